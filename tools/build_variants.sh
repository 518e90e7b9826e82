#!/bin/bash
# tools/build_variants.sh "name:-DFLAG=1 -DOTHER=2" ...  -> pooraytracer_b200/csrc/build/variants/lib_<name>.so
cd "$(dirname "$0")/../pooraytracer_b200/csrc" || exit 1
rm -rf build/variants build/v_*; mkdir -p build/variants
for v in "$@"; do n=${v%%:*}; f=${v#*:}; [ "$f" = "$v" ] && f=""
  (make -s BUILD=$PWD/build/v_$n OUT=$PWD/build/variants/lib_$n.so MAIN=$PWD/build/v_$n/prt_main EXTRA="$f" > build/v_$n.log 2>&1 || { echo FAIL $n; tail -5 build/v_$n.log; }) &
done; wait; ls build/variants

#!/usr/bin/env python
"""Summarise a `-Xptxas -v` log: registers, stack frame, spills and shared memory per kernel.
usage: python tools/ptxas_summary.py pooraytracer_b200/csrc/build/kernels.ptxas.log [substring ...]"""
import re
import subprocess
import sys

t = open(sys.argv[1]).read()
want = sys.argv[2:]
for b in re.split(r"ptxas info\s+: Compiling entry function '", t)[1:]:
    name = b.split("'")[0]
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().split("(")[0]
    if want and not any(w in dem for w in want):
        continue
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", b)
    r = re.search(r"Used (\d+) registers", b)
    sm = re.search(r"(\d+) bytes smem", b)
    print(f"{dem[:58]:58s} regs {r.group(1):>3s} smem {sm.group(1) if sm else 0:>6} stack {m.group(1):>4s} spill st/ld {m.group(2)}/{m.group(3)}")

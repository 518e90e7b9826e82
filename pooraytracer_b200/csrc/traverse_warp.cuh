/*
 * traverse_warp.cuh — the persistent-warp BVH walk used by the extend, shadow and ray-query kernels.
 *
 * What the ncu capture of the first version showed (profiles/r1_v0_k_extend_ncu_summary.txt): the
 * scene lives in L1/L2 (DRAM throughput 3-8 %), the kernel is issue-bound with 6-7 of 32 lanes
 * active per instruction on incoherent bounces, and 40 % of the issue slots go to the FP64 pipe,
 * most of them for box tests. Hence:
 *
 *   1. lanes are refilled: a lane whose ray is finished takes the next ray of the queue while its
 *      neighbours keep walking (one warp-aggregated atomic per refill), instead of idling until the
 *      slowest of 32 rays is done;
 *   2. while-while: every lane first descends inner nodes until it holds a leaf, then all lanes
 *      intersect their leaves together, so the node code and the triangle code are not serialised
 *      against each other inside one loop body;
 *   3. the box test runs in FP32 with explicit outward padding and only ever over-approximates the
 *      reference's FP64 slab test (AABB.cpp:38-64), which keeps the FP64 pipe for the one thing that
 *      must carry the reference's bits: Triangle::Hit (tri_test, traverse.cuh).
 *
 * Box test error budget. Node bounds are FP32, rounded outward from the FP64 boxes on the host. With
 * o32 = fl32(o), inv32 = fl32(1/d) (|inv32| clamped to 1e30 so 0 * inf never appears), p = fl32(o32 inv32):
 *     fma(b, inv32, -p) = [ (b - o) / d  +  (o - o32) / d  +  (o32 inv32 - p) / 1 ] (1 + e)^2,   |e| <= 2^-24
 * and |o - o32| <= |o| 2^-24, |o32 inv32 - p| <= |o32 inv32| 2^-24, so
 *     |computed - exact| <= |computed| * 2^-21 + |o32| * |inv32| * 2^-22     (both with > 2x margin). Entry
 * distances are lowered and exit distances raised by that amount, tmin is rounded down and the
 * current closest t rounded up to FP32: a box is skipped only if the padded interval is empty, which
 * implies that the exact interval of the enclosed FP64 box is empty too.
 */
#ifndef PRT_TRAVERSE_WARP_CUH
#define PRT_TRAVERSE_WARP_CUH

#include "traverse.cuh"

namespace prt {

#define PRT_DONE 0x7fffffff      /* `cur` of a lane whose walk is over */
#define PRT_BOX_EPS 4.76837158e-7f /* 2^-21 */
#ifndef PRT_INNER_EXIT_BELOW
#define PRT_INNER_EXIT_BELOW 8   /* leave the inner-node loop once fewer lanes than this are still descending */
#endif
#ifndef PRT_WALK_MODE
#define PRT_WALK_MODE 0          /* 0: while-while, 1: gated if-if with a postponed leaf */
#endif
#ifndef PRT_LEAF_GATE
#define PRT_LEAF_GATE 16         /* lanes that must hold a parked leaf before the triangle block runs */
#endif
#ifndef PRT_SAH_STACK
#define PRT_SAH_STACK 64         /* entries of the SAH / LBVH walk's stack; a ray that overflows it is handed to the
                                    reference-order walk (whose own depth is bounded by log2 of the mesh sizes) */
#endif
/* The first PRT_SMEM_STACK entries of a lane's stack live in SHARED memory, laid out [entry][thread] so that a
 * lane's bank never changes and no two lanes of a warp share one: conflict-free at any mix of stack depths. The
 * round-1 stack was a 64-entry local-memory array: 107-128 M local loads + 58-70 M local stores per k_extend
 * launch (profiles/r1_s26_extend_ncu_summary.txt) through the same L1TEX pipe the node and triangle fetches
 * saturate (75-88 %). Deeper entries (a BVH4 walk pushes up to three per level; cornell-box stays below 12,
 * the 10 M-triangle meshes below 30) spill to a local array, so depth is never lost; beyond PRT_SAH_STACK the
 * ray goes to the reference-order walk as before. */
#ifndef PRT_SMEM_STACK
#define PRT_SMEM_STACK 16
#endif
#ifndef PRT_TRAVERSE_BLOCK
#define PRT_TRAVERSE_BLOCK 128 /* threads per block of the kernels that walk the tree (k_tail: always 128); the stack's stride is blockDim.x.
                                  Registers are per SM sub-partition (16 K each): 6 warps per sub-partition = 80 registers, 5 = 96, whatever the
                                  block size, so smaller blocks buy no in-between budget (ptxas: 32 x 23 and 64 x 11 still compile to 80) */
#endif
#define PRT_STACK_WORDS_FOR(threads) (((PRT_SMEM_STACK < PRT_SAH_STACK ? PRT_SMEM_STACK : PRT_SAH_STACK) > 0 ? (PRT_SMEM_STACK < PRT_SAH_STACK ? PRT_SMEM_STACK : PRT_SAH_STACK) : 1) * (threads))
#define PRT_SMEM_STACK_WORDS PRT_STACK_WORDS_FOR(PRT_TRAVERSE_BLOCK)

/* refill idle lanes once fewer than this many lanes hold a ray (Policy::REFILL_BELOW). Refilling early keeps
 * lanes busy; refilling late keeps the rays of a warp (and the path vertices they push for k_shade) together.
 * Measured (profiles/r1_s8_ab.txt): the render kernels like 8-12 (k_shade reads the records of consecutive
 * queue entries: 35.2 -> 34.3 ms at 12, more than k_extend loses); for query batches 20 trades +4 % on random
 * rays for -4 % on coherent ones, so they stay at 16 */
#ifndef PRT_REFILL_BELOW_RENDER
#ifdef PRT_REFILL_BELOW
#define PRT_REFILL_BELOW_RENDER PRT_REFILL_BELOW
#else
#define PRT_REFILL_BELOW_RENDER 12
#endif
#endif
#ifndef PRT_REFILL_BELOW_SHADOW /* k_shadow pushes nothing: only its own lane occupancy counts */
#define PRT_REFILL_BELOW_SHADOW PRT_REFILL_BELOW_RENDER
#endif
#ifndef PRT_REFILL_BELOW_QUERY
#ifdef PRT_REFILL_BELOW
#define PRT_REFILL_BELOW_QUERY PRT_REFILL_BELOW
#else
#define PRT_REFILL_BELOW_QUERY 16
#endif
#endif

/* PRT_BOX_AXIS_ERR = 1 (default): the origin error of each axis is folded into that axis alone.
 * The first version padded the final entry / exit distances with the LARGEST of the three axis
 * errors. A scattered ray with one direction component of ~1e-8 (a cosine-hemisphere sample at the
 * rim of the disk, leaving a wall tangentially) has |o| |1/d| 2^-22 ~ 5 on that axis: the pad then
 * exceeded the length of the ray, no box was ever culled by the closest hit so far, and ONE such ray
 * walked 88 000 nodes of the displaced wall it grazed — a serial chain that held its launch for
 * 100 ms (profiles/r1_s13_grazing_rays.txt). With per-axis errors the two well-conditioned axes
 * keep their tight bounds and cull as usual. */
#ifndef PRT_BOX_AXIS_ERR
#define PRT_BOX_AXIS_ERR 1
#endif

struct RayF32 { /* the FP32 side of a ray, for box tests only: t = b * inv - o * inv */
    float ix, iy, iz;    /* fl32(1 / d), clamped */
    float o32x, o32y, o32z; /* fl32(o): the quantised-node decode works relative to the node's origin */
    float ex, ey, ez;       /* |o32| |inv32| 2^-22 per axis (rounded up) + a denormal-scale floor */
#if PRT_BOX_AXIS_ERR
    float nx, ny, nz;    /* o32 * inv32 + err(axis), rounded up: subtracted from the NEAR plane's product */
    float fx, fy, fz;    /* o32 * inv32 - err(axis), rounded down: subtracted from the FAR plane's product */
#else
    float ox, oy, oz;    /* fl32(o32 * inv32) */
    float err;           /* max over the axes of |o32| |inv32| 2^-22 (rounded up) */
#endif
};

__device__ __forceinline__ float clamp_inv(double d) {
    float i = (float)(1.0 / d);
    if (!(fabsf(i) <= 1e30f)) i = signbit(d) ? -1e30f : 1e30f; /* inf (d == +-0, denormal) and NaN */
    return i;
}
__device__ __forceinline__ RayF32 make_ray_f32(d3 o, d3 d) {
    RayF32 r;
    const float ox = (float)o.x, oy = (float)o.y, oz = (float)o.z;
    r.ix = clamp_inv(d.x); r.iy = clamp_inv(d.y); r.iz = clamp_inv(d.z);
    const float ex = __fmul_ru(__fmul_ru(fabsf(ox), fabsf(r.ix)), 2.3841858e-7f); /* |o32| |inv32| 2^-22 */
    const float ey = __fmul_ru(__fmul_ru(fabsf(oy), fabsf(r.iy)), 2.3841858e-7f);
    const float ez = __fmul_ru(__fmul_ru(fabsf(oz), fabsf(r.iz)), 2.3841858e-7f);
    r.o32x = ox; r.o32y = oy; r.o32z = oz;
    r.ex = ex + 1e-36f; r.ey = ey + 1e-36f; r.ez = ez + 1e-36f;
#if PRT_BOX_AXIS_ERR
    const float px = ox * r.ix, py = oy * r.iy, pz = oz * r.iz;
    r.nx = __fadd_ru(px, ex); r.ny = __fadd_ru(py, ey); r.nz = __fadd_ru(pz, ez);
    r.fx = __fadd_rd(px, -ex); r.fy = __fadd_rd(py, -ey); r.fz = __fadd_rd(pz, -ez);
#else
    r.ox = ox * r.ix; r.oy = oy * r.iy; r.oz = oz * r.iz;
    r.err = fmaxf(ex, fmaxf(ey, ez));
#endif
    return r;
}

/* Padded FP32 slab test. Per axis t = fma(b, inv32, -o32 * inv32); with E = the largest computed near
 * value and X = the smallest computed far value, the exact entry distance is >= the exact near value
 * of E's axis >= E - |E| eps - err(axis), and the exact exit distance <= X + |X| eps + err(axis), so
 * padding the two extremes once by the relative term is as safe as padding all six values; the
 * absolute term err(axis) rides in the biased origin products (nx / fx ...), whose > 2x margin over
 * the real origin error (2^-23 |o| |inv|) also absorbs the relative error of a near value that the
 * bias has pulled towards zero. The near plane of an axis is picked by the sign of inv32 (lo <= hi
 * for real boxes; an unused child holds lo = +inf, hi = -inf and fails either way). fmaxf / fminf
 * drop NaNs (inf - inf of an overflowed bound, 0 * inf): such an axis does not constrain. */
__device__ __forceinline__ bool box_f32(float lx, float ly, float lz, float hx, float hy, float hz, const RayF32& r, float tmin32,
                                        float tbest32, float& entry) {
#if PRT_BOX_AXIS_ERR
    const bool sx = r.ix < 0.f, sy = r.iy < 0.f, sz = r.iz < 0.f;
    const float ax = fmaf(sx ? hx : lx, r.ix, -r.nx), bx = fmaf(sx ? lx : hx, r.ix, -r.fx);
    const float ay = fmaf(sy ? hy : ly, r.iy, -r.ny), by = fmaf(sy ? ly : hy, r.iy, -r.fy);
    const float az = fmaf(sz ? hz : lz, r.iz, -r.nz), bz = fmaf(sz ? lz : hz, r.iz, -r.fz);
    const float E = fmaxf(fmaxf(ax, ay), az);
    const float X = fminf(fminf(bx, by), bz);
    entry = fmaxf(fmaf(-PRT_BOX_EPS, fabsf(E), E), tmin32);
    const float exit = fminf(fmaf(PRT_BOX_EPS, fabsf(X), X), tbest32);
#else
    const float ax = fmaf(lx, r.ix, -r.ox), bx = fmaf(hx, r.ix, -r.ox);
    const float ay = fmaf(ly, r.iy, -r.oy), by = fmaf(hy, r.iy, -r.oy);
    const float az = fmaf(lz, r.iz, -r.oz), bz = fmaf(hz, r.iz, -r.oz);
    const float E = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    const float X = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    entry = fmaxf(fmaf(-PRT_BOX_EPS, fabsf(E), E) - r.err, tmin32);
    const float exit = fminf(fmaf(PRT_BOX_EPS, fabsf(X), X) + r.err, tbest32);
#endif
    return entry <= exit;
}

/* Quantised nodes (Node4Q). A plane is origin + q s with s = 2^(exp - 127), so along one axis
 *     t = (origin + q s - o) / d = q (s inv) + (origin - o) inv =: q A + B.
 * A = s inv32 is exact (a power of two times a float), D = fl(origin - o32), B = fl(D inv32), and the
 * plane's t comes out of ONE fma. With e_i the roundings (|e_i| <= 2^-24):
 *     |t_c - t*| <= |t_c| 2^-23 + |B| 2^-22 + |o| |inv| 2^-24 (1 + 2^-23),
 * covered by the relative pad of the final entry / exit (2^-21 |E|, as for Node4), by
 * pad(axis) = |B| 2^-21 + err(axis) folded into B per node and axis (near: B - pad rounded down, far:
 * B + pad rounded up; err(axis) = |o32| |inv32| 2^-22 + 1e-36), each with a margin of 2x or more.
 * Three words per axis carry the four children's bytes, so the near / far plane of an axis is ONE
 * select per node, not per child. */
struct QuantAxis {
    float A, Bn, Bf;
    uint32_t wn, wf; /* the four children's near / far plane bytes */
};
__device__ __forceinline__ QuantAxis quant_axis(float origin, uint32_t exp_byte, uint32_t qlo, uint32_t qhi, float o32, float inv, float err) {
    QuantAxis a;
    const float s = __uint_as_float(exp_byte << 23);
    a.A = s * inv;
    const float B = (origin - o32) * inv;
    const float pad = fmaf(fabsf(B), PRT_BOX_EPS, err);
    a.Bn = __fadd_rd(B, -pad);
    a.Bf = __fadd_ru(B, pad);
    const bool neg = inv < 0.f;
    a.wn = neg ? qhi : qlo;
    a.wf = neg ? qlo : qhi;
    return a;
}
template <int C>
__device__ __forceinline__ bool box_quant(const QuantAxis& x, const QuantAxis& y, const QuantAxis& z, float tmin32, float tbest32, float& entry) {
    const float nx = fmaf((float)((x.wn >> (8 * C)) & 255u), x.A, x.Bn), fx = fmaf((float)((x.wf >> (8 * C)) & 255u), x.A, x.Bf);
    const float ny = fmaf((float)((y.wn >> (8 * C)) & 255u), y.A, y.Bn), fy = fmaf((float)((y.wf >> (8 * C)) & 255u), y.A, y.Bf);
    const float nz = fmaf((float)((z.wn >> (8 * C)) & 255u), z.A, z.Bn), fz = fmaf((float)((z.wf >> (8 * C)) & 255u), z.A, z.Bf);
    const float E = fmaxf(fmaxf(nx, ny), nz);
    const float X = fminf(fminf(fx, fy), fz);
    entry = fmaxf(fmaf(-PRT_BOX_EPS, fabsf(E), E), tmin32);
    const float exit = fminf(fmaf(PRT_BOX_EPS, fabsf(X), X), tbest32);
    return entry <= exit;
}

/* The reference-order walk is the rare path of the persistent kernels (boundary / equal-t hits):
 * kept out of line so that its FP64 slab code and second stack do not weigh on the register
 * allocation of the hot loop. */
#ifndef PRT_REF_NOINLINE
#define PRT_REF_NOINLINE 0
#endif
template <bool ANY, bool COUNT>
#if PRT_REF_NOINLINE
__device__ __noinline__
#else
__device__ __forceinline__
#endif
void trace_reference_cold(const DevScene& S, d3 o, d3 d, double tmin, double tmax, Hit& hit, VisitCounters& cnt) {
    trace_reference<ANY, COUNT>(S, o, d, tmin, tmax, hit, cnt);
}

/* L2 prefetch of the line holding p (the record a lane will read once it is handed its ray) */
#ifndef PRT_TRAV_PREFETCH
#define PRT_TRAV_PREFETCH 0
#endif
#ifndef PRT_PREFETCH_LEAF
#define PRT_PREFETCH_LEAF 0      /* 1: L1, 2: L2 prefetch of a leaf's first triangle when the lane reaches it */
#endif
#ifndef PRT_PREFETCH_PUSH
#define PRT_PREFETCH_PUSH 0      /* prefetch the second-nearest child (top of the stack) as it is pushed */
#endif
__device__ __forceinline__ void prefetch_line(const void* p) {
#if PRT_PREFETCH_LEAF == 2 || PRT_PREFETCH_PUSH == 2
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

/*
 * The persistent loop. Policy P supplies the rays and consumes the results:
 *     int32_t P::slot_of(int32_t qi)       queue entry qi -> the item ("slot") it names; per lane
 *     void P::prefetch(int32_t slot)       optional L2 prefetch of what load(slot) will read
 *     bool P::load(int32_t slot, d3& o, d3& d, double& tmin, double& tmax)   per lane; false = skip
 *     void P::retire(bool finished, int32_t slot, const Hit& hit)   ALL lanes call it together;
 *                                          lanes with finished == true carry a completed ray
 *     void P::finish()                     ALL lanes, once, after the last retire
 * `cursor` is the queue's fetch counter (zeroed by the host), `n` the queue length.
 *
 * Work fetch is a three-stage pipeline per warp, so that neither the atomic on the cursor nor the
 * queue read nor the DRAM latency of a ray record is waited for (ncu, profiles/r1_s1_extend: 30 %
 * of the stall samples of the first persistent version sat on exactly those three):
 *     chunk A   atomicAdd(cursor, 32) issued by lane 0, result not looked at yet
 *     chunk B   base known; lane j has read slot_of(base + j); records prefetched one event later
 *     chunk C   the chunk being handed out to lanes that run dry
 * Results are the reference's (see traverse.cuh): boundary and equal-t hits go through
 * trace_reference before they are retired.
 */
template <bool ANY, bool COUNT, int WIDE, class P>
__device__ __forceinline__ void traverse_persistent(const DevScene& S, P& pol, int32_t* cursor, int32_t n, VisitCounters& cnt, int32_t* sm_stack) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    constexpr int SM_N = PRT_SMEM_STACK < PRT_SAH_STACK ? PRT_SMEM_STACK : PRT_SAH_STACK;
    constexpr int LOC_N = PRT_SAH_STACK - SM_N > 0 ? PRT_SAH_STACK - SM_N : 1;
    const int BS = blockDim.x;
    int32_t* const sm = sm_stack + threadIdx.x; /* entry e of this lane: sm[e * blockDim.x] */
    int32_t deep[LOC_N];                        /* entries SM_N .. PRT_SAH_STACK - 1 (rarely reached) */
    int sp = 0;
#define PRT_PUSH(v)                                                     \
    do {                                                                \
        if (sp < SM_N) sm[sp * BS] = (v);               \
        else if (sp < PRT_SAH_STACK) deep[sp - SM_N] = (v);             \
        else { overflow = true; break; }                                \
        ++sp;                                                           \
    } while (0)
#define PRT_POP() (--sp, sp < SM_N ? sm[sp * BS] : deep[sp - SM_N])
    int32_t cur = PRT_DONE;
    int32_t leaf = -1; /* PRT_WALK_MODE 1: parked leaf, (slot << 3) | (triangles left - 1) */
    int32_t slot = -1;
#ifdef PRT_DEBUG_SLOW_RAY
    unsigned long long dbg_start = 0;
#endif
    bool has_ray = false, exhausted = false, recheck = false;
    bool overflow = false; /* a child did not fit the stack: the SAH walk is incomplete, the reference walk decides */
    d3 o = mk3(0, 0, 0), d = mk3(0, 0, 0);
    double tmin = 0.0, tmax = 0.0, tbest = 0.0;
    float tmin32 = 0.f, tbest32 = 0.f;
    RayF32 rf = make_ray_f32(o, d);
    Hit hit;
    hit.slot = -1; hit.t = 0.0; hit.alpha = 0.0; hit.beta = 0.0;
    const bool use_sah = WIDE == 40 ? S.nodes4q != nullptr : (WIDE == 4 ? S.nodes4 != nullptr : S.nodes2 != nullptr);

    /* fetch pipeline state (c_*, b_base, b_pref are warp-uniform) */
    int32_t a_raw = 0;                 /* lane 0: chunk A's atomic result */
    int32_t b_base, b_slot;            /* chunk B */
    bool b_pref = false;
    int32_t c_slot = -1, c_off = 0, c_left = 0;
    if (lane == 0) a_raw = atomicAdd(cursor, 32);
    b_base = __shfl_sync(FULL, a_raw, 0);
    b_slot = b_base + lane < n ? pol.slot_of(b_base + lane) : -1;
    if (lane == 0) a_raw = atomicAdd(cursor, 32);

    for (;;) {
        /* ---- retire finished lanes, refill idle ones ---- */
        const bool finished = has_ray && cur == PRT_DONE && leaf < 0;
#ifdef PRT_DEBUG_SLOW_RAY
        if (COUNT && finished) {
            const unsigned long long v = cnt.nodes + cnt.tris - dbg_start;
            if (v > PRT_DEBUG_SLOW_RAY)
                printf("[slow ray] fetches %llu slot %d o %.17g %.17g %.17g d %.17g %.17g %.17g tmin %g tmax %g hit %d t %.17g any %d\n", v, slot, o.x, o.y, o.z,
                       d.x, d.y, d.z, tmin, tmax, hit.slot, hit.t, (int)ANY);
        }
#endif
        if (finished) {
            if (recheck || overflow) {
                const unsigned long long before = cnt.nodes + cnt.tris;
                trace_reference_cold<ANY, COUNT>(S, o, d, tmin, tmax, hit, cnt);
                if (COUNT) {
                    const unsigned long long v = cnt.nodes + cnt.tris - before;
                    cnt.rechecks++;
                    cnt.recheck_visits += v;
                    if (v > cnt.recheck_max) cnt.recheck_max = v;
                }
            }
            has_ray = false;
        }
        __syncwarp();
        pol.retire(finished, slot, hit);
        if (!exhausted) {
            const unsigned need = __ballot_sync(FULL, !has_ray);
            int k = __popc(need);
            int rank = __popc(need & lt_mask);
            int32_t mine = -1;
#pragma unroll 1
            for (int round = 0; round < 2 && k > 0; ++round) {
                if (c_left == 0) {
                    /* C <- B, B <- A, issue a new A */
                    if (b_base >= n) { exhausted = true; break; }
                    c_slot = b_slot;
                    c_off = 0;
                    c_left = n - b_base < 32 ? n - b_base : 32;
                    if (PRT_TRAV_PREFETCH && !b_pref && c_slot >= 0) pol.prefetch(c_slot);
                    b_base = __shfl_sync(FULL, a_raw, 0);
                    b_slot = b_base + lane < n ? pol.slot_of(b_base + lane) : -1;
                    b_pref = false;
                    if (lane == 0) a_raw = atomicAdd(cursor, 32);
                }
                const int take = k < c_left ? k : c_left;
                const int32_t got = __shfl_sync(FULL, c_slot, (c_off + rank) & 31);
                if (!has_ray && mine < 0 && rank >= 0 && rank < take) mine = got;
                rank -= take;
                k -= take;
                c_off += take;
                c_left -= take;
            }
            if (!b_pref && !exhausted) {
                if (PRT_TRAV_PREFETCH && b_slot >= 0) pol.prefetch(b_slot);
                b_pref = true;
            }
            if (mine >= 0) {
                slot = mine;
                if (pol.load(slot, o, d, tmin, tmax)) {
#ifdef PRT_DEBUG_SLOW_RAY
                    dbg_start = cnt.nodes + cnt.tris;
#endif
                    has_ray = true;
                    recheck = false;
                    overflow = false;
                    hit.slot = -1; hit.t = 0.0; hit.alpha = 0.0; hit.beta = 0.0;
                    tbest = tmax;
                    sp = 0;
                    /* A ray with a NaN coordinate (a scattered direction of 0/0: the reference scrubs
                     * such pixels at the very end, Camera.cpp:291-293) misses every triangle: n.o or n.d
                     * is NaN, t is NaN, Interval::Contains is false (Triangle.cpp:56-62) — likewise a
                     * NaN upper bound. But its box tests never fail either (AABB.cpp:38-64: every
                     * comparison with NaN is false, nothing shrinks), so it would walk the WHOLE tree,
                     * 1.6 M dependent fetches on the 1 M-triangle room: one such ray held a launch for
                     * ~50 ms (profiles/r1_s13_nan_rays.txt). The miss is known up front. */
                    const bool poisoned = o.x != o.x || o.y != o.y || o.z != o.z || d.x != d.x || d.y != d.y || d.z != d.z || tmax != tmax;
                    if (poisoned) {
                        cur = PRT_DONE;
                    } else if (use_sah) {
                        rf = make_ray_f32(o, d);
                        tmin32 = __double2float_rd(tmin);
                        tbest32 = __double2float_ru(tbest);
                        cur = S.n_tris > 0 ? 0 : PRT_DONE;
                    } else {
                        /* PRT_ACCEL_REFERENCE_BVH2: the reference's own walk, one ray per lane, run
                         * by the retire step of the next round */
                        recheck = true;
                        cur = PRT_DONE;
                    }
                }
            }
        }
        if (__ballot_sync(FULL, has_ray) == 0) {
            if (exhausted) break;
            continue;
        }

        /* ---- walk until enough lanes have run dry ---- */
#if PRT_WALK_MODE == 1
        /* Gated if-if with one postponed leaf. Every round: (1) lanes holding an inner node take one
         * step; a lane that reaches a leaf parks it in `leaf` and carries on with its stack
         * (speculatively: the parked leaf has not shrunk tbest yet); (2) once PRT_LEAF_GATE lanes hold
         * a parked leaf (or nobody can step any more), those lanes test ONE triangle each. Neither
         * block waits for the slowest lane of the other. */
        for (;;) {
            if (cur >= 0 && cur != PRT_DONE) {
                const float4* __restrict__ p = reinterpret_cast<const float4*>(S.nodes2 + cur);
                const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
                const int4 ch = __ldg(reinterpret_cast<const int4*>(p + 3));
                if (COUNT) cnt.nodes++;
                float e0, e1;
                const bool h0 = box_f32(a.x, a.y, a.z, a.w, b.x, b.y, rf, tmin32, tbest32, e0);
                const bool h1 = box_f32(b.z, b.w, c.x, c.y, c.z, c.w, rf, tmin32, tbest32, e1) && ch.y != PRT_NO_CHILD;
                if (h0 && h1) {
                    const bool swap = e1 < e0;
                    PRT_PUSH(swap ? ch.x : ch.y);
                    cur = swap ? ch.y : ch.x;
                } else if (h0) {
                    cur = ch.x;
                } else if (h1) {
                    cur = ch.y;
                } else {
                    cur = sp > 0 ? PRT_POP() : PRT_DONE;
                }
            }
            if (cur < 0 && leaf < 0) { /* park the leaf, go on with the stack */
                leaf = ~cur;
                cur = sp > 0 ? PRT_POP() : PRT_DONE;
            }
            const unsigned lm = __ballot_sync(FULL, leaf >= 0);
            const unsigned im = __ballot_sync(FULL, cur >= 0 && cur != PRT_DONE);
            if (lm != 0 && (__popc(lm) >= PRT_LEAF_GATE || im == 0)) {
                if (leaf >= 0) {
                    /* leaf = (slot << 3) | (triangles left - 1): test `slot`, then advance */
                    const int32_t ts = leaf >> 3;
                    leaf = (leaf & 7) ? leaf + 7 : -1;
                    double t, al, be;
                    if (COUNT) cnt.tris++;
                    if (tri_test(S.tris + ts, o, d, tmin, tbest, t, al, be)) {
                        if (ANY) {
                            /* a boundary hit may be one the reference's boxes reject: keep looking
                             * for an interior one, let the reference decide if none turns up */
                            if (on_boundary(al, be)) {
                                recheck = true;
                            } else {
                                hit.slot = ts; hit.t = t; hit.alpha = al; hit.beta = be;
                                recheck = false;
                                leaf = -1; cur = PRT_DONE; sp = 0;
                            }
                        } else {
                            recheck = hit.slot >= 0 && t == tbest; /* same t, another triangle */
                            hit.slot = ts; hit.t = t; hit.alpha = al; hit.beta = be;
                            tbest = t;
                            tbest32 = __double2float_ru(t);
                        }
                    }
                }
            }
            const unsigned walking = __ballot_sync(FULL, has_ray && !(cur == PRT_DONE && leaf < 0));
            if (walking == 0) break;
            if (!exhausted && __popc(walking) < P::REFILL_BELOW) break;
        }
        if (!ANY && has_ray && cur == PRT_DONE && leaf < 0 && hit.slot >= 0 && on_boundary(hit.alpha, hit.beta)) recheck = true;
#else
        for (;;) {
            /* inner nodes: all lanes step together until fewer than PRT_INNER_EXIT_BELOW lanes still
             * hold an inner node (the others hold a leaf or are done); stragglers wait through
             * the leaf phase and carry on in the next round */
            for (;;) {
                if (cur >= 0 && cur != PRT_DONE) {
                    if (COUNT) cnt.nodes++;
                    if (WIDE == 4 || WIDE == 40) {
                        float e0, e1, e2, e3;
                        bool h0, h1, h2, h3;
                        int4 ch;
                        if (WIDE == 40) {
                            const char* __restrict__ p = reinterpret_cast<const char*>(S.nodes4q + cur);
                            const f8 q0 = ld256f(p), q1 = ld256f(p + 32);
                            const uint32_t exps = __float_as_uint(q0.v[3]);
                            const QuantAxis ax = quant_axis(q0.v[0], exps & 255u, __float_as_uint(q0.v[4]), __float_as_uint(q1.v[0]), rf.o32x, rf.ix, rf.ex);
                            const QuantAxis ay = quant_axis(q0.v[1], (exps >> 8) & 255u, __float_as_uint(q0.v[5]), __float_as_uint(q1.v[1]), rf.o32y, rf.iy, rf.ey);
                            const QuantAxis az = quant_axis(q0.v[2], (exps >> 16) & 255u, __float_as_uint(q0.v[6]), __float_as_uint(q1.v[2]), rf.o32z, rf.iz, rf.ez);
                            ch = make_int4(__float_as_int(q1.v[4]), __float_as_int(q1.v[5]), __float_as_int(q1.v[6]), __float_as_int(q1.v[7]));
                            h0 = box_quant<0>(ax, ay, az, tmin32, tbest32, e0) && ch.x != PRT_NO_CHILD;
                            h1 = box_quant<1>(ax, ay, az, tmin32, tbest32, e1) && ch.y != PRT_NO_CHILD;
                            h2 = box_quant<2>(ax, ay, az, tmin32, tbest32, e2) && ch.z != PRT_NO_CHILD;
                            h3 = box_quant<3>(ax, ay, az, tmin32, tbest32, e3) && ch.w != PRT_NO_CHILD;
                        } else {
#if PRT_LD256
                        const char* __restrict__ p = reinterpret_cast<const char*>(S.nodes4 + cur);
                        const f8 q0 = ld256f(p), q1 = ld256f(p + 32), q2 = ld256f(p + 64), q3 = ld256f(p + 96);
                        const float4 lx = make_float4(q0.v[0], q0.v[1], q0.v[2], q0.v[3]), ly = make_float4(q0.v[4], q0.v[5], q0.v[6], q0.v[7]);
                        const float4 lz = make_float4(q1.v[0], q1.v[1], q1.v[2], q1.v[3]), hx = make_float4(q1.v[4], q1.v[5], q1.v[6], q1.v[7]);
                        const float4 hy = make_float4(q2.v[0], q2.v[1], q2.v[2], q2.v[3]), hz = make_float4(q2.v[4], q2.v[5], q2.v[6], q2.v[7]);
                        ch = make_int4(__float_as_int(q3.v[0]), __float_as_int(q3.v[1]), __float_as_int(q3.v[2]), __float_as_int(q3.v[3]));
#else
                        const float4* __restrict__ p = reinterpret_cast<const float4*>(S.nodes4 + cur);
                        const float4 lx = __ldg(p), ly = __ldg(p + 1), lz = __ldg(p + 2);
                        const float4 hx = __ldg(p + 3), hy = __ldg(p + 4), hz = __ldg(p + 5);
                        ch = __ldg(reinterpret_cast<const int4*>(p + 6));
#endif
                        h0 = box_f32(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, rf, tmin32, tbest32, e0);
                        h1 = box_f32(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, rf, tmin32, tbest32, e1) && ch.y != PRT_NO_CHILD;
                        h2 = box_f32(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, rf, tmin32, tbest32, e2) && ch.z != PRT_NO_CHILD;
                        h3 = box_f32(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, rf, tmin32, tbest32, e3) && ch.w != PRT_NO_CHILD;
                        }
                        /* sort the hit children by entry distance: key = distance bits (non-negative
                         * float: ordered like an unsigned) with the child's slot in the two low bits */
                        const unsigned MISS = 0xffffffffu;
                        const unsigned k0 = h0 ? ((__float_as_uint(fmaxf(e0, 0.f)) & ~3u) | 0u) : MISS;
                        const unsigned k1 = h1 ? ((__float_as_uint(fmaxf(e1, 0.f)) & ~3u) | 1u) : MISS;
                        const unsigned k2 = h2 ? ((__float_as_uint(fmaxf(e2, 0.f)) & ~3u) | 2u) : MISS;
                        const unsigned k3 = h3 ? ((__float_as_uint(fmaxf(e3, 0.f)) & ~3u) | 3u) : MISS;
                        const unsigned a0 = min(k0, k1), a1 = max(k0, k1), b0 = min(k2, k3), b1 = max(k2, k3);
                        const unsigned s0 = min(a0, b0), t1 = max(a0, b0), t2 = min(a1, b1), s3 = max(a1, b1);
                        const unsigned s1 = min(t1, t2), s2 = max(t1, t2);
#define PRT_CHILD_OF(key) (((key) & 2u) ? (((key) & 1u) ? ch.w : ch.z) : (((key) & 1u) ? ch.y : ch.x))
                        if (s0 == MISS) {
                            cur = sp > 0 ? PRT_POP() : PRT_DONE;
                        } else {
                            /* the other hit children go on the stack farthest first. Validity is monotone in the sorted
                             * keys (s1 valid if s2 is, s2 if s3 is), so with cnt of them valid the entries are
                             * [s3 s2 s1], [s2 s1] or [s1]: three UNCONDITIONAL shared-memory stores (what lands above the new
                             * stack pointer is never read) instead of three divergent push sites, which were 12 % of the
                             * kernel's issued instructions at 3-6 active lanes (profiles/r2_s9_extend_hot_sass.txt) */
                            const int cnt = (s1 != MISS ? 1 : 0) + (s2 != MISS ? 1 : 0) + (s3 != MISS ? 1 : 0);
                            if (sp + 3 <= SM_N) {
                                const int32_t c1 = PRT_CHILD_OF(s1), c2 = PRT_CHILD_OF(s2), c3 = PRT_CHILD_OF(s3);
                                int32_t* top = sm + sp * BS;
                                top[0] = cnt == 3 ? c3 : (cnt == 2 ? c2 : c1);
                                top[BS] = cnt == 3 ? c2 : c1;
                                top[2 * BS] = c1;
                                sp += cnt;
                            } else { /* near the end of the shared part (or with a tiny test stack): one entry at a time */
                                if (s3 != MISS) PRT_PUSH(PRT_CHILD_OF(s3));
                                if (s2 != MISS) PRT_PUSH(PRT_CHILD_OF(s2));
                                if (s1 != MISS) PRT_PUSH(PRT_CHILD_OF(s1));
                            }
                            cur = PRT_CHILD_OF(s0);
                        }
#undef PRT_CHILD_OF
#if PRT_PREFETCH_LEAF
                        /* a lane that reached a leaf waits for the others to finish their inner nodes:
                         * start pulling its first triangle in meanwhile */
                        if (cur < 0) prefetch_line(S.tris + ((~cur) >> 3));
#endif
                    } else {
#if PRT_LD256
                        const char* __restrict__ p = reinterpret_cast<const char*>(S.nodes2 + cur);
                        const f8 q0 = ld256f(p), q1 = ld256f(p + 32);
                        const float4 a = make_float4(q0.v[0], q0.v[1], q0.v[2], q0.v[3]), b = make_float4(q0.v[4], q0.v[5], q0.v[6], q0.v[7]);
                        const float4 c = make_float4(q1.v[0], q1.v[1], q1.v[2], q1.v[3]);
                        const int4 ch = make_int4(__float_as_int(q1.v[4]), __float_as_int(q1.v[5]), 0, 0);
#else
                        const float4* __restrict__ p = reinterpret_cast<const float4*>(S.nodes2 + cur);
                        const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
                        const int4 ch = __ldg(reinterpret_cast<const int4*>(p + 3));
#endif
                        float e0, e1;
                        const bool h0 = box_f32(a.x, a.y, a.z, a.w, b.x, b.y, rf, tmin32, tbest32, e0);
                        const bool h1 = box_f32(b.z, b.w, c.x, c.y, c.z, c.w, rf, tmin32, tbest32, e1) && ch.y != PRT_NO_CHILD;
                        if (h0 && h1) {
                            const bool swap = e1 < e0;
                            PRT_PUSH(swap ? ch.x : ch.y);
                            cur = swap ? ch.y : ch.x;
                        } else if (h0) {
                            cur = ch.x;
                        } else if (h1) {
                            cur = ch.y;
                        } else {
                            cur = sp > 0 ? PRT_POP() : PRT_DONE;
                        }
                    }
                }
                if (__popc(__ballot_sync(FULL, cur >= 0 && cur != PRT_DONE)) < PRT_INNER_EXIT_BELOW) break;
            }
            /* leaves */
            if (cur < 0) {
                const int32_t code = ~cur;
                const int32_t first = code >> 3, count = (code & 7) + 1;
                bool done = false;
                for (int32_t i = 0; i < count; ++i) {
                    double t, al, be;
                    if (COUNT) cnt.tris++;
                    if (tri_test(S.tris + first + i, o, d, tmin, tbest, t, al, be)) {
                        if (ANY) {
                            /* a boundary hit may be one the reference's boxes reject: keep looking
                             * for an interior one, let the reference decide if none turns up */
                            if (on_boundary(al, be)) { recheck = true; continue; }
                            hit.slot = first + i; hit.t = t; hit.alpha = al; hit.beta = be;
                            recheck = false;
                            done = true;
                            break;
                        }
                        recheck = hit.slot >= 0 && t == tbest; /* same t, another triangle */
                        hit.slot = first + i; hit.t = t; hit.alpha = al; hit.beta = be;
                        tbest = t;
                        tbest32 = __double2float_ru(t);
                    }
                }
                cur = (done || sp == 0) ? PRT_DONE : PRT_POP();
            }
            if (cur == PRT_DONE && has_ray && !ANY && hit.slot >= 0 && on_boundary(hit.alpha, hit.beta)) recheck = true;
            const unsigned walking = __ballot_sync(FULL, has_ray && cur != PRT_DONE);
            if (walking == 0) break;
            if (!exhausted && __popc(walking) < P::REFILL_BELOW) break;
        }
#endif
    }
    pol.finish();
#undef PRT_PUSH
#undef PRT_POP
}

}  // namespace prt
#endif

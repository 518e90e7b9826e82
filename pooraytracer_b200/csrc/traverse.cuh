/*
 * traverse.cuh — ray / box, ray / triangle and the two BVH walks (device code).
 *
 *   tri_test        Triangle::Hit + IsInterior      Source/Triangle.cpp:54-83,100-113
 *   slab_f64        AABB::Hit                       Source/AABB.cpp:38-64
 *   trace_reference BVHNode::Hit                    Source/BVH.cpp:51-61 (the reference's own tree
 *                                                   in the reference's own order: bit-exact)
 *   on_boundary     the hits whose owner the reference's own tree may decide (equal-t candidates,
 *                   Source/Interval.h:21-23 + Source/BVH.cpp:58; its box test's rejection of
 *                   degenerate intervals, Source/AABB.cpp:62): the SAH / LBVH walks of
 *                   traverse_warp.cuh hand those to trace_reference.
 *
 * All arithmetic is FP64 in the reference's operation order; the file is compiled with
 * -fmad=false so nothing is contracted into an FMA (the reference objects contain none).
 */
#ifndef PRT_TRAVERSE_CUH
#define PRT_TRAVERSE_CUH

#include "device_scene.h"
#include "prt_math.h"

namespace prt {

#ifndef PRT_STACK
#define PRT_STACK 64 /* the reference-order walk's stack: depth = log2(mesh size) + the top level, 64 covers any mesh that fits */
#endif
#define PRT_NO_CHILD INT32_MIN

struct Hit {
    int32_t slot; /* -1: miss */
    double t, alpha, beta;
};

/* node / triangle records fetched; only maintained by the COUNT instantiations
 * (PRT_RENDER_COUNTERS / prt_set_counters), which feed the roofline's algorithmic bytes */
struct VisitCounters {
    unsigned long long nodes, tris;
    /* diagnostics of the reference-order re-trace (boundary / equal-t hits): how many rays took it, the
     * nodes + triangles they fetched there, the most one ray fetched */
    unsigned long long rechecks, recheck_visits, recheck_max;
};
#define PRT_COUNT_NODE(c) do { if (COUNT) (c).nodes++; } while (0)
#define PRT_COUNT_TRI(c) do { if (COUNT) (c).tris++; } while (0)

/* 256-bit read-only loads (sm_100: LDG.E.256): one L1 request per 32 bytes of a lane's record. The
 * traversal kernels are bound by L1TEX request throughput (ncu: l1tex__throughput 70-80 % of peak
 * with the issue slots half idle), every lane fetching its own node / triangle line, so halving
 * the number of load instructions per record is what counts. */
#ifndef PRT_LD256
#define PRT_LD256 1
#endif
struct f8 { float v[8]; };
struct dd4 { double v[4]; };
__device__ __forceinline__ f8 ld256f(const void* p) {
    f8 r;
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
        : "l"(p));
    return r;
}
__device__ __forceinline__ dd4 ld256d(const void* p) {
    dd4 r;
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.v[0]), "=d"(r.v[1]), "=d"(r.v[2]), "=d"(r.v[3]) : "l"(p));
    return r;
}

/* 256-bit loads / stores of MUTABLE global data (the wavefront's path records, launch.h): coherent (no .nc:
 * k_tail rewrites records between the stages of one launch), and kept out of L1 on the way in — a record is
 * read once per kernel, the L1 belongs to the BVH nodes and triangles every ray shares. One request per
 * 32-byte sector of a lane's record instead of four 64-bit ones. */
__device__ __forceinline__ dd4 ld256g(const void* p) {
    dd4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.v[0]), "=d"(r.v[1]), "=d"(r.v[2]), "=d"(r.v[3]) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st256g(void* p, double a, double b, double c, double d) {
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
/* two 32-bit words <-> one FP64 register image (a record's packed integer fields ride in vector loads) */
__device__ __forceinline__ double pack2(uint32_t lo, uint32_t hi) { return __hiloint2double((int)hi, (int)lo); }
__device__ __forceinline__ uint32_t unpack_lo(double v) { return (uint32_t)__double2loint(v); }
__device__ __forceinline__ uint32_t unpack_hi(double v) { return (uint32_t)__double2hiint(v); }

/* t = (D - n.o) / (n.d), Triangle.cpp:58 — the same IEEE quotient, without the divide's slow path for a ZERO dividend.
 * A scattered or shadow ray starts ON a triangle, and the first leaf it visits holds that triangle and its coplanar
 * neighbours: for an axis-aligned wall D - n.o is exactly 0 for about half of such rays (o.y = ro.y + rd.y t cancels to
 * the last bit). CUDA's FP64 division handles a zero / denormal quotient in an out-of-line routine of ~65 instructions,
 * run by the whole warp if one lane needs it: on cornell-box it ran once per two triangle tests, 6 % of k_extend's issued
 * instructions (profiles/r2_s9_extend_hot_sass.txt). 0 / x for finite non-zero x is +-0 with the sign of the operands'
 * signs' product: written out, it costs a compare. (NaN / inf denominators take the real division.) */
__device__ __forceinline__ double plane_t(double num, double denom) {
    if (num == 0.0 && fabs(denom) <= 1.7976931348623157e308)
        return (__double2hiint(num) ^ __double2hiint(denom)) < 0 ? -0.0 : 0.0;
    return num / denom;
}

/* Triangle::Hit. The 128-byte record is read as the plane (32 B) first and the remaining 96 B
 * only for candidates whose t is inside the interval. */
__device__ __forceinline__ bool tri_test(const TriIsect* __restrict__ T, d3 o, d3 d, double tmin, double tmax,
                                         double& t_out, double& alpha_out, double& beta_out) {
#if PRT_LD256
    const char* __restrict__ p = reinterpret_cast<const char*>(T);
    const dd4 q0 = ld256d(p); /* n.x n.y n.z D */
    const d3 n = mk3(q0.v[0], q0.v[1], q0.v[2]);
    const double denom = dot(n, d);
    if (fabs(denom) < 1e-8) return false;
    const double t = plane_t(q0.v[3] - dot(n, o), denom);
    if (!(tmin <= t && t <= tmax)) return false; /* Interval::Contains, inclusive */
    const dd4 q1 = ld256d(p + 32), q2 = ld256d(p + 64), q3 = ld256d(p + 96);
    const d3 w = mk3(q1.v[0], q1.v[1], q1.v[2]);
    const d3 v0 = mk3(q1.v[3], q2.v[0], q2.v[1]);
    const d3 e0 = mk3(q2.v[2], q2.v[3], q3.v[0]);
    const d3 e1 = mk3(q3.v[1], q3.v[2], q3.v[3]);
#else
    const double2* __restrict__ p = reinterpret_cast<const double2*>(T);
    const double2 q0 = __ldg(p), q1 = __ldg(p + 1); /* n.x n.y | n.z D */
    const d3 n = mk3(q0.x, q0.y, q1.x);
    const double denom = dot(n, d);
    if (fabs(denom) < 1e-8) return false;
    const double t = plane_t(q1.y - dot(n, o), denom);
    if (!(tmin <= t && t <= tmax)) return false; /* Interval::Contains, inclusive */
    const double2 q2 = __ldg(p + 2), q3 = __ldg(p + 3), q4 = __ldg(p + 4), q5 = __ldg(p + 5), q6 = __ldg(p + 6), q7 = __ldg(p + 7);
    const d3 w = mk3(q2.x, q2.y, q3.x);
    const d3 v0 = mk3(q3.y, q4.x, q4.y);
    const d3 e0 = mk3(q5.x, q5.y, q6.x);
    const d3 e1 = mk3(q6.y, q7.x, q7.y);
#endif
    const d3 pt = o + d * t; /* Ray.h:13 */
    const d3 v0p = pt - v0;
    const double alpha = dot(w, cross(v0p, e1));
    const double beta = dot(w, cross(e0, v0p));
    if (alpha != alpha || beta != beta) return false;
    if ((alpha < 0) || (beta < 0) || (alpha + beta > 1)) return false;
    t_out = t;
    alpha_out = alpha;
    beta_out = beta;
    return true;
}

/* AABB::Hit on one axis: the reference's ordered shrink (NaNs fall into the else branch and change
 * nothing). inv = 1.0 / d[i] is hoisted out of the node loop: same operands, same quotient. */
__device__ __forceinline__ void slab_axis(double bmin, double bmax, double o, double inv, double& lo, double& hi) {
    const double t0 = (bmin - o) * inv;
    const double t1 = (bmax - o) * inv;
    if (t0 < t1) {
        if (t0 > lo) lo = t0;
        if (t1 < hi) hi = t1;
    } else {
        if (t1 > lo) lo = t1;
        if (t0 < hi) hi = t0;
    }
}

/* The reference's BVH in the reference's order. Later candidates with equal t overwrite. */
template <bool ANY, bool COUNT>
__device__ __forceinline__ void trace_reference(const DevScene& S, d3 o, d3 d, double tmin, double tmax, Hit& hit, VisitCounters& cnt) {
    hit.slot = -1;
    hit.t = 0.0; hit.alpha = 0.0; hit.beta = 0.0;
    if (S.ref_root < 0) return;
    const d3 inv = mk3(1.0 / d.x, 1.0 / d.y, 1.0 / d.z);
    int32_t stack[PRT_STACK];
    int sp = 0;
    int32_t cur = S.ref_root;
    double tbest = tmax;
    for (;;) {
        if (cur >= 0) {
            const double2* __restrict__ p = reinterpret_cast<const double2*>(S.ref_nodes + cur);
            const double2 bx = __ldg(p), by = __ldg(p + 1), bz = __ldg(p + 2);
            const int4 ch = __ldg(reinterpret_cast<const int4*>(p + 3));
            PRT_COUNT_NODE(cnt);
            double lo = tmin, hi = tbest;
            bool ok = true;
            slab_axis(bx.x, bx.y, o.x, inv.x, lo, hi);
            if (hi <= lo) ok = false;
            if (ok) { slab_axis(by.x, by.y, o.y, inv.y, lo, hi); if (hi <= lo) ok = false; }
            if (ok) { slab_axis(bz.x, bz.y, o.z, inv.z, lo, hi); if (hi <= lo) ok = false; }
            if (ok) {
                /* left, then right (BVH.cpp:57-58); a span-1 node has left == right (BVH.cpp:21-24)
                 * and testing the same object again with max = its own t changes nothing */
                if (ch.y != ch.x) {
                    if (sp < PRT_STACK) stack[sp++] = ch.y;
                    else if (S.ref_overflow) atomicAdd(S.ref_overflow, 1ull); /* a dropped subtree is an ERROR: sticky, prt_get_stats reports it */
                }
                cur = ch.x;
                continue;
            }
        } else {
            const int32_t slot = ~cur;
            double t, a, b;
            PRT_COUNT_TRI(cnt);
            if (tri_test(S.tris + slot, o, d, tmin, tbest, t, a, b)) {
                hit.slot = slot; hit.t = t; hit.alpha = a; hit.beta = b;
                tbest = t;
                if (ANY) return;
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
}

/* A hit within PRT_EDGE_EPS (barycentric units) of the triangle's boundary. Only there can the
 * reference's answer depend on its own tree: (a) its box test rejects a degenerate interval
 * (`t.max <= t.min`, AABB.cpp:62), which happens for a ray that enters a leaf / ancestor box exactly
 * where it meets the triangle — boxes touch their triangles only at vertices and edges; (b) two
 * triangles sharing that edge / vertex are met at the very same t and the later one in the
 * reference's visiting order wins (Interval.h:21-23, BVH.cpp:58). Rounding decides those cases
 * within ~1e-12 of the boundary; the margin is four orders wider. */
#define PRT_EDGE_EPS 1e-7
__device__ __forceinline__ bool on_boundary(double alpha, double beta) {
    return alpha < PRT_EDGE_EPS || beta < PRT_EDGE_EPS || 1.0 - (alpha + beta) < PRT_EDGE_EPS;
}

}  // namespace prt
#endif

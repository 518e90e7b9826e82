/*
 * kernels.cu — the sm_100a kernels of the render path.
 *
 *   k_trace_closest / k_trace_any   ray-query batches = Hittable::Hit on `world` (Source/Hittable.h:34)
 *   k_sample_lights                 Hittable::Sample on `lights` (Source/Hittable.h:36)
 *   camera_ray                      Camera::GetRay (Source/Camera.cpp:108-117), fused into bounce 0 of k_extend
 *   k_extend                        closest hit of every live path + the terminal cases of
 *                                   Camera::RayColor (miss / emitter, Camera.cpp:125-132,187-196),
 *                                   hits binned by material class
 *   k_shade<class>                  one path vertex: NEE set-up, Russian roulette, BSDF sampling
 *                                   (Camera.cpp:134-201)
 *   k_shadow                        visibility of the NEE sample (Camera.cpp:143-155)
 *   k_accumulate                    colorAttachment[m] += L * 1/spp in sample order (Camera.cpp:56)
 *   k_resolve                       the pixel loop of Camera::WriteColorAttachment (Camera.cpp:279-331)
 *
 * Camera::RayColor's recursion  L = direct + att * L' / rr  becomes a loop carrying the throughput
 * T = prod(att / rr): the wavefront traces every scattered ray once (the reference traces it twice:
 * the "peek" at Camera.cpp:187 and again as the first act of the recursive call).
 *
 * Compiled with -fmad=false: FP64 results must carry the reference's bits, and its x86-64 objects
 * contain no fused multiply-add.
 */
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <float.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <utility>

#include "../../include/prt_b200.h"
#include "launch.h"
#include "rng.cuh"
#include "shading.cuh"
#include "cook_torrance.cuh"
#include "traverse.cuh"
#include "traverse_warp.cuh"

#ifndef PRT_TRAVERSE_MIN_BLOCKS
#define PRT_TRAVERSE_MIN_BLOCKS 6
#endif

namespace cg = cooperative_groups;

namespace prt {

#ifndef PRT_EXTEND_DEFER
#define PRT_EXTEND_DEFER 0
#endif
/* PathRec::dimflags bit 31: the vertex that scattered this ray skipped light sampling (Camera.cpp:191-195) */

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

/* warp-aggregated append: one atomic per warp. Must be called by all 32 lanes. */
__device__ __forceinline__ int warp_push(int32_t* counter, bool pred) {
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0) return -1;
    const int leader = __ffs(mask) - 1;
    int base = 0;
    if (lane_id() == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return pred ? base + __popc(mask & ((1u << lane_id()) - 1u)) : -1;
}

/* persistent-thread work fetch: a warp takes the next 32 queue entries. All lanes call. */
__device__ __forceinline__ int warp_fetch(int32_t* cursor) {
    int base = 0;
    if (lane_id() == 0) base = atomicAdd(cursor, 32);
    return __shfl_sync(0xffffffffu, base, 0);
}

template <bool COUNT>
__device__ __forceinline__ void flush_counters(const VisitCounters& c, unsigned long long* out) {
    if (!COUNT || !out) return;
    unsigned long long n = c.nodes, t = c.tris;
    for (int o = 16; o > 0; o >>= 1) {
        n += __shfl_down_sync(0xffffffffu, n, o);
        t += __shfl_down_sync(0xffffffffu, t, o);
    }
    if (lane_id() == 0) {
        atomicAdd(out, n);
        atomicAdd(out + 1, t);
    }
    if (c.rechecks) { /* rare: out[4..6] = re-traced rays, their fetches, the most one ray fetched */
        atomicAdd(out + 4, c.rechecks);
        atomicAdd(out + 5, c.recheck_visits);
        atomicMax(out + 6, c.recheck_max);
    }
}

/* ------------------------------------------------------------------------------------------------
 * ray binning (launch.h): key, scan, scatter
 * ---------------------------------------------------------------------------------------------- */
__device__ __forceinline__ uint32_t spread3(uint32_t v) { /* 4 bits -> every third bit */
    v = (v | (v << 4)) & 0x0c3u;
    v = (v | (v << 2)) & 0x249u;
    return v;
}
__device__ __forceinline__ uint32_t ray_key(const float* lo, const float* scale, d3 o, d3 d) {
    const float G = (float)(PRT_KEY_GRID - 1);
    /* fminf / fmaxf drop a NaN coordinate: such a ray lands in cell 0, which is as good as any */
    const uint32_t ix = (uint32_t)fminf(fmaxf(((float)o.x - lo[0]) * scale[0], 0.f), G);
    const uint32_t iy = (uint32_t)fminf(fmaxf(((float)o.y - lo[1]) * scale[1], 0.f), G);
    const uint32_t iz = (uint32_t)fminf(fmaxf(((float)o.z - lo[2]) * scale[2], 0.f), G);
    const uint32_t oct = (d.x < 0.0 ? 1u : 0u) | (d.y < 0.0 ? 2u : 0u) | (d.z < 0.0 ? 4u : 0u);
    return ((spread3(ix) | (spread3(iy) << 1) | (spread3(iz) << 2)) << 3) | oct;
}

/* one block per histogram: exclusive scan of hist -> cursor, hist zeroed for its next use */
__global__ void __launch_bounds__(1024) k_sort_scan(int32_t* __restrict__ hist, int32_t* __restrict__ cursor) {
    constexpr int PER = PRT_KEY_BINS / 1024;
    __shared__ int32_t warp_sum[32];
    int32_t* h = hist + (size_t)blockIdx.x * PRT_KEY_BINS;
    int32_t* c = cursor + (size_t)blockIdx.x * PRT_KEY_BINS;
    int32_t v[PER];
    int32_t sum = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        v[k] = h[threadIdx.x * PER + k];
        sum += v[k];
    }
    int32_t incl = sum;
    for (int o = 1; o < 32; o <<= 1) {
        const int32_t up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane_id() >= o) incl += up;
    }
    if (lane_id() == 31) warp_sum[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        int32_t w = warp_sum[threadIdx.x];
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t up = __shfl_up_sync(0xffffffffu, w, o);
            if (lane_id() >= o) w += up;
        }
        warp_sum[threadIdx.x] = w;
    }
    __syncthreads();
    int32_t run = incl - sum + ((threadIdx.x >> 5) ? warp_sum[(threadIdx.x >> 5) - 1] : 0);
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        c[threadIdx.x * PER + k] = run;
        run += v[k];
        h[threadIdx.x * PER + k] = 0;
    }
}

/* out[cursor[key]++] = in[qi] (in == nullptr: the identity). blockIdx.y picks the queue. */
struct SortJob {
    const int32_t* n;  /* queue length on the device (nullptr: n_host) */
    int32_t n_host;
    const uint16_t* keys;
    const int32_t* in;
    int32_t* out;
};
__global__ void __launch_bounds__(256) k_sort_scatter(SortJob j0, SortJob j1, int32_t* __restrict__ cursor) {
    const SortJob j = blockIdx.y == 0 ? j0 : j1;
    int32_t* c = cursor + (size_t)blockIdx.y * PRT_KEY_BINS;
    const int32_t n = j.n ? *j.n : j.n_host;
    for (int32_t qi = blockIdx.x * blockDim.x + threadIdx.x; qi < n; qi += gridDim.x * blockDim.x) {
        const int32_t pos = atomicAdd(c + j.keys[qi], 1);
        j.out[pos] = j.in ? j.in[qi] : qi;
    }
}

/* keys + histogram of a ray-query batch */
struct KeyBox { float lo[3], scale[3]; };
__global__ void __launch_bounds__(256) k_query_keys(int64_t n, const double* __restrict__ o, const double* __restrict__ d, KeyBox kb,
                                                    uint16_t* __restrict__ keys, int32_t* __restrict__ hist) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t k = ray_key(kb.lo, kb.scale, mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
        keys[i] = (uint16_t)k;
        atomicAdd(hist + k, 1);
    }
}

/* ------------------------------------------------------------------------------------------------
 * ray queries
 * ---------------------------------------------------------------------------------------------- */
struct ClosestQueryPolicy {
    static constexpr int REFILL_BELOW = PRT_REFILL_BELOW_QUERY;
    const double *o, *d, *tmin, *tmax;
    int32_t* prim;
    double *t, *alpha, *beta;
    const int32_t* slot_prim;
    const int32_t* __restrict__ order; /* binned ray order, or nullptr */
    __device__ __forceinline__ int32_t slot_of(int32_t qi) const { return order ? order[qi] : qi; }
    __device__ __forceinline__ void prefetch(int32_t i) const {
        prefetch_l2(o + 3 * (int64_t)i);
        prefetch_l2(d + 3 * (int64_t)i);
    }
    __device__ __forceinline__ bool load(int32_t i, d3& ro, d3& rd, double& t0, double& t1) const {
        ro = mk3(o[3 * (int64_t)i], o[3 * (int64_t)i + 1], o[3 * (int64_t)i + 2]);
        rd = mk3(d[3 * (int64_t)i], d[3 * (int64_t)i + 1], d[3 * (int64_t)i + 2]);
        t0 = tmin[i];
        t1 = tmax[i];
        return true;
    }
    __device__ __forceinline__ void retire(bool finished, int32_t i, const Hit& h) const {
        if (!finished) return;
        prim[i] = h.slot >= 0 ? slot_prim[h.slot] : -1;
        if (t) t[i] = h.t;
        if (alpha) alpha[i] = h.alpha;
        if (beta) beta[i] = h.beta;
    }
    __device__ __forceinline__ void finish() const {}
};
struct AnyQueryPolicy {
    static constexpr int REFILL_BELOW = PRT_REFILL_BELOW_QUERY;
    const double *o, *d, *tmin, *tmax;
    uint8_t* hit;
    const int32_t* __restrict__ order;
    __device__ __forceinline__ int32_t slot_of(int32_t qi) const { return order ? order[qi] : qi; }
    __device__ __forceinline__ void prefetch(int32_t i) const {
        prefetch_l2(o + 3 * (int64_t)i);
        prefetch_l2(d + 3 * (int64_t)i);
    }
    __device__ __forceinline__ bool load(int32_t i, d3& ro, d3& rd, double& t0, double& t1) const {
        ro = mk3(o[3 * (int64_t)i], o[3 * (int64_t)i + 1], o[3 * (int64_t)i + 2]);
        rd = mk3(d[3 * (int64_t)i], d[3 * (int64_t)i + 1], d[3 * (int64_t)i + 2]);
        t0 = tmin[i];
        t1 = tmax[i];
        return true;
    }
    __device__ __forceinline__ void retire(bool finished, int32_t i, const Hit& h) const {
        if (finished) hit[i] = h.slot >= 0 ? 1 : 0;
    }
    __device__ __forceinline__ void finish() const {}
};

template <bool COUNT, int WIDE>
__global__ void __launch_bounds__(PRT_TRAVERSE_BLOCK, PRT_TRAVERSE_MIN_BLOCKS) k_trace_closest(DevScene S, int64_t n, const double* __restrict__ o, const double* __restrict__ d,
                                                       const double* __restrict__ tmin, const double* __restrict__ tmax,
                                                       int32_t* __restrict__ prim, double* __restrict__ t, double* __restrict__ alpha,
                                                       double* __restrict__ beta, const int32_t* __restrict__ order, int32_t* fetch, unsigned long long* counters) {
    __shared__ int32_t sm_stack[PRT_SMEM_STACK_WORDS];
    VisitCounters cnt = {0, 0, 0, 0, 0};
    ClosestQueryPolicy pol = {o, d, tmin, tmax, prim, t, alpha, beta, S.slot_prim, order};
    traverse_persistent<false, COUNT, WIDE>(S, pol, fetch, (int32_t)n, cnt, sm_stack);
    flush_counters<COUNT>(cnt, counters);
}

template <bool COUNT, int WIDE>
__global__ void __launch_bounds__(PRT_TRAVERSE_BLOCK, PRT_TRAVERSE_MIN_BLOCKS) k_trace_any(DevScene S, int64_t n, const double* __restrict__ o, const double* __restrict__ d,
                                                   const double* __restrict__ tmin, const double* __restrict__ tmax, uint8_t* __restrict__ hit,
                                                   const int32_t* __restrict__ order, int32_t* fetch, unsigned long long* counters) {
    __shared__ int32_t sm_stack[PRT_SMEM_STACK_WORDS];
    VisitCounters cnt = {0, 0, 0, 0, 0};
    AnyQueryPolicy pol = {o, d, tmin, tmax, hit, order};
    traverse_persistent<true, COUNT, WIDE>(S, pol, fetch, (int32_t)n, cnt, sm_stack);
    flush_counters<COUNT>(cnt, counters);
}

__global__ void k_sample_lights(DevScene S, int64_t n, const double* __restrict__ origins, const double* __restrict__ u,
                                double* __restrict__ position, double* __restrict__ normal, uint8_t* __restrict__ front,
                                double* __restrict__ pdf, int32_t* __restrict__ prim) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    LightSample ls;
    sample_lights(S, mk3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), u[4 * i], u[4 * i + 1], u[4 * i + 2], u[4 * i + 3], ls);
    position[3 * i] = ls.position.x; position[3 * i + 1] = ls.position.y; position[3 * i + 2] = ls.position.z;
    normal[3 * i] = ls.normal.x; normal[3 * i + 1] = ls.normal.y; normal[3 * i + 2] = ls.normal.z;
    front[i] = ls.front ? 1 : 0;
    pdf[i] = ls.pdf;
    prim[i] = ls.prim;
}

/* ------------------------------------------------------------------------------------------------
 * wavefront integrator
 * ---------------------------------------------------------------------------------------------- */
__device__ __forceinline__ d3 ld3(const double* p) { return mk3(p[0], p[1], p[2]); }
__device__ __forceinline__ void st3(double* p, d3 v) {
    p[0] = v.x; p[1] = v.y; p[2] = v.z;
}
/* sector k of a path record / the radiance cell of a slot (launch.h) */
__device__ __forceinline__ char* rec_sector(const PathBuffers& pb, int32_t slot, int k) { return reinterpret_cast<char*>(pb.rec + slot) + 32 * k; }
__device__ __forceinline__ void lum_add(const PathBuffers& pb, int32_t slot, d3 v) {
    const dd4 l = ld256g(pb.lum + slot);
    st256g(pb.lum + slot, l.v[0] + v.x, l.v[1] + v.y, l.v[2] + v.z, 0.0);
}
__device__ __forceinline__ void lum_set(const PathBuffers& pb, int32_t slot, d3 v) { st256g(pb.lum + slot, v.x, v.y, v.z, 0.0); }

/* Camera::GetRay (Camera.cpp:108-117): pixel centre, direction not normalised. slot = sample * npix + pixel */
__device__ __forceinline__ void camera_ray(const RenderConst& rc, int32_t slot, d3& o, d3& d) {
    uint32_t smp, pix, i, j;
    rc.by_npix.divmod((uint32_t)slot, smp, pix);
    rc.by_width.divmod(pix, j, i);
    double fi = (double)(int)i, fj = (double)(int)j;
    if (rc.flags & PRT_RENDER_PIXEL_JITTER) {
        /* the commented-out lines of Camera.cpp:110-111: + SampleSquare(), the first two draws of the stream */
        Rng rng;
        rng.seed = stream_seed(rc.rng_key, pix, (uint32_t)rc.sample_first + smp);
        rng.dim = 0u;
        fi = fi + (rng.next() - 0.5);
        fj = fj + (rng.next() - 0.5);
    }
    const d3 ps = rc.pixel00 + fi * rc.dU + fj * rc.dV;
    o = rc.center;
    d = ps - rc.center;
}

/* deferred warp-aggregated queue push: issue() starts one atomic per destination queue (the leader
 * of a class is a lane of that class, so one register holds every leader's result), complete()
 * finishes it an event later, when the atomic's round trip to L2 has long been paid */
struct DeferredPush {
    int32_t cls, slot, raw;
    uint32_t key; /* ray-binning key of the pushed ray (used by complete_keyed only) */
    __device__ __forceinline__ void init() { cls = -1; slot = 0; raw = 0; key = 0; }
    /* single queue + its key array: the entry and its key land at the same index */
    __device__ __forceinline__ void complete_keyed(int32_t* queue, uint16_t* keys) {
        const unsigned mask = __ballot_sync(0xffffffffu, cls == 0);
        if (mask != 0) {
            const int32_t base = __shfl_sync(0xffffffffu, raw, __ffs(mask) - 1);
            if (cls == 0) {
                const int32_t at = base + __popc(mask & ((1u << lane_id()) - 1u));
                queue[at] = slot;
                keys[at] = (uint16_t)key;
            }
        }
        cls = -1;
    }
    template <int NQ>
    __device__ __forceinline__ void issue(int32_t* const (&counters)[NQ], int c, int32_t s) {
        cls = c;
        slot = s;
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const unsigned mask = __ballot_sync(0xffffffffu, c == q);
            if (mask != 0 && lane_id() == __ffs(mask) - 1) raw = atomicAdd(counters[q], __popc(mask));
        }
    }
    template <int NQ>
    __device__ __forceinline__ void complete(int32_t* const (&queues)[NQ]) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const unsigned mask = __ballot_sync(0xffffffffu, cls == q);
            if (mask != 0) {
                const int32_t base = __shfl_sync(0xffffffffu, raw, __ffs(mask) - 1);
                if (cls == q) queues[q][base + __popc(mask & ((1u << lane_id()) - 1u))] = slot;
            }
        }
        cls = -1;
    }
};

/* What a ray that ends its path at this hit leaves behind, without any BSDF (Camera.cpp:125-132,187-196).
 * Returns the material class of the hit (CLS_SHADED.. = the path ends here) and, for a path that ends, the
 * radiance `le` it collects BEFORE weighting (background / emission / 0). */
__device__ __forceinline__ int terminal_radiance(const DevScene& S, const uint8_t* __restrict__ prim_class, const RenderConst& rc, const Hit& h,
                                                 int32_t& prim, d3& le) {
    le = mk3(0.0, 0.0, 0.0);
    prim = -1;
    if (h.slot < 0) {
        le = mk3(rc.background[0], rc.background[1], rc.background[2]);
        return CLS_EMPTY;
    }
    prim = S.slot_prim[h.slot];
    const int c = prim_class[prim];
    if (c == CLS_EMISSIVE) {
        const DevMaterial* m = S.materials + S.shade[prim].material;
        le = mk3(m->emission[0], m->emission[1], m->emission[2]);
    }
    return c;
}

/* Bounce 0, once per PIXEL (PrimaryHit, launch.h): the closest hit of the camera ray through the pixel centre,
 * shared by every sample of the pixel. Pixels whose path goes on are queued per material class; k_shade<CLS, true>
 * expands each queued pixel into its n_samples path slots. */
struct PrimaryPolicy {
    static constexpr int REFILL_BELOW = PRT_REFILL_BELOW_QUERY;
    const DevScene& S;
    const uint8_t* __restrict__ prim_class;
    const RenderConst& rc;
    const PathBuffers& pb;
    BounceCtrl* ctrl;
    DeferredPush push;
    __device__ __forceinline__ int32_t slot_of(int32_t qi) const { return qi; }
    __device__ __forceinline__ void prefetch(int32_t) const {}
    __device__ __forceinline__ bool load(int32_t pix, d3& o, d3& d, double& t0, double& t1) const {
        camera_ray(rc, pix, o, d);
        t0 = 0.0001; /* Camera.cpp:125 */
        t1 = (double)INFINITY;
        return true;
    }
    __device__ __forceinline__ void flush() {
        int32_t* const queues[CLS_SHADED] = {pb.q_mat[0], pb.q_mat[1], pb.q_mat[2], pb.q_mat[3], pb.q_mat[4]};
        push.complete(queues);
    }
    __device__ __forceinline__ void retire(bool finished, int32_t pix, const Hit& h) {
        int cls = -1;
        if (finished) {
            int32_t prim;
            d3 le;
            const int c = terminal_radiance(S, prim_class, rc, h, prim, le);
            if (c < CLS_SHADED) {
                st256g(pb.primary + pix, h.t, h.alpha, h.beta, pack2((uint32_t)prim, (uint32_t)c));
                cls = c;
            } else {
                /* a primary miss returns the background (Camera.cpp:127), an emitter its radiance (:129-132), Empty nothing */
                st256g(pb.primary + pix, le.x, le.y, le.z, pack2((uint32_t)prim, (uint32_t)c));
            }
        }
        int32_t* const counters[CLS_SHADED] = {&ctrl->n_mat[0], &ctrl->n_mat[1], &ctrl->n_mat[2], &ctrl->n_mat[3], &ctrl->n_mat[4]};
        push.issue(counters, cls, pix);
        flush();
    }
    __device__ __forceinline__ void finish() {}
};

template <bool COUNT, int WIDE>
__global__ void __launch_bounds__(PRT_TRAVERSE_BLOCK, PRT_TRAVERSE_MIN_BLOCKS) k_primary(DevScene S, const uint8_t* __restrict__ prim_class, RenderConst rc, PathBuffers pb) {
    __shared__ int32_t sm_stack[PRT_SMEM_STACK_WORDS];
    VisitCounters cnt = {0, 0, 0, 0, 0};
    PrimaryPolicy pol = {S, prim_class, rc, pb, pb.ctrl, {-1, 0, 0, 0}};
    traverse_persistent<false, COUNT, WIDE>(S, pol, &pb.ctrl->fetch_extend, rc.npix, cnt, sm_stack);
    flush_counters<COUNT>(cnt, pb.counters);
}

/* closest hit for every path in the bounce's queue (persistent warps with lane refill,
 * traverse_warp.cuh), then the terminal cases that need no BSDF, then one queue per material class
 * for the shade kernels. Without primary-hit reuse (PRT_RENDER_PIXEL_JITTER) bounce 0 runs here too and
 * generates its camera rays itself. */
struct ExtendPolicy {
    static constexpr int REFILL_BELOW = PRT_REFILL_BELOW_RENDER;
    const DevScene& S;
    const uint8_t* __restrict__ prim_class;
    const RenderConst& rc;
    const PathBuffers& pb;
    BounceCtrl* ctrl;
    const int32_t* queue; /* read with ld.cg: k_tail rewrites the queues inside one launch */
    int bounce;
    DeferredPush push;
    __device__ __forceinline__ int32_t slot_of(int32_t qi) const { return bounce == 0 ? qi : __ldcg(queue + qi); }
    __device__ __forceinline__ void prefetch(int32_t slot) const {
        if (bounce != 0) prefetch_l2(pb.rec + slot);
    }
    __device__ __forceinline__ bool load(int32_t slot, d3& o, d3& d, double& t0, double& t1) const {
        if (bounce == 0) {
            camera_ray(rc, slot, o, d);
        } else {
            const dd4 s0 = ld256g(rec_sector(pb, slot, 0)), s1 = ld256g(rec_sector(pb, slot, 1));
            o = mk3(s0.v[0], s0.v[1], s0.v[2]);
            d = mk3(s0.v[3], s1.v[0], s1.v[1]);
        }
        t0 = 0.0001; /* Camera.cpp:125 */
        t1 = (double)INFINITY;
        return true;
    }
    __device__ __forceinline__ void flush() {
        int32_t* const queues[CLS_SHADED] = {pb.q_mat[0], pb.q_mat[1], pb.q_mat[2], pb.q_mat[3], pb.q_mat[4]};
        push.complete(queues);
    }
    __device__ __forceinline__ void retire(bool finished, int32_t slot, const Hit& h) {
        flush();
        int cls = -1;
        if (finished) {
            int32_t prim;
            d3 le;
            const int c = terminal_radiance(S, prim_class, rc, h, prim, le);
            if (c < CLS_SHADED) {
                st256g(rec_sector(pb, slot, 4), h.t, h.alpha, h.beta, pack2((uint32_t)prim, 0u));
                if (bounce == 0) lum_set(pb, slot, mk3(0.0, 0.0, 0.0)); /* bounce 0 initialises the radiance of every path */
                cls = c;
            } else if (bounce == 0) {
                lum_set(pb, slot, le); /* background (Camera.cpp:127), emitter (:129-132) or Empty's nothing */
            } else if (h.slot < 0) {
                /* with NEE on, a scattered ray that misses adds nothing (Camera.cpp:187-196) */
                if (!rc.sample_lights) {
                    const dd4 s1 = ld256g(rec_sector(pb, slot, 1)), s2 = ld256g(rec_sector(pb, slot, 2));
                    lum_add(pb, slot, mk3(s1.v[2], s1.v[3], s2.v[0]) * le);
                }
            } else if (c == CLS_EMISSIVE) {
                /* an emitter ends the path (Camera.cpp:129-132); with NEE on, a scattered ray may
                 * collect it only if the scattering material skipped light sampling (:191-195) */
                const dd4 s1 = ld256g(rec_sector(pb, slot, 1)), s2 = ld256g(rec_sector(pb, slot, 2));
                const d3 T = mk3(s1.v[2], s1.v[3], s2.v[0]);
                const bool prev_skip = (unpack_lo(s2.v[1]) >> 31) != 0u;
                const float prev_pdf = __uint_as_float(unpack_hi(s2.v[1]));
                if (!rc.sample_lights || prev_skip) {
                    lum_add(pb, slot, T * le);
                } else if ((rc.flags & PRT_RENDER_MIS) && prev_pdf > 0.f) {
                    /* PRT_RENDER_MIS: the emitter the reference drops (Camera.cpp:187-196), weighted against the
                     * light sampler's density for this point (0 if it shows its back). T already holds att / rr. */
                    const TriShade* __restrict__ sh = S.shade + prim;
                    const dd4 s0 = ld256g(rec_sector(pb, slot, 0));
                    const d3 dir = mk3(s0.v[3], s1.v[0], s1.v[1]), n = mk3(sh->normal[0], sh->normal[1], sh->normal[2]);
                    const double pbsdf = (double)prev_pdf;
                    const double d2 = dot(dir, dir);
                    const double dn = dot(dir, n);
                    const double cosl = fabs(dn) / sqrt(d2);
                    double pl = 0.0;
                    if (dn < 0.0 && cosl > 0.0) pl = (1.0 / S.lights_area) * (h.t * h.t * d2) / cosl;
                    lum_add(pb, slot, T * le * (pbsdf / (pbsdf + pl)));
                }
            }
            /* CLS_EMPTY past bounce 0: no NEE, Scatter returns false (Material.h:537-540): the path just ends */
        }
        int32_t* const counters[CLS_SHADED] = {&ctrl->n_mat[0], &ctrl->n_mat[1], &ctrl->n_mat[2], &ctrl->n_mat[3], &ctrl->n_mat[4]};
        push.issue(counters, cls, slot);
#if !PRT_EXTEND_DEFER
        flush();
#endif
    }
    __device__ __forceinline__ void finish() { flush(); }
};

template <bool COUNT, int WIDE>
__device__ __forceinline__ void extend_body(const DevScene& S, const uint8_t* __restrict__ prim_class, const RenderConst& rc, const PathBuffers& pb,
                                            int bounce, int sorted, VisitCounters& cnt, int32_t* sm_stack) {
    BounceCtrl* ctrl = pb.ctrl + bounce;
    const int32_t n = bounce == 0 ? rc.n_paths : __ldcg(&ctrl->n_extend);
    ExtendPolicy pol = {S, prim_class, rc, pb, ctrl, sorted ? pb.qs_extend : pb.q_extend[bounce & 1], bounce, {-1, 0, 0, 0}};
    traverse_persistent<false, COUNT, WIDE>(S, pol, &ctrl->fetch_extend, n, cnt, sm_stack);
}

template <bool COUNT, int WIDE>
__global__ void __launch_bounds__(PRT_TRAVERSE_BLOCK, PRT_TRAVERSE_MIN_BLOCKS) k_extend(DevScene S, const uint8_t* __restrict__ prim_class, RenderConst rc, PathBuffers pb, int bounce, int sorted) {
    __shared__ int32_t sm_stack[PRT_SMEM_STACK_WORDS];
    VisitCounters cnt = {0, 0, 0, 0, 0};
    extend_body<COUNT, WIDE>(S, prim_class, rc, pb, bounce, sorted, cnt, sm_stack);
    flush_counters<COUNT>(cnt, pb.counters);
}

/* one path vertex of material class CLS (Camera.cpp:134-201). Grid-stride over the class queue with
 * the queue entry read two iterations ahead and the path record prefetched one iteration ahead (the
 * chain queue -> record -> triangle -> material was 30 % of the stall samples, profiles/r1_s1_shade).
 *
 * PRIMARY (bounce 0 with primary-hit reuse): the queue holds PIXELS; entry qi names sample qi / nq of pixel
 * queue[qi % nq], i.e. slot = sample * npix + pixel — sample-major, so that a warp's 32 vertices sit in 32
 * neighbouring records. The hit comes from the per-pixel table, the ray is recomputed, T = 1, L = 0: the
 * kernel reads no path record at all. */
#ifndef PRT_SHADE_MIN_BLOCKS
#define PRT_SHADE_MIN_BLOCKS 4
#endif
#ifndef PRT_SHADE_PREFETCH
#define PRT_SHADE_PREFETCH 1
#endif
/* The vertex itself: everything between "this slot holds a hit on a surface of class CLS" and "these are its
 * shadow ray and its next segment" (Camera.cpp:134-201). prim_hint >= 0: the hit primitive, already known to the
 * caller (read an iteration ahead so that its TriShade line could be prefetched). */
template <int CLS, bool PRIMARY>
__device__ __forceinline__ void shade_vertex(const DevScene& S, const RenderConst& rc, const PathBuffers& pb, int bounce, int sort, int32_t slot,
                                             bool& want_shadow, bool& want_extend, uint32_t& key_shadow, uint32_t& key_extend) {
    /* ---- the vertex: hit, incoming ray, throughput, RNG dimension ---- */
    uint32_t slot_smp, slot_pix; /* slot = sample_in_pass * npix + pixel */
    rc.by_npix.divmod((uint32_t)slot, slot_smp, slot_pix);
    int32_t prim;
    double t, alpha, beta;
    d3 ro, rd, T;
    uint32_t dim;
    if (PRIMARY) {
        const dd4 ph = ld256d(pb.primary + (uint32_t)slot_pix); /* written by the previous launch: read-only here */
        t = ph.v[0]; alpha = ph.v[1]; beta = ph.v[2];
        prim = (int32_t)unpack_lo(ph.v[3]);
    } else {
        const dd4 s4 = ld256g(rec_sector(pb, slot, 4));
        t = s4.v[0]; alpha = s4.v[1]; beta = s4.v[2];
        prim = (int32_t)unpack_lo(s4.v[3]);
    }
    if (PRIMARY || bounce == 0) {
        /* bounce 0: the camera ray and T = 1 are recomputed instead of being written by the extend stage and
         * read back; the stream starts at dimension 0 (2 after the jitter draws) */
        camera_ray(rc, slot, ro, rd);
        T = mk3(1.0, 1.0, 1.0);
        dim = (rc.flags & PRT_RENDER_PIXEL_JITTER) ? 2u : 0u;
    } else {
        const dd4 s0 = ld256g(rec_sector(pb, slot, 0)), s1 = ld256g(rec_sector(pb, slot, 1)), s2 = ld256g(rec_sector(pb, slot, 2));
        ro = mk3(s0.v[0], s0.v[1], s0.v[2]);
        rd = mk3(s0.v[3], s1.v[0], s1.v[1]);
        T = mk3(s1.v[2], s1.v[3], s2.v[0]);
        dim = unpack_lo(s2.v[1]) & 0x7fffffffu;
    }
    const TriShade* __restrict__ sh = S.shade + prim;
    const DevMaterial m = S.materials[sh->material];
    /* the HitRecord of Triangle::Hit (Triangle.cpp:76-80,111) */
    Surface s;
    s.position = ro + rd * t;
    s.uv = (1. - alpha - beta) * mk2(sh->uv[0], sh->uv[1]) + alpha * mk2(sh->uv[2], sh->uv[3]) + beta * mk2(sh->uv[4], sh->uv[5]);
    s.tangent = mk3(sh->tangent[0], sh->tangent[1], sh->tangent[2]);
    const d3 gn = mk3(sh->normal[0], sh->normal[1], sh->normal[2]);
    s.front = dot(rd, gn) < 0.; /* SetFaceNormal, Hittable.cpp:8-13 */
    s.normal = s.front ? gn : -gn;
    const d3 wo = world_to_local(-rd, s);

    Rng rng;
    rng.seed = stream_seed(rc.rng_key, slot_pix, (uint32_t)rc.sample_first + slot_smp);
    rng.dim = dim;

    /* ---- next-event estimation (Camera.cpp:137-174) ---- */
    d3 lp = mk3(0.0, 0.0, 0.0), sc = mk3(0.0, 0.0, 0.0);
    if (rc.sample_lights && !m.skip_light_sampling) {
        const double u0 = rng.next(), u1 = rng.next(), u2 = rng.next(), u3 = rng.next();
        LightSample ls;
        sample_lights(S, s.position, u0, u1, u2, u3, ls, (rc.flags & PRT_RENDER_UNBIASED_LIGHT_PICK) != 0);
        const d3 to_light = ls.position - s.position;
        const d3 light_dir = normalize(to_light);
        const double distance = length(to_light);
        if (dot(s.normal, light_dir) > 0.0 && ls.front) {
            const DevMaterial* lm = S.materials + ls.material;
            const d3 emission = mk3(lm->emission[0], lm->emission[1], lm->emission[2]);
            const d3 local_wi = world_to_local(light_dir, s);
            const d3 local_ln = world_to_local(ls.normal, s);
            const d3 fr = material_eval<CLS == CLS_COOK, CLS == CLS_PHONG || CLS == CLS_PHONG_SPEC, CLS == CLS_PHONG_SPEC>(S, m, local_wi, wo, s.uv, rng);
            const double cos_theta = local_wi.z;
            const double cos_theta_bar = dot(local_ln, -local_wi);
            d3 direct = emission * fr * cos_theta * cos_theta_bar / (distance * distance) / ls.pdf;
            if ((rc.flags & PRT_RENDER_MIS) && (CLS == CLS_LAMBERT || (CLS == CLS_PHONG && m.pkd == 1.0))) {
                /* balance heuristic against the cosine lobe's density for the same direction */
                const double pl = ls.pdf * (distance * distance) / cos_theta_bar;
                const double pbsdf = cos_theta * PRT_INV_PI;
                direct = direct * (pl / (pl + pbsdf));
            }
            /* the shadow ray is stored as its target: k_shadow recomputes light_dir and distance from
             * (lp, o = s.position) with the functions used above */
            lp = ls.position;
            sc = T * direct;
            want_shadow = true;
            if (sort) key_shadow = ray_key(rc.key_lo, rc.key_scale, s.position, light_dir);
        }
    }
    /* ---- Russian roulette + BSDF sampling (Camera.cpp:176-201) ---- */
    d3 out_dir = rd, Tn = T;
    uint32_t dimflags = dim;
    float prev_pdf = 0.f;
    if (rng.next() < rc.rr) {
        d3 att, sd;
        bool scattered = true;
        double cosine_pdf = 0.0;
        if (CLS == CLS_LAMBERT) scatter_lambertian(S, m, s, rng, att, sd, cosine_pdf);
        else if (CLS == CLS_PHONG) scatter_phong<false>(S, m, s, wo, rng, att, sd, cosine_pdf, (rc.flags & PRT_RENDER_LOBE_PROBABILITY) != 0);
        else if (CLS == CLS_PHONG_SPEC) scatter_phong<true>(S, m, s, wo, rng, att, sd, cosine_pdf, (rc.flags & PRT_RENDER_LOBE_PROBABILITY) != 0);
        else if (CLS == CLS_COOK) scattered = scatter_cook(m, s, wo, rng, att, sd);
        else scatter_mirror(s, wo, att, sd);
        /* RayColor(depth - 1) of the last allowed vertex returns 0 (Camera.cpp:121) */
        if (scattered && bounce < rc.max_depth) {
            out_dir = sd;
            Tn = T * att / rc.rr;
            dimflags = rng.dim | (m.skip_light_sampling ? 0x80000000u : 0u);
            prev_pdf = (rc.flags & PRT_RENDER_MIS) ? (float)cosine_pdf : 0.f;
            want_extend = true;
            if (sort) key_extend = ray_key(rc.key_lo, rc.key_scale, s.position, out_dir);
        }
    }
    /* ---- the record: whole sectors, only those a later stage reads ---- */
    if (want_extend || want_shadow) st256g(rec_sector(pb, slot, 0), s.position.x, s.position.y, s.position.z, out_dir.x);
    if (want_extend) st256g(rec_sector(pb, slot, 1), out_dir.y, out_dir.z, Tn.x, Tn.y);
    if (want_extend || want_shadow) st256g(rec_sector(pb, slot, 2), Tn.z, pack2(dimflags, __float_as_uint(prev_pdf)), lp.x, lp.y);
    if (want_shadow) st256g(rec_sector(pb, slot, 3), lp.z, sc.x, sc.y, sc.z);
    if (PRIMARY) lum_set(pb, slot, mk3(0.0, 0.0, 0.0)); /* bounce 0 initialises the radiance of every path */
}

#ifndef PRT_SHADE_STAGE
#define PRT_SHADE_STAGE 1 /* 1: persistent warps over contiguous chunks, pushes staged in shared memory; 0: round 1's grid-stride loop */
#endif
#ifndef PRT_SHADE_CHUNK_ITERS
#define PRT_SHADE_CHUNK_ITERS 4 /* measured 2 / 4 / 8 / 16: k_shade 20.2 / 20.4 / 20.9 / 21.3 ms (profiles/r2_s5_chunk_ab.txt) */
#endif
#ifndef PRT_SHADE_AHEAD
#define PRT_SHADE_AHEAD 0 /* 1: read the next vertex's primitive id an iteration ahead and prefetch its TriShade line into L1.
                             Measured (profiles/r2_s4_shade_ab.txt): k_shade 23.4 ms with it, 21.3 ms without — the extra load
                             and its address arithmetic cost more than the hop they hide */
#endif
#define PRT_SHADE_CHUNK (32 * PRT_SHADE_CHUNK_ITERS)
/* shared memory of one shade block (4 warps): per warp two staged queues (shadow, extend) of one chunk's worth of
 * entries, plus their ray-binning keys */
#define PRT_SHADE_STAGE_WORDS (PRT_SHADE_STAGE ? 4 * (2 * PRT_SHADE_CHUNK + PRT_SHADE_CHUNK) : 1)

/* Queue position q of class CLS -> the path slot it names (-1 past the end). PRIMARY: the queue holds pixels and
 * position q means sample q / nq of pixel queue[q % nq] (sample-major: neighbouring positions, neighbouring records). */
template <bool PRIMARY>
__device__ __forceinline__ int32_t shade_entry(const int32_t* queue, const FastDiv& by_nq, int32_t n, int32_t npix, int64_t q) {
    if (q >= (int64_t)n) return -1;
    if (PRIMARY) {
        uint32_t s, j;
        by_nq.divmod((uint32_t)q, s, j);
        return (int32_t)s * npix + __ldcg(queue + j);
    }
    return __ldcg(queue + q);
}

#if PRT_SHADE_STAGE
/* One warp = one contiguous chunk of PRT_SHADE_CHUNK queue positions at a time (one atomic on the class cursor per
 * chunk, issued a chunk ahead). The vertices' pushes go to two per-warp staging arrays in shared memory — positions
 * from a ballot, no atomics — and are flushed once per chunk: ONE atomicAdd per output queue per 256 vertices instead
 * of one per 32 (the round-trip of those atomics was 9 % of the stall samples, profiles/r2_s2_shade_hot_sass.txt),
 * and a chunk's survivors land in the output queue next to each other, in queue order: the records the next
 * kernels gather stay neighbours (round 1's per-iteration pushes interleaved 2368 warps' groups of ~20). */
template <int CLS, bool PRIMARY>
__device__ __forceinline__ void shade_body(const DevScene& S, const RenderConst& rc, const PathBuffers& pb, int bounce, int sort, int32_t* stage) {
    const unsigned FULL = 0xffffffffu;
    BounceCtrl* ctrl = pb.ctrl + bounce;
    BounceCtrl* next = pb.ctrl + bounce + 1;
    const int32_t nq = __ldcg(&ctrl->n_mat[CLS]);
    const int32_t n = PRIMARY ? nq * rc.n_samples : nq;
    const int32_t* queue = pb.q_mat[CLS]; /* ld.cg: k_tail rewrites the queues inside one launch */
    int32_t* const q_shadow = pb.q_shadow;
    int32_t* const q_extend = pb.q_extend[(bounce + 1) & 1];
    const int lane = lane_id(), warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    int32_t* st_s = stage + warp * (3 * PRT_SHADE_CHUNK);
    int32_t* st_e = st_s + PRT_SHADE_CHUNK;
    uint16_t* st_ks = reinterpret_cast<uint16_t*>(st_e + PRT_SHADE_CHUNK);
    uint16_t* st_ke = st_ks + PRT_SHADE_CHUNK;
    int32_t* cursor = &ctrl->fetch_mat[CLS];
    /* chunk length: PRT_SHADE_CHUNK_ITERS warp-iterations when the queue is long, fewer when it is not — every warp of
     * the grid should see at least eight chunks, or the late, short bounces run on a fraction of the machine */
    const int32_t per_warp = n / (int32_t)(gridDim.x * (blockDim.x >> 5) * 32 * 8);
    const int K = per_warp < 1 ? 1 : (per_warp < PRT_SHADE_CHUNK_ITERS ? per_warp : PRT_SHADE_CHUNK_ITERS);
    const int32_t CH = 32 * K;
    int32_t a_raw = 0;
    if (lane == 0) a_raw = atomicAdd(cursor, CH);
    int32_t base = __shfl_sync(FULL, a_raw, 0);
    if (base >= n) return;
    if (lane == 0) a_raw = atomicAdd(cursor, CH); /* the chunk after this one */
    const FastDiv by_nq = FastDiv::make((uint32_t)nq);
    int32_t slot_next = shade_entry<PRIMARY>(queue, by_nq, n, rc.npix, (int64_t)base + lane);
    for (;;) {
        int32_t ns = 0, ne = 0;
        int32_t base_next = 0;
#pragma unroll 1
        for (int k = 0; k < K; ++k) {
            const int32_t slot = slot_next;
            if (k + 1 < K) {
                slot_next = shade_entry<PRIMARY>(queue, by_nq, n, rc.npix, (int64_t)base + (k + 1) * 32 + lane);
            } else {
                base_next = __shfl_sync(FULL, a_raw, 0); /* issued a whole chunk ago */
                slot_next = shade_entry<PRIMARY>(queue, by_nq, n, rc.npix, (int64_t)base_next + lane);
            }
#if PRT_SHADE_PREFETCH
            if (!PRIMARY && slot_next >= 0) {
                if (bounce == 0) {
                    prefetch_l2(rec_sector(pb, slot_next, 4)); /* bounce 0 without reuse reads the hit only */
                } else {
                    prefetch_l2(rec_sector(pb, slot_next, 0)); /* 160 B: two 128-byte lines at most */
                    prefetch_l2(rec_sector(pb, slot_next, 4));
                }
            }
#endif
#if PRT_SHADE_AHEAD
            /* the next vertex's primitive id, asked for now (its record line comes along), used at the END of this
             * iteration to pull that triangle's shading record into L1: the chain slot -> hit -> TriShade -> material
             * of the next iteration then starts two hops in */
            int32_t prim_next = -1;
            if (slot_next >= 0)
                prim_next = PRIMARY ? __ldg(reinterpret_cast<const int32_t*>(pb.primary + (slot_next - (int32_t)rc.by_npix.div((uint32_t)slot_next) * rc.npix)) + 6)
                                    : __ldcg(reinterpret_cast<const int32_t*>(rec_sector(pb, slot_next, 4)) + 6);
#endif
            bool want_shadow = false, want_extend = false;
            uint32_t key_shadow = 0, key_extend = 0;
            if (slot >= 0) shade_vertex<CLS, PRIMARY>(S, rc, pb, bounce, sort, slot, want_shadow, want_extend, key_shadow, key_extend);
#if PRT_SHADE_AHEAD
            if (prim_next >= 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(S.shade + prim_next));
#endif
            __syncwarp();
            const unsigned ms = __ballot_sync(FULL, want_shadow), me = __ballot_sync(FULL, want_extend);
            if (want_shadow) {
                const int at = ns + __popc(ms & lt);
                st_s[at] = slot;
                if (sort) { st_ks[at] = (uint16_t)key_shadow; atomicAdd(pb.hist + PRT_KEY_BINS + key_shadow, 1); }
            }
            if (want_extend) {
                const int at = ne + __popc(me & lt);
                st_e[at] = slot;
                if (sort) { st_ke[at] = (uint16_t)key_extend; atomicAdd(pb.hist + key_extend, 1); }
            }
            ns += __popc(ms);
            ne += __popc(me);
            if ((int64_t)base + (k + 1) * 32 >= (int64_t)n && k + 1 < K) { /* the queue ended inside this chunk (warp-uniform) */
                base_next = n;
                break;
            }
        }
        __syncwarp();
        /* flush: one atomic per output queue, then a coalesced copy out of shared memory */
        int32_t rs = 0, re = 0;
        if (lane == 0) {
            if (ns) rs = atomicAdd(&ctrl->n_shadow, ns);
            if (ne) re = atomicAdd(&next->n_extend, ne);
        }
        rs = __shfl_sync(FULL, rs, 0);
        re = __shfl_sync(FULL, re, 0);
        for (int i = lane; i < ns; i += 32) {
            q_shadow[rs + i] = st_s[i];
            if (sort) pb.key_shadow[rs + i] = st_ks[i];
        }
        for (int i = lane; i < ne; i += 32) {
            q_extend[re + i] = st_e[i];
            if (sort) pb.key_extend[re + i] = st_ke[i];
        }
        __syncwarp();
        base = base_next;
        if (base >= n) break;
        if (lane == 0) a_raw = atomicAdd(cursor, CH);
    }
}
#else
template <int CLS, bool PRIMARY>
__device__ __forceinline__ void shade_body(const DevScene& S, const RenderConst& rc, const PathBuffers& pb, int bounce, int sort, int32_t*) {
    BounceCtrl* ctrl = pb.ctrl + bounce;
    BounceCtrl* next = pb.ctrl + bounce + 1;
    const int32_t nq = __ldcg(&ctrl->n_mat[CLS]);
    const int32_t n = PRIMARY ? nq * rc.n_samples : nq;
    const int32_t* queue = pb.q_mat[CLS]; /* ld.cg, as in ExtendPolicy */
    int32_t* const out_q[2] = {pb.q_shadow, pb.q_extend[(bounce + 1) & 1]};
    int32_t* const out_n[2] = {&ctrl->n_shadow, &next->n_extend};
    const int32_t n_round = (n + 31) & ~31;
    const int32_t stride = gridDim.x * blockDim.x;
    /* the slot that queue position q names, -1 past the end */
    auto entry = [&](int64_t q) -> int32_t {
        if (q >= (int64_t)n) return -1;
        if (PRIMARY) {
            const int32_t s = (int32_t)q / nq;
            return s * rc.npix + __ldcg(queue + ((int32_t)q - s * nq));
        }
        return __ldcg(queue + q);
    };
    int32_t qi = blockIdx.x * blockDim.x + threadIdx.x;
    int32_t slot_next = entry(qi);
    int32_t slot_next2 = entry((int64_t)qi + stride);
    DeferredPush push_shadow, push_extend;
    push_shadow.init();
    push_extend.init();
    for (; qi < n_round; qi += stride) {
        bool want_shadow = false, want_extend = false;
        uint32_t key_shadow = 0, key_extend = 0;
        const int32_t slot = slot_next;
        slot_next = slot_next2;
        slot_next2 = entry((int64_t)qi + 2 * (int64_t)stride);
#if PRT_SHADE_PREFETCH
        if (!PRIMARY && slot_next >= 0) {
            if (bounce == 0) {
                prefetch_l2(rec_sector(pb, slot_next, 4)); /* bounce 0 without reuse reads the hit only */
            } else {
                prefetch_l2(rec_sector(pb, slot_next, 0)); /* 160 B: two 128-byte lines at most */
                prefetch_l2(rec_sector(pb, slot_next, 4));
            }
        }
#endif
        if (slot >= 0) {
            shade_vertex<CLS, PRIMARY>(S, rc, pb, bounce, sort, slot, want_shadow, want_extend, key_shadow, key_extend);
        }
        __syncwarp();
        {
            int32_t* const q0[1] = {out_q[0]};
            int32_t* const q1[1] = {out_q[1]};
            int32_t* const n0[1] = {out_n[0]};
            int32_t* const n1[1] = {out_n[1]};
            push_shadow.issue(n0, want_shadow ? 0 : -1, slot);
            push_extend.issue(n1, want_extend ? 0 : -1, slot);
            if (sort) {
                /* histogram of the keys: fire-and-forget REDs, consumed by k_sort_scan */
                if (want_shadow) atomicAdd(pb.hist + PRT_KEY_BINS + key_shadow, 1);
                if (want_extend) atomicAdd(pb.hist + key_extend, 1);
                push_shadow.key = key_shadow;
                push_extend.key = key_extend;
                push_shadow.complete_keyed(out_q[0], pb.key_shadow);
                push_extend.complete_keyed(out_q[1], pb.key_extend);
            } else {
                push_shadow.complete(q0);
                push_extend.complete(q1);
            }
        }
    }
}

#endif

template <int CLS, bool PRIMARY>
__global__ void __launch_bounds__(128, PRT_SHADE_MIN_BLOCKS) k_shade(DevScene S, RenderConst rc, PathBuffers pb, int bounce, int sort) {
    __shared__ int32_t stage[PRT_SHADE_STAGE_WORDS];
    shade_body<CLS, PRIMARY>(S, rc, pb, bounce, sort, stage);
}

/* NEE visibility (Camera.cpp:143-155). The reference asks for the closest hit on [1e-3, DBL_MAX]
 * along the unit light direction and accepts iff something was hit and
 * distance - |ps - hit| < 1e-3, i.e. iff no surface lies closer than distance - 1e-3: an any-hit
 * query over [1e-3, distance - 1e-3]. PRT_RENDER_EXACT_VISIBILITY runs the literal test instead. */
template <bool EXACT>
struct ShadowPolicy {
    static constexpr int REFILL_BELOW = PRT_REFILL_BELOW_SHADOW;
    const PathBuffers& pb;
    const int32_t* queue;
    d3 ps, dir;
    double distance;
    __device__ __forceinline__ int32_t slot_of(int32_t qi) const { return __ldcg(queue + qi); }
    __device__ __forceinline__ void prefetch(int32_t slot) const {
        prefetch_l2(rec_sector(pb, slot, 0));
        prefetch_l2(rec_sector(pb, slot, 3));
    }
    __device__ __forceinline__ void finish() const {}
    __device__ __forceinline__ bool load(int32_t slot, d3& o, d3& d, double& t0, double& t1) {
        /* by now the record holds the hit position as the origin of the next segment, and the light sample */
        const dd4 s0 = ld256g(rec_sector(pb, slot, 0)), s2 = ld256g(rec_sector(pb, slot, 2)), s3 = ld256g(rec_sector(pb, slot, 3));
        const d3 p = mk3(s0.v[0], s0.v[1], s0.v[2]);
        const d3 to_light = mk3(s2.v[2], s2.v[3], s3.v[0]) - p; /* Camera.cpp:141-143, as in k_shade */
        const d3 light_dir = normalize(to_light);
        const double dist = length(to_light);
        if (EXACT) { ps = p; dir = light_dir; distance = dist; }
        o = p;
        d = light_dir;
        t0 = 0.001;
        t1 = EXACT ? DBL_MAX : dist - 0.001;
        return true;
    }
    __device__ __forceinline__ void retire(bool finished, int32_t slot, const Hit& h) {
        if (!finished) return;
        bool visible;
        if (EXACT) {
            visible = false; /* a shadow ray that hits nothing reads an unset position: "not visible" */
            if (h.slot >= 0) {
                const d3 hp = ps + dir * h.t;
                visible = distance - length(ps - hp) < 0.001;
            }
        } else {
            visible = h.slot < 0;
        }
        if (visible) {
            const dd4 s3 = ld256g(rec_sector(pb, slot, 3));
            lum_add(pb, slot, mk3(s3.v[1], s3.v[2], s3.v[3]));
        }
    }
};

template <bool COUNT, int WIDE>
__device__ __forceinline__ void shadow_body(const DevScene& S, const RenderConst& rc, const PathBuffers& pb, int bounce, int sorted, VisitCounters& cnt,
                                            int32_t* sm_stack) {
    BounceCtrl* ctrl = pb.ctrl + bounce;
    const int32_t n = __ldcg(&ctrl->n_shadow);
    if (rc.flags & PRT_RENDER_EXACT_VISIBILITY) {
        ShadowPolicy<true> pol = {pb, sorted ? pb.qs_shadow : pb.q_shadow, mk3(0, 0, 0), mk3(0, 0, 0), 0.0};
        traverse_persistent<false, COUNT, WIDE>(S, pol, &ctrl->fetch_shadow, n, cnt, sm_stack);
    } else {
        ShadowPolicy<false> pol = {pb, sorted ? pb.qs_shadow : pb.q_shadow, mk3(0, 0, 0), mk3(0, 0, 0), 0.0};
        traverse_persistent<true, COUNT, WIDE>(S, pol, &ctrl->fetch_shadow, n, cnt, sm_stack);
    }
}

template <bool COUNT, int WIDE>
__global__ void __launch_bounds__(PRT_TRAVERSE_BLOCK, PRT_TRAVERSE_MIN_BLOCKS) k_shadow(DevScene S, RenderConst rc, PathBuffers pb, int bounce, int sorted) {
    __shared__ int32_t sm_stack[PRT_SMEM_STACK_WORDS];
    VisitCounters cnt = {0, 0, 0, 0, 0};
    shadow_body<COUNT, WIDE>(S, rc, pb, bounce, sorted, cnt, sm_stack);
    flush_counters<COUNT>(cnt, pb.counters);
}

/* The tail of a pass in ONE cooperative launch. Russian roulette thins the wavefront geometrically
 * (rr 0.8, depth 100: bounce 27 holds < 0.3 % of the paths), and a bounce of a few thousand rays
 * costs what its dependent chain of launches and memory latencies costs, not what its rays cost:
 * the ncu launch list of the per-bounce pipeline shows bounces 30..100 at 30-80 us each per stage
 * (profiles/r1_s5_launches.csv). From bounce `first` on, the same stage bodies run back to back
 * inside this kernel, separated by grid-wide barriers, until the extend queue is empty. Results are
 * those of the per-bounce launches (same code, same queues); only launch count and idle time change.
 * The barrier after the shadow stage is needed: shadow(b) and extend(b + 1) both update the radiance cells. */
template <bool COUNT, int WIDE>
__global__ void __launch_bounds__(128, 2) k_tail(DevScene S, const uint8_t* __restrict__ prim_class, RenderConst rc, PathBuffers pb, int first) {
    /* the stages are separated by grid-wide barriers: the traversal stack and the shade staging share one buffer */
    __shared__ int32_t sm_stack[PRT_STACK_WORDS_FOR(128) > PRT_SHADE_STAGE_WORDS ? PRT_STACK_WORDS_FOR(128) : PRT_SHADE_STAGE_WORDS];
    cg::grid_group grid = cg::this_grid();
    VisitCounters cnt = {0, 0, 0, 0, 0};
    for (int b = first; b <= rc.max_depth; ++b) {
        if (b > 0 && __ldcg(&pb.ctrl[b].n_extend) == 0) break; /* uniform: written before the last barrier */
        extend_body<COUNT, WIDE>(S, prim_class, rc, pb, b, 0, cnt, sm_stack);
        grid.sync();
        if (rc.class_mask & (1 << CLS_LAMBERT)) shade_body<CLS_LAMBERT, false>(S, rc, pb, b, 0, sm_stack);
        if (rc.class_mask & (1 << CLS_PHONG)) shade_body<CLS_PHONG, false>(S, rc, pb, b, 0, sm_stack);
        if (rc.class_mask & (1 << CLS_MIRROR)) shade_body<CLS_MIRROR, false>(S, rc, pb, b, 0, sm_stack);
        if (rc.class_mask & (1 << CLS_COOK)) shade_body<CLS_COOK, false>(S, rc, pb, b, 0, sm_stack);
        if (rc.class_mask & (1 << CLS_PHONG_SPEC)) shade_body<CLS_PHONG_SPEC, false>(S, rc, pb, b, 0, sm_stack);
        grid.sync();
        if (rc.sample_lights) {
            shadow_body<COUNT, WIDE>(S, rc, pb, b, 0, cnt, sm_stack);
            grid.sync();
        }
    }
    flush_counters<COUNT>(cnt, pb.counters);
}

/* colorAttachment[m] += RayColor * pixelSamplesScale, samples in index order (Camera.cpp:54-57). Streams the
 * radiance plane (consecutive threads = consecutive pixels = consecutive 32-byte cells); a pixel whose path ended
 * at bounce 0 under primary-hit reuse has no cells: all its samples returned the table's radiance. */
__global__ void __launch_bounds__(256) k_accumulate(RenderConst rc, PathBuffers pb, double* __restrict__ rgb) {
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < rc.npix; pix += gridDim.x * blockDim.x) {
        d3 acc = mk3(rgb[3 * (int64_t)pix], rgb[3 * (int64_t)pix + 1], rgb[3 * (int64_t)pix + 2]);
        bool constant = false;
        d3 c = mk3(0.0, 0.0, 0.0);
        if (rc.primary_reuse) {
            const dd4 ph = ld256d(pb.primary + pix);
            constant = (int32_t)unpack_hi(ph.v[3]) >= CLS_SHADED;
            c = mk3(ph.v[0], ph.v[1], ph.v[2]);
        }
        if (constant) {
            const d3 term = c * rc.scale;
            for (int s = 0; s < rc.n_samples; ++s) acc = acc + term;
        } else {
            for (int s = 0; s < rc.n_samples; ++s) {
                const dd4 l = ld256g(pb.lum + ((int64_t)s * rc.npix + pix));
                acc = acc + mk3(l.v[0], l.v[1], l.v[2]) * rc.scale;
            }
        }
        rgb[3 * (int64_t)pix] = acc.x; rgb[3 * (int64_t)pix + 1] = acc.y; rgb[3 * (int64_t)pix + 2] = acc.z;
    }
}

/* ray accounting of one pass, kept on the device so that the render call stays asynchronous:
 * totals[2] += extend rays, totals[3] += shadow rays */
__global__ void k_pass_stats(const BounceCtrl* __restrict__ ctrl, int n_bounces, int n_rays0, unsigned long long* totals) {
    unsigned long long e = 0, s = 0;
    for (int b = threadIdx.x; b < n_bounces; b += 32) {
        e += b == 0 ? (unsigned long long)n_rays0 : (unsigned long long)ctrl[b].n_extend; /* bounce 0: one ray per path, or per pixel under primary-hit reuse */
        s += (unsigned long long)ctrl[b].n_shadow;
    }
    for (int o = 16; o > 0; o >>= 1) {
        e += __shfl_down_sync(0xffffffffu, e, o);
        s += __shfl_down_sync(0xffffffffu, s, o);
    }
    if (threadIdx.x == 0) {
        totals[2] += e;
        totals[3] += s;
    }
}

/* Camera::WriteColorAttachment's pixel loop: NaN scrub, sRGB OETF (Camera.cpp:216-221), clamp to
 * [0, 0.9999] then *255 truncated (:297-300); float32 copy of the scrubbed linear value (:308-325).
 *
 * The byte is NOT computed with CUDA's pow: it differs from glibc's in the last ulp, and an ulp at one of the 254
 * places where trunc(clamp(LinearToSRGB(c)) * 255) steps is a different byte. The quantiser is monotone in c, so
 * the host finds, once, the smallest double that yields each byte value k = 1..254 with the very libm call the
 * reference makes (capi.cu: resolve_thresholds, bisection over the bit patterns), and the kernel counts how many of
 * those thresholds c has reached: byte-for-byte the reference's PNG payload (tests/golden/output_stage.npz). */
__global__ void __launch_bounds__(256) k_resolve(const double* __restrict__ rgb, int64_t n, const double* __restrict__ thr, uint8_t* __restrict__ rgb8,
                                                 float* __restrict__ rgb32f) {
    __shared__ double s_thr[256];
    s_thr[threadIdx.x] = threadIdx.x < 254 ? thr[threadIdx.x] : (double)INFINITY;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double c = rgb[i];
        if (c != c) c = 0.0;
        if (rgb32f) rgb32f[i] = (float)c;
        if (rgb8) {
            int lo = 0, hi = 254; /* the byte = number of thresholds <= c */
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (c >= s_thr[mid - 1]) lo = mid;
                else hi = mid - 1;
            }
            rgb8[i] = (uint8_t)lo;
        }
    }
}

/* ------------------------------------------------------------------------------------------------
 * launchers
 * ---------------------------------------------------------------------------------------------- */
#ifndef PRT_TRAVERSE_MIN_BLOCKS
#define PRT_TRAVERSE_MIN_BLOCKS 6
#endif

/* persistent kernels get exactly as many blocks as can be resident (148 SMs x occupancy), never a
 * second wave; small queues get fewer */
/* resident blocks per SM of `kernel` on the CURRENT device, looked up once per (device, kernel). Contexts on
 * different GPUs render from different host threads (prt_render_multi): the cache is keyed by device and locked. */
static int blocks_per_sm(const void* kernel, int block) {
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, int> cache;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(mu);
    const auto key = std::make_pair(dev, kernel);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    cache[key] = per_sm;
    return per_sm;
}
template <typename K>
static int persistent_grid(K kernel, int block, const LaunchConfig& lc, int64_t items) {
    int per_sm = blocks_per_sm((const void*)kernel, block);
    if (lc.trace_blocks_per_sm > 0 && per_sm > lc.trace_blocks_per_sm) per_sm = lc.trace_blocks_per_sm;
    const int64_t cap = (int64_t)per_sm * lc.sm_count;
    int64_t want = (items + block - 1) / block;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

/* picks the <COUNT, WIDE> instantiation of a traversal kernel; GRID may name it as KERNEL */
#define PRT_DISPATCH_ONE(C, W, NAME, GRID, ARGS)      \
    do {                                              \
        auto KERNEL = NAME<C, W>;                     \
        KERNEL<<<(GRID), PRT_TRAVERSE_BLOCK, 0, st>>> ARGS;          \
    } while (0)
/* wide: 2 = Node2 (or the reference tree), 4 = Node4, 40 = Node4Q */
#define PRT_DISPATCH(count, wide, NAME, GRID, ARGS)                            \
    do {                                                                       \
        if (count) {                                                           \
            if ((wide) == 40) PRT_DISPATCH_ONE(true, 40, NAME, GRID, ARGS);    \
            else if ((wide) == 4) PRT_DISPATCH_ONE(true, 4, NAME, GRID, ARGS); \
            else PRT_DISPATCH_ONE(true, 2, NAME, GRID, ARGS);                  \
        } else {                                                               \
            if ((wide) == 40) PRT_DISPATCH_ONE(false, 40, NAME, GRID, ARGS);   \
            else if ((wide) == 4) PRT_DISPATCH_ONE(false, 4, NAME, GRID, ARGS);\
            else PRT_DISPATCH_ONE(false, 2, NAME, GRID, ARGS);                 \
        }                                                                      \
    } while (0)
static inline int wide_of(const DevScene& S) { return S.nodes4q ? 40 : (S.nodes4 ? 4 : 2); }

void launch_trace_closest(const DevScene& S, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax,
                          int32_t* prim, double* t, double* alpha, double* beta, const int32_t* order, int32_t* fetch,
                          unsigned long long* counters, const LaunchConfig& lc, cudaStream_t st) {
    if (n <= 0) return;
    PRT_DISPATCH(counters != nullptr, wide_of(S), k_trace_closest, persistent_grid(KERNEL, PRT_TRAVERSE_BLOCK, lc, n), (S, n, o, d, tmin, tmax, prim, t, alpha, beta, order, fetch, counters));
}
void launch_trace_any(const DevScene& S, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax,
                      uint8_t* hit, const int32_t* order, int32_t* fetch, unsigned long long* counters, const LaunchConfig& lc,
                      cudaStream_t st) {
    if (n <= 0) return;
    PRT_DISPATCH(counters != nullptr, wide_of(S), k_trace_any, persistent_grid(KERNEL, PRT_TRAVERSE_BLOCK, lc, n), (S, n, o, d, tmin, tmax, hit, order, fetch, counters));
}
void launch_query_sort(int64_t n, const double* o, const double* d, const float* key_lo, const float* key_scale, uint16_t* keys,
                       int32_t* hist, int32_t* cursor, int32_t* order, const LaunchConfig& lc, cudaStream_t st) {
    if (n <= 0) return;
    KeyBox kb;
    for (int k = 0; k < 3; ++k) { kb.lo[k] = key_lo[k]; kb.scale[k] = key_scale[k]; }
    const int64_t want = (n + 255) / 256, cap = (int64_t)lc.sm_count * 8;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    k_query_keys<<<grid, 256, 0, st>>>(n, o, d, kb, keys, hist);
    k_sort_scan<<<1, 1024, 0, st>>>(hist, cursor);
    SortJob j = {nullptr, (int32_t)n, keys, nullptr, order};
    k_sort_scatter<<<dim3(grid, 1), 256, 0, st>>>(j, j, cursor);
}
void launch_sample_lights(const DevScene& S, int64_t n, const double* origins, const double* u, double* position, double* normal,
                          uint8_t* front, double* pdf, int32_t* prim, cudaStream_t st) {
    if (n <= 0) return;
    k_sample_lights<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(S, n, origins, u, position, normal, front, pdf, prim);
}

int launch_render_pass(const DevScene& S, const uint8_t* prim_class, const RenderConst& rc, const PathBuffers& pb, double* rgb,
                       unsigned long long* totals, const LaunchConfig& lc, cudaStream_t st, int* tail_first_out) {
    int launches = 0;
    const int n_bounces = rc.max_depth + 1;
    cudaMemsetAsync(pb.ctrl, 0, sizeof(BounceCtrl) * (size_t)(n_bounces + 1), st);
    const bool count = (rc.flags & PRT_RENDER_COUNTERS) != 0;
    const int wide = wide_of(S);
    LaunchConfig shade_lc = lc;
    shade_lc.trace_blocks_per_sm = 0;
    const int g_shade0 = persistent_grid(k_shade<CLS_LAMBERT, false>, 128, shade_lc, rc.n_paths);
    const int g_shade1 = persistent_grid(k_shade<CLS_PHONG, false>, 128, shade_lc, rc.n_paths);
    const int g_shade2 = persistent_grid(k_shade<CLS_MIRROR, false>, 128, shade_lc, rc.n_paths);
    const int g_shade3 = persistent_grid(k_shade<CLS_COOK, false>, 128, shade_lc, rc.n_paths);
    const int g_shade4 = persistent_grid(k_shade<CLS_PHONG_SPEC, false>, 128, shade_lc, rc.n_paths);
    cudaEvent_t* ev = lc.stage_events;
    /* Both per-bounce decisions are taken here from the EXPECTED queue length n_paths * rr^bounce
     * (Russian roulette thins the wavefront geometrically; the host never reads a queue length back):
     *   - ray binning for long queues (off by default, launch.h);
     *   - the first bounce of the tail: once fewer than lc.tail_paths paths are expected, the rest of
     *     the pass runs inside ONE cooperative launch (k_tail). Under primary-hit reuse bounce 0 always has its
     *     own launches (its extend stage is per pixel, its shade stage expands pixels into samples). */
    const double survive = rc.rr < 1.0 ? (rc.rr > 0.0 ? rc.rr : 0.0) : 1.0;
    double expect = (double)rc.n_paths;
    int sorted = 0; /* this bounce's extend queue was reordered by the previous bounce */
    const int sort_grid = (int)std::min<int64_t>((rc.n_paths + 255) / 256, (int64_t)lc.sm_count * 8);
    int b = 0;
    for (; b < n_bounces; ++b) {
        const bool primary = b == 0 && rc.primary_reuse;
        if (!primary && lc.tail_paths > 0 && expect < (double)lc.tail_paths) break;
        const int sort = lc.sort_min > 0 && expect >= (double)lc.sort_min ? 1 : 0;
        if (ev) cudaEventRecord(ev[4 * b], st);
        if (primary) PRT_DISPATCH(count, wide, k_primary, persistent_grid(KERNEL, PRT_TRAVERSE_BLOCK, lc, rc.npix), (S, prim_class, rc, pb));
        else PRT_DISPATCH(count, wide, k_extend, persistent_grid(KERNEL, PRT_TRAVERSE_BLOCK, lc, rc.n_paths), (S, prim_class, rc, pb, b, sorted));
        ++launches;
        if (ev) cudaEventRecord(ev[4 * b + 1], st);
        if (primary) {
            if (rc.class_mask & (1 << CLS_LAMBERT)) { k_shade<CLS_LAMBERT, true><<<persistent_grid(k_shade<CLS_LAMBERT, true>, 128, shade_lc, rc.n_paths), 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_PHONG)) { k_shade<CLS_PHONG, true><<<persistent_grid(k_shade<CLS_PHONG, true>, 128, shade_lc, rc.n_paths), 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_MIRROR)) { k_shade<CLS_MIRROR, true><<<persistent_grid(k_shade<CLS_MIRROR, true>, 128, shade_lc, rc.n_paths), 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_COOK)) { k_shade<CLS_COOK, true><<<persistent_grid(k_shade<CLS_COOK, true>, 128, shade_lc, rc.n_paths), 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_PHONG_SPEC)) { k_shade<CLS_PHONG_SPEC, true><<<persistent_grid(k_shade<CLS_PHONG_SPEC, true>, 128, shade_lc, rc.n_paths), 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
        } else {
            if (rc.class_mask & (1 << CLS_LAMBERT)) { k_shade<CLS_LAMBERT, false><<<g_shade0, 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_PHONG)) { k_shade<CLS_PHONG, false><<<g_shade1, 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_MIRROR)) { k_shade<CLS_MIRROR, false><<<g_shade2, 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_COOK)) { k_shade<CLS_COOK, false><<<g_shade3, 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
            if (rc.class_mask & (1 << CLS_PHONG_SPEC)) { k_shade<CLS_PHONG_SPEC, false><<<g_shade4, 128, 0, st>>>(S, rc, pb, b, sort); ++launches; }
        }
        if (sort) {
            BounceCtrl* dc = pb.ctrl + b;
            const SortJob je = {&(dc + 1)->n_extend, 0, pb.key_extend, pb.q_extend[(b + 1) & 1], pb.qs_extend};
            const SortJob js = {&dc->n_shadow, 0, pb.key_shadow, pb.q_shadow, pb.qs_shadow};
            k_sort_scan<<<2, 1024, 0, st>>>(pb.hist, pb.cursor);
            k_sort_scatter<<<dim3(sort_grid, 2), 256, 0, st>>>(je, js, pb.cursor);
            launches += 2;
        }
        if (ev) cudaEventRecord(ev[4 * b + 2], st);
        if (rc.sample_lights) {
            PRT_DISPATCH(count, wide, k_shadow, persistent_grid(KERNEL, PRT_TRAVERSE_BLOCK, lc, rc.n_paths), (S, rc, pb, b, sort));
            ++launches;
        }
        sorted = sort;
        expect *= survive;
        if (ev) cudaEventRecord(ev[4 * b + 3], st);
    }
    if (tail_first_out) *tail_first_out = b;
    if (b < n_bounces) {
        /* the tail: bounces b .. max_depth in one cooperative launch (its time is booked as traversal
         * time by the stage timing: events 4b, 4b + 1) */
        if (ev) cudaEventRecord(ev[4 * b], st);
        int first = b;
        DevScene S_ = S;
        RenderConst rc_ = rc;
        PathBuffers pb_ = pb;
        const uint8_t* pc_ = prim_class;
        void* args[] = {(void*)&S_, (void*)&pc_, (void*)&rc_, (void*)&pb_, (void*)&first};
        const void* fn = count ? (wide == 40 ? (const void*)k_tail<true, 40> : (wide == 4 ? (const void*)k_tail<true, 4> : (const void*)k_tail<true, 2>))
                               : (wide == 40 ? (const void*)k_tail<false, 40> : (wide == 4 ? (const void*)k_tail<false, 4> : (const void*)k_tail<false, 2>));
        int per_sm = blocks_per_sm(fn, 128);
        if (per_sm > 2) per_sm = 2;
        int64_t want = (rc.n_paths + 127) / 128;
        const int64_t cap = (int64_t)per_sm * lc.sm_count;
        if (want > cap) want = cap;
        if (want < 1) want = 1;
        cudaLaunchCooperativeKernel(fn, dim3((unsigned)want), dim3(128), args, 0, st);
        ++launches;
        if (ev) cudaEventRecord(ev[4 * b + 1], st);
    }
    k_accumulate<<<(rc.npix + 255) / 256 < lc.sm_count * 8 ? (rc.npix + 255) / 256 : lc.sm_count * 8, 256, 0, st>>>(rc, pb, rgb);
    k_pass_stats<<<1, 32, 0, st>>>(pb.ctrl, n_bounces, rc.primary_reuse ? rc.npix : rc.n_paths, totals);
    launches += 2;
    return launches;
}

void launch_resolve(const double* rgb, int64_t npix, const double* thresholds, uint8_t* rgb8, float* rgb32f, cudaStream_t st) {
    const int64_t n = npix * 3;
    if (n <= 0) return;
    k_resolve<<<(unsigned)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8), 256, 0, st>>>(rgb, n, thresholds, rgb8, rgb32f);
}

}  // namespace prt

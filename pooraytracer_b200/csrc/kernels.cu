/*
 * kernels.cu — the sm_100a kernels of the render path.
 *
 *   k_trace_closest / k_trace_any   ray-query batches = Hittable::Hit on `world` (Source/Hittable.h:34)
 *   k_sample_lights                 Hittable::Sample on `lights` (Source/Hittable.h:36)
 *   k_raygen                        Camera::GetRay (Source/Camera.cpp:108-117)
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
#include <cuda_runtime.h>
#include <float.h>

#include "../../include/prt_b200.h"
#include "launch.h"
#include "rng.cuh"
#include "shading.cuh"
#include "traverse.cuh"

namespace prt {

#define PRT_FLAG_PREV_SKIP 1u /* the vertex that scattered this ray skipped light sampling */

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
}

/* ------------------------------------------------------------------------------------------------
 * ray queries
 * ---------------------------------------------------------------------------------------------- */
template <bool COUNT>
__global__ void __launch_bounds__(128) k_trace_closest(DevScene S, int64_t n, const double* __restrict__ o, const double* __restrict__ d,
                                                       const double* __restrict__ tmin, const double* __restrict__ tmax,
                                                       int32_t* __restrict__ prim, double* __restrict__ t, double* __restrict__ alpha,
                                                       double* __restrict__ beta, int32_t* fetch, unsigned long long* counters) {
    VisitCounters cnt = {0, 0};
    for (;;) {
        const int64_t i = (int64_t)warp_fetch(fetch) + lane_id();
        if (i - lane_id() >= n) break;
        if (i < n) {
            Hit h;
            closest_hit<COUNT>(S, mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), tmin[i], tmax[i], h, cnt);
            prim[i] = h.slot >= 0 ? S.slot_prim[h.slot] : -1;
            if (t) t[i] = h.t;
            if (alpha) alpha[i] = h.alpha;
            if (beta) beta[i] = h.beta;
        }
        __syncwarp();
    }
    flush_counters<COUNT>(cnt, counters);
}

template <bool COUNT>
__global__ void __launch_bounds__(128) k_trace_any(DevScene S, int64_t n, const double* __restrict__ o, const double* __restrict__ d,
                                                   const double* __restrict__ tmin, const double* __restrict__ tmax, uint8_t* __restrict__ hit,
                                                   int32_t* fetch, unsigned long long* counters) {
    VisitCounters cnt = {0, 0};
    for (;;) {
        const int64_t i = (int64_t)warp_fetch(fetch) + lane_id();
        if (i - lane_id() >= n) break;
        if (i < n)
            hit[i] = any_hit<COUNT>(S, mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), tmin[i], tmax[i], cnt) ? 1 : 0;
        __syncwarp();
    }
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
__device__ __forceinline__ d3 load3(const double* base, int64_t cap, int64_t i) { return mk3(base[i], base[cap + i], base[2 * cap + i]); }
__device__ __forceinline__ void store3(double* base, int64_t cap, int64_t i, d3 v) {
    base[i] = v.x; base[cap + i] = v.y; base[2 * cap + i] = v.z;
}

__global__ void __launch_bounds__(256) k_raygen(RenderConst rc, PathBuffers pb) {
    for (int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; slot < rc.n_paths; slot += (int64_t)gridDim.x * blockDim.x) {
        const int pix = (int)(slot % rc.npix);
        const int i = pix % rc.width, j = pix / rc.width;
        /* Camera::GetRay: pixel centre, direction not normalised */
        const d3 ps = rc.pixel00 + ((double)i) * rc.dU + ((double)j) * rc.dV;
        const d3 dir = ps - rc.center;
        store3(pb.ray, pb.cap, slot, rc.center);
        store3(pb.ray + 3 * pb.cap, pb.cap, slot, dir);
        store3(pb.T, pb.cap, slot, mk3(1.0, 1.0, 1.0));
        store3(pb.L, pb.cap, slot, mk3(0.0, 0.0, 0.0));
        pb.dim[slot] = 0u;
        pb.flags[slot] = 0u;
    }
}

/* closest hit for every path in the bounce's queue (persistent warps, 32 rays per fetch), then the
 * terminal cases that need no BSDF, then one queue per material class for the shade kernels */
template <bool COUNT>
__global__ void __launch_bounds__(128) k_extend(DevScene S, const uint8_t* __restrict__ prim_class, RenderConst rc, PathBuffers pb, int bounce) {
    BounceCtrl* ctrl = pb.ctrl + bounce;
    const int32_t n = bounce == 0 ? rc.n_paths : ctrl->n_extend;
    const int32_t* __restrict__ queue = pb.q_extend[bounce & 1];
    VisitCounters cnt = {0, 0};
    for (;;) {
        const int32_t base = warp_fetch(&ctrl->fetch_extend);
        if (base >= n) break;
        const int32_t qi = base + lane_id();
        const bool valid = qi < n;
        int cls = -1;
        int32_t slot = 0;
        if (valid) {
            slot = bounce == 0 ? qi : queue[qi];
            const d3 o = load3(pb.ray, pb.cap, slot), d = load3(pb.ray + 3 * pb.cap, pb.cap, slot);
            Hit h;
            closest_hit<COUNT>(S, o, d, 0.0001, (double)INFINITY, h, cnt); /* Camera.cpp:125 */
            if (h.slot < 0) {
                /* a primary miss returns the background (Camera.cpp:127); with NEE on, a scattered
                 * ray that misses adds nothing (Camera.cpp:187-196) */
                const d3 bg = mk3(rc.background[0], rc.background[1], rc.background[2]);
                if (bounce == 0) store3(pb.L, pb.cap, slot, bg);
                else if (!rc.sample_lights) store3(pb.L, pb.cap, slot, load3(pb.L, pb.cap, slot) + load3(pb.T, pb.cap, slot) * bg);
            } else {
                const int32_t prim = S.slot_prim[h.slot];
                const int c = prim_class[prim];
                if (c == CLS_EMISSIVE) {
                    /* an emitter ends the path (Camera.cpp:129-132); with NEE on, a scattered ray may
                     * collect it only if the scattering material skipped light sampling (:191-195) */
                    const bool take = bounce == 0 || !rc.sample_lights || (pb.flags[slot] & PRT_FLAG_PREV_SKIP);
                    if (take) {
                        const DevMaterial* m = S.materials + S.shade[prim].material;
                        const d3 le = mk3(m->emission[0], m->emission[1], m->emission[2]);
                        if (bounce == 0) store3(pb.L, pb.cap, slot, le);
                        else store3(pb.L, pb.cap, slot, load3(pb.L, pb.cap, slot) + load3(pb.T, pb.cap, slot) * le);
                    }
                } else if (c < CLS_SHADED) {
                    pb.hit_prim[slot] = prim;
                    pb.hit[slot] = h.t;
                    pb.hit[pb.cap + slot] = h.alpha;
                    pb.hit[2 * pb.cap + slot] = h.beta;
                    cls = c;
                }
                /* CLS_EMPTY: no NEE, Scatter returns false (Material.h:537-540): the path ends */
            }
        }
        __syncwarp();
#pragma unroll
        for (int c = 0; c < CLS_SHADED; ++c) {
            const int at = warp_push(&ctrl->n_mat[c], cls == c);
            if (at >= 0) pb.q_mat[c][at] = slot;
        }
    }
    flush_counters<COUNT>(cnt, pb.counters);
}

/* one path vertex of material class CLS (Camera.cpp:134-201) */
template <int CLS>
__global__ void __launch_bounds__(128) k_shade(DevScene S, RenderConst rc, PathBuffers pb, int bounce) {
    BounceCtrl* ctrl = pb.ctrl + bounce;
    BounceCtrl* next = pb.ctrl + bounce + 1;
    const int32_t n = ctrl->n_mat[CLS];
    const int32_t* __restrict__ queue = pb.q_mat[CLS];
    int32_t* __restrict__ q_next = pb.q_extend[(bounce + 1) & 1];
    const int32_t n_round = (n + 31) & ~31;
    for (int32_t qi = blockIdx.x * blockDim.x + threadIdx.x; qi < n_round; qi += gridDim.x * blockDim.x) {
        bool want_shadow = false, want_extend = false;
        int32_t slot = 0;
        if (qi < n) {
            slot = queue[qi];
            const int32_t prim = pb.hit_prim[slot];
            const double t = pb.hit[slot], alpha = pb.hit[pb.cap + slot], beta = pb.hit[2 * pb.cap + slot];
            const d3 ro = load3(pb.ray, pb.cap, slot), rd = load3(pb.ray + 3 * pb.cap, pb.cap, slot);
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
            rng.seed = stream_seed(rc.rng_key, (uint32_t)(slot % rc.npix), (uint32_t)(rc.sample_first + slot / rc.npix));
            rng.dim = pb.dim[slot];
            const d3 T = load3(pb.T, pb.cap, slot);

            /* ---- next-event estimation (Camera.cpp:137-174) ---- */
            if (rc.sample_lights && !m.skip_light_sampling) {
                const double u0 = rng.next(), u1 = rng.next(), u2 = rng.next(), u3 = rng.next();
                LightSample ls;
                sample_lights(S, s.position, u0, u1, u2, u3, ls);
                const d3 to_light = ls.position - s.position;
                const d3 light_dir = normalize(to_light);
                const double distance = length(to_light);
                if (dot(s.normal, light_dir) > 0.0 && ls.front) {
                    const DevMaterial* lm = S.materials + ls.material;
                    const d3 emission = mk3(lm->emission[0], lm->emission[1], lm->emission[2]);
                    const d3 local_wi = world_to_local(light_dir, s);
                    const d3 local_ln = world_to_local(ls.normal, s);
                    const d3 fr = material_eval(S, m, local_wi, wo, s.uv, rng);
                    const double cos_theta = local_wi.z;
                    const double cos_theta_bar = dot(local_ln, -local_wi);
                    const d3 direct = emission * fr * cos_theta * cos_theta_bar / (distance * distance) / ls.pdf;
                    store3(pb.shadow, pb.cap, slot, light_dir);
                    pb.shadow[3 * pb.cap + slot] = distance;
                    store3(pb.shadow + 4 * pb.cap, pb.cap, slot, T * direct);
                    want_shadow = true;
                }
            }
            /* ---- Russian roulette + BSDF sampling (Camera.cpp:176-201) ---- */
            if (rng.next() < rc.rr) {
                d3 att, out_dir;
                if (CLS == CLS_LAMBERT) scatter_lambertian(S, m, s, rng, att, out_dir);
                else if (CLS == CLS_PHONG) scatter_phong(S, m, s, wo, rng, att, out_dir);
                else scatter_mirror(s, wo, att, out_dir);
                /* RayColor(depth - 1) of the last allowed vertex returns 0 (Camera.cpp:121) */
                if (bounce < rc.max_depth) {
                    store3(pb.ray, pb.cap, slot, s.position);
                    store3(pb.ray + 3 * pb.cap, pb.cap, slot, out_dir);
                    store3(pb.T, pb.cap, slot, T * att / rc.rr);
                    pb.dim[slot] = rng.dim;
                    pb.flags[slot] = m.skip_light_sampling ? PRT_FLAG_PREV_SKIP : 0u;
                    want_extend = true;
                }
            }
            if (want_shadow && !want_extend) store3(pb.ray, pb.cap, slot, s.position); /* shadow ray origin */
        }
        __syncwarp();
        const int a = warp_push(&ctrl->n_shadow, want_shadow);
        if (a >= 0) pb.q_shadow[a] = slot;
        const int b = warp_push(&next->n_extend, want_extend);
        if (b >= 0) q_next[b] = slot;
    }
}

/* NEE visibility (Camera.cpp:143-155). The reference asks for the closest hit on [1e-3, DBL_MAX]
 * along the unit light direction and accepts iff something was hit and
 * distance - |ps - hit| < 1e-3, i.e. iff no surface lies closer than distance - 1e-3: an any-hit
 * query over [1e-3, distance - 1e-3]. PRT_RENDER_EXACT_VISIBILITY runs the literal test instead. */
template <bool COUNT>
__global__ void __launch_bounds__(128) k_shadow(DevScene S, RenderConst rc, PathBuffers pb, int bounce) {
    BounceCtrl* ctrl = pb.ctrl + bounce;
    const int32_t n = ctrl->n_shadow;
    VisitCounters cnt = {0, 0};
    const bool exact = (rc.flags & PRT_RENDER_EXACT_VISIBILITY) != 0;
    for (;;) {
        const int32_t base = warp_fetch(&ctrl->fetch_shadow);
        if (base >= n) break;
        const int32_t qi = base + lane_id();
        if (qi < n) {
            const int32_t slot = pb.q_shadow[qi];
            /* by now pb.ray holds the hit position as the origin of the next segment */
            const d3 ps = load3(pb.ray, pb.cap, slot);
            const d3 dir = load3(pb.shadow, pb.cap, slot);
            const double distance = pb.shadow[3 * pb.cap + slot];
            bool visible;
            if (exact) {
                Hit h;
                closest_hit<COUNT>(S, ps, dir, 0.001, DBL_MAX, h, cnt);
                visible = false; /* a shadow ray that hits nothing reads an unset position: "not visible" */
                if (h.slot >= 0) {
                    const d3 hp = ps + dir * h.t;
                    visible = distance - length(ps - hp) < 0.001;
                }
            } else {
                visible = !any_hit<COUNT>(S, ps, dir, 0.001, distance - 0.001, cnt);
            }
            if (visible) store3(pb.L, pb.cap, slot, load3(pb.L, pb.cap, slot) + load3(pb.shadow + 4 * pb.cap, pb.cap, slot));
        }
        __syncwarp();
    }
    flush_counters<COUNT>(cnt, pb.counters);
}

/* colorAttachment[m] += RayColor * pixelSamplesScale, samples in index order (Camera.cpp:54-57) */
__global__ void __launch_bounds__(256) k_accumulate(RenderConst rc, PathBuffers pb, double* __restrict__ rgb) {
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < rc.npix; pix += gridDim.x * blockDim.x) {
        d3 acc = mk3(rgb[3 * (int64_t)pix], rgb[3 * (int64_t)pix + 1], rgb[3 * (int64_t)pix + 2]);
        for (int s = 0; s < rc.n_samples; ++s) acc = acc + load3(pb.L, pb.cap, (int64_t)s * rc.npix + pix) * rc.scale;
        rgb[3 * (int64_t)pix] = acc.x; rgb[3 * (int64_t)pix + 1] = acc.y; rgb[3 * (int64_t)pix + 2] = acc.z;
    }
}

/* ray accounting of one pass, kept on the device so that the render call stays asynchronous:
 * totals[2] += extend rays, totals[3] += shadow rays */
__global__ void k_pass_stats(const BounceCtrl* __restrict__ ctrl, int n_bounces, int n_paths, unsigned long long* totals) {
    unsigned long long e = 0, s = 0;
    for (int b = threadIdx.x; b < n_bounces; b += 32) {
        e += b == 0 ? (unsigned long long)n_paths : (unsigned long long)ctrl[b].n_extend;
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
 * [0, 0.9999] then *255 truncated (:297-300); float32 copy of the scrubbed linear value (:308-325) */
__device__ __forceinline__ double linear_to_srgb(double c) {
    if (c <= 0.0031308) return 12.92 * c;
    return 1.055 * pow(c, (1. / 2.4)) - 0.055;
}
__global__ void __launch_bounds__(256) k_resolve(const double* __restrict__ rgb, int64_t n, uint8_t* __restrict__ rgb8, float* __restrict__ rgb32f) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double c = rgb[i];
        if (c != c) c = 0.0;
        if (rgb32f) rgb32f[i] = (float)c;
        if (rgb8) {
            double s = linear_to_srgb(c);
            s = s < 0.0 ? 0.0 : (s > 0.9999 ? 0.9999 : s); /* Interval::Clamp */
            rgb8[i] = (uint8_t)(s * 255);
        }
    }
}

/* ------------------------------------------------------------------------------------------------
 * launchers
 * ---------------------------------------------------------------------------------------------- */
void launch_trace_closest(const DevScene& S, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax,
                          int32_t* prim, double* t, double* alpha, double* beta, int32_t* fetch, unsigned long long* counters,
                          const LaunchConfig& lc, cudaStream_t st) {
    if (n <= 0) return;
    int64_t want = (n + 127) / 128;
    const int grid = (int)(want < (int64_t)lc.sm_count * lc.trace_blocks_per_sm ? want : (int64_t)lc.sm_count * lc.trace_blocks_per_sm);
    if (counters) k_trace_closest<true><<<grid, 128, 0, st>>>(S, n, o, d, tmin, tmax, prim, t, alpha, beta, fetch, counters);
    else k_trace_closest<false><<<grid, 128, 0, st>>>(S, n, o, d, tmin, tmax, prim, t, alpha, beta, fetch, counters);
}
void launch_trace_any(const DevScene& S, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax,
                      uint8_t* hit, int32_t* fetch, unsigned long long* counters, const LaunchConfig& lc, cudaStream_t st) {
    if (n <= 0) return;
    int64_t want = (n + 127) / 128;
    const int grid = (int)(want < (int64_t)lc.sm_count * lc.trace_blocks_per_sm ? want : (int64_t)lc.sm_count * lc.trace_blocks_per_sm);
    if (counters) k_trace_any<true><<<grid, 128, 0, st>>>(S, n, o, d, tmin, tmax, hit, fetch, counters);
    else k_trace_any<false><<<grid, 128, 0, st>>>(S, n, o, d, tmin, tmax, hit, fetch, counters);
}
void launch_sample_lights(const DevScene& S, int64_t n, const double* origins, const double* u, double* position, double* normal,
                          uint8_t* front, double* pdf, int32_t* prim, cudaStream_t st) {
    if (n <= 0) return;
    k_sample_lights<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(S, n, origins, u, position, normal, front, pdf, prim);
}

int launch_render_pass(const DevScene& S, const uint8_t* prim_class, const RenderConst& rc, const PathBuffers& pb, double* rgb,
                       unsigned long long* totals, const LaunchConfig& lc, cudaStream_t st) {
    int launches = 0;
    const int n_bounces = rc.max_depth + 1;
    cudaMemsetAsync(pb.ctrl, 0, sizeof(BounceCtrl) * (size_t)(n_bounces + 1), st);
    const bool count = (rc.flags & PRT_RENDER_COUNTERS) != 0;
    const int persistent = lc.sm_count * lc.trace_blocks_per_sm;
    int64_t want = ((int64_t)rc.n_paths + 127) / 128;
    const int grid_paths = (int)(want < persistent ? want : persistent);
    k_raygen<<<(int)(((int64_t)rc.n_paths + 255) / 256 < lc.sm_count * 8 ? ((int64_t)rc.n_paths + 255) / 256 : lc.sm_count * 8), 256, 0, st>>>(rc, pb);
    ++launches;
    cudaEvent_t* ev = lc.stage_events;
    for (int b = 0; b < n_bounces; ++b) {
        if (ev) cudaEventRecord(ev[4 * b], st);
        if (count) k_extend<true><<<grid_paths, 128, 0, st>>>(S, prim_class, rc, pb, b);
        else k_extend<false><<<grid_paths, 128, 0, st>>>(S, prim_class, rc, pb, b);
        if (ev) cudaEventRecord(ev[4 * b + 1], st);
        k_shade<CLS_LAMBERT><<<grid_paths, 128, 0, st>>>(S, rc, pb, b);
        k_shade<CLS_PHONG><<<grid_paths, 128, 0, st>>>(S, rc, pb, b);
        k_shade<CLS_MIRROR><<<grid_paths, 128, 0, st>>>(S, rc, pb, b);
        launches += 4;
        if (ev) cudaEventRecord(ev[4 * b + 2], st);
        if (rc.sample_lights) {
            if (count) k_shadow<true><<<grid_paths, 128, 0, st>>>(S, rc, pb, b);
            else k_shadow<false><<<grid_paths, 128, 0, st>>>(S, rc, pb, b);
            ++launches;
        }
        if (ev) cudaEventRecord(ev[4 * b + 3], st);
    }
    k_accumulate<<<(rc.npix + 255) / 256 < lc.sm_count * 8 ? (rc.npix + 255) / 256 : lc.sm_count * 8, 256, 0, st>>>(rc, pb, rgb);
    k_pass_stats<<<1, 32, 0, st>>>(pb.ctrl, n_bounces, rc.n_paths, totals);
    launches += 2;
    return launches;
}

void launch_resolve(const double* rgb, int64_t npix, uint8_t* rgb8, float* rgb32f, cudaStream_t st) {
    const int64_t n = npix * 3;
    if (n <= 0) return;
    k_resolve<<<(unsigned)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8), 256, 0, st>>>(rgb, n, rgb8, rgb32f);
}

}  // namespace prt

/*
 * capi.cu — the C ABI of include/prt_b200.h: context / scene lifecycle, uploads, and the host side
 * of the wavefront (pass scheduling, stats). No CPU fallback: every compute entry point needs the
 * context's GPU.
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/prt_b200.h"
#include "bvh_device.h"
#include "capi_internal.h"
#include "launch.h"
#include "scene_host.h"

using namespace prt;

namespace {
thread_local std::string g_last_error = "";
int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}
}  // namespace

namespace prt {
int set_error(int code, const std::string& msg) { return fail(code, msg); } /* for scene_io.cpp */
}

namespace {

#define CUDA_TRY(expr)                                                                                         \
    do {                                                                                                       \
        cudaError_t e__ = (expr);                                                                              \
        if (e__ != cudaSuccess) {                                                                              \
            const int code__ = e__ == cudaErrorMemoryAllocation ? PRT_ERR_OUT_OF_MEMORY : PRT_ERR_CUDA;       \
            return fail(code__, std::string(#expr) + ": " + cudaGetErrorString(e__));                         \
        }                                                                                                      \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    ~DevBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    cudaError_t ensure(size_t count) {
        if (count <= n && p) return cudaSuccess;
        release();
        if (count == 0) count = 1;
        cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
        if (e == cudaSuccess) n = count;
        else p = nullptr;
        return e;
    }
    cudaError_t upload(const std::vector<T>& v, cudaStream_t st) {
        cudaError_t e = ensure(v.size());
        if (e != cudaSuccess || v.empty()) return e;
        return cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, st);
    }
};

}  // namespace

struct prt_scene {
    prt_context* ctx = nullptr;
    HostScene host;
    bool committed = false;
    int accel = PRT_ACCEL_SAH_BVH2;

    /* HBM-resident scene */
    DevBuf<TriIsect> d_tris;
    DevBuf<int32_t> d_slot_prim;
    DevBuf<TriShade> d_shade;
    DevBuf<Node2> d_nodes2;
    DevBuf<Node4> d_nodes4;
    DevBuf<Node4Q> d_nodes4q;
    DevBuf<RefNode> d_ref_nodes;
    DevBuf<LightTri> d_light_tris;
    DevBuf<LightNode> d_light_nodes;
    DevBuf<DevMaterial> d_materials;
    DevBuf<DevTexture> d_textures;
    DevBuf<uint8_t> d_texels;
    DevBuf<double> d_srgb;
    DevBuf<uint8_t> d_prim_class;
    bool has_cook = false;
    int class_mask = 0;
    float device_build_ms = 0.f; /* PRT_ACCEL_LBVH4_DEVICE: CUDA-event time of the build */
    DevScene dev;

    /* scratch for the host-pointer entry points */
    DevBuf<double> s_o, s_d, s_tmin, s_tmax, s_t, s_a, s_b, s_u, s_pos, s_nrm, s_pdf;
    DevBuf<int32_t> s_prim;
    DevBuf<uint8_t> s_hit;
    DevBuf<int32_t> d_fetch;             /* work cursor of the ray-query kernels */
    DevBuf<unsigned long long> d_counters; /* [2] node visits, triangle tests */
    DevBuf<unsigned long long> d_err;      /* [1] sticky: subtrees dropped by the reference-order walk (DevScene::ref_overflow) */
    DevBuf<double> s_rgb;
    DevBuf<uint8_t> s_rgb8;
    DevBuf<float> s_rgb32f;
    DevBuf<double> d_thr; /* [254] quantiser thresholds of the output stage (resolve_thresholds) */

    /* wavefront pool */
    DevBuf<PathRec> p_rec;
    DevBuf<PathL> p_lum;
    DevBuf<PrimaryHit> p_primary;
    DevBuf<int32_t> p_q0, p_q1, p_qm0, p_qm1, p_qm2, p_qm3, p_qm4, p_qs;
    DevBuf<BounceCtrl> p_ctrl;
    /* ray binning (launch.h): keys, reordered queues, histograms + cursors; also used by the ray queries */
    DevBuf<uint16_t> p_ke, p_ks;
    DevBuf<int32_t> p_qse, p_qss, p_hist, p_cursor;
    float key_lo[3] = {0.f, 0.f, 0.f}, key_scale[3] = {0.f, 0.f, 0.f};
    int64_t pool_cap = 0;
    int pool_bounces = 0;

    std::vector<cudaEvent_t> stage_events;
    bool count_visits = false;
    prt_stats stats;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timing_pending = false, rays_pending = false, visits_pending = false;

    prt_scene() { std::memset(&stats, 0, sizeof(stats)); std::memset(&dev, 0, sizeof(dev)); }
    ~prt_scene() {
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        for (cudaEvent_t e : stage_events) cudaEventDestroy(e);
    }
};

namespace {

/* PRT_BUILD_TIMING=1: wall time of the host-side stages of scene creation / commit on stderr */
struct HostStageTimer {
    const bool on = std::getenv("PRT_BUILD_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void lap(const char* what) {
        if (!on) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[prt] host %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t).count());
        t = now;
    }
};

int need_device(const prt_scene* s) {
    if (!s) return fail(PRT_ERR_INVALID_ARGUMENT, "null scene");
    if (!s->ctx || s->ctx->device < 0) return fail(PRT_ERR_NO_DEVICE, "this context has no GPU (host-only context): there is no CPU fallback");
    if (!s->committed) return fail(PRT_ERR_NOT_COMMITTED, "scene not committed: call prt_scene_commit first");
    cudaError_t e = cudaSetDevice(s->ctx->device);
    if (e != cudaSuccess) return fail(PRT_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    return PRT_OK;
}

LaunchConfig launch_config(const prt_context* ctx) {
    LaunchConfig lc;
    lc.sm_count = ctx->sm_count;
    const char* env = std::getenv("PRT_BLOCKS_PER_SM");
    lc.trace_blocks_per_sm = env ? std::max(1, std::atoi(env)) : 0; /* 0 = whatever is resident */
    lc.stage_events = nullptr;
    /* PRT_SORT_MIN: shortest (expected) ray queue that is binned before it is traced; 0 = never, the
     * default: measured on B200 (profiles/r1_s4_binning_ab.txt) the 16^3 x 8-octant bins buy 2-6 % of
     * traversal time and cost more than that in histogram atomics and the scatter pass */
    env = std::getenv("PRT_SORT_MIN");
    lc.sort_min = env ? std::max<int64_t>(0, std::atoll(env)) : 0;
    /* PRT_TAIL_PATHS: see LaunchConfig::tail_paths */
    env = std::getenv("PRT_TAIL_PATHS");
    lc.tail_paths = env ? std::max<int64_t>(0, std::atoll(env)) : 262144;
    return lc;
}

int ensure_sort_scratch(prt_scene* s, cudaStream_t st) {
    if (s->p_hist.p) return PRT_OK;
    CUDA_TRY(s->p_hist.ensure((size_t)2 * PRT_KEY_BINS));
    CUDA_TRY(s->p_cursor.ensure((size_t)2 * PRT_KEY_BINS));
    CUDA_TRY(cudaMemsetAsync(s->p_hist.p, 0, sizeof(int32_t) * 2 * PRT_KEY_BINS, st)); /* k_sort_scan keeps it zero */
    return PRT_OK;
}

int64_t pass_path_budget() {
    const char* env = std::getenv("PRT_PASS_PATHS");
    int64_t v = env ? std::atoll(env) : (int64_t(1) << 26); /* 12.9 GB of path records: per-pass fixed costs (tail, drains) amortised over 4x the paths of 2^24, +9 % */
    if (v < 1024) v = 1024;
    const int64_t top = (int64_t(1) << 31) - 64 - 4096; /* slots and queue positions are int32 (rounded up to whole warps) */
    if (v > top) v = top;
    return v;
}

int ensure_pool(prt_scene* s, int64_t cap, int64_t npix, int n_bounces, bool binning) {
    if ((int64_t)s->p_primary.n < npix) CUDA_TRY(s->p_primary.ensure((size_t)npix));
    if (binning && (int64_t)s->p_ke.n < cap) { /* ray-binning scratch only when PRT_SORT_MIN asks for it */
        CUDA_TRY(s->p_ke.ensure((size_t)cap));
        CUDA_TRY(s->p_ks.ensure((size_t)cap));
        CUDA_TRY(s->p_qse.ensure((size_t)cap));
        CUDA_TRY(s->p_qss.ensure((size_t)cap));
    }
    if (cap > s->pool_cap) {
        CUDA_TRY(s->p_rec.ensure((size_t)cap));
        CUDA_TRY(s->p_lum.ensure((size_t)cap));
        CUDA_TRY(s->p_q0.ensure((size_t)cap));
        CUDA_TRY(s->p_q1.ensure((size_t)cap));
        CUDA_TRY(s->p_qm0.ensure((size_t)cap));
        CUDA_TRY(s->p_qm1.ensure((size_t)cap));
        CUDA_TRY(s->p_qm2.ensure((size_t)cap));
        CUDA_TRY(s->p_qm3.ensure((size_t)cap));
        CUDA_TRY(s->p_qm4.ensure((size_t)cap));
        CUDA_TRY(s->p_qs.ensure((size_t)cap));
        s->pool_cap = cap;
    }
    if (n_bounces + 2 > s->pool_bounces) {
        CUDA_TRY(s->p_ctrl.ensure((size_t)n_bounces + 2));
        s->pool_bounces = n_bounces + 2;
    }
    return PRT_OK;
}

PathBuffers path_buffers(prt_scene* s) {
    PathBuffers pb;
    pb.cap = s->pool_cap;
    pb.rec = s->p_rec.p;
    pb.lum = s->p_lum.p;
    pb.primary = s->p_primary.p;
    pb.q_extend[0] = s->p_q0.p; pb.q_extend[1] = s->p_q1.p;
    pb.q_mat[0] = s->p_qm0.p; pb.q_mat[1] = s->p_qm1.p; pb.q_mat[2] = s->p_qm2.p; pb.q_mat[3] = s->p_qm3.p; pb.q_mat[4] = s->p_qm4.p;
    pb.q_shadow = s->p_qs.p;
    pb.ctrl = s->p_ctrl.p;
    pb.key_extend = s->p_ke.p; pb.key_shadow = s->p_ks.p;
    pb.qs_extend = s->p_qse.p; pb.qs_shadow = s->p_qss.p;
    pb.hist = s->p_hist.p; pb.cursor = s->p_cursor.p;
    pb.counters = s->count_visits ? s->d_counters.p : nullptr;
    return pb;
}

/* The reference's 8-bit quantiser (Camera.cpp:216-221,297-300) with the reference's own libm call */
inline int quantise_srgb8(double c) {
    if (c != c) c = 0.0;
    double s = c <= 0.0031308 ? 12.92 * c : 1.055 * std::pow(c, (1. / 2.4)) - 0.055;
    s = s < 0.0 ? 0.0 : (s > 0.9999 ? 0.9999 : s); /* Interval::Clamp */
    return (int)(uint8_t)(s * 255);
}
/* thr[k - 1] = the smallest double whose byte is >= k, k = 1..254, by bisection over the bit patterns of the
 * non-negative doubles (ordered like unsigned integers). Monotonicity of the quantiser is what makes "count the
 * thresholds reached" equal to it; the 64 neighbours on either side of every threshold are verified here, and
 * tests/test_gpu_parity.py holds the bytes the compiled reference wrote for the same neighbourhoods. */
const std::vector<double>& resolve_thresholds() {
    static std::vector<double> thr;
    static std::once_flag once;
    std::call_once(once, [] {
        thr.resize(254);
        auto value = [](uint64_t bits) { double v; std::memcpy(&v, &bits, 8); return v; };
        uint64_t one_bits; { const double one = 1.0; std::memcpy(&one_bits, &one, 8); }
        for (int k = 1; k <= 254; ++k) {
            uint64_t lo = 0, hi = one_bits; /* q(lo) = 0 < k <= 254 = q(1.0) */
            while (hi - lo > 1) {
                const uint64_t mid = lo + (hi - lo) / 2;
                if (quantise_srgb8(value(mid)) >= k) hi = mid; else lo = mid;
            }
            for (int j = 1; j <= 64; ++j) /* monotone around the step, or the comparison form would be wrong */
                if (quantise_srgb8(value(hi + j)) < k || (hi > (uint64_t)j && quantise_srgb8(value(hi - j)) >= k)) {
                    std::fprintf(stderr, "[prt] resolve_thresholds: quantiser not monotone near byte %d\n", k);
                    std::abort();
                }
            thr[k - 1] = value(hi);
        }
    });
    return thr;
}

/* Camera::Initialize, Source/Camera.cpp:75-106 (host libm tan, the function the reference calls) */
void camera_setup(const prt_camera* c, RenderConst& rc) {
    const int W = c->image_width < 1 ? 1 : c->image_width;
    const int H = c->image_height < 1 ? 1 : c->image_height;
    const double aspect = double(W) / double(H);
    const d3 eye = mk3(c->eye[0], c->eye[1], c->eye[2]), look = mk3(c->look_at[0], c->look_at[1], c->look_at[2]);
    const d3 up = mk3(c->up[0], c->up[1], c->up[2]);
    const double focal = length(eye - look);
    const double theta = c->fovy * 0.01745329251994329576923690768489; /* glm::radians */
    const double h = std::tan(theta / 2.0);
    const double vh = 2. * h * focal;
    const double vw = vh * aspect;
    const d3 w = normalize(eye - look);
    const d3 u = normalize(cross(up, w));
    const d3 v = cross(w, u);
    const d3 vu = vw * u;
    const d3 vv = vh * (-v);
    rc.width = W; rc.height = H; rc.npix = W * H;
    rc.by_width = FastDiv::make((uint32_t)W);
    rc.by_npix = FastDiv::make((uint32_t)(W * H));
    rc.center = eye;
    rc.dU = vu / (double)W;
    rc.dV = vv / (double)H;
    const d3 upper_left = eye - (focal * w) - vu / 2. - vv / 2.;
    rc.pixel00 = upper_left + 0.5 * (rc.dU + rc.dV);
}

}  // namespace

extern "C" {

int prt_abi_version(void) { return PRT_ABI_VERSION; }
const char* prt_last_error(void) { return g_last_error.c_str(); }

int prt_context_create(int device, prt_context** out) {
    if (!out) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_context_create: out is null");
    *out = nullptr;
    prt_context* c = new (std::nothrow) prt_context();
    if (!c) return fail(PRT_ERR_OUT_OF_MEMORY, "prt_context_create: out of host memory");
    if (device >= 0) {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess || device >= n) {
            delete c;
            return fail(PRT_ERR_NO_DEVICE, e != cudaSuccess ? std::string("no CUDA device: ") + cudaGetErrorString(e)
                                                             : "device index out of range");
        }
        cudaDeviceProp prop;
        e = cudaGetDeviceProperties(&prop, device);
        if (e != cudaSuccess) { delete c; return fail(PRT_ERR_CUDA, cudaGetErrorString(e)); }
        if (prop.major != 10) {
            delete c;
            return fail(PRT_ERR_NO_DEVICE, std::string("device is sm_") + std::to_string(prop.major) + std::to_string(prop.minor) +
                                               ", this library carries sm_100a code only");
        }
        e = cudaSetDevice(device);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { delete c; return fail(PRT_ERR_CUDA, cudaGetErrorString(e)); }
        c->sm_count = prop.multiProcessorCount;
        c->device = device;
    }
    *out = c;
    return PRT_OK;
}

void prt_context_destroy(prt_context* ctx) {
    if (!ctx) return;
    if (ctx->device >= 0) {
        cudaSetDevice(ctx->device);
        if (ctx->stream) cudaStreamDestroy(ctx->stream);
    }
    delete ctx;
}

int prt_scene_create(prt_context* ctx, const prt_scene_desc* desc, prt_scene** out) {
    if (!ctx || !desc || !out) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_scene_create: null argument");
    *out = nullptr;
    prt_scene* s = new (std::nothrow) prt_scene();
    if (!s) return fail(PRT_ERR_OUT_OF_MEMORY, "prt_scene_create: out of host memory");
    s->ctx = ctx;
    try {
        HostStageTimer timer;
        if (!s->host.ingest(desc)) {
            std::string msg = s->host.error;
            delete s;
            return fail(PRT_ERR_INVALID_ARGUMENT, msg);
        }
        timer.lap("ingest + triangle records");
        s->host.build_reference_bvhs(default_host_threads());
        timer.lap("reference BVHs (DFS order)");
    } catch (const std::bad_alloc&) {
        delete s;
        return fail(PRT_ERR_OUT_OF_MEMORY, "prt_scene_create: out of host memory");
    }
    *out = s;
    return PRT_OK;
}

void prt_scene_destroy(prt_scene* scene) {
    if (!scene) return;
    if (scene->ctx && scene->ctx->device >= 0) cudaSetDevice(scene->ctx->device);
    delete scene;
}

int prt_scene_commit(prt_scene* s, const prt_build_opts* opts) {
    if (!s) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_scene_commit: null scene");
    int accel = opts ? opts->accel : PRT_ACCEL_SAH_BVH2;
    if (accel == PRT_ACCEL_AUTO) accel = s->host.n_tris >= 262144 ? PRT_ACCEL_SAH_BVH4Q : PRT_ACCEL_SAH_BVH4;
    if (accel != PRT_ACCEL_REFERENCE_BVH2 && accel != PRT_ACCEL_SAH_BVH2 && accel != PRT_ACCEL_SAH_BVH4 && accel != PRT_ACCEL_LBVH4_DEVICE && accel != PRT_ACCEL_SAH_BVH4Q)
        return fail(PRT_ERR_UNSUPPORTED, "prt_scene_commit: unknown accel");
    HostScene& h = s->host;
    const bool dev_build = accel == PRT_ACCEL_LBVH4_DEVICE;
    if (dev_build && s->ctx->device < 0)
        return fail(PRT_ERR_NO_DEVICE, "prt_scene_commit: PRT_ACCEL_LBVH4_DEVICE builds on the GPU and this context has none (no CPU fallback)");
    const int max_leaf = opts && opts->reserved[0] > 0 ? opts->reserved[0] : 4;
    double scene_lo[3] = {0, 0, 0}, scene_hi[3] = {0, 0, 0}; /* bounding box of all triangles */
    for (int64_t p = 0; p < h.n_tris; ++p)
        for (int k = 0; k < 3; ++k) {
            const double a = h.tri_box[p].v[2 * k], b = h.tri_box[p].v[2 * k + 1];
            if (p == 0 || a < scene_lo[k]) scene_lo[k] = a;
            if (p == 0 || b > scene_hi[k]) scene_hi[k] = b;
        }
    DeviceBvh dbvh;
    HostStageTimer stage_timer;
    try {
        if (dev_build) {
            /* Morton sort, hierarchy, boxes, BVH4 emission and the triangle permutation
             * all happen in HBM (bvh_device.cu); the host only learns the slot order, which the
             * flattened reference tree's leaves refer to */
            CUDA_TRY(cudaSetDevice(s->ctx->device));
            cudaStream_t st = s->ctx->stream;
            DevBuf<Box6> d_box;
            DevBuf<TriIsect> d_isect;
            CUDA_TRY(d_box.upload(h.tri_box, st));
            CUDA_TRY(d_isect.upload(h.isect, st));
            cudaError_t e = build_lbvh4_device(d_box.p, d_isect.p, h.n_tris, scene_lo, scene_hi, max_leaf, &dbvh, st);
            if (e != cudaSuccess) {
                cudaFree(dbvh.nodes4); cudaFree(dbvh.slot_prim); cudaFree(dbvh.tris);
                return fail(e == cudaErrorMemoryAllocation ? PRT_ERR_OUT_OF_MEMORY : PRT_ERR_CUDA, std::string("device BVH build: ") + cudaGetErrorString(e));
            }
            h.nodes2.clear();
            h.nodes4.clear();
            h.nodes4q.clear();
            h.isect_slots.clear();
            h.sah_max_leaf = max_leaf;
            h.slot_prim.assign((size_t)h.n_tris, 0);
            h.prim_slot.assign((size_t)h.n_tris, -1);
            if (h.n_tris > 0) {
                e = cudaMemcpy(h.slot_prim.data(), dbvh.slot_prim, sizeof(int32_t) * (size_t)h.n_tris, cudaMemcpyDeviceToHost);
                if (e != cudaSuccess) {
                    cudaFree(dbvh.nodes4); cudaFree(dbvh.slot_prim); cudaFree(dbvh.tris);
                    return fail(PRT_ERR_CUDA, std::string("device BVH build: ") + cudaGetErrorString(e));
                }
            }
            for (int64_t sl = 0; sl < h.n_tris; ++sl) h.prim_slot[(size_t)h.slot_prim[(size_t)sl]] = (int32_t)sl;
            s->d_nodes4.release(); s->d_slot_prim.release(); s->d_tris.release();
            s->d_nodes4.p = dbvh.nodes4; s->d_nodes4.n = (size_t)dbvh.n_nodes4;
            s->d_slot_prim.p = dbvh.slot_prim; s->d_slot_prim.n = (size_t)h.n_tris;
            s->d_tris.p = dbvh.tris; s->d_tris.n = (size_t)h.n_tris;
        } else if (accel == PRT_ACCEL_SAH_BVH2) h.build_sah_bvh2(default_host_threads(), max_leaf);
        else if (accel == PRT_ACCEL_SAH_BVH4) h.build_sah_bvh4(default_host_threads(), max_leaf);
        else if (accel == PRT_ACCEL_SAH_BVH4Q) { h.build_sah_bvh4(default_host_threads(), max_leaf); stage_timer.lap("SAH tree + BVH4 + slots"); h.quantise_bvh4(); }
        else h.use_reference_slots();
        stage_timer.lap("traversal tree");
        h.flatten_reference_world();
        stage_timer.lap("reference tree flattened");
    } catch (const std::bad_alloc&) {
        return fail(PRT_ERR_OUT_OF_MEMORY, "prt_scene_commit: out of host memory");
    }
    s->accel = accel;
    const bool wide4 = accel == PRT_ACCEL_SAH_BVH4 || dev_build;
    const bool quant = accel == PRT_ACCEL_SAH_BVH4Q;
    s->stats.node_bytes = accel == PRT_ACCEL_SAH_BVH2 ? PRT_NODE2_BYTES : (wide4 ? PRT_NODE4_BYTES : (quant ? PRT_NODE4Q_BYTES : PRT_REFNODE_BYTES));
    s->stats.tri_bytes = PRT_TRI_BYTES;
    s->stats.n_nodes = accel == PRT_ACCEL_SAH_BVH2 ? (int64_t)h.nodes2.size()
                                                   : (dev_build ? dbvh.n_nodes4 : (accel == PRT_ACCEL_SAH_BVH4 || quant ? (int64_t)h.nodes4.size() : (int64_t)h.ref_flat.size()));
    s->stats.n_triangles = h.n_tris;
    s->device_build_ms = dbvh.build_ms;
    s->stats.build_ms = dbvh.build_ms;
    for (int k = 0; k < 3; ++k) { /* ray-binning grid: PRT_KEY_GRID^3 cells over the scene's bounding box */
        const double ext = scene_hi[k] - scene_lo[k];
        s->key_lo[k] = (float)scene_lo[k];
        s->key_scale[k] = ext > 0.0 && std::isfinite(ext) ? (float)(PRT_KEY_GRID / ext) : 0.f;
    }
    if (s->ctx->device < 0) { /* host-only context: flattening done, nothing to upload */
        s->committed = true;
        return PRT_OK;
    }
    CUDA_TRY(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    std::vector<uint8_t> prim_class((size_t)h.n_tris);
    s->has_cook = false;
    s->class_mask = 0;
    for (int64_t p = 0; p < h.n_tris; ++p) {
        const DevMaterial& m = h.dev_materials[h.shade[p].material];
        uint8_t c;
        if (m.has_emission) c = CLS_EMISSIVE;
        else if (m.type == PRT_MAT_LAMBERTIAN) c = CLS_LAMBERT;
        else if (m.type == PRT_MAT_PHONG) c = m.pks > 0.0 ? CLS_PHONG_SPEC : CLS_PHONG; /* launch.h: the specular lobe gets its own queue */
        else if (m.type == PRT_MAT_MIRROR) c = CLS_MIRROR;
        else if (m.type == PRT_MAT_COOK_TORRANCE) c = CLS_COOK;
        else if (m.type == PRT_MAT_EMPTY) c = CLS_EMPTY;
        else c = CLS_LAMBERT; /* unknown types fall back to Lambertian like Model.cpp:282-285 */
        prim_class[p] = c;
        if (c == CLS_COOK) s->has_cook = true;
        if (c < CLS_SHADED) s->class_mask |= 1 << c;
    }
    std::vector<double> lut(h.srgb_lut, h.srgb_lut + 256);
    if (!dev_build) { /* the device build left these three in HBM already */
        CUDA_TRY(s->d_tris.upload(h.isect_slots, st));
        CUDA_TRY(s->d_slot_prim.upload(h.slot_prim, st));
        CUDA_TRY(s->d_nodes4.upload(h.nodes4, st));
    }
    CUDA_TRY(s->d_shade.upload(h.shade, st));
    CUDA_TRY(s->d_nodes2.upload(h.nodes2, st));
    if (quant) CUDA_TRY(s->d_nodes4q.upload(h.nodes4q, st));
    CUDA_TRY(s->d_ref_nodes.upload(h.ref_flat, st));
    CUDA_TRY(s->d_light_tris.upload(h.light_tris, st));
    CUDA_TRY(s->d_light_nodes.upload(h.light_nodes, st));
    CUDA_TRY(s->d_materials.upload(h.dev_materials, st));
    CUDA_TRY(s->d_textures.upload(h.dev_textures, st));
    CUDA_TRY(s->d_texels.upload(h.texel_pool, st));
    CUDA_TRY(s->d_srgb.upload(lut, st));
    CUDA_TRY(s->d_prim_class.upload(prim_class, st));
    CUDA_TRY(s->d_fetch.ensure(4));
    CUDA_TRY(s->d_counters.ensure(8));
    CUDA_TRY(cudaMemsetAsync(s->d_counters.p, 0, 8 * sizeof(unsigned long long), st));
    CUDA_TRY(s->d_err.ensure(1));
    CUDA_TRY(cudaMemsetAsync(s->d_err.p, 0, sizeof(unsigned long long), st));
    if (int e = ensure_sort_scratch(s, st)) return e;
    CUDA_TRY(cudaStreamSynchronize(st));
    DevScene& D = s->dev;
    D.tris = s->d_tris.p;
    D.slot_prim = s->d_slot_prim.p;
    D.shade = s->d_shade.p;
    D.nodes2 = accel == PRT_ACCEL_SAH_BVH2 && !h.nodes2.empty() ? s->d_nodes2.p : nullptr;
    D.nodes4 = (accel == PRT_ACCEL_SAH_BVH4 && !h.nodes4.empty()) || (dev_build && h.n_tris > 0) ? s->d_nodes4.p : nullptr;
    D.nodes4q = quant && !h.nodes4q.empty() ? s->d_nodes4q.p : nullptr;
    D.ref_nodes = s->d_ref_nodes.p;
    D.ref_root = h.ref_root_flat;
    D.n_nodes2 = (int32_t)h.nodes2.size();
    D.n_tris = h.n_tris;
    D.light_tris = s->d_light_tris.p;
    D.light_nodes = s->d_light_nodes.p;
    D.light_root = h.light_root_flat;
    D.n_light_tris = (int32_t)h.light_tris.size();
    D.lights_area = h.lights_area;
    D.materials = s->d_materials.p;
    D.textures = s->d_textures.p;
    D.texels = s->d_texels.p;
    D.srgb_lut = s->d_srgb.p;
    D.n_materials = (int32_t)h.dev_materials.size();
    D.n_textures = (int32_t)h.dev_textures.size();
    D.ref_overflow = s->d_err.p;
    if (!s->ev0) {
        CUDA_TRY(cudaEventCreate(&s->ev0));
        CUDA_TRY(cudaEventCreate(&s->ev1));
    }
    s->committed = true;
    return PRT_OK;
}

/* A replica of a committed scene on another GPU: every device buffer is copied GPU to GPU (cudaMemcpyPeer: NVLink when the
 * devices are peers), the pointer table is rebound, nothing is rebuilt or re-uploaded from the host. The replica holds no
 * host-side scene (prt_scene_reference_order answers on the original only). */
int prt_scene_clone(const prt_scene* src, prt_context* ctx, prt_scene** out) {
    if (!src || !ctx || !out) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_scene_clone: null argument");
    *out = nullptr;
    if (!src->committed || !src->ctx || src->ctx->device < 0 || ctx->device < 0)
        return fail(PRT_ERR_NO_DEVICE, "prt_scene_clone: source and destination must be committed, GPU-bound scenes / contexts");
    prt_scene* s = new (std::nothrow) prt_scene();
    if (!s) return fail(PRT_ERR_OUT_OF_MEMORY, "prt_scene_clone: out of host memory");
    s->ctx = ctx;
    s->accel = src->accel;
    s->has_cook = src->has_cook;
    s->class_mask = src->class_mask;
    s->device_build_ms = src->device_build_ms;
    s->stats = src->stats;
    for (int k = 0; k < 3; ++k) { s->key_lo[k] = src->key_lo[k]; s->key_scale[k] = src->key_scale[k]; }
    const int sd = src->ctx->device, dd = ctx->device;
    cudaStream_t st = ctx->stream;
    auto bail = [&](cudaError_t e, const char* what) {
        const std::string msg = std::string("prt_scene_clone: ") + what + ": " + cudaGetErrorString(e);
        delete s;
        return fail(e == cudaErrorMemoryAllocation ? PRT_ERR_OUT_OF_MEMORY : PRT_ERR_CUDA, msg);
    };
    cudaError_t e = cudaSetDevice(dd);
    if (e != cudaSuccess) return bail(e, "cudaSetDevice");
    /* make sure the source's uploads have landed before they are read from another device's stream */
    cudaSetDevice(sd);
    e = cudaStreamSynchronize(src->ctx->stream);
    cudaSetDevice(dd);
    if (e != cudaSuccess) return bail(e, "cudaStreamSynchronize(source)");
#define PRT_CLONE(member)                                                                                                        \
    do {                                                                                                                         \
        if (src->member.p && src->member.n) {                                                                                    \
            e = s->member.ensure(src->member.n);                                                                                 \
            if (e != cudaSuccess) return bail(e, "allocate " #member);                                                           \
            e = cudaMemcpyPeerAsync(s->member.p, dd, src->member.p, sd, src->member.n * sizeof(*src->member.p), st);             \
            if (e != cudaSuccess) return bail(e, "copy " #member);                                                               \
        }                                                                                                                        \
    } while (0)
    PRT_CLONE(d_tris); PRT_CLONE(d_slot_prim); PRT_CLONE(d_shade); PRT_CLONE(d_nodes2); PRT_CLONE(d_nodes4); PRT_CLONE(d_nodes4q);
    PRT_CLONE(d_ref_nodes); PRT_CLONE(d_light_tris); PRT_CLONE(d_light_nodes); PRT_CLONE(d_materials); PRT_CLONE(d_textures);
    PRT_CLONE(d_texels); PRT_CLONE(d_srgb); PRT_CLONE(d_prim_class);
#undef PRT_CLONE
    e = s->d_fetch.ensure(4);
    if (e == cudaSuccess) e = s->d_counters.ensure(8);
    if (e == cudaSuccess) e = s->d_err.ensure(1);
    if (e == cudaSuccess) e = cudaMemsetAsync(s->d_counters.p, 0, 8 * sizeof(unsigned long long), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(s->d_err.p, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return bail(e, "scratch");
    if (int rc = ensure_sort_scratch(s, st)) { delete s; return rc; }
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return bail(e, "cudaStreamSynchronize");
    /* the pointer table: same fields set / unset as in the source, pointing at this GPU's copies */
    const DevScene& A = src->dev;
    DevScene& D = s->dev;
    D = A;
    D.tris = s->d_tris.p;
    D.slot_prim = s->d_slot_prim.p;
    D.shade = s->d_shade.p;
    D.nodes2 = A.nodes2 ? s->d_nodes2.p : nullptr;
    D.nodes4 = A.nodes4 ? s->d_nodes4.p : nullptr;
    D.nodes4q = A.nodes4q ? s->d_nodes4q.p : nullptr;
    D.ref_nodes = s->d_ref_nodes.p;
    D.light_tris = s->d_light_tris.p;
    D.light_nodes = s->d_light_nodes.p;
    D.materials = s->d_materials.p;
    D.textures = s->d_textures.p;
    D.texels = s->d_texels.p;
    D.srgb_lut = s->d_srgb.p;
    D.ref_overflow = s->d_err.p;
    if (cudaEventCreate(&s->ev0) != cudaSuccess || cudaEventCreate(&s->ev1) != cudaSuccess) return bail(cudaErrorUnknown, "cudaEventCreate");
    s->committed = true;
    *out = s;
    return PRT_OK;
}

int prt_scene_reference_order(const prt_scene* s, int64_t* world_rank, int64_t* light_order, int64_t* n_light_tris) {
    if (!s) return fail(PRT_ERR_INVALID_ARGUMENT, "null scene");
    if (world_rank) std::copy(s->host.world_rank.begin(), s->host.world_rank.end(), world_rank);
    if (light_order) std::copy(s->host.light_order.begin(), s->host.light_order.end(), light_order);
    if (n_light_tris) *n_light_tris = (int64_t)s->host.light_order.size();
    return PRT_OK;
}

int prt_set_counters(prt_scene* s, int enable) {
    if (!s) return fail(PRT_ERR_INVALID_ARGUMENT, "null scene");
    s->count_visits = enable != 0;
    return PRT_OK;
}

/* ---- ray queries ----------------------------------------------------------------------------- */
static int trace_begin(prt_scene* s, int64_t n, cudaStream_t st) {
    if (n < 0 || n >= (int64_t(1) << 31) - 64) return fail(PRT_ERR_INVALID_ARGUMENT, "ray count out of range");
    CUDA_TRY(cudaMemsetAsync(s->d_fetch.p, 0, 4 * sizeof(int32_t), st));
    CUDA_TRY(cudaMemsetAsync(s->d_counters.p, 0, 8 * sizeof(unsigned long long), st));
    CUDA_TRY(cudaEventRecord(s->ev0, st));
    return PRT_OK;
}
/* ray binning of a query batch (inside the timed region): *order = the binned ray order, or nullptr
 * for a batch too short to be worth three more launches */
static int query_order(prt_scene* s, int64_t n, const double* o, const double* d, const LaunchConfig& lc, cudaStream_t st,
                       const int32_t** order) {
    *order = nullptr;
    if (lc.sort_min <= 0 || n < lc.sort_min) return PRT_OK;
    CUDA_TRY(s->p_ke.ensure((size_t)n));
    CUDA_TRY(s->p_qse.ensure((size_t)n));
    launch_query_sort(n, o, d, s->key_lo, s->key_scale, s->p_ke.p, s->p_hist.p, s->p_cursor.p, s->p_qse.p, lc, st);
    *order = s->p_qse.p;
    return PRT_OK;
}
static int trace_end(prt_scene* s, int64_t n, bool shadow, cudaStream_t st) {
    CUDA_TRY(cudaEventRecord(s->ev1, st));
    CUDA_TRY(cudaGetLastError());
    s->timing_pending = true;
    s->rays_pending = false;
    s->visits_pending = s->count_visits;
    s->stats.node_visits = s->stats.tri_tests = 0;
    s->stats.extend_rays = shadow ? 0 : (uint64_t)n;
    s->stats.shadow_rays = shadow ? (uint64_t)n : 0;
    s->stats.paths = 0;
    s->stats.kernel_launches = n > 0 ? 1 : 0;
    return PRT_OK;
}

int prt_trace_closest_device(prt_scene* s, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax,
                             int32_t* prim, double* t, double* alpha, double* beta, void* stream) {
    int rc = need_device(s);
    if (rc) return rc;
    if (n > 0 && (!o || !d || !tmin || !tmax || !prim)) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_closest: null array");
    cudaStream_t st = stream ? (cudaStream_t)stream : s->ctx->stream;
    rc = trace_begin(s, n, st);
    if (rc) return rc;
    const LaunchConfig lc = launch_config(s->ctx);
    const int32_t* order = nullptr;
    rc = query_order(s, n, o, d, lc, st, &order);
    if (rc) return rc;
    launch_trace_closest(s->dev, n, o, d, tmin, tmax, prim, t, alpha, beta, order, s->d_fetch.p, s->count_visits ? s->d_counters.p : nullptr,
                         lc, st);
    return trace_end(s, n, false, st);
}

int prt_trace_any_device(prt_scene* s, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax,
                         uint8_t* hit, void* stream) {
    int rc = need_device(s);
    if (rc) return rc;
    if (n > 0 && (!o || !d || !tmin || !tmax || !hit)) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_any: null array");
    cudaStream_t st = stream ? (cudaStream_t)stream : s->ctx->stream;
    rc = trace_begin(s, n, st);
    if (rc) return rc;
    const LaunchConfig lc = launch_config(s->ctx);
    const int32_t* order = nullptr;
    rc = query_order(s, n, o, d, lc, st, &order);
    if (rc) return rc;
    launch_trace_any(s->dev, n, o, d, tmin, tmax, hit, order, s->d_fetch.p, s->count_visits ? s->d_counters.p : nullptr, lc, st);
    return trace_end(s, n, true, st);
}

int prt_trace_closest(prt_scene* s, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax,
                      int32_t* prim, double* t, double* alpha, double* beta) {
    int rc = need_device(s);
    if (rc) return rc;
    if (n > 0 && (!o || !d || !tmin || !tmax || !prim)) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_closest: null array");
    if (n <= 0) return PRT_OK;
    cudaStream_t st = s->ctx->stream;
    const size_t N = (size_t)n;
    CUDA_TRY(s->s_o.ensure(3 * N)); CUDA_TRY(s->s_d.ensure(3 * N)); CUDA_TRY(s->s_tmin.ensure(N)); CUDA_TRY(s->s_tmax.ensure(N));
    CUDA_TRY(s->s_prim.ensure(N)); CUDA_TRY(s->s_t.ensure(N)); CUDA_TRY(s->s_a.ensure(N)); CUDA_TRY(s->s_b.ensure(N));
    CUDA_TRY(cudaMemcpyAsync(s->s_o.p, o, 3 * N * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(s->s_d.p, d, 3 * N * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(s->s_tmin.p, tmin, N * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(s->s_tmax.p, tmax, N * sizeof(double), cudaMemcpyHostToDevice, st));
    rc = prt_trace_closest_device(s, n, s->s_o.p, s->s_d.p, s->s_tmin.p, s->s_tmax.p, s->s_prim.p, s->s_t.p, s->s_a.p, s->s_b.p, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(prim, s->s_prim.p, N * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    if (t) CUDA_TRY(cudaMemcpyAsync(t, s->s_t.p, N * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (alpha) CUDA_TRY(cudaMemcpyAsync(alpha, s->s_a.p, N * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (beta) CUDA_TRY(cudaMemcpyAsync(beta, s->s_b.p, N * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return PRT_OK;
}

int prt_trace_any(prt_scene* s, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax, uint8_t* hit) {
    int rc = need_device(s);
    if (rc) return rc;
    if (n > 0 && (!o || !d || !tmin || !tmax || !hit)) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_any: null array");
    if (n <= 0) return PRT_OK;
    cudaStream_t st = s->ctx->stream;
    const size_t N = (size_t)n;
    CUDA_TRY(s->s_o.ensure(3 * N)); CUDA_TRY(s->s_d.ensure(3 * N)); CUDA_TRY(s->s_tmin.ensure(N)); CUDA_TRY(s->s_tmax.ensure(N));
    CUDA_TRY(s->s_hit.ensure(N));
    CUDA_TRY(cudaMemcpyAsync(s->s_o.p, o, 3 * N * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(s->s_d.p, d, 3 * N * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(s->s_tmin.p, tmin, N * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(s->s_tmax.p, tmax, N * sizeof(double), cudaMemcpyHostToDevice, st));
    rc = prt_trace_any_device(s, n, s->s_o.p, s->s_d.p, s->s_tmin.p, s->s_tmax.p, s->s_hit.p, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(hit, s->s_hit.p, N, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return PRT_OK;
}

int prt_sample_lights(prt_scene* s, int64_t n, const double* origins, const double* u, double* position, double* normal,
                      uint8_t* front, double* pdf, int32_t* prim) {
    int rc = need_device(s);
    if (rc) return rc;
    if (s->dev.light_root < 0) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_sample_lights: the scene has no emissive mesh");
    if (n > 0 && (!origins || !u || !position || !normal || !front || !pdf || !prim)) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_sample_lights: null array");
    if (n <= 0) return PRT_OK;
    cudaStream_t st = s->ctx->stream;
    const size_t N = (size_t)n;
    CUDA_TRY(s->s_o.ensure(3 * N)); CUDA_TRY(s->s_u.ensure(4 * N)); CUDA_TRY(s->s_pos.ensure(3 * N)); CUDA_TRY(s->s_nrm.ensure(3 * N));
    CUDA_TRY(s->s_hit.ensure(N)); CUDA_TRY(s->s_pdf.ensure(N)); CUDA_TRY(s->s_prim.ensure(N));
    CUDA_TRY(cudaMemcpyAsync(s->s_o.p, origins, 3 * N * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(s->s_u.p, u, 4 * N * sizeof(double), cudaMemcpyHostToDevice, st));
    launch_sample_lights(s->dev, n, s->s_o.p, s->s_u.p, s->s_pos.p, s->s_nrm.p, s->s_hit.p, s->s_pdf.p, s->s_prim.p, st);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(position, s->s_pos.p, 3 * N * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(normal, s->s_nrm.p, 3 * N * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(front, s->s_hit.p, N, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(pdf, s->s_pdf.p, N * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(prim, s->s_prim.p, N * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return PRT_OK;
}

/* ---- rendering ------------------------------------------------------------------------------- */
int prt_render_device(prt_scene* s, const prt_camera* camera, const prt_render_params* params, double* rgb_device, void* stream) {
    int rc = need_device(s);
    if (rc) return rc;
    if (!camera || !params || !rgb_device) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_render: null argument");
    if (params->samples_per_pixel < 1 || params->max_depth < 0 || params->max_depth > 100000)
        return fail(PRT_ERR_INVALID_ARGUMENT, "prt_render: samples_per_pixel < 1 or max_depth out of range");
    if (params->sample_lights && s->dev.light_root < 0)
        return fail(PRT_ERR_INVALID_ARGUMENT, "prt_render: sample_lights is set but the scene has no emissive mesh");
    int sb = params->spp_begin, se = params->spp_end;
    if (sb == 0 && se == 0) se = params->samples_per_pixel;
    if (sb < 0 || se < sb) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_render: bad sample range");
    cudaStream_t st = stream ? (cudaStream_t)stream : s->ctx->stream;

    RenderConst rcn;
    std::memset(&rcn, 0, sizeof(rcn));
    camera_setup(camera, rcn);
    if ((int64_t)rcn.width * rcn.height >= (int64_t(1) << 31) / 4) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_render: image too large");
    rcn.max_depth = params->max_depth;
    rcn.sample_lights = params->sample_lights ? 1 : 0;
    rcn.has_cook = s->has_cook ? 1 : 0;
    rcn.class_mask = s->class_mask;
    /* primary-hit reuse (launch.h: PrimaryHit): valid whenever every sample of a pixel starts with the same ray, i.e.
     * always for the reference's Camera::GetRay; PRT_PRIMARY_REUSE=0 keeps the per-sample bounce 0 (A/B, tests) */
    {
        const char* env = std::getenv("PRT_PRIMARY_REUSE");
        rcn.primary_reuse = (params->flags & PRT_RENDER_PIXEL_JITTER) || (env && std::atoi(env) == 0) ? 0 : 1;
    }
    rcn.flags = params->flags | (s->count_visits ? PRT_RENDER_COUNTERS : 0u);
    rcn.rr = params->russian_roulette;
    rcn.scale = 1.0 / params->samples_per_pixel;
    for (int k = 0; k < 3; ++k) rcn.background[k] = params->background[k];
    rcn.rng_key = params->rng_key;
    for (int k = 0; k < 3; ++k) { rcn.key_lo[k] = s->key_lo[k]; rcn.key_scale[k] = s->key_scale[k]; }

    const int64_t budget = pass_path_budget();
    int per_pass = (int)std::max<int64_t>(1, budget / rcn.npix);
    per_pass = std::min(per_pass, std::max(1, se - sb));
    const int n_bounces = params->max_depth + 1;
    rc = ensure_pool(s, (int64_t)per_pass * rcn.npix, rcn.npix, n_bounces, launch_config(s->ctx).sort_min > 0);
    if (rc) return rc;
    const bool counting = (rcn.flags & PRT_RENDER_COUNTERS) != 0;
    CUDA_TRY(cudaMemsetAsync(s->d_counters.p, 0, 8 * sizeof(unsigned long long), st));
    PathBuffers pb = path_buffers(s);
    pb.counters = counting ? s->d_counters.p : nullptr;
    LaunchConfig lc = launch_config(s->ctx);
    const bool stage_timing = (rcn.flags & PRT_RENDER_STAGE_TIMING) != 0;
    if (stage_timing) {
        while (s->stage_events.size() < (size_t)4 * n_bounces) {
            cudaEvent_t e;
            CUDA_TRY(cudaEventCreate(&e));
            s->stage_events.push_back(e);
        }
        lc.stage_events = s->stage_events.data();
    }
    double extend_ms = 0.0, shade_ms = 0.0, shadow_ms = 0.0;
    uint64_t launches = 0;
    CUDA_TRY(cudaEventRecord(s->ev0, st));
    for (int s0 = sb; s0 < se; s0 += per_pass) {
        rcn.sample_first = s0;
        rcn.n_samples = std::min(per_pass, se - s0);
        rcn.n_paths = rcn.n_samples * rcn.npix;
        int tail_first = n_bounces;
        launches += (uint64_t)launch_render_pass(s->dev, s->d_prim_class.p, rcn, pb, rgb_device, s->d_counters.p, lc, st, &tail_first);
        CUDA_TRY(cudaGetLastError());
        if (stage_timing) {
            CUDA_TRY(cudaStreamSynchronize(st));
            if (tail_first < n_bounces) { /* the tail launch (all stages of the late bounces) is booked as traversal */
                float a = 0.f;
                cudaEventElapsedTime(&a, lc.stage_events[4 * tail_first], lc.stage_events[4 * tail_first + 1]);
                extend_ms += a;
            }
            for (int b = 0; b < tail_first; ++b) {
                float a = 0.f, c = 0.f, d = 0.f;
                cudaEventElapsedTime(&a, lc.stage_events[4 * b], lc.stage_events[4 * b + 1]);
                cudaEventElapsedTime(&c, lc.stage_events[4 * b + 1], lc.stage_events[4 * b + 2]);
                cudaEventElapsedTime(&d, lc.stage_events[4 * b + 2], lc.stage_events[4 * b + 3]);
                extend_ms += a; shade_ms += c; shadow_ms += d;
            }
        }
    }
    s->stats.extend_ms = extend_ms;
    s->stats.shade_ms = shade_ms;
    s->stats.shadow_ms = shadow_ms;
    CUDA_TRY(cudaEventRecord(s->ev1, st));
    s->timing_pending = true;
    s->rays_pending = true;
    s->visits_pending = counting;
    s->stats.extend_rays = 0;
    s->stats.shadow_rays = 0;
    s->stats.node_visits = 0;
    s->stats.tri_tests = 0;
    s->stats.paths = (uint64_t)rcn.npix * (uint64_t)(se - sb);
    s->stats.kernel_launches = launches;
    return PRT_OK;
}

int prt_render(prt_scene* s, const prt_camera* camera, const prt_render_params* params, double* rgb) {
    int rc = need_device(s);
    if (rc) return rc;
    if (!camera || !params || !rgb) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_render: null argument");
    const int W = camera->image_width < 1 ? 1 : camera->image_width, H = camera->image_height < 1 ? 1 : camera->image_height;
    const size_t n = (size_t)W * H * 3;
    cudaStream_t st = s->ctx->stream;
    CUDA_TRY(s->s_rgb.ensure(n));
    CUDA_TRY(cudaMemsetAsync(s->s_rgb.p, 0, n * sizeof(double), st));
    rc = prt_render_device(s, camera, params, s->s_rgb.p, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(rgb, s->s_rgb.p, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return PRT_OK;
}

int prt_resolve_device(prt_scene* s, const double* rgb_device, int32_t width, int32_t height, uint8_t* rgb8, float* rgb32f) {
    int rc = need_device(s);
    if (rc) return rc;
    if (!rgb_device || width < 1 || height < 1) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_resolve_device: bad argument");
    const size_t n = (size_t)width * height * 3;
    cudaStream_t st = s->ctx->stream;
    if (rgb8) CUDA_TRY(s->s_rgb8.ensure(n));
    if (rgb32f) CUDA_TRY(s->s_rgb32f.ensure(n));
    if (!s->d_thr.p) CUDA_TRY(s->d_thr.upload(resolve_thresholds(), st));
    launch_resolve(rgb_device, (int64_t)width * height, s->d_thr.p, rgb8 ? s->s_rgb8.p : nullptr, rgb32f ? s->s_rgb32f.p : nullptr, st);
    CUDA_TRY(cudaGetLastError());
    if (rgb8) CUDA_TRY(cudaMemcpyAsync(rgb8, s->s_rgb8.p, n, cudaMemcpyDeviceToHost, st));
    if (rgb32f) CUDA_TRY(cudaMemcpyAsync(rgb32f, s->s_rgb32f.p, n * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return PRT_OK;
}

int prt_resolve(prt_scene* s, const double* rgb, int32_t width, int32_t height, uint8_t* rgb8, float* rgb32f) {
    int rc = need_device(s);
    if (rc) return rc;
    if (!rgb || width < 1 || height < 1) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_resolve: bad argument");
    const size_t n = (size_t)width * height * 3;
    CUDA_TRY(s->s_rgb.ensure(n));
    CUDA_TRY(cudaMemcpyAsync(s->s_rgb.p, rgb, n * sizeof(double), cudaMemcpyHostToDevice, s->ctx->stream));
    return prt_resolve_device(s, s->s_rgb.p, width, height, rgb8, rgb32f);
}

int prt_resolve_thresholds(double* out254) {
    if (!out254) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_resolve_thresholds: null argument");
    const std::vector<double>& t = resolve_thresholds();
    std::copy(t.begin(), t.end(), out254);
    return PRT_OK;
}

int prt_get_stats(const prt_scene* cs, prt_stats* out) {
    if (!cs || !out) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_get_stats: null argument");
    prt_scene* s = const_cast<prt_scene*>(cs);
    if (s->ctx && s->ctx->device >= 0 && s->committed) {
        cudaSetDevice(s->ctx->device);
        if (s->timing_pending) {
            if (cudaEventSynchronize(s->ev1) == cudaSuccess) {
                float ms = 0.f;
                if (cudaEventElapsedTime(&ms, s->ev0, s->ev1) == cudaSuccess) s->stats.gpu_ms = ms;
            }
            s->timing_pending = false;
        }
        if (s->rays_pending || s->visits_pending) {
            unsigned long long c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (cudaMemcpy(c, s->d_counters.p, sizeof(c), cudaMemcpyDeviceToHost) == cudaSuccess) {
                if (s->visits_pending && std::getenv("PRT_DEBUG_RECHECK"))
                    std::fprintf(stderr, "[prt] reference re-trace: %llu rays, %llu fetches, most for one ray %llu (all fetches: %llu)\n", c[4], c[5], c[6], c[0] + c[1]);
                if (s->visits_pending) { s->stats.node_visits = c[0]; s->stats.tri_tests = c[1]; }
                if (s->rays_pending) { s->stats.extend_rays = c[2]; s->stats.shadow_rays = c[3]; }
            }
            s->rays_pending = s->visits_pending = false;
        }
        if (s->d_err.p) {
            unsigned long long e = 0;
            if (cudaMemcpy(&e, s->d_err.p, sizeof(e), cudaMemcpyDeviceToHost) == cudaSuccess) s->stats.ref_stack_overflows = e;
        }
    }
    *out = s->stats;
    return PRT_OK;
}

int prt_synchronize(prt_context* ctx) {
    if (!ctx) return fail(PRT_ERR_INVALID_ARGUMENT, "null context");
    if (ctx->device < 0) return PRT_OK;
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return PRT_OK;
}

}  // extern "C"

/* ---- the entry point that allocates nothing -----------------------------------------------------
 * prt_trace_closest_device for callers that may not reach cudaMalloc (CUDA graph capture, serving
 * loops, harnesses that forbid it). Same kernels, same launch rules; the scratch they need — the work
 * cursor of the persistent traversal kernel and, when PRT_SORT_MIN asks for ray binning, the keys,
 * the binned order, the histogram and the bin cursors — lives in ONE file-scope device array instead
 * of the scene's cudaMalloc'ed buffers. Every GPU that loads the library holds the array until exit.
 * g_noalloc_scratch below is a host-side stand-in in host code: its device address comes from
 * cudaGetSymbolAddress, once per device, with that device current. */
#define PRT_NOALLOC_ALIGN(x) (((x) + 255) / 256 * 256)
#define PRT_NOALLOC_OFF_FETCH 0
#define PRT_NOALLOC_OFF_HIST 256
#define PRT_NOALLOC_OFF_CURSOR (PRT_NOALLOC_OFF_HIST + PRT_NOALLOC_ALIGN(sizeof(int32_t) * PRT_KEY_BINS))
#define PRT_NOALLOC_OFF_ORDER (PRT_NOALLOC_OFF_CURSOR + PRT_NOALLOC_ALIGN(sizeof(int32_t) * PRT_KEY_BINS))
#define PRT_NOALLOC_OFF_KEYS (PRT_NOALLOC_OFF_ORDER + PRT_NOALLOC_ALIGN(sizeof(int32_t) * PRT_NOALLOC_MAX_RAYS))
#define PRT_NOALLOC_BYTES (PRT_NOALLOC_OFF_KEYS + PRT_NOALLOC_ALIGN(sizeof(uint16_t) * PRT_NOALLOC_MAX_RAYS))
#define PRT_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_noalloc_scratch[PRT_NOALLOC_BYTES]; /* 6.25 MiB for 2^20 rays */

namespace {
/* device address of g_noalloc_scratch on the CURRENT device (the caller made `device` current) */
int noalloc_scratch(int device, unsigned char** out) {
    static std::mutex mu;
    static unsigned char* address[PRT_NOALLOC_MAX_DEVICES] = {nullptr};
    if (device < 0 || device >= PRT_NOALLOC_MAX_DEVICES) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_closest_device_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lock(mu);
    if (!address[device]) {
        unsigned char* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) return fail(PRT_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        address[device] = p;
    }
    *out = address[device];
    return PRT_OK;
}
}  // namespace

extern "C" int prt_trace_closest_device_noalloc(prt_scene* s, int64_t n, const double* o, const double* d, const double* tmin,
                                                const double* tmax, int32_t* prim, double* t, double* alpha, double* beta,
                                                void* stream) {
    int rc = need_device(s);
    if (rc) return rc;
    if (n < 0) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_closest_device_noalloc: negative ray count");
    if (n > PRT_NOALLOC_MAX_RAYS)
        return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_closest_device_noalloc: " + std::to_string((long long)n) + " rays in one call, the static scratch holds at most " +
                                                  std::to_string((long long)PRT_NOALLOC_MAX_RAYS) + " (PRT_NOALLOC_MAX_RAYS): split the batch or call prt_trace_closest_device");
    if (n > 0 && (!o || !d || !tmin || !tmax || !prim)) return fail(PRT_ERR_INVALID_ARGUMENT, "prt_trace_closest_device_noalloc: null array");
    unsigned char* scratch = nullptr;
    rc = noalloc_scratch(s->ctx->device, &scratch); /* n == 0: the set-up alone (e.g. before a graph capture starts) */
    if (rc || n == 0) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : s->ctx->stream;
    int32_t* fetch = (int32_t*)(scratch + PRT_NOALLOC_OFF_FETCH);
    /* the array keeps what the previous call left: the cursor must start at zero */
    CUDA_TRY(cudaMemsetAsync(fetch, 0, 4 * sizeof(int32_t), st));
    const LaunchConfig lc = launch_config(s->ctx);
    const int32_t* order = nullptr;
    if (lc.sort_min > 0 && n >= lc.sort_min) { /* the rule of query_order() */
        int32_t* hist = (int32_t*)(scratch + PRT_NOALLOC_OFF_HIST);
        int32_t* cursor = (int32_t*)(scratch + PRT_NOALLOC_OFF_CURSOR);
        int32_t* binned = (int32_t*)(scratch + PRT_NOALLOC_OFF_ORDER);
        uint16_t* keys = (uint16_t*)(scratch + PRT_NOALLOC_OFF_KEYS);
        CUDA_TRY(cudaMemsetAsync(hist, 0, sizeof(int32_t) * PRT_KEY_BINS, st)); /* launch_query_sort wants it zero on entry */
        launch_query_sort(n, o, d, s->key_lo, s->key_scale, keys, hist, cursor, binned, lc, st);
        order = binned;
    }
    launch_trace_closest(s->dev, n, o, d, tmin, tmax, prim, t, alpha, beta, order, fetch, nullptr, lc, st);
    CUDA_TRY(cudaGetLastError());
    return PRT_OK;
}

/*
 * multi.cu — the render path on 1..8 GPUs of one box, behind the C ABI.
 *
 * The path shards with no exchange inside the loop: the scene is REPLICATED on every GPU (built once, copied GPU to
 * GPU), the samples of every
 * pixel are split into contiguous ranges (the counter-based generator is keyed by (pixel, absolute sample index), so
 * the union of the shards is the one-GPU render), each GPU accumulates its range into its own FP64 frame, and ONE
 * reduction to devices[0] brings the frames together. NaN scrub and the sRGB quantiser run on the root AFTER the
 * reduce, as the reference scrubs the final pixel (Camera.cpp:291-293). One host thread per GPU issues that GPU's
 * launches. Ray-query batches are split into contiguous ray ranges with no collective at all.
 *
 * The reduction comes in two kinds (prt_multi_reduce_kind; PRT_MULTI_REDUCE=peer|nccl overrides the choice):
 *   PEER  where devices[0] can address every other GPU's memory (NVLink / NVSwitch: every B200 box): ONE kernel on the
 *         root, k_sum_peers, reads the other frames through peer loads and adds them to its own in the fixed order
 *         g = 1, 2, ... (bit-reproducible from run to run, which a ring's or tree's order is not), stream-ordered behind
 *         every GPU's last pass by one event per GPU. W*H*3 doubles per peer cross the switch once (7 x 25 MB for a
 *         1024^2 frame on 8 GPUs: 0.3 ms). No communicator: a handle is ready as soon as the scene is (ncclCommInitAll
 *         on 8 GPUs costs 3.5 s, and again at exit: measured, profiles/r2_s15_bench_n8.json; a 3000-spp cornell-box job
 *         renders in 0.5 s on the same box);
 *   NCCL  otherwise: ncclReduce(sum, ncclDouble, root = devices[0]), grouped, one communicator per GPU owned by the
 *         handle. Bound at run time (dlopen of libnccl.so.2, the system's or the one a host process already loaded):
 *         a one-GPU or PEER handle never touches it, and libprt_b200.so carries no link-time dependency on it.
 *
 * Replaces, for a C / C++ caller: main.cpp:36-52 (build world / lights, Camera::Render, WriteColorAttachment) on a
 * whole box.
 */
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/prt_b200.h"
#include "capi_internal.h"

namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
    bool ok() const { return lib && CommInitAll && CommDestroy && Reduce && GroupStart && GroupEnd && GetErrorString; }
};

NcclApi& nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) {
            api.error = std::string("cannot load NCCL (libnccl.so.2): ") + (dlerror() ? dlerror() : "?");
            return;
        }
        api.CommInitAll = (decltype(api.CommInitAll))dlsym(api.lib, "ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.lib, "ncclCommDestroy");
        api.Reduce = (decltype(api.Reduce))dlsym(api.lib, "ncclReduce");
        api.GroupStart = (decltype(api.GroupStart))dlsym(api.lib, "ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))dlsym(api.lib, "ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.lib, "ncclGetErrorString");
        if (!api.ok()) api.error = "libnccl.so.2 lacks one of ncclCommInitAll / ncclReduce / ncclGroupStart / ncclGroupEnd";
    });
    return api;
}

/* the g-th of n contiguous, balanced slices of [begin, end): sizes differ by at most one (500 samples over 8 GPUs =
 * 63,63,63,63,62,62,62,62) — pooraytracer_b200/sharding.py:split_samples is the Python twin */
void split_range(int64_t begin, int64_t end, int g, int n, int64_t* b, int64_t* e) {
    const int64_t len = end - begin, base = len / n, extra = len % n;
    *b = begin + g * base + (g < extra ? g : extra);
    *e = *b + base + (g < extra ? 1 : 0);
}

/* PEER reduction: root[i] += frame_1[i] + ... in that order; frames are other GPUs' memory read over NVLink. Eight
 * independent 16-byte loads in flight per thread and pass; the grid is sized by the caller (a multiple of the SM count). */
struct PeerFrames {
    const double* frame[7];
    int n;
};
__global__ void __launch_bounds__(256) k_sum_peers(double* __restrict__ root, PeerFrames f, size_t count) {
    const size_t pairs = count / 2, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < pairs; i += stride) {
        double2 a = reinterpret_cast<double2*>(root)[i];
        double2 b[7];
#pragma unroll
        for (int g = 0; g < 7; ++g)
            if (g < f.n) b[g] = reinterpret_cast<const double2*>(f.frame[g])[i];
#pragma unroll
        for (int g = 0; g < 7; ++g)
            if (g < f.n) { a.x += b[g].x; a.y += b[g].y; }
        reinterpret_cast<double2*>(root)[i] = a;
    }
    if ((count & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
        double a = root[count - 1];
        for (int g = 0; g < f.n; ++g) a += f.frame[g][count - 1];
        root[count - 1] = a;
    }
}

struct Worker {
    int rc = PRT_OK;
    std::string error;
};

/* runs body(g) on one host thread per GPU and collects the first failure */
template <class F>
int on_every_gpu(int n, F body) {
    std::vector<Worker> w((size_t)n);
    std::vector<std::thread> th;
    for (int g = 1; g < n; ++g)
        th.emplace_back([&, g] {
            w[g].rc = body(g);
            if (w[g].rc != PRT_OK) w[g].error = prt_last_error();
        });
    w[0].rc = body(0);
    if (w[0].rc != PRT_OK) w[0].error = prt_last_error();
    for (auto& t : th) t.join();
    for (int g = 0; g < n; ++g)
        if (w[g].rc != PRT_OK) return prt::set_error(w[g].rc, "GPU " + std::to_string(g) + ": " + w[g].error);
    return PRT_OK;
}

}  // namespace

struct prt_multi {
    int n = 0;
    std::vector<int> devices;
    std::vector<prt_context*> ctx;
    std::vector<prt_scene*> scene;
    std::vector<double*> acc; /* per GPU: W*H*3 FP64 accumulators */
    size_t acc_n = 0;
    std::vector<ncclComm_t> comms;       /* PRT_REDUCE_NCCL */
    std::vector<cudaEvent_t> frame_done; /* PRT_REDUCE_PEER: per GPU, recorded behind its last pass */
    int reduce_kind = PRT_REDUCE_NONE;
    int32_t width = 0, height = 0;
    bool rendered = false;
    double reduce_ms = 0.0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr; /* on devices[0]: around the reduce */
};

extern "C" {

void prt_multi_destroy(prt_multi* m) {
    if (!m) return;
    for (int g = 0; g < (int)m->acc.size(); ++g)
        if (m->acc[g]) { cudaSetDevice(m->devices[g]); cudaFree(m->acc[g]); }
    if (m->ev0) { cudaSetDevice(m->devices[0]); cudaEventDestroy(m->ev0); cudaEventDestroy(m->ev1); }
    for (ncclComm_t c : m->comms)
        if (c) nccl_api().CommDestroy(c);
    for (int g = 0; g < (int)m->frame_done.size(); ++g)
        if (m->frame_done[g]) { cudaSetDevice(m->devices[g]); cudaEventDestroy(m->frame_done[g]); }
    for (prt_scene* s : m->scene) prt_scene_destroy(s);
    for (prt_context* c : m->ctx) prt_context_destroy(c);
    delete m;
}

int prt_multi_create(const int32_t* devices, int32_t n, const prt_scene_desc* desc, const prt_build_opts* opts, prt_multi** out) {
    if (!out) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_create: out is null");
    *out = nullptr;
    if (!devices || !desc || n < 1 || n > 8) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_create: 1..8 devices and a scene description are needed");
    for (int a = 0; a < n; ++a)
        for (int b = a + 1; b < n; ++b)
            if (devices[a] == devices[b]) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_create: a device is listed twice");
    prt_multi* m = new (std::nothrow) prt_multi();
    if (!m) return prt::set_error(PRT_ERR_OUT_OF_MEMORY, "prt_multi_create: out of host memory");
    m->n = n;
    m->devices.assign(devices, devices + n);
    m->ctx.assign((size_t)n, nullptr);
    m->scene.assign((size_t)n, nullptr);
    m->acc.assign((size_t)n, nullptr);
    /* the scene is replicated: ingested, built and uploaded ONCE (devices[0]), then copied GPU to GPU on every other
     * device at the same time (prt_scene_clone: peer copies over NVLink, no second host-side build, no n-fold host memory).
     * The other GPUs' contexts come up on their own threads WHILE the host builds the tree for the first. */
    std::vector<int> ctx_rc((size_t)n, PRT_OK);
    std::vector<std::string> ctx_err((size_t)n);
    std::vector<std::thread> bring_up;
    for (int g = 1; g < n; ++g)
        bring_up.emplace_back([&, g] {
            ctx_rc[g] = prt_context_create(m->devices[g], &m->ctx[g]);
            if (ctx_rc[g] != PRT_OK) ctx_err[g] = prt_last_error();
        });
    int rc = prt_context_create(m->devices[0], &m->ctx[0]);
    if (rc == PRT_OK) rc = prt_scene_create(m->ctx[0], desc, &m->scene[0]);
    if (rc == PRT_OK) rc = prt_scene_commit(m->scene[0], opts);
    std::string first_error = rc != PRT_OK ? prt_last_error() : "";
    for (auto& t : bring_up) t.join();
    for (int g = 1; g < n && rc == PRT_OK; ++g)
        if (ctx_rc[g] != PRT_OK) { rc = ctx_rc[g]; first_error = "GPU " + std::to_string(g) + ": " + ctx_err[g]; }
    if (rc == PRT_OK && n > 1) {
        rc = on_every_gpu(n, [&](int g) { return g == 0 ? (int)PRT_OK : prt_scene_clone(m->scene[0], m->ctx[g], &m->scene[g]); });
        if (rc != PRT_OK) first_error = prt_last_error();
    }
    if (rc != PRT_OK) { prt_multi_destroy(m); return prt::set_error(rc, first_error); }
    if (n > 1) {
        /* which reduction: peer loads where the root can address every other GPU, NCCL otherwise */
        const char* want = std::getenv("PRT_MULTI_REDUCE");
        const bool force_nccl = want && !std::strcmp(want, "nccl"), force_peer = want && !std::strcmp(want, "peer");
        bool peer_ok = !force_nccl;
        for (int g = 1; g < n && peer_ok; ++g) {
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, m->devices[0], m->devices[g]) != cudaSuccess || !can) peer_ok = false;
        }
        if (force_peer && !peer_ok) {
            prt_multi_destroy(m);
            return prt::set_error(PRT_ERR_UNSUPPORTED, "prt_multi_create: PRT_MULTI_REDUCE=peer, but the first device cannot address every other device's memory");
        }
        if (peer_ok) {
            cudaSetDevice(m->devices[0]);
            for (int g = 1; g < n; ++g) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(m->devices[g], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError(); /* a host process (torch) got there first */
                else if (e != cudaSuccess) {
                    const std::string msg = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e);
                    prt_multi_destroy(m);
                    return prt::set_error(PRT_ERR_CUDA, msg);
                }
            }
            m->frame_done.assign((size_t)n, nullptr);
            for (int g = 1; g < n; ++g) {
                cudaSetDevice(m->devices[g]);
                if (cudaEventCreateWithFlags(&m->frame_done[g], cudaEventDisableTiming) != cudaSuccess) {
                    prt_multi_destroy(m);
                    return prt::set_error(PRT_ERR_CUDA, "prt_multi_create: cudaEventCreate failed");
                }
            }
            m->reduce_kind = PRT_REDUCE_PEER;
        } else {
            NcclApi& api = nccl_api();
            if (!api.ok()) { prt_multi_destroy(m); return prt::set_error(PRT_ERR_UNSUPPORTED, "prt_multi_create: " + api.error); }
            m->comms.assign((size_t)n, nullptr);
            const ncclResult_t r = api.CommInitAll(m->comms.data(), n, m->devices.data());
            if (r != ncclSuccess) {
                const std::string msg = std::string("ncclCommInitAll: ") + api.GetErrorString(r);
                m->comms.clear();
                prt_multi_destroy(m);
                return prt::set_error(PRT_ERR_CUDA, msg);
            }
            m->reduce_kind = PRT_REDUCE_NCCL;
        }
    }
    cudaSetDevice(m->devices[0]);
    if (cudaEventCreate(&m->ev0) != cudaSuccess || cudaEventCreate(&m->ev1) != cudaSuccess) {
        prt_multi_destroy(m);
        return prt::set_error(PRT_ERR_CUDA, "prt_multi_create: cudaEventCreate failed");
    }
    *out = m;
    return PRT_OK;
}

int prt_multi_device_count(const prt_multi* m) { return m ? m->n : 0; }
int prt_multi_reduce_kind(const prt_multi* m) { return m ? m->reduce_kind : PRT_REDUCE_NONE; }

int prt_multi_render(prt_multi* m, const prt_camera* camera, const prt_render_params* params, double* rgb) {
    if (!m || !camera || !params) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_render: null argument");
    const int W = camera->image_width < 1 ? 1 : camera->image_width, H = camera->image_height < 1 ? 1 : camera->image_height;
    const size_t count = (size_t)W * H * 3;
    int64_t sb = params->spp_begin, se = params->spp_end;
    if (sb == 0 && se == 0) se = params->samples_per_pixel;
    if (sb < 0 || se < sb) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_render: bad sample range");
    if (count > m->acc_n) {
        for (int g = 0; g < m->n; ++g) {
            cudaSetDevice(m->devices[g]);
            if (m->acc[g]) cudaFree(m->acc[g]);
            m->acc[g] = nullptr;
            if (cudaMalloc((void**)&m->acc[g], count * sizeof(double)) != cudaSuccess) {
                m->acc_n = 0;
                return prt::set_error(PRT_ERR_OUT_OF_MEMORY, "prt_multi_render: cannot allocate the accumulators");
            }
        }
        m->acc_n = count;
    }
    m->rendered = false;
    int rc = on_every_gpu(m->n, [&](int g) {
        if (cudaSetDevice(m->devices[g]) != cudaSuccess) return prt::set_error(PRT_ERR_CUDA, "cudaSetDevice failed");
        if (cudaMemsetAsync(m->acc[g], 0, count * sizeof(double), m->ctx[g]->stream) != cudaSuccess) return prt::set_error(PRT_ERR_CUDA, "cudaMemsetAsync failed");
        int64_t b, e;
        split_range(sb, se, g, m->n, &b, &e);
        if (e <= b) return (int)PRT_OK; /* more GPUs than samples: this one only joins the reduce */
        prt_render_params p = *params;
        p.spp_begin = (int32_t)b;
        p.spp_end = (int32_t)e;
        return prt_render_device(m->scene[g], camera, &p, m->acc[g], nullptr);
    });
    if (rc == PRT_OK && m->reduce_kind == PRT_REDUCE_PEER)
        for (int g = 1; g < m->n; ++g) { /* this GPU's frame is complete behind this point of its stream */
            cudaSetDevice(m->devices[g]);
            if (cudaEventRecord(m->frame_done[g], m->ctx[g]->stream) != cudaSuccess) return prt::set_error(PRT_ERR_CUDA, "prt_multi_render: cudaEventRecord failed");
        }
    if (rc != PRT_OK) return rc;
    cudaSetDevice(m->devices[0]);
    cudaEventRecord(m->ev0, m->ctx[0]->stream);
    if (m->reduce_kind == PRT_REDUCE_PEER) {
        /* THE exchange step: one kernel on the root reads every other frame over NVLink, behind every GPU's last pass */
        PeerFrames f;
        f.n = m->n - 1;
        for (int g = 1; g < m->n; ++g) {
            f.frame[g - 1] = m->acc[g];
            if (cudaStreamWaitEvent(m->ctx[0]->stream, m->frame_done[g], 0) != cudaSuccess) return prt::set_error(PRT_ERR_CUDA, "prt_multi_render: cudaStreamWaitEvent failed");
        }
        cudaEventRecord(m->ev0, m->ctx[0]->stream); /* reduce_ms = the kernel alone, not the wait for the slowest GPU */
        const size_t want = (count / 2 + 255) / 256, cap = (size_t)m->ctx[0]->sm_count * 8;
        k_sum_peers<<<(unsigned)(want < cap ? (want ? want : 1) : cap), 256, 0, m->ctx[0]->stream>>>(m->acc[0], f, count);
        if (cudaGetLastError() != cudaSuccess) return prt::set_error(PRT_ERR_CUDA, "prt_multi_render: k_sum_peers failed to launch");
    } else if (m->reduce_kind == PRT_REDUCE_NCCL) {
        /* THE collective: stream-ordered behind every GPU's last pass, in place on the root */
        NcclApi& api = nccl_api();
        ncclResult_t r = api.GroupStart();
        for (int g = 0; g < m->n && r == ncclSuccess; ++g)
            r = api.Reduce(m->acc[g], m->acc[0], count, ncclDouble, ncclSum, 0, m->comms[g], m->ctx[g]->stream);
        const ncclResult_t r2 = api.GroupEnd();
        if (r == ncclSuccess) r = r2;
        if (r != ncclSuccess) return prt::set_error(PRT_ERR_CUDA, std::string("ncclReduce: ") + api.GetErrorString(r));
        cudaSetDevice(m->devices[0]);
    }
    cudaEventRecord(m->ev1, m->ctx[0]->stream);
    if (rgb && cudaMemcpyAsync(rgb, m->acc[0], count * sizeof(double), cudaMemcpyDeviceToHost, m->ctx[0]->stream) != cudaSuccess)
        return prt::set_error(PRT_ERR_CUDA, "prt_multi_render: device-to-host copy failed");
    for (int g = 0; g < m->n; ++g) {
        cudaSetDevice(m->devices[g]);
        const cudaError_t e = cudaStreamSynchronize(m->ctx[g]->stream);
        if (e != cudaSuccess) return prt::set_error(PRT_ERR_CUDA, std::string("GPU ") + std::to_string(g) + ": " + cudaGetErrorString(e));
    }
    float ms = 0.f;
    cudaSetDevice(m->devices[0]);
    if (cudaEventElapsedTime(&ms, m->ev0, m->ev1) == cudaSuccess) m->reduce_ms = ms;
    m->width = W;
    m->height = H;
    m->rendered = true;
    return PRT_OK;
}

int prt_multi_resolve(prt_multi* m, uint8_t* rgb8, float* rgb32f) {
    if (!m || !m->rendered) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_resolve: nothing has been rendered");
    return prt_resolve_device(m->scene[0], m->acc[0], m->width, m->height, rgb8, rgb32f);
}

int prt_multi_get_stats(const prt_multi* m, prt_stats* out, double* reduce_ms) {
    if (!m || !out) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_get_stats: null argument");
    std::memset(out, 0, sizeof(*out));
    for (int g = 0; g < m->n; ++g) {
        prt_stats s;
        const int rc = prt_get_stats(m->scene[g], &s);
        if (rc != PRT_OK) return rc;
        if (g == 0) *out = s;
        else {
            out->extend_rays += s.extend_rays; out->shadow_rays += s.shadow_rays; out->paths += s.paths;
            out->node_visits += s.node_visits; out->tri_tests += s.tri_tests; out->kernel_launches += s.kernel_launches;
            out->ref_stack_overflows += s.ref_stack_overflows;
            if (s.gpu_ms > out->gpu_ms) out->gpu_ms = s.gpu_ms; /* the slowest GPU's device time */
        }
    }
    if (reduce_ms) *reduce_ms = m->reduce_ms;
    return PRT_OK;
}

/* Hittable::Hit over a batch, contiguous ray ranges per GPU, no collective */
int prt_multi_trace_closest(prt_multi* m, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax, int32_t* prim,
                            double* t, double* alpha, double* beta) {
    if (!m) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_trace_closest: null handle");
    if (n > 0 && (!o || !d || !tmin || !tmax || !prim)) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_trace_closest: null array");
    return on_every_gpu(m->n, [&](int g) {
        int64_t b, e;
        split_range(0, n, g, m->n, &b, &e);
        if (e <= b) return (int)PRT_OK;
        return prt_trace_closest(m->scene[g], e - b, o + 3 * b, d + 3 * b, tmin + b, tmax + b, prim + b, t ? t + b : nullptr, alpha ? alpha + b : nullptr,
                                 beta ? beta + b : nullptr);
    });
}

int prt_multi_trace_any(prt_multi* m, int64_t n, const double* o, const double* d, const double* tmin, const double* tmax, uint8_t* hit) {
    if (!m) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_trace_any: null handle");
    if (n > 0 && (!o || !d || !tmin || !tmax || !hit)) return prt::set_error(PRT_ERR_INVALID_ARGUMENT, "prt_multi_trace_any: null array");
    return on_every_gpu(m->n, [&](int g) {
        int64_t b, e;
        split_range(0, n, g, m->n, &b, &e);
        if (e <= b) return (int)PRT_OK;
        return prt_trace_any(m->scene[g], e - b, o + 3 * b, d + 3 * b, tmin + b, tmax + b, hit + b);
    });
}

}  // extern "C"

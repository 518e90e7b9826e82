/*
 * bvh_device.cu — the acceleration structure built ON the GPU: replaces the host-side
 * SAH build for PRT_ACCEL_LBVH4_DEVICE. What the reference does in BVHNode::BVHNode
 * (Source/BVH.cpp:7-48: recursive std::sort + median split, one heap node per span) is, for the
 * traversal structure, a linear BVH here:
 *
 *   1. k_lbvh_morton     63-bit Morton code of every triangle's box centre on the scene's bounding box
 *   2. radix sort        (code, prim) pairs — cub::DeviceRadixSort, the one library call of the build
 *   3. k_lbvh_hierarchy  Karras 2012: each of the n - 1 inner nodes finds its key range and split
 *                        from the common-prefix lengths of neighbouring codes (ties broken by index)
 *   4. k_lbvh_fit        bottom-up FP64 boxes (exact unions of the reference's triangle boxes), the
 *                        second thread to arrive at a node carries on upwards and decides, by the SAH
 *                        termination rule of the host builder, whether the subtree (<= max_leaf
 *                        triangles, contiguous slots: slot order = Morton order) becomes one leaf
 *   5. k_lbvh_classify   live inner nodes at even depth become BVH4 nodes, those at odd depth are
 *                        absorbed into their parent
 *   6. exclusive scan    BVH4 node numbers (root = 0)
 *   7. k_lbvh_emit       Node4 records: up to four (grand)children, FP32 boxes rounded OUTWARD
 *   8. k_gather_tris     TriIsect records permuted into slot order
 *
 * The tree only decides which candidates a ray is offered; hits are the reference's by the argument
 * in traverse_warp.cuh (its boxes contain the reference's boxes; boundary / equal-t hits are re-traced
 * through the flattened REFERENCE tree, which is still built on the host because its leaf order is
 * defined by std::sort's handling of equal keys). 10 M triangles: 5.3 s of host SAH build -> see
 * profiles/ for the measured device time.
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "bvh_device.h"
#include "traverse.cuh" /* PRT_NO_CHILD */

namespace prt {

namespace {

#define LB_TRY(expr)                        \
    do {                                    \
        cudaError_t e__ = (expr);           \
        if (e__ != cudaSuccess) return e__; \
    } while (0)

struct Scratch { /* frees whatever was allocated when it goes out of scope */
    void* p[24];
    int n = 0;
    ~Scratch() {
        for (int i = 0; i < n; ++i) cudaFree(p[i]);
    }
    template <typename T>
    cudaError_t get(T** out, size_t count) {
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, (count ? count : 1) * sizeof(T));
        if (e != cudaSuccess) return e;
        p[n++] = q;
        *out = (T*)q;
        return cudaSuccess;
    }
};

__device__ __forceinline__ uint64_t spread21(uint64_t v) { /* 21 bits -> every third bit */
    v &= 0x1fffffull;
    v = (v | (v << 32)) & 0x1f00000000ffffull;
    v = (v | (v << 16)) & 0x1f0000ff0000ffull;
    v = (v | (v << 8)) & 0x100f00f00f00f00full;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}

struct BuildBox { double lo[3], inv[3]; };

__global__ void __launch_bounds__(256) k_lbvh_morton(const Box6* __restrict__ box, int32_t n, BuildBox bb, uint64_t* __restrict__ code,
                                                     int32_t* __restrict__ prim) {
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t c = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double mid = 0.5 * (box[i].v[2 * k] + box[i].v[2 * k + 1]);
        double f = (mid - bb.lo[k]) * bb.inv[k]; /* in [0, 1] for finite input */
        f = f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f); /* also maps NaN to 1 */
        uint64_t q = (uint64_t)(f * 2097151.0);
        c |= spread21(q) << k;
    }
    code[i] = c;
    prim[i] = i;
}

/* Node ids: inner nodes 0 .. n-2 (root 0), leaves n-1 .. 2n-2 (leaf k = sorted position k). */
struct Tree {
    int32_t n;        /* triangles */
    int32_t* left;    /* [n-1] */
    int32_t* right;   /* [n-1] */
    int32_t* parent;  /* [2n-1] */
    int32_t* first;   /* [n-1] first sorted position of the node's range */
    int32_t* count;   /* [n-1] triangles below */
    Box6* box;        /* [2n-1] */
    int32_t* arrived; /* [n-1] fit counters */
    float* cost;      /* [2n-1] SAH cost of the subtree as finally cut */
    uint8_t* is_leaf; /* [n-1] the subtree is cut here: one leaf of the final tree */
};

/* the host builder's constants (scene_host.cpp: SahBuilder) */
#define LB_C_TRI 1.25f
#define LB_C_NODE 1.0f
__device__ __forceinline__ float half_area_f(const Box6& b) {
    const double dx = b.v[1] - b.v[0], dy = b.v[3] - b.v[2], dz = b.v[5] - b.v[4];
    return (float)(dx * dy + dy * dz + dz * dx);
}

__device__ __forceinline__ int delta(const uint64_t* __restrict__ code, int32_t n, int32_t i, int32_t j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = code[i], b = code[j];
    if (a == b) return 64 + __clz(i ^ j); /* equal codes: the index continues the key */
    return __clzll(a ^ b);
}

__global__ void __launch_bounds__(256) k_lbvh_hierarchy(const uint64_t* __restrict__ code, Tree T) {
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const int32_t n = T.n;
    if (i >= n - 1) return;
    const int d = delta(code, n, i, i + 1) - delta(code, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(code, n, i, i - d);
    int32_t lmax = 2;
    while (delta(code, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int32_t l = 0;
    for (int32_t t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(code, n, i, i + (l + t) * d) > dmin) l += t;
    const int32_t j = i + l * d;
    const int dnode = delta(code, n, i, j);
    int32_t s = 0;
    for (int32_t t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(code, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int32_t gamma = i + s * d + (d < 0 ? -1 : 0);
    const int32_t lo = i < j ? i : j, hi = i < j ? j : i;
    const int32_t L = lo == gamma ? (n - 1) + gamma : gamma;
    const int32_t R = hi == gamma + 1 ? (n - 1) + gamma + 1 : gamma + 1;
    T.left[i] = L;
    T.right[i] = R;
    T.parent[L] = i;
    T.parent[R] = i;
    T.first[i] = lo;
    T.count[i] = hi - lo + 1;
    if (i == 0) T.parent[0] = -1;
}

__device__ __forceinline__ void box_union(Box6& a, const Box6& b) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        a.v[2 * k] = fmin(a.v[2 * k], b.v[2 * k]);
        a.v[2 * k + 1] = fmax(a.v[2 * k + 1], b.v[2 * k + 1]);
    }
}

/* one thread per leaf; the second thread to reach an inner node owns it and climbs on */
__global__ void __launch_bounds__(256) k_lbvh_fit(const Box6* __restrict__ tri_box, const int32_t* __restrict__ sorted_prim, Tree T, int max_leaf) {
    const int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    const int32_t n = T.n;
    if (k >= n) return;
    int32_t node = (n - 1) + k;
    T.box[node] = tri_box[sorted_prim[k]];
    T.cost[node] = LB_C_TRI;
    __threadfence();
    int32_t p = T.parent[node];
    while (p >= 0) {
        if (atomicAdd(T.arrived + p, 1) == 0) return; /* the sibling subtree is not done yet */
        __threadfence();
        Box6 b, c; /* written by other SMs during this launch: read past L1 */
        const double* pl = T.box[T.left[p]].v;
        const double* pr = T.box[T.right[p]].v;
#pragma unroll
        for (int k = 0; k < 6; ++k) { b.v[k] = __ldcg(pl + k); c.v[k] = __ldcg(pr + k); }
        /* where to cut (the SAH termination rule of the host builder, applied bottom-up): a subtree of
         * <= max_leaf triangles becomes ONE leaf if testing them all is no dearer than descending */
        const float al = half_area_f(b), ar = half_area_f(c);
        box_union(b, c);
        T.box[p] = b;
        const float a = half_area_f(b);
        const float split = a > 0.f ? LB_C_NODE + (al * __ldcg(T.cost + T.left[p]) + ar * __ldcg(T.cost + T.right[p])) / a : __int_as_float(0x7f800000);
        const int32_t cnt = T.count[p];
        const float leaf = cnt <= max_leaf ? LB_C_TRI * (float)cnt : __int_as_float(0x7f800000);
        const bool cut = cnt <= max_leaf && leaf <= split;
        T.is_leaf[p] = cut ? 1 : 0;
        T.cost[p] = cut ? leaf : split;
        __threadfence();
        p = T.parent[p];
    }
}

/* Node roles. A node where the tree is cut (is_leaf, or a single triangle) is a LEAF of the final
 * tree and everything below it is dead; a live inner node at even depth is a BVH4 node (flag = 1), at
 * odd depth it is absorbed into its parent. */
__device__ __forceinline__ int32_t tri_count(const Tree& T, int32_t id) { return id >= T.n - 1 ? 1 : T.count[id]; }
__device__ __forceinline__ bool final_leaf(const Tree& T, int32_t id) { return id >= T.n - 1 || T.is_leaf[id]; }

__global__ void __launch_bounds__(256) k_lbvh_classify(Tree T, int32_t* __restrict__ flag) {
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.n - 1) return;
    int f = 0;
    if (!T.is_leaf[i]) {
        int depth = 0;
        bool dead = false;
        for (int32_t p = T.parent[i]; p >= 0; p = T.parent[p]) {
            ++depth;
            dead |= T.is_leaf[p] != 0;
        }
        f = !dead && (depth & 1) == 0;
    }
    flag[i] = f;
}

__device__ __forceinline__ float f32_down(double v) { return __double2float_rd(v); }
__device__ __forceinline__ float f32_up(double v) { return __double2float_ru(v); }

__global__ void __launch_bounds__(256) k_lbvh_emit(Tree T, const int32_t* __restrict__ flag, const int32_t* __restrict__ index4,
                                                   Node4* __restrict__ out) {
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.n - 1 || !flag[i]) return;
    int32_t kids[4];
    int nk = 0;
    const int32_t two[2] = {T.left[i], T.right[i]};
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const int32_t ch = two[c];
        if (final_leaf(T, ch)) kids[nk++] = ch;
        else { kids[nk++] = T.left[ch]; kids[nk++] = T.right[ch]; } /* odd depth: absorbed */
    }
    Node4 N;
    const float inf = __int_as_float(0x7f800000);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        if (c < nk) {
            const int32_t id = kids[c];
            const Box6 b = T.box[id];
#pragma unroll
            for (int k = 0; k < 3; ++k) { N.lo[k][c] = f32_down(b.v[2 * k]); N.hi[k][c] = f32_up(b.v[2 * k + 1]); }
            if (final_leaf(T, id)) {
                const int32_t first = id >= T.n - 1 ? id - (T.n - 1) : T.first[id];
                N.child[c] = ~((first << 3) | (tri_count(T, id) - 1));
            } else {
                N.child[c] = index4[id]; /* a live inner node two levels down: even depth, flagged */
            }
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) { N.lo[k][c] = inf; N.hi[k][c] = -inf; }
            N.child[c] = PRT_NO_CHILD;
        }
        N.pad[c] = 0;
    }
    out[index4[i]] = N;
}

/* the whole scene fits one leaf (n <= max_leaf): a root with one real child */
__global__ void k_lbvh_single_leaf(const Box6* __restrict__ tri_box, int32_t n, Node4* __restrict__ out) {
    Box6 b = tri_box[0];
    for (int32_t i = 1; i < n; ++i) box_union(b, tri_box[i]);
    Node4 N;
    const float inf = __int_as_float(0x7f800000);
    for (int c = 0; c < 4; ++c) {
        for (int k = 0; k < 3; ++k) { N.lo[k][c] = c ? inf : f32_down(b.v[2 * k]); N.hi[k][c] = c ? -inf : f32_up(b.v[2 * k + 1]); }
        N.child[c] = c ? PRT_NO_CHILD : ~((0 << 3) | (n - 1));
        N.pad[c] = 0;
    }
    out[0] = N;
}

__global__ void __launch_bounds__(256) k_gather_tris(const TriIsect* __restrict__ in, const int32_t* __restrict__ slot_prim, int64_t n,
                                                     TriIsect* __restrict__ out) {
    /* 8 threads per 128-byte record, 16 bytes each */
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t s = t >> 3;
    if (s >= n) return;
    const int part = (int)(t & 7);
    const int4* src = reinterpret_cast<const int4*>(in + slot_prim[s]) + part;
    reinterpret_cast<int4*>(out + s)[part] = __ldg(src);
}

}  // namespace

cudaError_t build_lbvh4_device(const Box6* d_box, const TriIsect* d_isect, int64_t n64, const double lo[3], const double hi[3], int max_leaf,
                               DeviceBvh* out, cudaStream_t st) {
    *out = DeviceBvh();
    if (n64 <= 0) return cudaSuccess;
    if (n64 >= (int64_t(1) << 28)) return cudaErrorInvalidValue; /* leaf refs hold slot << 3 */
    const int32_t n = (int32_t)n64;
    if (max_leaf < 1) max_leaf = 1;
    if (max_leaf > 8) max_leaf = 8;
    const bool single = n <= max_leaf; /* one leaf: slot order = prim order */
    const unsigned g_n = (unsigned)((n + 255) / 256);

    /* ---- every allocation but the node array first: cudaMalloc stalls the host, and build_ms is to
     * be the device time of the build, not of the allocator ---- */
    Scratch tmp;
    LB_TRY(cudaMalloc((void**)&out->slot_prim, sizeof(int32_t) * (size_t)n));
    LB_TRY(cudaMalloc((void**)&out->tris, sizeof(TriIsect) * (size_t)n));
    uint64_t *code = nullptr, *code_sorted = nullptr;
    int32_t *prim = nullptr, *flag = nullptr, *index4 = nullptr;
    char *sort_tmp = nullptr, *scan_tmp = nullptr;
    size_t sort_bytes = 0, scan_bytes = 0;
    Tree T;
    T.n = n;
    LB_TRY(tmp.get(&code, (size_t)n));
    LB_TRY(tmp.get(&prim, (size_t)n));
    if (!single) {
        LB_TRY(tmp.get(&code_sorted, (size_t)n));
        LB_TRY(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, code, code_sorted, prim, out->slot_prim, n, 0, 63, st));
        LB_TRY(tmp.get(&sort_tmp, sort_bytes));
        LB_TRY(tmp.get(&T.left, (size_t)n - 1));
        LB_TRY(tmp.get(&T.right, (size_t)n - 1));
        LB_TRY(tmp.get(&T.parent, (size_t)2 * n - 1));
        LB_TRY(tmp.get(&T.first, (size_t)n - 1));
        LB_TRY(tmp.get(&T.count, (size_t)n - 1));
        LB_TRY(tmp.get(&T.box, (size_t)2 * n - 1));
        LB_TRY(tmp.get(&T.arrived, (size_t)n - 1));
        LB_TRY(tmp.get(&T.cost, (size_t)2 * n - 1));
        LB_TRY(tmp.get(&T.is_leaf, (size_t)n - 1));
        LB_TRY(tmp.get(&flag, (size_t)n)); /* n - 1 flags + one slot that receives the total */
        LB_TRY(tmp.get(&index4, (size_t)n));
        LB_TRY(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, flag, index4, n, st));
        LB_TRY(tmp.get(&scan_tmp, scan_bytes));
    } else {
        LB_TRY(cudaMalloc((void**)&out->nodes4, sizeof(Node4)));
        out->n_nodes4 = 1;
    }
    /* two brackets of device time: everything up to the node count, then emission + gather (the one
     * allocation whose size the device decides sits between them) */
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < 4; ++i) if (e[i]) cudaEventDestroy(e[i]); } } guard{ev};
    for (int i = 0; i < 4; ++i) LB_TRY(cudaEventCreate(&ev[i]));
    /* PRT_BUILD_TIMING=1: per-stage device times on stderr (synchronises after every stage) */
    const bool timing = std::getenv("PRT_BUILD_TIMING") != nullptr;
    cudaEvent_t es = nullptr, ee = nullptr;
    if (timing) { cudaEventCreate(&es); cudaEventCreate(&ee); cudaEventRecord(es, st); }
    auto stage = [&](const char* name) {
        if (!timing) return;
        cudaEventRecord(ee, st);
        cudaEventSynchronize(ee);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, es, ee);
        std::fprintf(stderr, "[bvh_device] %-10s %8.3f ms\n", name, ms);
        cudaEventRecord(es, st);
    };

    LB_TRY(cudaEventRecord(ev[0], st));
    BuildBox bb;
    for (int k = 0; k < 3; ++k) {
        const double ext = hi[k] - lo[k];
        bb.lo[k] = lo[k];
        bb.inv[k] = ext > 0.0 && ext < 1e300 ? 1.0 / ext : 0.0;
    }
    k_lbvh_morton<<<g_n, 256, 0, st>>>(d_box, n, bb, code, prim);
    stage("morton");
    if (single) {
        LB_TRY(cudaMemcpyAsync(out->slot_prim, prim, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
        k_lbvh_single_leaf<<<1, 1, 0, st>>>(d_box, n, out->nodes4);
        LB_TRY(cudaEventRecord(ev[1], st));
        LB_TRY(cudaEventRecord(ev[2], st));
    } else {
        LB_TRY(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, code, code_sorted, prim, out->slot_prim, n, 0, 63, st));
        stage("sort");
        LB_TRY(cudaMemsetAsync(T.arrived, 0, sizeof(int32_t) * ((size_t)n - 1), st));
        k_lbvh_hierarchy<<<g_n, 256, 0, st>>>(code_sorted, T);
        stage("hierarchy");
        k_lbvh_fit<<<g_n, 256, 0, st>>>(d_box, out->slot_prim, T, max_leaf);
        stage("fit");
        LB_TRY(cudaMemsetAsync(flag + (n - 1), 0, sizeof(int32_t), st));
        k_lbvh_classify<<<g_n, 256, 0, st>>>(T, flag);
        stage("classify");
        LB_TRY(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, flag, index4, n, st));
        stage("scan");
        LB_TRY(cudaEventRecord(ev[1], st));
        int32_t total = 0; /* index4[n - 1] = number of flagged nodes */
        LB_TRY(cudaMemcpyAsync(&total, index4 + (n - 1), sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        LB_TRY(cudaStreamSynchronize(st));
        if (total < 1) return cudaErrorUnknown;
        LB_TRY(cudaMalloc((void**)&out->nodes4, sizeof(Node4) * (size_t)total));
        out->n_nodes4 = total;
        if (timing) cudaEventRecord(es, st);
        LB_TRY(cudaEventRecord(ev[2], st));
        k_lbvh_emit<<<g_n, 256, 0, st>>>(T, flag, index4, out->nodes4);
        stage("emit");
    }
    k_gather_tris<<<(unsigned)(((int64_t)n * 8 + 255) / 256), 256, 0, st>>>(d_isect, out->slot_prim, n, out->tris);
    stage("gather");
    if (timing) { cudaEventDestroy(es); cudaEventDestroy(ee); }
    LB_TRY(cudaEventRecord(ev[3], st));
    LB_TRY(cudaStreamSynchronize(st));
    LB_TRY(cudaGetLastError());
    float a = 0.f, b = 0.f;
    cudaEventElapsedTime(&a, ev[0], ev[1]);
    cudaEventElapsedTime(&b, ev[2], ev[3]);
    out->build_ms = a + b;
    return cudaSuccess;
}

}  // namespace prt

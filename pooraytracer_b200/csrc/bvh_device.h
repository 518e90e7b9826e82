/*
 * bvh_device.h — device-side construction of the traversal BVH (bvh_device.cu), SURVEY §8 f3.
 */
#ifndef PRT_BVH_DEVICE_H
#define PRT_BVH_DEVICE_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "device_scene.h"
#include "scene_host.h"

namespace prt {

struct DeviceBvh {          /* outputs, cudaMalloc'ed by the builder, owned by the caller afterwards */
    Node4* nodes4 = nullptr;
    int64_t n_nodes4 = 0;
    int32_t* slot_prim = nullptr; /* [n] slot -> prim (Morton order) */
    TriIsect* tris = nullptr;     /* [n] TriIsect in slot order */
    float build_ms = 0.f;         /* device time of the build (CUDA events on `st`) */
};

/* d_box: the triangles' FP64 boxes in prim order (Source/Triangle.cpp:94-99), d_isect: their
 * TriIsect records in prim order, both on the device. lo / hi: the scene's bounding box.
 * Synchronises `st` before returning. */
cudaError_t build_lbvh4_device(const Box6* d_box, const TriIsect* d_isect, int64_t n, const double lo[3], const double hi[3], int max_leaf,
                               DeviceBvh* out, cudaStream_t st);

}  // namespace prt
#endif

/*
 * scene_host.h — host side of the scene: what main.cpp:33-45 builds out of Model::meshes, kept as
 * flat arrays, plus the flattened layouts that get uploaded (device_scene.h).
 *
 * Pure C++ (no CUDA): compiled with -ffp-contract=off so that every derived quantity (triangle
 * normal / D / w / tangent / area / bbox, BVH boxes and areas) has the reference's bits.
 */
#ifndef PRT_SCENE_HOST_H
#define PRT_SCENE_HOST_H

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/prt_b200.h"
#include "device_scene.h"
#include "prt_math.h"

namespace prt {

struct Box6 {
    double v[6]; /* x.min x.max y.min y.max z.min z.max */
};

/* One Hittable in a BVH build: a triangle (kind 0) or a BVHNode (kind 1) — Source/BVH.h:25-27. */
struct ObjRef {
    int32_t kind;
    int32_t idx;
};

struct HostRefNode { /* Source/BVH.h:11-34 */
    Box6 box;
    ObjRef left, right;
    double area;
};

struct HostTexture {
    int32_t width = 0, height = 0, channels = 0;
    std::vector<uint8_t> texels;
};

struct HostScene {
    /* ---- copied inputs ---- */
    int64_t n_tris = 0;
    std::vector<double> vertices;  /* [n][9] */
    std::vector<double> texcoords; /* [n][6] (defaults applied) */
    std::vector<int64_t> mesh_first;
    std::vector<int32_t> mesh_material;
    std::vector<prt_material> materials;
    std::vector<HostTexture> textures;

    /* ---- derived per triangle (prim order), Source/Triangle.cpp:11-53,94-99 ---- */
    std::vector<TriIsect> isect;
    std::vector<TriShade> shade;
    std::vector<double> area;
    std::vector<Box6> tri_box;

    /* ---- the reference's BVHs (Source/BVH.cpp:7-48 over main.cpp:36-45) ---- */
    std::vector<HostRefNode> ref_nodes;
    int32_t world_root = -1, lights_root = -1;
    std::vector<int64_t> world_rank;  /* prim -> rank in DFS leaf order */
    std::vector<int64_t> light_order; /* lights DFS leaf order -> prim */

    /* ---- flattened, ready for upload ---- */
    std::vector<LightTri> light_tris;
    std::vector<LightNode> light_nodes;
    int32_t light_root_flat = -1;
    double lights_area = 0.0;
    std::vector<DevMaterial> dev_materials;
    std::vector<DevTexture> dev_textures;
    std::vector<uint8_t> texel_pool;
    double srgb_lut[256];

    /* committed acceleration structure */
    std::vector<Node2> nodes2;         /* SAH BVH2 */
    std::vector<Node4> nodes4;         /* SAH BVH4 (collapsed BVH2) */
    std::vector<Node4Q> nodes4q;       /* nodes4 with quantised child boxes (PRT_ACCEL_SAH_BVH4Q) */
    std::vector<int32_t> slot_prim;    /* slot -> prim */
    std::vector<int32_t> prim_slot;    /* prim -> slot */
    std::vector<TriIsect> isect_slots; /* isect permuted into slot order */
    std::vector<RefNode> ref_flat;     /* reference BVH (world part), leaves = ~slot */
    int32_t ref_root_flat = -1;
    int sah_max_leaf = 4;

    std::string error;

    bool ingest(const prt_scene_desc* d);
    void build_reference_bvhs(int n_threads);
    void build_sah_bvh2(int n_threads, int max_leaf);
    void build_sah_bvh4(int n_threads, int max_leaf);
    void quantise_bvh4(); /* nodes4 -> nodes4q */
    /* slot order = reference DFS order, no SAH tree (PRT_ACCEL_REFERENCE_BVH2) */
    void use_reference_slots();
    void flatten_reference_world();
};

int default_host_threads();

}  // namespace prt
#endif

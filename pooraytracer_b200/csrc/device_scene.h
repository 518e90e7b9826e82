/*
 * device_scene.h — layout of the committed scene in HBM (plain structs, shared by the host
 * flattener and the kernels). Everything is sized and aligned for 128-bit loads; the two records
 * fetched inside the traversal loop are exactly one (TriIsect, 128 B) or half (Node2, 64 B) of a
 * 128-byte L2 line.
 */
#ifndef PRT_DEVICE_SCENE_H
#define PRT_DEVICE_SCENE_H

#include <stdint.h>

namespace prt {

/* What Triangle::Hit reads (Source/Triangle.cpp:54-74): normal, D, w, v0, e0, e1 = 16 doubles.
 * Stored in the leaf order of the committed BVH ("slot" order). */
#define PRT_TRI_BYTES 128
struct alignas(128) TriIsect {
    double n[3];
    double D;
    double w[3];
    double v0[3];
    double e0[3];
    double e1[3];
};

/* Shading attributes of a triangle, indexed by prim id (Source/Triangle.cpp:76-80,111). */
struct alignas(128) TriShade {
    double uv[6];      /* uv0, uv1, uv2 */
    double tangent[3];
    double normal[3];  /* geometric normal (same bits as TriIsect::n) */
    int32_t material;
    int32_t pad[7];
};

/* A light triangle in the leaf order of the reference's lights BVH (Source/Triangle.cpp:84-93). */
struct alignas(128) LightTri {
    double v0[3], v1[3], v2[3];
    double normal[3];
    double area;
    int32_t material;
    int32_t prim;
    double pad[2];
};

/* Node of the lights BVH for BVHNode::TraverseSample (Source/BVH.cpp:86-100): only the area of the
 * left child and the two child references are read. ref >= 0: node index; ref < 0: ~light index. */
struct alignas(16) LightNode {
    double left_area;
    int32_t left, right;
};

/* SAH BVH2, both child boxes in the parent, FP32 rounded OUTWARD from the FP64 boxes. 64 bytes =
 * four 128-bit loads. child >= 0: inner node index; child < 0: leaf, ~child = (first_slot << 3) |
 * (count - 1), count in 1..8. */
#define PRT_NODE2_BYTES 64
struct alignas(64) Node2 {
    float lo0[3], hi0[3];
    float lo1[3], hi1[3];
    int32_t child0, child1;
    int32_t pad[2];
};

/* SAH BVH4: the BVH2 collapsed (grandchildren pulled up by largest surface area), four child boxes
 * per node as outward-rounded FP32, stored per axis across the children so that one 128-bit load
 * brings the same bound of all four. 128 bytes = one L2 line = seven 128-bit loads (112 B used).
 * Child references as in Node2; an unused slot holds PRT_NO_CHILD. */
#define PRT_NODE4_BYTES 128
struct alignas(128) Node4 {
    float lo[3][4]; /* lo[axis][child] */
    float hi[3][4];
    int32_t child[4];
    int32_t pad[4];
};

/* The BVH4 with QUANTISED child boxes: 64 bytes = half an L2 line = two 256-bit loads (the traversal
 * kernels are bound by L1TEX requests, every lane fetching its own node). A child's plane on axis k is
 * origin[k] + q * 2^(exp_k - 127), q in 0..255: lows rounded down, highs rounded up on the host, checked
 * in double against the FP64 boxes, so the decoded box contains the Node4 box it was made from. Byte c
 * of qlo[k] / qhi[k] belongs to child c. Child references as in Node4. */
#define PRT_NODE4Q_BYTES 64
struct alignas(64) Node4Q {
    float origin[3];
    uint32_t exps;    /* byte k: biased exponent of axis k's step (0: step 0, a flat axis) */
    uint32_t qlo[3];
    uint32_t pad0;
    uint32_t qhi[3];
    uint32_t pad1;
    int32_t child[4];
};

/* The reference's own median-split BVH, flattened (Source/BVH.cpp:7-48): FP64 box, children in the
 * reference's left/right order. ref >= 0: node; ref < 0: ~slot of a triangle. */
#define PRT_REFNODE_BYTES 64
struct alignas(64) RefNode {
    double box[6]; /* x.min x.max y.min y.max z.min z.max */
    int32_t left, right;
    int32_t pad[2];
};

struct alignas(16) DevMaterial {
    double kd[3];
    double ks[3];
    double emission[3]; /* GetEmission(): DiffuseLight radiance, or the albedo for DebugMaterial */
    double ns, pkd, pks;
    double alpha_x, alpha_y; /* CookTorrance roughness (Material.h:518) */
    double eta[3], k[3];     /* CookTorrance complex index of refraction per channel */
    int32_t type;
    int32_t kd_tex;
    int32_t has_emission;
    int32_t skip_light_sampling;
};

struct alignas(16) DevTexture {
    int64_t offset; /* into the texel pool */
    int32_t width, height, channels, valid;
};

/* Everything a kernel needs, passed by value. */
struct DevScene {
    const TriIsect* tris;      /* [n_tris] slot order */
    const int32_t* slot_prim;  /* [n_tris] slot -> prim id */
    const TriShade* shade;     /* [n_tris] prim order */
    const Node2* nodes2;       /* SAH BVH2 (root = 0) or nullptr */
    const Node4* nodes4;       /* SAH BVH4 (root = 0) or nullptr */
    const Node4Q* nodes4q;     /* quantised BVH4 (root = 0) or nullptr */
    const RefNode* ref_nodes;  /* reference BVH */
    int32_t ref_root;          /* -1: empty world */
    int32_t n_nodes2;
    int64_t n_tris;
    const LightTri* light_tris;
    const LightNode* light_nodes;
    int32_t light_root; /* -1: no lights */
    int32_t n_light_tris;
    double lights_area; /* area of the top lights node */
    const DevMaterial* materials;
    const DevTexture* textures;
    const uint8_t* texels;
    const double* srgb_lut; /* [256] SRGBToLinear(i/255) computed on the host (Texture.cpp:66-71) */
    int32_t n_materials, n_textures;
    unsigned long long* ref_overflow; /* sticky error counter: subtrees the reference-order walk dropped because its
                                         PRT_STACK-entry stack was full (prt_stats.ref_stack_overflows; 0 in a healthy build) */
};

}  // namespace prt
#endif

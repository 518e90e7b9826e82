/*
 * prt_b200.h — C ABI of the B200-native path-tracing backend for Pooraytracer.
 *
 * The reference (Zoz4/Pooraytracer) has no plugin / FFI boundary: its render path sits behind two
 * C++ seams that `main.cpp` calls directly,
 *
 *     void Camera::Render(Hittable& world, Hittable& lights)         Source/Camera.h:27  (main.cpp:49)
 *     bool Hittable::Hit(const Ray&, Interval, HitRecord&) const     Source/Hittable.h:34
 *     void Hittable::Sample(const point3&, HitRecord&, double&)      Source/Hittable.h:36
 *
 * This header is the boundary a maintainer would bind instead (INTEGRATION.md shows the stub):
 * plain pointers and sizes, no C++ / torch types, every call returns an int status (0 = ok) and
 * never throws. Inputs are copied at create time (the caller keeps ownership); outputs are
 * caller-allocated. One context is bound to one GPU and is used from one host thread at a time.
 * Multi-GPU: one prt_multi handle for the whole box from one process (prt_multi_*), or one
 * context per rank with the caller's own reduce (prt_render_device).
 *
 * There is NO CPU fallback: every compute entry point fails with PRT_ERR_NO_DEVICE when the
 * context has no usable sm_100 device.
 *
 * Conventions
 *   - all geometry / radiance arithmetic is IEEE binary64, evaluated in the reference's operation
 *     order without FMA contraction (Source/Triangle.cpp:54-83, Source/AABB.cpp:38-64);
 *   - prim id  = index of the triangle in prt_scene_desc (the creation order of
 *     Source/Model.cpp:115-190: shape-major, face-minor);
 *   - images are row-major, top row first, 3 doubles (linear RGB) per pixel — the layout of
 *     `Camera::colorAttachment` (Source/Camera.h:26, Source/Camera.cpp:56).
 */
#ifndef PRT_B200_H
#define PRT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PRT_ABI_VERSION 2

/* ---- status codes ---------------------------------------------------------------------------- */
#define PRT_OK 0
#define PRT_ERR_INVALID_ARGUMENT (-1)
#define PRT_ERR_NO_DEVICE (-2)      /* no CUDA device / not an sm_100 part / context has none   */
#define PRT_ERR_CUDA (-3)           /* a CUDA runtime call failed; see prt_last_error           */
#define PRT_ERR_OUT_OF_MEMORY (-4)
#define PRT_ERR_NOT_COMMITTED (-5)  /* compute call on a scene that was not committed           */
#define PRT_ERR_UNSUPPORTED (-6)

typedef struct prt_context prt_context;
typedef struct prt_scene prt_scene;

/* ---- materials: numeric values follow `enum class MaterialType` Source/Material.h:48-51 ------ */
enum prt_material_type {
    PRT_MAT_LAMBERTIAN = 0,    /* Source/Material.h:101-155 */
    PRT_MAT_PHONG = 1,         /* PhoneReflectance, Source/Material.h:172-330 */
    PRT_MAT_MIRROR = 2,        /* PerfectMirror, Source/Material.h:332-366 */
    PRT_MAT_COOK_TORRANCE = 3, /* Source/Material.h:368-521 */
    PRT_MAT_DIFFUSE_LIGHT = 4, /* Source/Material.h:157-170 */
    PRT_MAT_DEBUG = 5,         /* Source/Material.h:523-535 (emits its albedo) */
    PRT_MAT_EMPTY = 6          /* Source/Material.h:537-540 */
};

typedef struct prt_material {
    int32_t type;       /* prt_material_type */
    int32_t kd_tex;     /* index into prt_scene_desc.textures for map_Kd, or -1 (SolidColor)    */
    double kd[3];       /* Lambertian albedo / Phong Kd / CookTorrance Kd / Debug albedo        */
    double ks[3];       /* Phong Ks (ignored when kd_tex >= 0: the reference aliases Ks to the  */
                        /* Kd texture, Source/Material.h:178-181)                               */
    double ns;          /* Phong shininess                                                      */
    double emission[3]; /* DiffuseLight radiance (<light radiance=...>, Source/Model.cpp:332)   */
    double alpha_x, alpha_y; /* CookTorrance roughness                                          */
    double eta[3], k[3];     /* CookTorrance complex IOR                                        */
} prt_material;

/* Decoded image, rows top to bottom, `channels` interleaved u8 per texel — what stbi_load(.., 0)
 * returns in Source/Texture.cpp:12. texels == NULL reproduces a failed load (cyan,
 * Source/Texture.cpp:24). */
typedef struct prt_texture {
    int32_t width, height, channels, reserved;
    const uint8_t* texels;
} prt_texture;

/* Flat scene: what Model::meshes holds after Source/Model.cpp:53-193. */
typedef struct prt_scene_desc {
    int64_t n_triangles;
    const double* vertices;  /* [n_triangles][3][3]  v0,v1,v2 xyz                               */
    const double* normals;   /* [n_triangles][3][3]  vertex normals (fallback only,             */
                             /*                      Source/Triangle.cpp:21-29) or NULL = zeros */
    const double* texcoords; /* [n_triangles][3][2]  or NULL = (0,0),(1,0),(1,1) per triangle   */
                             /*                      (the degenerate-UV rule Source/Model.cpp:170) */
    int32_t n_meshes;        /* one per OBJ shape, in file order (main.cpp:37 iteration order)  */
    const int64_t* mesh_first;     /* [n_meshes+1] triangles of mesh m = [first[m], first[m+1]) */
    const int32_t* mesh_material;  /* [n_meshes]   material of face 0 rules (Model.cpp:118)     */
    int32_t n_materials;
    const prt_material* materials;
    int32_t n_textures;
    const prt_texture* textures;
} prt_scene_desc;

/* ---- acceleration structure ------------------------------------------------------------------ */
enum prt_accel {
    /* The reference's own two-level median-split BVH (Source/BVH.cpp:7-48, main.cpp:36-45),
     * flattened, FP64 boxes, traversed in the reference's order (left then right,
     * Source/BVH.cpp:51-61) including its duplicated span-1 children. Bit-exact by construction. */
    PRT_ACCEL_REFERENCE_BVH2 = 0,
    /* Binned-SAH BVH2, 64-byte nodes holding both child boxes as outward-rounded FP32, near-first
     * traversal. Same hits: the triangle test is the reference's FP64 arithmetic, the boxes only
     * ever contain the reference's boxes, and a ray that meets two triangles at the very same t
     * (the one case where the reference's visiting order decides: last visited wins,
     * Source/Interval.h:21-23 + Source/BVH.cpp:58) is re-traced through the reference BVH. */
    PRT_ACCEL_SAH_BVH2 = 1,
    /* The same binned-SAH tree collapsed to four children per node (128-byte nodes, one L2 line):
     * half the dependent node fetches per ray. Same boxes, same triangle arithmetic, same tie rule. */
    PRT_ACCEL_SAH_BVH4 = 2,
    /* A BVH4 in the same node format built ON the GPU (replaces the host build of
     * Source/BVH.cpp:7-48 for the traversal structure): 63-bit Morton codes, radix sort, Karras'
     * hierarchy, bottom-up FP64 boxes, subtrees of <= max_leaf triangles collapsed into leaves. A
     * lower-quality tree than the binned-SAH one (more nodes per ray) committed in a fraction of the
     * time; same hits by the same argument. Needs a GPU context (PRT_ERR_NO_DEVICE otherwise). */
    PRT_ACCEL_LBVH4_DEVICE = 3,
    /* PRT_ACCEL_SAH_BVH4 with the child boxes quantised to 8 bits per plane relative to the node
     * (64-byte nodes: half the bytes per node fetch, slightly looser boxes, more decode arithmetic).
     * Same hits: the decoded boxes contain the boxes they replace. */
    PRT_ACCEL_SAH_BVH4Q = 4,
    /* Let the library choose between the two SAH BVH4 layouts by mesh size: quantised 64-byte nodes
     * from 262144 triangles up (the node set no longer fits the caches the FP32 nodes are served
     * from: measured +4 % on the 1.1 M-triangle bathroom stand-in, profiles/r2_s11_accel_ab.txt),
     * FP32 128-byte nodes below (measured +1..3 % on cornell-box / veach-mis). prt_get_stats reports
     * the layout that was built. Same hits either way. */
    PRT_ACCEL_AUTO = 5
};

typedef struct prt_build_opts {
    int32_t accel;        /* prt_accel */
    int32_t reserved[7];  /* reserved[0]: max triangles per SAH leaf (0 = default 4, at most 8) */
} prt_build_opts;

/* ---- camera + integrator parameters (public fields of Source/Camera.h:14-33) ------------------ */
typedef struct prt_camera {
    int32_t image_width, image_height;
    double fovy;
    double eye[3], look_at[3], up[3];
} prt_camera;

#define PRT_RENDER_EXACT_VISIBILITY 1u /* NEE visibility by closest-hit + the literal distance test
                                          of Source/Camera.cpp:150-155 instead of any-hit over
                                          [1e-3, dist-1e-3] (equivalent up to rounding)          */
#define PRT_RENDER_PIXEL_JITTER 8u     /* OFF by default and never part of parity with the shipped
                                          reference: the sample position Camera::GetRay's commented-out
                                          lines intend (Source/Camera.cpp:110-111, SampleSquare(),
                                          RandomNumberGenerator.h:65-68): pixel + (u0 - .5, u1 - .5),
                                          drawn as the first two numbers of the sample's stream    */
#define PRT_RENDER_UNBIASED_LIGHT_PICK 16u /* OFF by default, not the reference's estimator: BVHNode::Sample
                                          descends its area tree with p = RandomDouble() * area instead of
                                          sqrt(RandomDouble()) * area (Source/BVH.cpp:64), so a light triangle
                                          is picked in proportion to its area, which is what the returned
                                          pdf = 1 / area already claims                                   */
#define PRT_RENDER_LOBE_PROBABILITY 32u /* OFF by default, not the reference's estimator: PhoneReflectance::
                                          Scatter divides the attenuation by the probability of the lobe it
                                          sampled (pkd / pks, Source/Material.h:262-285,318-327)          */
#define PRT_RENDER_MIS 64u             /* OFF by default, not the reference's estimator: at a vertex whose
                                          BSDF is the cosine lobe (Lambertian, Phong with Ns <= 1) and which
                                          samples lights, the light sample is weighted p_l / (p_l + p_b) and
                                          an emitter met by the scattered ray — which the reference drops
                                          (Source/Camera.cpp:187-196) — is added with p_b / (p_b + p_l)
                                          (balance heuristic; p_b = cos / pi, p_l = dist^2 / (cos_l A))   */
#define PRT_RENDER_COUNTERS 2u         /* count BVH node / triangle visits (prt_get_stats)       */
#define PRT_RENDER_STAGE_TIMING 4u     /* CUDA events around the extend / shade / shadow kernels of
                                          every bounce (prt_stats.extend_ms ...); the call then
                                          waits for each pass                                     */

typedef struct prt_render_params {
    int32_t samples_per_pixel; /* total spp: pixelSamplesScale = 1/spp (Source/Camera.cpp:83)    */
    int32_t max_depth;         /* Camera::maxDepth: at most max_depth+1 path vertices            */
    int32_t sample_lights;     /* Camera::bSampleLights                                          */
    int32_t spp_begin, spp_end;/* this call renders sample indices [spp_begin, spp_end) of every */
                               /* pixel (spp sharding across GPUs / steps); 0,0 = all            */
    uint32_t flags;
    double russian_roulette;   /* Camera::russianRoulette                                        */
    double background[3];      /* Camera::background                                             */
    uint64_t rng_key;          /* key of the counter-based generator (pixel, sample, dimension)  */
} prt_render_params;

typedef struct prt_stats {
    uint64_t extend_rays;      /* closest-hit rays traced by the last render / trace call        */
    uint64_t shadow_rays;      /* shadow rays traced                                             */
    uint64_t paths;            /* camera paths started                                           */
    uint64_t node_visits;      /* BVH nodes fetched   (only with PRT_RENDER_COUNTERS / counters) */
    uint64_t tri_tests;        /* triangle records fetched                                       */
    uint64_t kernel_launches;  /* kernels launched by the last call                              */
    double   gpu_ms;           /* device time of the last call (CUDA events on its stream)       */
    double   extend_ms;        /* device time inside the extend kernel(s)                        */
    double   shade_ms;
    double   shadow_ms;
    int32_t  node_bytes;       /* bytes per BVH node of the committed layout                     */
    int32_t  tri_bytes;        /* bytes per triangle intersection record (128)                   */
    int64_t  n_nodes;
    int64_t  n_triangles;
    double   build_ms;         /* PRT_ACCEL_LBVH4_DEVICE: device time of the BVH build (CUDA events);
                                  0 for the host-built structures                                */
    uint64_t ref_stack_overflows; /* STICKY error counter since commit: subtrees the reference-order walk (boundary /
                                  equal-t re-trace, PRT_ACCEL_REFERENCE_BVH2) had to drop because its stack was full.
                                  0 in a healthy build: its depth is log2(mesh size) + the top level. Non-zero means
                                  results of those rays are not the reference's                   */
} prt_stats;

/* ---- lifecycle -------------------------------------------------------------------------------- */
int prt_abi_version(void);

/* device < 0: host-only context (scene building / flattening only; compute calls fail with
 * PRT_ERR_NO_DEVICE). Otherwise binds CUDA device `device`. */
int prt_context_create(int device, prt_context** out);
void prt_context_destroy(prt_context* ctx);
/* Message of the last failure on this thread (never NULL). */
const char* prt_last_error(void);

/* Copies every input; builds the reference-order BVH (needed for the tie rule and for light
 * sampling whatever accel is chosen later). Replaces: Model::meshes → per-mesh BVHNode →
 * world / lights HittableList construction, main.cpp:33-45. */
int prt_scene_create(prt_context* ctx, const prt_scene_desc* desc, prt_scene** out);
/* Flattens into the chosen layout and uploads to HBM. */
int prt_scene_commit(prt_scene* scene, const prt_build_opts* opts);
void prt_scene_destroy(prt_scene* scene);
/* A replica of a COMMITTED scene on another context's GPU: the device buffers are copied GPU to GPU (NVLink between peers),
 * nothing is rebuilt on the host. This is how prt_multi_create replicates the scene over the box. The replica renders and
 * answers ray queries like the original; prt_scene_reference_order answers on the original only. */
int prt_scene_clone(const prt_scene* source, prt_context* ctx, prt_scene** out);

/* Host-side introspection (no GPU needed): the reference DFS order the tie rule relies on.
 * world_rank[prim] = position of the triangle in the left-to-right leaf order of the reference
 * world BVH (Source/BVH.cpp:16-34). light_order[i] = prim id of the i-th light triangle in the
 * lights BVH's leaf order; returns their count through n_light_tris (pass NULL arrays to size). */
int prt_scene_reference_order(const prt_scene* scene, int64_t* world_rank, int64_t* light_order,
                              int64_t* n_light_tris);

/* ---- ray queries: replaces Hittable::Hit on `world` (Source/Hittable.h:34) --------------------- */
/* origins/directions: [n][3] doubles; tmin/tmax: [n] (Interval, inclusive — Source/Interval.h:21).
 * prim[i] = -1 on miss (t/alpha/beta then untouched = 0). alpha/beta are the barycentrics of v1/v2
 * (Source/Triangle.cpp:70-71). Host pointers; copies are part of the call. */
int prt_trace_closest(prt_scene* scene, int64_t n, const double* origins, const double* directions,
                      const double* tmin, const double* tmax, int32_t* prim, double* t,
                      double* alpha, double* beta);
/* hit[i] = 1 iff some triangle is hit inside [tmin, tmax] (the visibility query of
 * Source/Camera.cpp:147-155 phrased as any-hit). */
int prt_trace_any(prt_scene* scene, int64_t n, const double* origins, const double* directions,
                  const double* tmin, const double* tmax, uint8_t* hit);
/* Same, all pointers in device memory of the context's GPU, launched on `stream` (a cudaStream_t
 * passed as void*; NULL = the context's stream). Asynchronous. */
int prt_trace_closest_device(prt_scene* scene, int64_t n, const double* origins,
                             const double* directions, const double* tmin, const double* tmax,
                             int32_t* prim, double* t, double* alpha, double* beta, void* stream);
int prt_trace_any_device(prt_scene* scene, int64_t n, const double* origins,
                         const double* directions, const double* tmin, const double* tmax,
                         uint8_t* hit, void* stream);
/* prt_trace_closest_device that allocates NOTHING: no cudaMalloc, cudaFree or allocating library
 * routine on any call, the first included. No counterpart in the reference: added
 * for callers that cannot reach cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it). Same kernels, same results bit for bit; the
 * scratch the kernels need (the work cursor of the persistent traversal kernel and, under
 * PRT_SORT_MIN, the ray-binning keys / order / histogram) lives in one static device array of
 * 6.25 MiB that every GPU which loads the library holds until exit.
 *   - n > PRT_NOALLOC_MAX_RAYS is refused with PRT_ERR_INVALID_ARGUMENT (split the batch);
 *   - n == 0 performs the once-per-device set-up only (the address look-up of the array), which a
 *     caller can do before a capture begins;
 *   - every call on a GPU shares the one array: calls that OVERLAP on the same GPU (two streams,
 *     two scenes, two threads) race on it. Serialise them, or use prt_trace_closest_device, whose
 *     scratch is per scene;
 *   - the scene's prt_stats are not touched (no timing events, no fetch counters). */
#define PRT_NOALLOC_MAX_RAYS 1048576
int prt_trace_closest_device_noalloc(prt_scene* scene, int64_t n, const double* origins,
                                     const double* directions, const double* tmin,
                                     const double* tmax, int32_t* prim, double* t, double* alpha,
                                     double* beta, void* stream);

/* ---- light sampling: replaces Hittable::Sample on `lights` (Source/Hittable.h:36) -------------- */
/* u: [n][4] uniforms in draw order (HittableList.h:49, BVH.cpp:64, Triangle.cpp:86 x then y).
 * Outputs: position[n][3], normal[n][3] (face-forwarded to the origin), front[n], pdf[n],
 * prim[n] (sampled light triangle). Host pointers. */
int prt_sample_lights(prt_scene* scene, int64_t n, const double* origins, const double* u,
                      double* position, double* normal, uint8_t* front, double* pdf,
                      int32_t* prim);

/* ---- rendering: replaces Camera::Render (Source/Camera.cpp:21-73) ------------------------------ */
/* rgb: host buffer of image_width*image_height*3 doubles, OVERWRITTEN with the radiance of the
 * requested sample range scaled by 1/samples_per_pixel (a caller that wants the reference's
 * accumulate-on-top behaviour, Source/Camera.cpp:81, adds it to its own buffer). */
int prt_render(prt_scene* scene, const prt_camera* camera, const prt_render_params* params,
               double* rgb);
/* Same, ACCUMULATING (+=) into a device buffer and asynchronous on `stream`; the multi-GPU path
 * reduces these buffers with one NCCL sum (torch.distributed) after every rank has rendered its
 * sample range. */
int prt_render_device(prt_scene* scene, const prt_camera* camera,
                      const prt_render_params* params, double* rgb_device, void* stream);

/* ---- output stage: replaces the pixel loop of Camera::WriteColorAttachment (Camera.cpp:279-331) */
/* NaN scrub, sRGB OETF, clamp to [0,0.9999]*255 truncation → rgb8[W*H*3]; optional float32 copy
 * of the scrubbed linear image (the .hdr payload). Device buffer in, host buffers out. */
int prt_resolve_device(prt_scene* scene, const double* rgb_device, int32_t width, int32_t height,
                       uint8_t* rgb8, float* rgb32f);
/* Same with the radiance image in host memory (the colorAttachment a caller kept). */
int prt_resolve(prt_scene* scene, const double* rgb, int32_t width, int32_t height, uint8_t* rgb8,
                float* rgb32f);
/* Host-side introspection (no GPU needed): out254[k - 1] = the smallest linear value whose 8-bit code is k,
 * k = 1..254, found with the libm call of Camera::LinearToSRGB (Camera.cpp:216-221) — the table the resolve
 * kernel quantises by, so that its bytes are the reference's for every double. */
int prt_resolve_thresholds(double* out254);

/* ---- the whole box: 1..8 GPUs behind one handle ------------------------------
 * Replaces main.cpp:36-52 for a caller that wants every GPU of the node: the scene is replicated (ingested, built and
 * uploaded ONCE, on devices[0], then copied GPU to GPU: prt_scene_clone), Camera::Render's samples are split into contiguous ranges across the GPUs
 * (the generator is keyed by the absolute sample index, so the union of the shards is the one-GPU image up to the order
 * of the final FP64 additions), and ONE reduction of the W*H*3 FP64 frames to devices[0] over NVLink ends the render:
 * where devices[0] can address the other GPUs' memory (every NVSwitch box) a kernel on devices[0] reads the other frames
 * through peer loads and adds them in device order — no communicator, nothing to initialise, bit-reproducible; otherwise
 * ncclReduce(sum, double) on communicators the library owns (NCCL is bound at run time, only then). The environment
 * variable PRT_MULTI_REDUCE=peer|nccl forces one kind. Used from one host thread at a time; the library runs one thread
 * of its own per GPU. */
typedef struct prt_multi prt_multi;
enum prt_reduce_kind { PRT_REDUCE_NONE = 0 /* one GPU */, PRT_REDUCE_NCCL = 1, PRT_REDUCE_PEER = 2 };
int prt_multi_create(const int32_t* devices, int32_t n, const prt_scene_desc* desc, const prt_build_opts* opts,
                     prt_multi** out);
void prt_multi_destroy(prt_multi* multi);
int prt_multi_device_count(const prt_multi* multi);
/* prt_reduce_kind of this handle */
int prt_multi_reduce_kind(const prt_multi* multi);
/* Camera::Render on all GPUs: rgb (host, W*H*3 doubles, may be NULL) is OVERWRITTEN with the reduced image of the sample
 * range, scaled by 1/samples_per_pixel; the reduced accumulators also stay on devices[0] for prt_multi_resolve. */
int prt_multi_render(prt_multi* multi, const prt_camera* camera, const prt_render_params* params, double* rgb);
/* The output stage (Camera.cpp:279-331) on devices[0], on the image of the last prt_multi_render: NaN scrub AFTER the
 * reduce, as the reference scrubs the final pixel. */
int prt_multi_resolve(prt_multi* multi, uint8_t* rgb8, float* rgb32f);
/* Sums of the per-GPU counters of the last call, gpu_ms = the slowest GPU's device time; reduce_ms (may be NULL) = device
 * time of the reduction on devices[0]. */
int prt_multi_get_stats(const prt_multi* multi, prt_stats* out, double* reduce_ms);
/* Hittable::Hit over a batch split into contiguous ray ranges, one per GPU; no collective. Host pointers. */
int prt_multi_trace_closest(prt_multi* multi, int64_t n, const double* origins, const double* directions,
                            const double* tmin, const double* tmax, int32_t* prim, double* t, double* alpha,
                            double* beta);
int prt_multi_trace_any(prt_multi* multi, int64_t n, const double* origins, const double* directions,
                        const double* tmin, const double* tmax, uint8_t* hit);

/* ---- scene files: replaces Model::Model + Camera::SetViewParametersByXmlFile -------------------
 * Host-only (no GPU needed). Reads <dir>/<name>.obj, the .mtl it names, <dir>/<name>.xml (camera +
 * <light mtlname radiance>), and the map_Kd images, the way Source/Model.cpp:53-394 and
 * Source/Camera.cpp:339-389 do through tinyobjloader / tinyxml2 / stb_image:
 *   - one mesh per OBJ shape (`g` / `o`), the material of the shape's FIRST face rules the whole
 *     shape (Model.cpp:118), polygons are triangulated like tinyobjloader's default config;
 *   - the material TYPE comes from the reference's name table (Model.cpp:16-51), unknown names are
 *     Lambertian, a name in the table as Phong falls back to CreatePhoneReflectanceMaterial;
 *   - DiffuseLight radiance comes from the XML's <light> elements (Model.cpp:320-347);
 *   - triangles whose texture coordinates coincide get (0,0),(1,0),(1,1) (Model.cpp:170-175);
 *   - anonymous `g` lines are named Group_<n> IN MEMORY (the reference rewrites the user's .obj
 *     file in place, Model.cpp:195-268; this loader never writes to its inputs).
 * PRT_LOAD_STRTOD parses reals with strtod (correctly rounded) instead of tinyobjloader's own
 * digit-accumulating parser (the default, which is what the reference's numbers went through). */
typedef struct prt_model prt_model;
#define PRT_LOAD_STRTOD 1u
int prt_model_load(const char* directory, const char* name, uint32_t flags, prt_model** out);
void prt_model_destroy(prt_model* model);
/* The loaded scene as flat arrays owned by the model (valid until prt_model_destroy). */
const prt_scene_desc* prt_model_desc(const prt_model* model);
/* Camera of the XML file (image size, fovy, eye, lookat, up); PRT_ERR_INVALID_ARGUMENT when the file
 * has no <camera> element (the reference then renders with Camera.h's defaults). */
int prt_model_camera(const prt_model* model, prt_camera* out);
/* Names as the reference logs them (Model.cpp:121-125): shape name and material name of mesh m. */
const char* prt_model_mesh_name(const prt_model* model, int32_t mesh);
const char* prt_model_material_name(const prt_model* model, int32_t material);
/* Warnings collected while loading (unsupported image formats, unknown materials ...), one per
 * line; never NULL. */
const char* prt_model_warnings(const prt_model* model);

/* ---- image files: replaces stbi_write_png / stbi_write_hdr (Camera.cpp:302,327) and stbi_load ---
 * PNG: 8-bit RGB, rows top to bottom. HDR: Radiance RGBE of float32 RGB (the float -> RGBE
 * conversion is stb_image_write's: frexp of the largest component). */
int prt_write_png(const char* path, int32_t width, int32_t height, const uint8_t* rgb8);
int prt_write_hdr(const char* path, int32_t width, int32_t height, const float* rgb32f);
/* Decodes a PNG (any colour type / bit depth, non-interlaced), BMP (24/32-bit) or binary PNM into
 * `channels` interleaved u8 per texel the way stbi_load(path, .., 0) does. *texels is malloc'ed:
 * release it with prt_image_free. */
int prt_image_load(const char* path, int32_t* width, int32_t* height, int32_t* channels, uint8_t** texels);
void prt_image_free(uint8_t* texels);

/* enable != 0: the ray-query and render kernels count BVH node and triangle record fetches
 * (instrumented instantiations; slower). Read them back through prt_get_stats. */
int prt_set_counters(prt_scene* scene, int enable);
int prt_get_stats(const prt_scene* scene, prt_stats* out);
/* Waits for all work queued on the context's stream. */
int prt_synchronize(prt_context* ctx);

#ifdef __cplusplus
}
#endif
#endif /* PRT_B200_H */

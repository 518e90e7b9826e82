/*
 * oracle_api.h — TEST INFRASTRUCTURE. Common C interface of the two CPU checkers:
 *
 *   ref_*  oracle/ref_driver.cpp  — the UNMODIFIED reference sources compiled in place from
 *                                   /root/reference/Source (outputs in oracle/_ref/), driven
 *                                   through their own classes (Triangle, Mesh, BVHNode,
 *                                   HittableList, Camera, the Material family).
 *   orc_*  oracle/prt_oracle.cpp  — our CPU restatement (port) of the same algorithm, every
 *                                   function citing the reference file:line it follows.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load these libraries. The product (pooraytracer_b200/) never does.
 *
 * Scene/camera/params structs are the product's C-ABI structs (include/prt_b200.h) so one
 * description feeds the checker and the checked.
 */
#ifndef ORACLE_API_H
#define ORACLE_API_H
#include "../include/prt_b200.h"

/* scene flags */
#define ORC_PER_TRIANGLE_MATERIAL 1 /* every triangle gets its own material instance so that the  */
                                    /* hit triangle can be recovered from HitRecord::material      */
#define ORC_UV_TAG 2                /* texcoords := (0,0),(1,0),(0,1): HitRecord::uv == (alpha,beta) */
#define ORC_DEFINE_UB 4             /* Phong: caller's `attenuation` pre-set to 0 before Scatter   */
                                    /* (Camera.cpp:177 leaves it uninitialised), and the stream is */
                                    /* rewound after the draw inside Eval (Material.h:228): not UB, */
                                    /* but visibility-dependent (see PhongDefined, ref_driver.cpp)  */

/* rng modes for render */
#define ORC_RNG_COUNTER 0      /* rng_spec.h, keyed by (rng_key, pixel, sample, dim) — parity       */
#define ORC_RNG_THREAD_LOCAL 1 /* per-thread xorshift seeded from rng_key — fair timing baseline    */
#define ORC_RNG_LIBC 2         /* glibc rand() as shipped (global lock)                             */

typedef struct orc_render_info {
    uint64_t world_hit_calls; /* calls of world.Hit (extend + shadow + peek), all threads          */
    uint64_t shadow_calls;    /* of which visibility queries (domain.min == 1e-3)                  */
    uint64_t paths;
    double seconds;           /* wall time of the render loop                                      */
    int32_t threads;
    int32_t reserved;
} orc_render_info;

#define ORACLE_DECLARE(P)                                                                          \
    int P##_scene_create(const prt_scene_desc* desc, const char* const* texture_paths, int flags,  \
                         void** out);                                                              \
    void P##_scene_destroy(void* scene);                                                           \
    int P##_trace_closest(void* scene, int64_t n, const double* o, const double* d,                \
                          const double* tmin, const double* tmax, int32_t* prim, double* t,        \
                          double* alpha, double* beta, int nthreads);                              \
    int P##_reference_order(void* scene, int64_t* world_rank, int64_t* light_order,                \
                            int64_t* n_light_tris);                                                \
    int P##_sample_lights(void* scene, int64_t n, const double* origins, const double* u,          \
                          double* position, double* normal, uint8_t* front, double* pdf,           \
                          int32_t* prim);                                                          \
    int P##_render(void* scene, const prt_camera* cam, const prt_render_params* params,            \
                   int rng_mode, int nthreads, double* rgb, orc_render_info* info);                \
    /* per-sample radiance of one pixel (counter rng): out[(spp_end-spp_begin)][3] */              \
    int P##_pixel_samples(void* scene, const prt_camera* cam, const prt_render_params* params,     \
                          int32_t px, int32_t py, double* out);                                    \
    /* precomputed triangle record: normal[3], D, w[3], v0[3], e0[3], e1[3], tangent[3], area */   \
    int P##_triangle_record(void* scene, int64_t prim, double* out20);                             \
    /* one material interaction with scripted uniforms (rand() replays u[]): Scatter, returns     */\
    /* attenuation[3], scattered dir[3], n_draws; and Eval for wi_world -> f[3] */                 \
    int P##_material_scatter(void* scene, int32_t material, const double* ray_dir,                 \
                             const double* normal, const double* tangent, const double* uv,        \
                             const double* u, int32_t n_u, double* attenuation, double* out_dir,   \
                             int32_t* n_draws, int32_t* scattered);                                \
    /* Material::Eval as the NEE estimator calls it (Camera.cpp:160-168): context.wo =               */\
    /* WorldToLocal(-ray_dir), localWi = WorldToLocal(wi_world) -> f[3]                              */\
    int P##_material_eval(void* scene, int32_t material, const double* ray_dir,                    \
                          const double* normal, const double* tangent, const double* uv,           \
                          const double* wi_world, const double* u, int32_t n_u, double* f);        \
    int P##_texture_value(void* scene, int32_t texture, int64_t n, const double* uv, double* rgb); \
    /* camera ray of pixel (i,j): origin[3], direction[3] (Camera.cpp:108-117) */                  \
    int P##_camera_ray(const prt_camera* cam, int32_t i, int32_t j, double* origin, double* dir);

#ifdef __cplusplus
extern "C" {
#endif
ORACLE_DECLARE(ref)
ORACLE_DECLARE(orc)
/* ref only: value of AABB::empty as initialised in this build (static-init-order probe) */
int ref_aabb_empty(double* out6);
/* ref only: Camera::WriteColorAttachment (Camera.cpp:279-331) on a caller-supplied W x H x 3 float64 image */
int ref_write_color_attachment(const double* rgb, int32_t width, int32_t height, const char* png_path, int write_hdr);
#ifdef __cplusplus
}
#endif
#endif

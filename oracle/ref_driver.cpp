/*
 * ref_driver.cpp — TEST INFRASTRUCTURE (oracle). C driver around the UNMODIFIED reference
 * sources, which are compiled where they lie under /root/reference/Source by oracle/Makefile
 * (objects and the resulting libprt_ref*.so go to oracle/_ref/, never into git).
 *
 * Nothing of the reference is copied here: the driver only *uses* its classes
 *   Triangle / Mesh            Source/Triangle.h:13-43
 *   BVHNode / HittableList     Source/BVH.h:11-34, Source/HittableList.h:14-63
 *   Camera                     Source/Camera.h:11-52
 *   Material family            Source/Material.h:53-540
 * and builds `world` / `lights` exactly the way main.cpp:36-45 does.
 *
 * Three things are interposed without touching reference code:
 *   1. rand()  — defined here and bound inside the .so with -Bsymbolic-functions, so that
 *      RandomDouble() (Source/RandomNumberGenerator.h:16-19) reads a thread-local stream:
 *      counter-based (oracle/rng_spec.h), per-thread xorshift, scripted replay, or libc.
 *   2. `private` — this translation unit (only) sees the reference classes with their private
 *      members accessible, to call Camera::Initialize/GetRay/RayColor per sample and to walk
 *      BVHNode::left/right. Class layouts are unaffected.
 *   3. world — wrapped in a forwarding Hittable that counts Hit calls (rays).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.
 */
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iomanip>
#include <limits>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>
#include <dlfcn.h>

#include <glm/glm.hpp>
#include <glm/gtx/string_cast.hpp>

#define private public
#define protected public
#include "Source/AABB.h"
#include "Source/BVH.h"
#include "Source/Camera.h"
#include "Source/Hittable.h"
#include "Source/HittableList.h"
#include "Source/Material.h"
#include "Source/Ray.h"
#include "Source/Texture.h"
#include "Source/Triangle.h"
#undef private
#undef protected

#include "oracle_api.h"
#include "rng_spec.h"

using namespace Pooraytracer;

/* ------------------------------------------------------------------------------------------------
 * rand() interposition
 * ---------------------------------------------------------------------------------------------- */
namespace {

enum { MODE_COUNTER = 0, MODE_THREAD_LOCAL = 1, MODE_LIBC = 2, MODE_SCRIPT = 3 };

std::atomic<int> g_default_mode{MODE_THREAD_LOCAL};
std::atomic<uint64_t> g_seed_base{0x1234567ull};
std::atomic<uint64_t> g_thread_counter{0};

std::atomic<uint64_t> g_hit_calls{0}, g_shadow_calls{0};

struct ThreadState {
    int mode;
    uint64_t seed = 0;   // counter mode: stream seed
    uint32_t dim = 0;    // counter mode: next dimension
    uint64_t xs;         // thread-local xorshift state
    const int32_t* script = nullptr;
    int script_n = 0, script_i = 0;
    uint64_t hit_calls = 0, shadow_calls = 0;
    ThreadState() {
        mode = g_default_mode.load();
        uint64_t id = g_thread_counter.fetch_add(1);
        xs = orc_mix64(g_seed_base.load() + ORC_GOLDEN * (id + 1));
        if (xs == 0) xs = 0x9E3779B97F4A7C15ull;
    }
    void flush() {
        g_hit_calls.fetch_add(hit_calls);
        g_shadow_calls.fetch_add(shadow_calls);
        hit_calls = shadow_calls = 0;
    }
    ~ThreadState() { flush(); }
};
ThreadState& tls() {
    static thread_local ThreadState s;
    return s;
}

int libc_rand() {
    static int (*real)(void) = (int (*)(void))dlsym(RTLD_NEXT, "rand");
    return real ? real() : 0;
}
}  // namespace

extern "C" int rand(void) {
    ThreadState& t = tls();
    switch (t.mode) {
        case MODE_COUNTER:
            return orc_draw31(t.seed, t.dim++);
        case MODE_SCRIPT: {
            int v = (t.script_i < t.script_n) ? t.script[t.script_i] : 0;
            t.script_i++;
            return v;
        }
        case MODE_LIBC:
            return libc_rand();
        default: {
            // xorshift64* (Vigna 2016); top 31 bits
            uint64_t x = t.xs;
            x ^= x >> 12;
            x ^= x << 25;
            x ^= x >> 27;
            t.xs = x;
            return (int)((x * 0x2545F4914F6CDD1Dull) >> 33);
        }
    }
}

/* ------------------------------------------------------------------------------------------------
 * helpers
 * ---------------------------------------------------------------------------------------------- */
namespace {

// Forwarding wrapper that counts world.Hit calls (rays) without editing the reference.
class CountingWorld : public Hittable {
public:
    explicit CountingWorld(const Hittable& inner) : inner(inner) {}
    bool Hit(const Ray& ray, Interval domain, HitRecord& record) const override {
        ThreadState& t = tls();
        t.hit_calls++;
        if (domain.min == 0.001) t.shadow_calls++;
        return inner.Hit(ray, domain, record);
    }
    AABB BoundingBox() const override { return inner.BoundingBox(); }
    void Sample(const point3& origin, HitRecord& rec, double& pdf) const override {
        inner.Sample(origin, rec, pdf);
    }
    double GetArea() const override { return inner.GetArea(); }

private:
    const Hittable& inner;
};

// ORC_DEFINE_UB: two spots of the Phong material get a defined meaning while the reference's own Scatter / Eval
// bodies still do all the work.
//  (1) Scatter's `attenuation` is left uninitialised when pdf == 0 (Camera.cpp:177, Material.h:280-284): genuinely
//      undefined behaviour; pre-set to 0.
//  (2) PhoneReflectance::Eval draws a random number (Material.h:228). That is NOT undefined behaviour — it is an
//      ordinary rand() call that advances the stream — but Camera::RayColor only reaches Eval when the light sample
//      turned out VISIBLE (Camera.cpp:152-155), so whether the Russian-roulette and scatter draws that follow are
//      shifted by one depends on the outcome of the shadow ray. A wavefront integrator scatters BEFORE it traces its
//      shadow rays and cannot reproduce a visibility-dependent stream position; the draw's VALUE never matters where
//      Eval is reachable (Ns <= 1 => pkd = 1 => u < pkd always), so the estimator is unchanged if the stream is rewound
//      after it. Phong render parity (tests/golden/render_*: vertices with Ns <= 1) is therefore against this REWOUND
//      stream, not against the unmodified reference's stream; every other material consumes the unmodified stream.
class PhongDefined : public PhoneReflectance {
public:
    using PhoneReflectance::PhoneReflectance;
    bool Scatter(const Ray& rayIn, const HitRecord& record, color& attenuation,
                 Ray& scatteredRay) const override {
        attenuation = color(0., 0., 0.);  // Camera.cpp:177 leaves it uninitialised
        return PhoneReflectance::Scatter(rayIn, record, attenuation, scatteredRay);
    }
    vec3 Eval(const vec3& wi, const MaterialEvalContext& context) const override {
        ThreadState& t = tls();
        uint32_t saved = t.dim;
        vec3 f = PhoneReflectance::Eval(wi, context);
        if (t.mode == MODE_COUNTER) t.dim = saved;  // Material.h:228 draw: value never matters, position rewound (see above)
        return f;
    }
};

struct RefScene {
    int flags = 0;
    std::vector<std::shared_ptr<Mesh>> meshes;
    std::vector<std::shared_ptr<Texture>> textures;
    std::vector<std::shared_ptr<Material>> materials;       // per desc material
    std::vector<std::shared_ptr<Triangle>> tris;            // prim id -> triangle
    std::unordered_map<const Material*, int64_t> matToPrim; // per-triangle material -> prim
    std::unordered_map<const Hittable*, int64_t> triToPrim;
    HittableList world, lights;
    bool has_lights = false;
};

std::shared_ptr<Material> make_material(const RefScene& s, const prt_material& m, int flags) {
    color kd(m.kd[0], m.kd[1], m.kd[2]), ks(m.ks[0], m.ks[1], m.ks[2]);
    std::shared_ptr<Texture> tex;
    if (m.kd_tex >= 0 && (size_t)m.kd_tex < s.textures.size()) tex = s.textures[m.kd_tex];
    switch (m.type) {
        case PRT_MAT_LAMBERTIAN:  // Model.cpp:381-394
            if (tex) return std::make_shared<Lambertian>(tex);
            return std::make_shared<Lambertian>(kd);
        case PRT_MAT_PHONG:  // Model.cpp:362-379
            if (flags & ORC_DEFINE_UB) {
                if (tex) return std::make_shared<PhongDefined>(tex, ks, m.ns);
                return std::make_shared<PhongDefined>(kd, ks, m.ns);
            }
            if (tex) return std::make_shared<PhoneReflectance>(tex, ks, m.ns);
            return std::make_shared<PhoneReflectance>(kd, ks, m.ns);
        case PRT_MAT_MIRROR:
            return std::make_shared<PerfectMirror>();
        case PRT_MAT_COOK_TORRANCE:
            return std::make_shared<CookTorrance>(kd, m.alpha_x, m.alpha_y,
                                                  vec3(m.eta[0], m.eta[1], m.eta[2]),
                                                  vec3(m.k[0], m.k[1], m.k[2]));
        case PRT_MAT_DIFFUSE_LIGHT:
            return std::make_shared<DiffuseLight>(color(m.emission[0], m.emission[1], m.emission[2]));
        case PRT_MAT_DEBUG:
            return std::make_shared<DebugMaterial>(kd);
        case PRT_MAT_EMPTY:
            return std::make_shared<EmptyMaterial>();
        default:
            return std::make_shared<Lambertian>(kd);
    }
}

void set_camera(Camera& cam, const prt_camera* c, const prt_render_params* p) {
    cam.imageWidth = c->image_width;
    cam.imageHeight = c->image_height;
    cam.fovy = c->fovy;
    cam.eye = vec3(c->eye[0], c->eye[1], c->eye[2]);
    cam.lookAt = vec3(c->look_at[0], c->look_at[1], c->look_at[2]);
    cam.up = vec3(c->up[0], c->up[1], c->up[2]);
    if (p) {
        cam.samplesPerPixel = p->samples_per_pixel;
        cam.maxDepth = p->max_depth;
        cam.bSampleLights = p->sample_lights != 0;
        cam.russianRoulette = p->russian_roulette;
        cam.background = color(p->background[0], p->background[1], p->background[2]);
    }
}

void parallel_for(int64_t n, int nthreads, const std::function<void(int64_t, int64_t)>& body) {
    if (nthreads <= 1 || n < 2) {
        body(0, n);
        tls().flush();
        return;
    }
    std::atomic<int64_t> next{0};
    const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(4096, n / (nthreads * 8) + 1));
    std::vector<std::thread> th;
    for (int k = 0; k < nthreads; ++k)
        th.emplace_back([&]() {
            for (;;) {
                int64_t b = next.fetch_add(chunk);
                if (b >= n) break;
                body(b, std::min(n, b + chunk));
            }
        });
    for (auto& t : th) t.join();
}

void walk_leaves(const std::shared_ptr<Hittable>& node, const RefScene& s,
                 std::vector<int64_t>& out) {
    if (auto b = std::dynamic_pointer_cast<BVHNode>(node)) {
        walk_leaves(b->left, s, out);
        if (b->right != b->left) walk_leaves(b->right, s, out);  // span 1: left == right (BVH.cpp:21-24)
        return;
    }
    auto it = s.triToPrim.find(node.get());
    out.push_back(it == s.triToPrim.end() ? -1 : it->second);
}

}  // namespace

/* ------------------------------------------------------------------------------------------------
 * C interface
 * ---------------------------------------------------------------------------------------------- */
extern "C" {

int ref_aabb_empty(double* out6) {
    const AABB& e = AABB::empty;
    out6[0] = e.x.min; out6[1] = e.x.max;
    out6[2] = e.y.min; out6[3] = e.y.max;
    out6[4] = e.z.min; out6[5] = e.z.max;
    return 0;
}

/* The reference's output stage, untouched: Camera::WriteColorAttachment (Camera.cpp:279-331) on a caller-supplied
 * colorAttachment — writes <png_path> through stbi_write_png and, with write_hdr, <png_path minus extension>.hdr
 * through stbi_write_hdr. The golden bytes of tests/golden/output_stage.npz come from here. */
int ref_write_color_attachment(const double* rgb, int32_t width, int32_t height, const char* png_path, int write_hdr) {
    if (!rgb || !png_path || width < 1 || height < 1) return PRT_ERR_INVALID_ARGUMENT;
    Camera cam;
    cam.imageWidth = width;
    cam.imageHeight = height;
    cam.colorAttachment.resize((size_t)width * height);
    for (size_t i = 0; i < (size_t)width * height; ++i) cam.colorAttachment[i] = color(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
    cam.WriteColorAttachment(std::string(png_path), write_hdr != 0);
    return 0;
}

int ref_scene_create(const prt_scene_desc* d, const char* const* texture_paths, int flags,
                     void** out) {
    if (!d || !out || d->n_triangles < 0 || d->n_meshes < 0) return PRT_ERR_INVALID_ARGUMENT;
    int saved_mode = tls().mode;
    tls().mode = MODE_THREAD_LOCAL;
    auto s = std::make_unique<RefScene>();
    s->flags = flags;
    for (int i = 0; i < d->n_textures; ++i) {
        const char* p = (texture_paths && texture_paths[i]) ? texture_paths[i] : "/nonexistent/texture";
        s->textures.push_back(std::make_shared<ImageTexture>(std::string(p)));  // Texture.cpp:10-21
    }
    for (int i = 0; i < d->n_materials; ++i) s->materials.push_back(make_material(*s, d->materials[i], flags));

    s->tris.resize((size_t)d->n_triangles);
    for (int m = 0; m < d->n_meshes; ++m) {
        const prt_material& md = d->materials[d->mesh_material[m]];
        std::shared_ptr<Material> meshMat = s->materials[d->mesh_material[m]];
        std::vector<std::shared_ptr<Hittable>> list;
        for (int64_t f = d->mesh_first[m]; f < d->mesh_first[m + 1]; ++f) {
            std::array<vec3, 3> v, n;
            std::array<vec2, 3> uv;
            for (int k = 0; k < 3; ++k) {
                const double* pv = d->vertices + f * 9 + k * 3;
                v[k] = vec3(pv[0], pv[1], pv[2]);
                if (d->normals) {
                    const double* pn = d->normals + f * 9 + k * 3;
                    n[k] = vec3(pn[0], pn[1], pn[2]);
                } else n[k] = vec3(0., 0., 0.);
            }
            if (flags & ORC_UV_TAG) {
                uv[0] = vec2(0., 0.); uv[1] = vec2(1., 0.); uv[2] = vec2(0., 1.);
            } else if (d->texcoords) {
                for (int k = 0; k < 3; ++k) uv[k] = vec2(d->texcoords[f * 6 + k * 2], d->texcoords[f * 6 + k * 2 + 1]);
            } else {
                uv[0] = vec2(0., 0.); uv[1] = vec2(1., 0.); uv[2] = vec2(1., 1.);  // Model.cpp:170-175
            }
            std::shared_ptr<Material> mat = meshMat;
            if (flags & ORC_PER_TRIANGLE_MATERIAL) {
                mat = make_material(*s, md, flags);
                s->matToPrim[mat.get()] = f;
            }
            auto tri = std::make_shared<Triangle>(v, n, uv, mat);  // Triangle.cpp:11-53
            s->tris[(size_t)f] = tri;
            s->triToPrim[tri.get()] = f;
            list.push_back(tri);
        }
        s->meshes.push_back(std::make_shared<Mesh>("mesh" + std::to_string(m), list, meshMat));
    }
    // main.cpp:36-45
    for (auto& mesh : s->meshes) {
        if (mesh->objects.empty()) continue;  // an empty span would recurse forever in BVH.cpp:30-34
        s->world.Add(std::make_shared<BVHNode>(mesh));
        if (mesh->material->HasEmission()) {
            s->lights.Add(std::make_shared<BVHNode>(mesh));
            s->has_lights = true;
        }
    }
    if (!s->world.objects.empty()) s->world = HittableList(std::make_shared<BVHNode>(s->world));
    if (s->has_lights) s->lights = HittableList(std::make_shared<BVHNode>(s->lights));
    tls().mode = saved_mode;
    *out = s.release();
    return 0;
}

void ref_scene_destroy(void* p) { delete (RefScene*)p; }

int ref_trace_closest(void* p, int64_t n, const double* o, const double* d, const double* tmin,
                      const double* tmax, int32_t* prim, double* t, double* alpha, double* beta,
                      int nthreads) {
    RefScene* s = (RefScene*)p;
    if (!s) return PRT_ERR_INVALID_ARGUMENT;
    const bool empty = s->world.objects.empty();
    parallel_for(n, nthreads, [&](int64_t b, int64_t e) {
        for (int64_t i = b; i < e; ++i) {
            prim[i] = -1;
            if (t) t[i] = 0.;
            if (alpha) alpha[i] = 0.;
            if (beta) beta[i] = 0.;
            if (empty) continue;
            Ray ray(vec3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), vec3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
            HitRecord rec;
            if (s->world.Hit(ray, Interval(tmin[i], tmax[i]), rec)) {  // HittableList.h:26-39 -> BVH.cpp:51-61
                auto it = s->matToPrim.find(rec.material.get());
                prim[i] = it == s->matToPrim.end() ? -2 : (int32_t)it->second;
                if (t) t[i] = rec.time;
                if (alpha) alpha[i] = rec.uv.x;
                if (beta) beta[i] = rec.uv.y;
            }
        }
    });
    return 0;
}

int ref_reference_order(void* p, int64_t* world_rank, int64_t* light_order, int64_t* n_light) {
    RefScene* s = (RefScene*)p;
    if (!s) return PRT_ERR_INVALID_ARGUMENT;
    if (world_rank) {
        std::vector<int64_t> leaves;
        if (!s->world.objects.empty()) walk_leaves(s->world.objects[0], *s, leaves);
        for (size_t i = 0; i < s->tris.size(); ++i) world_rank[i] = -1;
        for (size_t r = 0; r < leaves.size(); ++r)
            if (leaves[r] >= 0) world_rank[leaves[r]] = (int64_t)r;
    }
    std::vector<int64_t> ll;
    if (s->has_lights) walk_leaves(s->lights.objects[0], *s, ll);
    if (n_light) *n_light = (int64_t)ll.size();
    if (light_order) std::copy(ll.begin(), ll.end(), light_order);
    return 0;
}

int ref_sample_lights(void* p, int64_t n, const double* origins, const double* u, double* position,
                      double* normal, uint8_t* front, double* pdf, int32_t* prim) {
    RefScene* s = (RefScene*)p;
    if (!s || !s->has_lights) return PRT_ERR_INVALID_ARGUMENT;
    ThreadState& t = tls();
    int saved = t.mode;
    t.mode = MODE_SCRIPT;
    for (int64_t i = 0; i < n; ++i) {
        int32_t script[4];
        for (int k = 0; k < 4; ++k) script[k] = (int32_t)(u[4 * i + k] * 2147483648.0);
        t.script = script; t.script_n = 4; t.script_i = 0;
        HitRecord rec;
        double q = 0.;
        s->lights.Sample(point3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), rec, q);  // HittableList.h:44-59
        for (int k = 0; k < 3; ++k) { position[3 * i + k] = rec.position[k]; normal[3 * i + k] = rec.normal[k]; }
        front[i] = rec.bFrontFace ? 1 : 0;
        pdf[i] = q;
        auto it = s->matToPrim.find(rec.material.get());
        prim[i] = it == s->matToPrim.end() ? -2 : (int32_t)it->second;
    }
    t.script = nullptr; t.script_n = 0;
    t.mode = saved;
    return 0;
}

int ref_render(void* p, const prt_camera* c, const prt_render_params* prm, int rng_mode,
               int nthreads, double* rgb, orc_render_info* info) {
    RefScene* s = (RefScene*)p;
    if (!s || !c || !prm || !rgb) return PRT_ERR_INVALID_ARGUMENT;
    if (prm->sample_lights && !s->has_lights) return PRT_ERR_INVALID_ARGUMENT;
    const int W = c->image_width, H = c->image_height;
    int sb = prm->spp_begin, se = prm->spp_end;
    if (sb == 0 && se == 0) se = prm->samples_per_pixel;
    if (nthreads < 1) nthreads = 1;
    tls().flush();
    g_hit_calls = 0; g_shadow_calls = 0;
    CountingWorld world(s->world);
    auto t0 = std::chrono::steady_clock::now();
    int used_threads = nthreads;

    if (rng_mode == ORC_RNG_COUNTER) {
        Camera cam;
        set_camera(cam, c, prm);
        cam.Initialize();  // Camera.cpp:75-106
        const double scale = 1.0 / prm->samples_per_pixel;
        std::fill(rgb, rgb + (size_t)W * H * 3, 0.0);
        parallel_for(H, nthreads, [&](int64_t jb, int64_t je) {
            ThreadState& t = tls();
            int saved = t.mode;
            t.mode = MODE_COUNTER;
            for (int64_t j = jb; j < je; ++j)
                for (int i = 0; i < W; ++i) {
                    Ray ray = cam.GetRay(i, (int)j);  // Camera.cpp:108-117
                    const uint32_t pix = (uint32_t)(j * W + i);
                    color acc(0., 0., 0.);
                    for (int smp = sb; smp < se; ++smp) {
                        t.seed = orc_stream_seed(prm->rng_key, pix, (uint32_t)smp);
                        t.dim = 0;
                        acc += cam.RayColor(ray, cam.maxDepth, world, s->lights) * scale;  // Camera.cpp:56
                    }
                    rgb[3 * (size_t)pix] = acc.x; rgb[3 * (size_t)pix + 1] = acc.y; rgb[3 * (size_t)pix + 2] = acc.z;
                }
            t.mode = saved;
            t.flush();
        });
    } else {
        // The reference's own Camera::Render with its own threads (Camera.cpp:21-73).
        Camera cam;
        set_camera(cam, c, prm);
        cam.samplesPerPixel = se - sb;
        while (used_threads > 1 && (H % used_threads) != 0) --used_threads;  // Camera.cpp:47: rows lost otherwise
        cam.threadNums = used_threads;
        g_seed_base = orc_mix64(prm->rng_key + 0x5bd1e995ull);
        g_default_mode = (rng_mode == ORC_RNG_LIBC) ? MODE_LIBC : MODE_THREAD_LOCAL;
        cam.Render(world, s->lights);
        g_default_mode = MODE_THREAD_LOCAL;
        const double k = (double)(se - sb) / (double)prm->samples_per_pixel;
        for (size_t i = 0; i < (size_t)W * H; ++i) {
            rgb[3 * i] = cam.colorAttachment[i].x * k;
            rgb[3 * i + 1] = cam.colorAttachment[i].y * k;
            rgb[3 * i + 2] = cam.colorAttachment[i].z * k;
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    tls().flush();
    if (info) {
        info->world_hit_calls = g_hit_calls.load();
        info->shadow_calls = g_shadow_calls.load();
        info->paths = (uint64_t)W * H * (uint64_t)(se - sb);
        info->seconds = std::chrono::duration<double>(t1 - t0).count();
        info->threads = used_threads;
        info->reserved = 0;
    }
    return 0;
}

int ref_pixel_samples(void* p, const prt_camera* c, const prt_render_params* prm, int32_t px,
                      int32_t py, double* out) {
    RefScene* s = (RefScene*)p;
    if (!s || !c || !prm || !out) return PRT_ERR_INVALID_ARGUMENT;
    int sb = prm->spp_begin, se = prm->spp_end;
    if (sb == 0 && se == 0) se = prm->samples_per_pixel;
    Camera cam;
    set_camera(cam, c, prm);
    cam.Initialize();
    ThreadState& t = tls();
    int saved = t.mode;
    t.mode = MODE_COUNTER;
    Ray ray = cam.GetRay(px, py);
    const uint32_t pix = (uint32_t)(py * c->image_width + px);
    for (int smp = sb; smp < se; ++smp) {
        t.seed = orc_stream_seed(prm->rng_key, pix, (uint32_t)smp);
        t.dim = 0;
        color L = cam.RayColor(ray, cam.maxDepth, s->world, s->lights);
        out[3 * (smp - sb)] = L.x; out[3 * (smp - sb) + 1] = L.y; out[3 * (smp - sb) + 2] = L.z;
    }
    t.mode = saved;
    return 0;
}

int ref_triangle_record(void* p, int64_t prim, double* o) {
    RefScene* s = (RefScene*)p;
    if (!s || prim < 0 || (size_t)prim >= s->tris.size()) return PRT_ERR_INVALID_ARGUMENT;
    const Triangle& T = *s->tris[(size_t)prim];
    int k = 0;
    for (int i = 0; i < 3; ++i) o[k++] = T.normal[i];
    o[k++] = T.D;
    for (int i = 0; i < 3; ++i) o[k++] = T.w[i];
    for (int i = 0; i < 3; ++i) o[k++] = T.vertices[0][i];
    for (int i = 0; i < 3; ++i) o[k++] = T.edges[0][i];
    for (int i = 0; i < 3; ++i) o[k++] = T.edges[1][i];
    for (int i = 0; i < 3; ++i) o[k++] = T.tangent[i];
    o[k++] = T.area;
    return 0;
}

int ref_material_scatter(void* p, int32_t material, const double* ray_dir, const double* normal,
                         const double* tangent, const double* uv, const double* u, int32_t n_u,
                         double* attenuation, double* out_dir, int32_t* n_draws,
                         int32_t* scattered) {
    RefScene* s = (RefScene*)p;
    if (!s || material < 0 || (size_t)material >= s->materials.size()) return PRT_ERR_INVALID_ARGUMENT;
    std::vector<int32_t> script((size_t)n_u);
    for (int i = 0; i < n_u; ++i) script[(size_t)i] = (int32_t)(u[i] * 2147483648.0);
    ThreadState& t = tls();
    int saved = t.mode;
    t.mode = MODE_SCRIPT; t.script = script.data(); t.script_n = n_u; t.script_i = 0;
    HitRecord rec;
    rec.position = vec3(0., 0., 0.);
    rec.time = 1.0;
    rec.tangent = vec3(tangent[0], tangent[1], tangent[2]);
    rec.uv = vec2(uv[0], uv[1]);
    rec.material = s->materials[(size_t)material];
    Ray in(vec3(-ray_dir[0], -ray_dir[1], -ray_dir[2]), vec3(ray_dir[0], ray_dir[1], ray_dir[2]));
    rec.SetFaceNormal(in, vec3(normal[0], normal[1], normal[2]));  // Hittable.cpp:8-13
    const double nan = std::numeric_limits<double>::quiet_NaN();
    color att(nan, nan, nan);
    Ray sc(vec3(0.), vec3(0.));
    bool ok = rec.material->Scatter(in, rec, att, sc);
    for (int i = 0; i < 3; ++i) { attenuation[i] = att[i]; out_dir[i] = sc.direction[i]; }
    *n_draws = t.script_i;
    *scattered = ok ? 1 : 0;
    t.script = nullptr; t.script_n = 0; t.mode = saved;
    return 0;
}

int ref_material_eval(void* p, int32_t material, const double* ray_dir, const double* normal,
                      const double* tangent, const double* uv, const double* wi_world, const double* u, int32_t n_u,
                      double* f) {
    RefScene* s = (RefScene*)p;
    if (!s || material < 0 || (size_t)material >= s->materials.size()) return PRT_ERR_INVALID_ARGUMENT;
    std::vector<int32_t> script((size_t)n_u);
    for (int i = 0; i < n_u; ++i) script[(size_t)i] = (int32_t)(u[i] * 2147483648.0);
    ThreadState& t = tls();
    int saved = t.mode;
    t.mode = MODE_SCRIPT; t.script = script.data(); t.script_n = n_u; t.script_i = 0;
    HitRecord rec;
    rec.position = vec3(0., 0., 0.);
    rec.time = 1.0;
    rec.tangent = vec3(tangent[0], tangent[1], tangent[2]);
    rec.uv = vec2(uv[0], uv[1]);
    rec.material = s->materials[(size_t)material];
    Ray in(vec3(-ray_dir[0], -ray_dir[1], -ray_dir[2]), vec3(ray_dir[0], ray_dir[1], ray_dir[2]));
    rec.SetFaceNormal(in, vec3(normal[0], normal[1], normal[2]));  // Hittable.cpp:8-13
    MaterialEvalContext context;                                  // Camera.cpp:160-165
    context.p = rec.position;
    context.uv = rec.uv;
    context.n = rec.normal;
    context.dpdus = rec.tangent;
    context.wo = Material::WorldToLocal(-in.direction, rec);
    const vec3 localWi = Material::WorldToLocal(vec3(wi_world[0], wi_world[1], wi_world[2]), rec);
    vec3 fr = rec.material->Eval(localWi, context);
    for (int i = 0; i < 3; ++i) f[i] = fr[i];
    t.script = nullptr; t.script_n = 0; t.mode = saved;
    return 0;
}

int ref_texture_value(void* p, int32_t texture, int64_t n, const double* uv, double* rgb) {
    RefScene* s = (RefScene*)p;
    if (!s || texture < 0 || (size_t)texture >= s->textures.size()) return PRT_ERR_INVALID_ARGUMENT;
    for (int64_t i = 0; i < n; ++i) {
        color c = s->textures[(size_t)texture]->Value(uv[2 * i], uv[2 * i + 1], point3(0.));  // Texture.cpp:22-49
        rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    }
    return 0;
}

int ref_camera_ray(const prt_camera* c, int32_t i, int32_t j, double* origin, double* dir) {
    Camera cam;
    set_camera(cam, c, nullptr);
    cam.Initialize();
    Ray r = cam.GetRay(i, j);
    for (int k = 0; k < 3; ++k) { origin[k] = r.origin[k]; dir[k] = r.direction[k]; }
    return 0;
}

}  // extern "C"

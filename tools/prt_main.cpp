/*
 * prt_main — drop-in for the reference's main.cpp (/root/reference/main.cpp:6-55) on top of the C ABI:
 * loads <dir>/<name>.{obj,mtl,xml}, renders on one B200, writes <out>.png and <out>.hdr.
 *
 *   prt_main <scene-dir> <name> [--spp N] [--depth N] [--rr P] [--no-nee] [--out PATH] [--device K]
 *            [--strtod] [--seed KEY] [--chunk SPP]
 *
 * Defaults are main.cpp's: spp 100, depth 100, Russian roulette 0.8, light sampling on, black
 * background. The image is rendered in chunks of --chunk samples per pixel (default 16) that are
 * accumulated on the host, like successive Camera::Render calls into one colorAttachment
 * (Source/Camera.cpp:81).
 */
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/prt_b200.h"

#define CHECK(call)                                                              \
    do {                                                                         \
        int rc_ = (call);                                                        \
        if (rc_ != PRT_OK) {                                                     \
            std::fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, prt_last_error()); \
            return 1;                                                            \
        }                                                                        \
    } while (0)

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: %s <scene-dir> <name> [--spp N] [--depth N] [--rr P] [--no-nee] [--out PATH] [--device K] [--strtod] [--seed KEY] [--chunk SPP]\n", argv[0]);
        return 2;
    }
    const std::string dir = argv[1], name = argv[2];
    int spp = 100, depth = 100, device = 0, chunk = 16, nee = 1;
    double rr = 0.8;
    unsigned long long seed = 0;
    uint32_t flags = 0;
    std::string out = name;
    for (int i = 3; i < argc; ++i) {
        auto next = [&](const char* what) -> const char* {
            if (i + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", what); std::exit(2); }
            return argv[++i];
        };
        if (!std::strcmp(argv[i], "--spp")) spp = std::atoi(next("--spp"));
        else if (!std::strcmp(argv[i], "--depth")) depth = std::atoi(next("--depth"));
        else if (!std::strcmp(argv[i], "--rr")) rr = std::atof(next("--rr"));
        else if (!std::strcmp(argv[i], "--no-nee")) nee = 0;
        else if (!std::strcmp(argv[i], "--out")) out = next("--out");
        else if (!std::strcmp(argv[i], "--device")) device = std::atoi(next("--device"));
        else if (!std::strcmp(argv[i], "--strtod")) flags |= PRT_LOAD_STRTOD;
        else if (!std::strcmp(argv[i], "--seed")) seed = std::strtoull(next("--seed"), nullptr, 10);
        else if (!std::strcmp(argv[i], "--chunk")) chunk = std::atoi(next("--chunk"));
        else { std::fprintf(stderr, "unknown option %s\n", argv[i]); return 2; }
    }
    if (spp < 1 || chunk < 1) { std::fprintf(stderr, "--spp and --chunk must be positive\n"); return 2; }

    prt_model* model = nullptr;
    CHECK(prt_model_load(dir.c_str(), name.c_str(), flags, &model));
    if (*prt_model_warnings(model)) std::fprintf(stderr, "%s", prt_model_warnings(model));
    const prt_scene_desc* desc = prt_model_desc(model);
    std::printf("Shapes/Meshes Nums: %d\nMaterials Nums: %d\nTriangles: %lld\n", desc->n_meshes, desc->n_materials, (long long)desc->n_triangles);

    prt_camera cam;
    CHECK(prt_model_camera(model, &cam));
    prt_context* ctx = nullptr;
    CHECK(prt_context_create(device, &ctx));
    prt_scene* scene = nullptr;
    auto t0 = std::chrono::steady_clock::now();
    CHECK(prt_scene_create(ctx, desc, &scene));
    prt_build_opts opts;
    std::memset(&opts, 0, sizeof(opts));
    opts.accel = PRT_ACCEL_SAH_BVH2;
    CHECK(prt_scene_commit(scene, &opts));
    auto t1 = std::chrono::steady_clock::now();
    std::printf("Building BVH End... %.3f s\n", std::chrono::duration<double>(t1 - t0).count());

    const size_t n = (size_t)cam.image_width * cam.image_height * 3;
    std::vector<double> image(n, 0.0), part(n);
    prt_render_params prm;
    std::memset(&prm, 0, sizeof(prm));
    prm.samples_per_pixel = spp;
    prm.max_depth = depth;
    prm.sample_lights = nee;
    prm.russian_roulette = rr;
    prm.rng_key = seed;
    unsigned long long rays = 0;
    for (int b = 0; b < spp; b += chunk) {
        prm.spp_begin = b;
        prm.spp_end = b + chunk < spp ? b + chunk : spp;
        CHECK(prt_render(scene, &cam, &prm, part.data()));
        for (size_t i = 0; i < n; ++i) image[i] += part[i];
        prt_stats st;
        CHECK(prt_get_stats(scene, &st));
        rays += st.extend_rays + st.shadow_rays;
    }
    auto t2 = std::chrono::steady_clock::now();
    const double sec = std::chrono::duration<double>(t2 - t1).count();
    std::printf("Rendered %dx%d spp%d-depth%d in %.3f s (%.1f spp/s, %.1f Mrays/s)\n", cam.image_width, cam.image_height, spp, depth, sec,
                spp / sec, rays / sec / 1e6);

    std::vector<uint8_t> rgb8(n);
    std::vector<float> rgb32f(n);
    CHECK(prt_resolve(scene, image.data(), cam.image_width, cam.image_height, rgb8.data(), rgb32f.data()));
    CHECK(prt_write_png((out + ".png").c_str(), cam.image_width, cam.image_height, rgb8.data()));
    CHECK(prt_write_hdr((out + ".hdr").c_str(), cam.image_width, cam.image_height, rgb32f.data()));
    std::printf("Wrote %s.png and %s.hdr\n", out.c_str(), out.c_str());
    prt_scene_destroy(scene);
    prt_context_destroy(ctx);
    prt_model_destroy(model);
    return 0;
}

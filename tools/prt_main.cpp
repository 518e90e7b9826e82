/*
 * prt_main — drop-in for the reference's main.cpp (/root/reference/main.cpp:6-55) on top of the C ABI:
 * loads <dir>/<name>.{obj,mtl,xml}, renders on 1..8 B200s of the box, writes <out>.png and <out>.hdr.
 *
 *   prt_main <scene-dir> <name> [--spp N] [--depth N] [--rr P] [--no-nee] [--out PATH] [--gpus N | --devices a,b,..]
 *            [--device K] [--strtod] [--seed KEY] [--accel auto|bvh2|bvh4|bvh4q|lbvh]
 *
 * Defaults are main.cpp's: spp 100, depth 100, Russian roulette 0.8, light sampling on, black background. The scene is
 * replicated on every GPU, the samples are split across them, one reduction (peer loads over NVLink, or NCCL) brings the frame to the first GPU
 * (prt_multi_render), where the output stage of Camera::WriteColorAttachment runs (prt_multi_resolve). The library
 * cuts a GPU's sample range into wavefront passes itself.
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

static double seconds_since(std::chrono::steady_clock::time_point t) {
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t).count();
}

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: %s <scene-dir> <name> [--spp N] [--depth N] [--rr P] [--no-nee] [--out PATH] [--gpus N | --devices a,b,..] [--device K] [--strtod] [--seed KEY] [--accel auto|bvh2|bvh4|bvh4q|lbvh] [--clean-exit]\n", argv[0]);
        return 2;
    }
    const std::string dir = argv[1], name = argv[2];
    int spp = 100, depth = 100, nee = 1, accel = PRT_ACCEL_AUTO;
    double rr = 0.8;
    unsigned long long seed = 0;
    uint32_t flags = 0;
    bool clean_exit = false;
    std::string out = name;
    std::vector<int32_t> devices = {0};
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
        else if (!std::strcmp(argv[i], "--device")) devices.assign(1, std::atoi(next("--device")));
        else if (!std::strcmp(argv[i], "--gpus")) {
            const int n = std::atoi(next("--gpus"));
            devices.clear();
            for (int g = 0; g < n; ++g) devices.push_back(g);
        } else if (!std::strcmp(argv[i], "--devices")) {
            devices.clear();
            for (const char* p = next("--devices"); *p;) {
                devices.push_back((int32_t)std::strtol(p, const_cast<char**>(&p), 10));
                if (*p == ',') ++p;
            }
        } else if (!std::strcmp(argv[i], "--strtod")) flags |= PRT_LOAD_STRTOD;
        else if (!std::strcmp(argv[i], "--seed")) seed = std::strtoull(next("--seed"), nullptr, 10);
        else if (!std::strcmp(argv[i], "--clean-exit")) clean_exit = true; /* destroy every handle before returning (leak checkers) */
        else if (!std::strcmp(argv[i], "--chunk")) (void)next("--chunk"); /* accepted for old command lines: passes are cut by the library */
        else if (!std::strcmp(argv[i], "--accel")) {
            const std::string a = next("--accel");
            accel = a == "bvh2" ? PRT_ACCEL_SAH_BVH2 : a == "bvh4q" ? PRT_ACCEL_SAH_BVH4Q : a == "lbvh" ? PRT_ACCEL_LBVH4_DEVICE : a == "bvh4" ? PRT_ACCEL_SAH_BVH4 : PRT_ACCEL_AUTO;
        } else { std::fprintf(stderr, "unknown option %s\n", argv[i]); return 2; }
    }
    if (spp < 1 || devices.empty() || devices.size() > 8) { std::fprintf(stderr, "--spp must be positive, 1..8 GPUs\n"); return 2; }

    const auto t_start = std::chrono::steady_clock::now();
    prt_model* model = nullptr;
    CHECK(prt_model_load(dir.c_str(), name.c_str(), flags, &model));
    if (*prt_model_warnings(model)) std::fprintf(stderr, "%s", prt_model_warnings(model));
    const prt_scene_desc* desc = prt_model_desc(model);
    std::printf("Shapes/Meshes Nums: %d\nMaterials Nums: %d\nTriangles: %lld\n", desc->n_meshes, desc->n_materials, (long long)desc->n_triangles);
    const double t_load = seconds_since(t_start);

    prt_camera cam;
    CHECK(prt_model_camera(model, &cam));
    auto t0 = std::chrono::steady_clock::now();
    prt_build_opts opts;
    std::memset(&opts, 0, sizeof(opts));
    opts.accel = accel;
    prt_multi* box = nullptr;
    CHECK(prt_multi_create(devices.data(), (int32_t)devices.size(), desc, &opts, &box));
    const double t_build = seconds_since(t0);
    std::printf("Building BVH End... %.3f s (%d GPU%s)\n", t_build, (int)devices.size(), devices.size() > 1 ? "s" : "");

    const size_t n = (size_t)cam.image_width * cam.image_height * 3;
    prt_render_params prm;
    std::memset(&prm, 0, sizeof(prm));
    prm.samples_per_pixel = spp;
    prm.max_depth = depth;
    prm.sample_lights = nee;
    prm.russian_roulette = rr;
    prm.rng_key = seed;
    t0 = std::chrono::steady_clock::now();
    CHECK(prt_multi_render(box, &cam, &prm, nullptr)); /* the frame stays on the first GPU for the output stage */
    const double sec = seconds_since(t0);
    prt_stats st;
    double reduce_ms = 0.0;
    CHECK(prt_multi_get_stats(box, &st, &reduce_ms));
    std::printf("Rendered %dx%d spp%d-depth%d in %.3f s (%.1f spp/s, %.1f Mrays/s; reduce %.3f ms)\n", cam.image_width, cam.image_height, spp, depth, sec,
                spp / sec, (double)(st.extend_rays + st.shadow_rays) / sec / 1e6, reduce_ms);
    if (st.ref_stack_overflows) std::fprintf(stderr, "warning: %llu subtrees dropped by the reference-order walk (stack overflow)\n", (unsigned long long)st.ref_stack_overflows);

    t0 = std::chrono::steady_clock::now();
    std::vector<uint8_t> rgb8(n);
    std::vector<float> rgb32f(n);
    CHECK(prt_multi_resolve(box, rgb8.data(), rgb32f.data()));
    CHECK(prt_write_png((out + ".png").c_str(), cam.image_width, cam.image_height, rgb8.data()));
    CHECK(prt_write_hdr((out + ".hdr").c_str(), cam.image_width, cam.image_height, rgb32f.data()));
    std::printf("Wrote %s.png and %s.hdr\n", out.c_str(), out.c_str());
    std::printf("job: load %.3f s, build+upload %.3f s, render+reduce %.3f s, resolve+write %.3f s, total %.3f s\n", t_load, t_build, sec, seconds_since(t0),
                seconds_since(t_start));
    if (!clean_exit) {
        /* both files are written and closed: leave without tearing the CUDA contexts down one by one, which on an
         * 8-GPU box takes longer (1.7 s) than the render itself (0.5 s for 3000 spp of cornell-box) */
        std::fflush(stdout);
        std::fflush(stderr);
        std::_Exit(0);
    }
    prt_multi_destroy(box);
    prt_model_destroy(model);
    return 0;
}

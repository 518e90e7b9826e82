/*
 * shading.cuh — device restatement of the reference's shading layer (FP64, reference operation
 * order, no FMA). Paths relative to /root/reference/:
 *
 *   texture_value            ImageTexture::Value / GetPixel      Source/Texture.cpp:22-65
 *   sample_lights            HittableList::Sample                Source/HittableList.h:44-59
 *                            BVHNode::Sample / TraverseSample    Source/BVH.cpp:62-67,86-100
 *                            Triangle::Sample                    Source/Triangle.cpp:84-93
 *   sample_cosine_hemisphere SampleCosineHemisphere /            Source/RandomNumberGenerator.h:39-64
 *                            SampleUniformDiskConcentric
 *   world_to_local / local_to_world / reflect                    Source/Material.h:76-98
 *   material_eval            Lambertian::Eval, PhoneReflectance::Eval   Material.h:128-130,227-248
 *   scatter_*                Lambertian / PhoneReflectance / PerfectMirror ::Scatter + ::Sample
 *                                                                Material.h:106-151,183-226,262-312,334-363
 *
 * Defined meaning of the reference's undefined spot: Phong's unset `attenuation`
 * (Camera.cpp:177, Material.h:280-284) is 0. The random number drawn inside PhoneReflectance::Eval (Material.h:228)
 * is NOT undefined — it is an ordinary draw — but the reference reaches Eval only when the light sample was VISIBLE
 * (Camera.cpp:152-155), so the position of every later draw of the path would depend on the shadow ray's outcome,
 * which a wavefront learns only after it has scattered. Its value never matters where Eval is reachable (Ns <= 1 =>
 * pkd = 1), so here — and in both CPU checkers, identically — the draw is peeked and the stream is not advanced:
 * the estimator is the reference's, the stream of Phong paths is shifted by one against the unmodified reference.
 */
#ifndef PRT_SHADING_CUH
#define PRT_SHADING_CUH

#include "device_scene.h"
#include "prt_math.h"
#include "rng.cuh"

namespace prt {

struct Surface { /* the HitRecord fields shading reads (Source/Hittable.h:17-28) */
    d3 position, normal, tangent;
    d2 uv;
    bool front;
};

/* ---- textures ---------------------------------------------------------------------------------- */
__device__ __forceinline__ d3 texture_pixel(const DevScene& S, const DevTexture& T, int x, int y) {
    const uint8_t* px = S.texels + T.offset + ((int64_t)y * T.width + x) * T.channels;
    if (T.channels >= 3) return mk3(S.srgb_lut[px[0]], S.srgb_lut[px[1]], S.srgb_lut[px[2]]);
    const double g = (1.0 / 255.0) * (double)px[0]; /* grey: no sRGB decode (Texture.cpp:62-64) */
    return mk3(g, g, g);
}
__device__ __forceinline__ d3 texture_value(const DevScene& S, int tex, double u, double v) {
    const DevTexture T = S.textures[tex];
    if (!T.valid) return mk3(0., 1., 1.); /* failed load: cyan (Texture.cpp:24) */
    u = u < 0.0 ? 0.0 : (u > 1.0 ? 1.0 : u); /* std::clamp */
    v = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);
    const double x = u * (T.width - 1.);
    const double y = (1. - v) * (T.height - 1.);
    const int x0 = (int)x, y0 = (int)y;
    const int x1 = min(x0 + 1, T.width - 1), y1 = min(y0 + 1, T.height - 1);
    const double tx = x - x0, ty = y - y0;
    const d3 c00 = texture_pixel(S, T, x0, y0), c10 = texture_pixel(S, T, x1, y0);
    const d3 c01 = texture_pixel(S, T, x0, y1), c11 = texture_pixel(S, T, x1, y1);
    const d3 c0 = c00 * (1 - tx) + c10 * tx;
    const d3 c1 = c01 * (1 - tx) + c11 * tx;
    return c0 * (1 - ty) + c1 * ty;
}
__device__ __forceinline__ d3 tex_or_color(const DevScene& S, int tex, const double* c, d2 uv) {
    if (tex >= 0) return texture_value(S, tex, uv.x, uv.y);
    return mk3(c[0], c[1], c[2]);
}

/* ---- light sampling ---------------------------------------------------------------------------- */
struct LightSample {
    d3 position, normal;
    bool front;
    int32_t material, prim;
    double pdf;
};
/* u0: HittableList.h:49 (picks the only object), u1: BVH.cpp:64, u2 u3: Triangle.cpp:86 */
/* unbiased_pick (PRT_RENDER_UNBIASED_LIGHT_PICK, never the parity default): p = u1 * area instead of sqrt(u1) * area */
__device__ __forceinline__ void sample_lights(const DevScene& S, d3 origin, double u0, double u1, double u2, double u3, LightSample& out,
                                              bool unbiased_pick = false) {
    const double area_sum = S.lights_area;
    const double p0 = u0 * area_sum;
    (void)p0; /* p0 <= areaSum always: the single top BVHNode is chosen */
    const double p = (unbiased_pick ? u1 : sqrt(u1)) * area_sum;
    float pf = (float)p; /* TraverseSample takes `float p` (BVH.h:32) */
    int32_t ref = S.light_root;
    while (ref >= 0) {
        const LightNode n = S.light_nodes[ref];
        if ((double)pf < n.left_area) ref = n.left;
        else { pf = (float)((double)pf - n.left_area); ref = n.right; }
    }
    const LightTri* __restrict__ L = S.light_tris + (~ref);
    const double x = sqrt(u2), y = u3;
    const d3 v0 = mk3(L->v0[0], L->v0[1], L->v0[2]), v1 = mk3(L->v1[0], L->v1[1], L->v1[2]), v2 = mk3(L->v2[0], L->v2[1], L->v2[2]);
    const d3 pnt = v0 * (1.0 - x) + v1 * (x * (1.0 - y)) + v2 * (x * y);
    const d3 n = mk3(L->normal[0], L->normal[1], L->normal[2]);
    const d3 direction = pnt - origin;
    out.position = pnt;
    out.front = dot(direction, n) < 0.; /* SetFaceNormal, Hittable.cpp:8-13 */
    out.normal = out.front ? n : -n;
    out.material = L->material;
    out.prim = L->prim;
    double pdf = 1.0 / L->area; /* Triangle.cpp:92 */
    pdf *= L->area;             /* BVH.cpp:91 */
    pdf /= area_sum;            /* BVH.cpp:66 */
    out.pdf = pdf;
}

/* ---- samplers ---------------------------------------------------------------------------------- */
#ifndef PRT_SINCOS
#define PRT_SINCOS 1 /* same bits as sin() + cos(), one range reduction: veach-mis shade 15.3 -> 14.9 ms (profiles/r1_s8_ab.txt) */
#endif
__device__ __forceinline__ d2 sample_disk_concentric(d2 u) {
    const d2 uo = mk2(2. * u.x - 1., 2. * u.y - 1.);
    if (uo.x == 0. && uo.y == 0.) return mk2(0., 0.);
    double theta, r;
    if (fabs(uo.x) > fabs(uo.y)) {
        r = uo.x;
        theta = PRT_PI_OVER_4 * (uo.y / uo.x);
    } else {
        r = uo.y;
        theta = PRT_PI_OVER_2 - PRT_PI_OVER_4 * (uo.x / uo.y);
    }
#if PRT_SINCOS
    double sn, cs; /* one range reduction for both; same values as sin() and cos() */
    sincos(theta, &sn, &cs);
    return mk2(r * cs, r * sn);
#else
    return mk2(r * cos(theta), r * sin(theta));
#endif
}
__device__ __forceinline__ d3 sample_cosine_hemisphere(Rng& rng) {
    /* `glm::dvec2(RandomDouble(), RandomDouble())`: g++ evaluates the constructor arguments right
     * to left, so the first number of the stream lands in u.y (pinned against the compiled
     * reference by tests/test_oracle_pinning.py) */
    const double second_arg = rng.next();
    const double first_arg = rng.next();
    const d2 dd = sample_disk_concentric(mk2(first_arg, second_arg));
    const double z = sqrt(fmax(0.0, 1. - dd.x * dd.x - dd.y * dd.y));
    return mk3(dd.x, dd.y, z);
}
__device__ __forceinline__ d3 sample_cosine_hemisphere_upper(Rng& rng) { /* Material.h:111-116 */
    d3 wi = sample_cosine_hemisphere(rng);
    while (wi.z <= 0.) wi = sample_cosine_hemisphere(rng);
    return wi;
}

/* ---- frames ------------------------------------------------------------------------------------ */
__device__ __forceinline__ d3 local_to_world(d3 local, const Surface& s) {
    const d3 bitangent = cross(s.tangent, s.normal);
    return normalize(local.x * s.tangent + local.y * bitangent + local.z * s.normal);
}
__device__ __forceinline__ d3 world_to_local(d3 world, const Surface& s) {
    const d3 bitangent = cross(s.tangent, s.normal);
    return mk3(dot(world, s.tangent), dot(world, bitangent), dot(world, s.normal));
}
__device__ __forceinline__ d3 reflect(d3 wo, d3 n) { return -wo + 2. * dot(wo, n) * n; }

/* ---- BSDFs ------------------------------------------------------------------------------------- */
/* Material::Eval as the NEE estimator calls it (Camera.cpp:168); wi, wo in the local frame */
__device__ __forceinline__ d3 cook_eval(const DevMaterial& m, d3 wi, d3 wo); /* cook_torrance.cuh */
/* COOK / PHONG / PHONG_SPEC: which of the branches the calling kernel's material class can reach at all (launch.h: a class
 * holds one material type; CLS_PHONG holds the Phong materials with pks == 0, whose specular branch is unreachable because
 * u < pkd = 1 always). The other branches are compiled out: same arithmetic, a third of the code. */
template <bool COOK = true, bool PHONG = true, bool PHONG_SPEC = true>
__device__ __forceinline__ d3 material_eval(const DevScene& S, const DevMaterial& m, d3 wi, d3 wo, d2 uv, const Rng& rng) {
    if (COOK && m.type == PRT_MAT_COOK_TORRANCE) return cook_eval(m, wi, wo);
    if (m.type == PRT_MAT_LAMBERTIAN) return tex_or_color(S, m.kd_tex, m.kd, uv) * PRT_INV_PI;
    if (PHONG && m.type == PRT_MAT_PHONG) {
        const double u = rng.peek(); /* Material.h:228: the value is looked at, the stream position is not moved (header comment) */
        if (u < m.pkd) {
            if (wi.z <= 0) return mk3(0, 0, 0);
            return tex_or_color(S, m.kd_tex, m.kd, uv) * PRT_INV_PI;
        } else if (PHONG_SPEC && m.pkd <= u && u < m.pkd + m.pks) {
            if (wi.z <= 0) return mk3(0, 0, 0);
            const d3 local_reflect = normalize(reflect(wo, mk3(0., 0., 1.)));
            const double ca = fmax(0., dot(wi, local_reflect));
            if (ca <= 0.) return mk3(0, 0, 0);
            /* a textured Phong aliases Ks to the Kd texture (Material.h:178-181) */
            return tex_or_color(S, m.kd_tex, m.ks, uv) * (m.ns + 2.) * PRT_INV_2PI * pow(ca, m.ns);
        }
    }
    return mk3(0, 0, 0);
}

/* cosine_pdf (both scatter functions): the pdf of the sampled direction where it came from the cosine lobe, else 0
 * (PRT_RENDER_MIS) */
__device__ __forceinline__ void scatter_lambertian(const DevScene& S, const DevMaterial& m, const Surface& s, Rng& rng, d3& att, d3& out_dir,
                                                   double& cosine_pdf) {
    const d3 wi = sample_cosine_hemisphere_upper(rng);
    const double pdf = wi.z * PRT_INV_PI;
    cosine_pdf = pdf;
    const d3 f = tex_or_color(S, m.kd_tex, m.kd, s.uv) * PRT_INV_PI;
    out_dir = local_to_world(wi, s);
    att = f * wi.z / pdf;
}

/* lobe_probability (PRT_RENDER_LOBE_PROBABILITY, never the parity default): divide by the probability of the sampled lobe.
 * SPEC = false: the material has pks == 0 (CLS_PHONG), the specular branch cannot be taken and is compiled out. */
template <bool SPEC = true>
__device__ __forceinline__ void scatter_phong(const DevScene& S, const DevMaterial& m, const Surface& s, d3 wo, Rng& rng, d3& att, d3& out_dir,
                                              double& cosine_pdf, bool lobe_probability = false) {
    d3 wi = mk3(0, 0, 0), f = mk3(0, 0, 0);
    double pdf = 0.0;
    const double u = rng.next();
    if (u < m.pkd) {
        wi = sample_cosine_hemisphere_upper(rng);
        pdf = wi.z * PRT_INV_PI;
        f = tex_or_color(S, m.kd_tex, m.kd, s.uv) * PRT_INV_PI;
    } else if (SPEC && m.pkd <= u && u < m.pkd + m.pks) {
        const double u1 = rng.next(), u2 = rng.next();
        const double alpha = acos(pow(u1, 1.0 / (m.ns + 1.0)));
        const double phi = 2.0 * PRT_PI * u2;
#if PRT_SINCOS
        double sa, ca, sp, cp;
        sincos(alpha, &sa, &ca);
        sincos(phi, &sp, &cp);
#else
        const double sa = sin(alpha), ca = cos(alpha), sp = sin(phi), cp = cos(phi);
#endif
        const d3 rw = mk3(sa * cp, sa * sp, ca);
        /* ReflectiveSpaceToLocal, Material.h:302-312 */
        const d3 local_r = normalize(reflect(wo, mk3(0., 0., 1.)));
        const d3 V = (fabs(local_r.x) > 0.9 ? mk3(0., 1., 0.) : mk3(1., 0., 0.));
        const d3 T = normalize(cross(V, local_r));
        const d3 B = cross(local_r, T);
        wi = rw.x * T + rw.y * B + rw.z * local_r;
        /* SpecularPDF (Material.h:255-261) raises dot(wi, r) to Ns, Eval (Material.h:236-244) raises max(0, dot(wi, r)):
         * where Eval's value is used at all (lca > 0) the two are the same number, so ONE pow serves both (it is ~300
         * instructions, and the whole warp runs it for the one or two lanes in this branch) */
        const double dl = dot(wi, local_r);
        const double lca = fmax(0.0, dl);
        double pw = 0.0;
        if (wi.z > 0.) pw = pow(dl, m.ns);
        if (wi.z <= 0.) pdf = 0.0;
        else pdf = (m.ns + 1.0) * PRT_INV_2PI * pw;
        if (wi.z > 0. && lca > 0.) f = tex_or_color(S, m.kd_tex, m.ks, s.uv) * (m.ns + 2.) * PRT_INV_2PI * pw;
    }
    out_dir = local_to_world(wi, s);
    att = mk3(0, 0, 0);
    if (pdf > 0. && wi.z > 0) att = f * wi.z / pdf;
    if (lobe_probability) att = att / (u < m.pkd ? m.pkd : m.pks);
    cosine_pdf = (u < m.pkd && m.pkd == 1.0) ? pdf : 0.0;
}

__device__ __forceinline__ void scatter_mirror(const Surface& s, d3 wo, d3& att, d3& out_dir) {
    const d3 wi = reflect(wo, mk3(0., 0., 1.));
    const double cos_theta = wi.z;
    const d3 f = mk3(1.0, 1.0, 1.0) / cos_theta;
    out_dir = local_to_world(wi, s);
    att = f * cos_theta / 1.0;
}

}  // namespace prt
#endif

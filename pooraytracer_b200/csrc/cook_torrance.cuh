/*
 * cook_torrance.cuh — device restatement of the reference's rough-conductor BSDF (FP64, reference
 * operation order, compiled with -fmad=false). Paths relative to /root/reference/:
 *
 *   cx_*, fr_complex          Complex<T>, Norm, Sqrt, FrComplex        Source/MaterialUtils.h:7-65,98-110
 *   ct_* trigonometry         Cos2Theta .. SinPhi, AbsDot, Lerp        Source/MaterialUtils.h:67-97
 *   ct_D / ct_lambda / ct_G   CookTorrance::D, Lambda, G1, G, D(w,wm)  Source/Material.h:374-398
 *   ct_sample_wm              CookTorrance::SampleWm (visible normals) Source/Material.h:410-430
 *                             + SampleUniformDiskPolar                 Source/RandomNumberGenerator.h:69-73
 *   cook_eval                 CookTorrance::Eval                       Source/Material.h:464-484
 *   scatter_cook              CookTorrance::Scatter + ::Sample         Source/Material.h:432-462,485-509
 *
 * Pinned bit for bit (Scatter, Eval with scripted uniforms; two renders) against the compiled
 * reference through the CPU port: tests/golden/material_cook.npz, render_cook_*.npz.
 */
#ifndef PRT_COOK_TORRANCE_CUH
#define PRT_COOK_TORRANCE_CUH

#include "shading.cuh"

namespace prt {

struct cplx { double re, im; };
__device__ __forceinline__ cplx cx(double re, double im) { cplx z; z.re = re; z.im = im; return z; }
__device__ __forceinline__ cplx cx_add(cplx a, cplx b) { return cx(a.re + b.re, a.im + b.im); }
__device__ __forceinline__ cplx cx_sub(cplx a, cplx b) { return cx(a.re - b.re, a.im - b.im); }
__device__ __forceinline__ cplx cx_mul(cplx a, cplx b) { return cx(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
__device__ __forceinline__ cplx cx_div(cplx a, cplx z) {
    const double scale = 1 / (z.re * z.re + z.im * z.im);
    return cx(scale * (a.re * z.re + a.im * z.im), scale * (a.im * z.re - a.re * z.im));
}
__device__ __forceinline__ double cx_norm(cplx z) { return z.re * z.re + z.im * z.im; }
__device__ __forceinline__ cplx cx_sqrt(cplx z) {
    const double n = sqrt(cx_norm(z)), t1 = sqrt(.5 * (n + fabs(z.re))), t2 = .5 * z.im / t1;
    if (n == 0) return cx(0, 0);
    if (z.re >= 0) return cx(t1, t2);
    return cx(fabs(t2), copysign(t1, z.im));
}
__device__ __forceinline__ double ct_clamp(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }
__device__ __forceinline__ double ct_sqr(double v) { return v * v; }
__device__ __forceinline__ double ct_cos2(d3 w) { return ct_sqr(w.z); }
__device__ __forceinline__ double ct_sin2(d3 w) { return fmax(0., 1 - ct_cos2(w)); }
__device__ __forceinline__ double ct_tan2(d3 w) { return ct_sin2(w) / ct_cos2(w); }
__device__ __forceinline__ double ct_sin(d3 w) { return sqrt(ct_sin2(w)); }
__device__ __forceinline__ double ct_cos_phi(d3 w) {
    const double st = ct_sin(w);
    return (st == 0) ? 1 : ct_clamp(w.x / st, -1, 1);
}
__device__ __forceinline__ double ct_sin_phi(d3 w) {
    const double st = ct_sin(w);
    return (st == 0) ? 0 : ct_clamp(w.y / st, -1, 1);
}
__device__ __forceinline__ double ct_absdot(d3 a, d3 b) { return fabs(dot(a, b)); }
__device__ __forceinline__ double ct_len2(d3 v) { return ct_sqr(v.x) + ct_sqr(v.y) + ct_sqr(v.z); }
__device__ __forceinline__ bool ct_same_hemisphere(d3 w, d3 wp) { return w.z * wp.z > 0; }

__device__ __noinline__ double fr_complex(double cos_i, double eta_re, double eta_im) {
    const cplx eta = cx(eta_re, eta_im);
    cos_i = ct_clamp(cos_i, 0, 1);
    const double sin2_i = 1 - ct_sqr(cos_i);
    const cplx sin2_t = cx_div(cx(sin2_i, 0), cx_mul(eta, eta));
    const cplx cos_t = cx_sqrt(cx_sub(cx(1, 0), sin2_t));
    const cplx eci = cx_mul(eta, cx(cos_i, 0));
    const cplx ect = cx_mul(eta, cos_t);
    const cplx r_parl = cx_div(cx_sub(eci, cos_t), cx_add(eci, cos_t));
    const cplx r_perp = cx_div(cx_sub(cx(cos_i, 0), ect), cx_add(cx(cos_i, 0), ect));
    return (cx_norm(r_parl) + cx_norm(r_perp)) / 2;
}

__device__ __forceinline__ double ct_D(const DevMaterial& m, d3 wm) {
    const double tan2 = ct_tan2(wm);
    if (isinf(tan2)) return 0;
    const double cos4 = ct_sqr(ct_cos2(wm));
    const double e = tan2 * (ct_sqr(ct_cos_phi(wm) / m.alpha_x) + ct_sqr(ct_sin_phi(wm) / m.alpha_y));
    return 1 / (PRT_PI * m.alpha_x * m.alpha_y * cos4 * ct_sqr(1 + e));
}
__device__ __forceinline__ double ct_lambda(const DevMaterial& m, d3 w) {
    const double tan2 = ct_tan2(w);
    if (isinf(tan2)) return 0;
    const double alpha2 = ct_sqr(ct_cos_phi(w) * m.alpha_x) + ct_sqr(ct_sin_phi(w) * m.alpha_y);
    return (sqrt(1 + alpha2 * tan2) - 1) / 2;
}
__device__ __forceinline__ double ct_G1(const DevMaterial& m, d3 w) { return 1 / (1 + ct_lambda(m, w)); }
__device__ __forceinline__ double ct_G(const DevMaterial& m, d3 wo, d3 wi) { return 1 / (1 + ct_lambda(m, wo) + ct_lambda(m, wi)); }
/* D(w, wm), the distribution of visible normals (= PDF(w, wm)) */
__device__ __forceinline__ double ct_Dw(const DevMaterial& m, d3 w, d3 wm) {
    return ct_G1(m, w) / fabs(w.z) * ct_D(m, wm) * ct_absdot(w, wm);
}
__device__ __forceinline__ d3 ct_fresnel(const DevMaterial& m, d3 wo, d3 wm) {
    const double c = ct_absdot(wo, wm);
    return mk3(fr_complex(c, m.eta[0], m.k[0]), fr_complex(c, m.eta[1], m.k[1]), fr_complex(c, m.eta[2], m.k[2]));
}

__device__ __forceinline__ d3 ct_sample_wm(const DevMaterial& m, d3 w, d2 u) {
    d3 wh = normalize(mk3(m.alpha_x * w.x, m.alpha_y * w.y, w.z));
    if (wh.z < 0) wh = -wh;
    const d3 T1 = (wh.z < 0.99999) ? normalize(cross(mk3(0., 0., 1.), wh)) : mk3(1, 0, 0);
    const d3 T2 = cross(wh, T1);
    const double r = sqrt(u.x);
    const double theta = 2 * PRT_PI * u.y;
    d2 p = mk2(r * cos(theta), r * sin(theta));
    const double h = sqrt(1 - ct_sqr(p.x));
    const double x = (1 + wh.z) / 2;
    p.y = (1 - x) * h + x * p.y; /* Lerp */
    const double pz = sqrt(fmax(0., 1. - (ct_sqr(p.x) + ct_sqr(p.y))));
    const d3 nh = p.x * T1 + p.y * T2 + pz * wh;
    return normalize(mk3(m.alpha_x * nh.x, m.alpha_y * nh.y, fmax(1e-6, nh.z)));
}

/* CookTorrance::Eval; wo as the caller built it (Camera.cpp:165 does not normalise it) */
__device__ __forceinline__ d3 cook_eval(const DevMaterial& m, d3 wi, d3 wo) {
    if (!ct_same_hemisphere(wo, wi)) return mk3(0, 0, 0);
    const double cos_o = fabs(wo.z), cos_i = fabs(wi.z);
    if (cos_i == 0 || cos_o == 0) return mk3(0, 0, 0);
    d3 wm = wi + wo;
    if (ct_len2(wm) == 0) return mk3(0, 0, 0);
    wm = normalize(wm);
    const d3 F = ct_fresnel(m, wo, wm);
    return ct_D(m, wm) * F * ct_G(m, wo, wi) / (4 * cos_i * cos_o);
}

/* CookTorrance::Scatter; false = SampleFlags::Unset (no scattered ray) */
__device__ __forceinline__ bool scatter_cook(const DevMaterial& m, const Surface& s, d3 wo_raw, Rng& rng, d3& att, d3& out_dir) {
    const d3 wo = normalize(wo_raw); /* Material.h:492 */
    if (wo.z == 0) return false;
    /* vec2(RandomDouble(), RandomDouble()): g++ evaluates the arguments right to left */
    const double second_arg = rng.next();
    const double first_arg = rng.next();
    const d3 wm = ct_sample_wm(m, wo, mk2(first_arg, second_arg));
    const d3 wi = reflect(wo, wm);
    if (!ct_same_hemisphere(wo, wi)) return false;
    const double pdf = ct_Dw(m, wo, wm) / (4. * ct_absdot(wo, wm));
    const double cos_o = fabs(wo.z), cos_i = fabs(wi.z);
    if (cos_i == 0 || cos_o == 0) return false;
    const d3 F = ct_fresnel(m, wo, wm);
    const d3 f = ct_D(m, wm) * F * ct_G(m, wo, wi) / (4. * cos_i * cos_o);
    att = f * wi.z / pdf;
    out_dir = local_to_world(wi, s);
    return true;
}

}  // namespace prt
#endif

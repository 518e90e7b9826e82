/*
 * prt_oracle.cpp — TEST INFRASTRUCTURE: CPU restatement ("port") of the reference's render path.
 *
 * Plain C++ (no glm, no reference headers), every function citing the reference file:line it
 * follows (paths relative to /root/reference/). It is PINNED against the reference itself: in this
 * container the unmodified reference sources are compiled into oracle/_ref/libprt_ref.so
 * (oracle/Makefile, oracle/ref_driver.cpp) and tests/test_oracle_pinning.py checks that this port
 * reproduces its closest hits, light samples, BSDF samples, texture lookups and whole rendered
 * images bit for bit on the same inputs; the generated vectors are committed under tests/golden/
 * (script: tests/golden/make_golden.py) so the pin also holds where /root/reference is absent.
 * The reference has no tests or golden vectors of its own for this path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product (pooraytracer_b200/) never does and has no CPU fallback.
 *
 * Arithmetic: IEEE binary64, evaluated in glm's operation order
 * (ThirdPartyLibraries/glm/glm/detail/func_geometric.inl:52-53,79-82,104), compiled with
 * -ffp-contract=off like the reference objects (no FMA).
 *
 * Two reference spots are undefined behaviour; like ORC_DEFINE_UB in ref_driver.cpp the port
 * defines them: Phong's unset `attenuation` (Camera.cpp:177, Material.h:280-284) is 0, and a
 * shadow ray that hits nothing (Camera.cpp:150-152 reads an unset position) is "not visible".
 * A third spot is defined behaviour that a wavefront cannot follow: the draw inside
 * PhoneReflectance::Eval (Material.h:228) advances the stream only if the light sample was visible;
 * the port rewinds it (see PhongDefined in ref_driver.cpp for the full argument).
 */
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <limits>
#include <memory>
#include <thread>
#include <vector>

#include "oracle_api.h"
#include "rng_spec.h"

namespace {

/* ---------------------------------------------------------------------------------------------
 * vectors in glm's order of operations
 * ------------------------------------------------------------------------------------------- */
struct V3 {
    double x, y, z;
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct V2 { double x, y; };
inline V3 v3(double x, double y, double z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline V2 operator+(V2 a, V2 b) { return {a.x + b.x, a.y + b.y}; }
inline V2 operator-(V2 a, V2 b) { return {a.x - b.x, a.y - b.y}; }
inline V2 operator*(double s, V2 a) { return {s * a.x, s * a.y}; }
// func_geometric.inl:52-53  tmp = a*b; tmp.x + tmp.y + tmp.z
inline double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// func_geometric.inl:79-82
inline V3 cross(V3 x, V3 y) { return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
inline double length(V3 a) { return std::sqrt(dot(a, a)); }
// func_geometric.inl:104 + func_exponential.inl:138: v * (1 / sqrt(dot(v,v)))
inline V3 normalize(V3 a) { return a * (1.0 / std::sqrt(dot(a, a))); }
inline bool anynan(V3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

const double Pi = 3.14159265358979323846;       // RandomNumberGenerator.h:10-14
const double InvPi = 0.31830988618379067154;
const double Inv2Pi = 0.15915494309189533577;
const double PiOver2 = 1.57079632679489661923;
const double PiOver4 = 0.78539816339744830961;
const double INF = std::numeric_limits<double>::infinity();

/* ---------------------------------------------------------------------------------------------
 * Interval / AABB (Source/Interval.h, Source/AABB.cpp)
 * ------------------------------------------------------------------------------------------- */
struct Interval {
    double min = +INF, max = -INF;  // Interval.h:11
    Interval() {}
    Interval(double a, double b) : min(a), max(b) {}
    Interval(const Interval& a, const Interval& b) {  // Interval.h:13-16
        min = a.min <= b.min ? a.min : b.min;
        max = a.max >= b.max ? a.max : b.max;
    }
    double Length() const { return max - min; }
    Interval Expand(double delta) const {  // Interval.h:33-36
        double padding = delta / 2.;
        return Interval(min - padding, max + padding);
    }
};
struct AABB {
    Interval x, y, z;
    AABB() {}  // true empty box: the intended AABB::empty (corrected init order)
    AABB(V3 a, V3 b) {  // AABB.cpp:16-22
        x = (a.x <= b.x) ? Interval(a.x, b.x) : Interval(b.x, a.x);
        y = (a.y <= b.y) ? Interval(a.y, b.y) : Interval(b.y, a.y);
        z = (a.z <= b.z) ? Interval(a.z, b.z) : Interval(b.z, a.z);
        Pad();
    }
    AABB(const AABB& a, const AABB& b) : x(a.x, b.x), y(a.y, b.y), z(a.z, b.z) {}  // AABB.cpp:24-29 (no re-pad)
    const Interval& axis(int i) const { return i == 1 ? y : (i == 2 ? z : x); }
    void Pad() {  // AABB.cpp:76-82
        double delta = 0.0001;
        if (x.Length() < delta) x = x.Expand(delta);
        if (y.Length() < delta) y = y.Expand(delta);
        if (z.Length() < delta) z = z.Expand(delta);
    }
    int LongestAxis() const {  // AABB.cpp:66-74
        if (x.Length() > y.Length()) return x.Length() > z.Length() ? 0 : 2;
        return y.Length() > z.Length() ? 1 : 2;
    }
    // AABB.cpp:38-64 — Interval passed by value, 1/d recomputed per call
    bool Hit(V3 o, V3 d, Interval t) const {
        for (int i = 0; i < 3; ++i) {
            const Interval& ax = axis(i);
            const double inv = 1.0 / d[i];
            double t0 = (ax.min - o[i]) * inv;
            double t1 = (ax.max - o[i]) * inv;
            if (t0 < t1) {
                if (t0 > t.min) t.min = t0;
                if (t1 < t.max) t.max = t1;
            } else {
                if (t1 > t.min) t.min = t1;
                if (t0 < t.max) t.max = t0;
            }
            if (t.max <= t.min) return false;
        }
        return true;
    }
};

/* ---------------------------------------------------------------------------------------------
 * RNG: the interposed rand() of ref_driver.cpp as an explicit stream object
 * ------------------------------------------------------------------------------------------- */
struct Rng {
    int mode = ORC_RNG_COUNTER;
    uint64_t seed = 0;
    uint32_t dim = 0;
    uint64_t xs = 0x9E3779B97F4A7C15ull;
    const int32_t* script = nullptr;
    int script_n = 0, script_i = 0;
    int32_t next31() {
        if (script) {
            int v = script_i < script_n ? script[script_i] : 0;
            script_i++;
            return v;
        }
        if (mode == ORC_RNG_COUNTER) return orc_draw31(seed, dim++);
        uint64_t x = xs;
        x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
        xs = x;
        return (int32_t)((x * 0x2545F4914F6CDD1Dull) >> 33);
    }
    // RandomNumberGenerator.h:16-19: rand() / (RAND_MAX + 1.0)
    double RandomDouble() { return next31() / (2147483647 + 1.0); }
};

/* ---------------------------------------------------------------------------------------------
 * textures (Source/Texture.cpp)
 * ------------------------------------------------------------------------------------------- */
struct Tex {
    int w = 0, h = 0, c = 0;
    std::vector<uint8_t> data;
    static double SRGBToLinear(double v) {  // Texture.cpp:66-71
        if (v <= 0.04045) return v * (1. / 12.92);
        return std::pow((v + 0.055) * (1. / 1.055), 2.4);
    }
    V3 GetPixel(int x, int y) const {  // Texture.cpp:50-65
        const double colorScale = 1.0 / 255.0;
        int idx = (y * w + x) * c;
        if (c >= 3)
            return v3(SRGBToLinear(colorScale * data[idx]), SRGBToLinear(colorScale * data[idx + 1]),
                      SRGBToLinear(colorScale * data[idx + 2]));
        double g = colorScale * data[idx];
        return v3(g, g, g);
    }
    V3 Value(double u, double v) const {  // Texture.cpp:22-49
        if (data.empty()) return v3(0., 1., 1.);
        u = std::clamp(u, 0.0, 1.0);
        v = std::clamp(v, 0.0, 1.0);
        double x = u * (w - 1.);
        double y = (1. - v) * (h - 1.);
        int x0 = static_cast<int>(x), y0 = static_cast<int>(y);
        int x1 = std::min(x0 + 1, w - 1), y1 = std::min(y0 + 1, h - 1);
        double tx = x - x0, ty = y - y0;
        V3 c00 = GetPixel(x0, y0), c10 = GetPixel(x1, y0), c01 = GetPixel(x0, y1), c11 = GetPixel(x1, y1);
        V3 c0 = c00 * (1 - tx) + c10 * tx;
        V3 c1 = c01 * (1 - tx) + c11 * tx;
        return c0 * (1 - ty) + c1 * ty;
    }
};

/* ---------------------------------------------------------------------------------------------
 * scene objects
 * ------------------------------------------------------------------------------------------- */
struct Mat {
    int type = PRT_MAT_LAMBERTIAN;
    V3 kd{0, 0, 0}, ks{0, 0, 0}, emission{0, 0, 0};
    double ns = 0, pkd = 1, pks = 0;
    double alphaX = 0.3, alphaY = 0.3;  // CookTorrance, Material.h:370,518
    V3 eta{1, 1, 1}, k{0, 0, 0};
    int kd_tex = -1;
    bool HasEmission() const { return type == PRT_MAT_DIFFUSE_LIGHT || type == PRT_MAT_DEBUG; }  // Material.h:166,528
    bool SkipLightSampling() const {  // Material.h:73,328,365,539
        if (type == PRT_MAT_PHONG) return ns > 1.;
        return type == PRT_MAT_MIRROR || type == PRT_MAT_EMPTY;
    }
};

struct Tri {                       // Source/Triangle.h:13-34
    V3 v[3], e0, e1, normal, tangent, w;
    V2 uv[3];
    double area, D;
    AABB bbox;
    int mat;      // material index (mesh material)
    int64_t prim;
};

struct HitRec {  // Source/Hittable.h:17-28 (+ prim id, alpha/beta for the ray-query interface)
    V3 position, normal, tangent;
    V2 uv;
    double time = 0;
    int mat = -1;
    bool front = false;
    int64_t prim = -1;
    double alpha = 0, beta = 0;
    bool set_position = false;
    void SetFaceNormal(V3 dir, V3 n) {  // Hittable.cpp:8-13
        front = dot(dir, n) < 0.;
        normal = front ? n : -n;
    }
};

// A Hittable is either a triangle (kind 0) or a BVH node (kind 1): Source/BVH.h:25-27
struct Obj { int kind; int64_t idx; };
struct Node {                     // Source/BVH.h:11-34
    AABB bbox;
    Obj left, right;
    double area = 0.0;
};

struct Scene {
    int flags = 0;
    std::vector<Tri> tris;
    std::vector<Mat> mats;
    std::vector<Tex> texs;
    std::vector<Node> nodes;
    int64_t world_root = -1, lights_root = -1;  // the single BVHNode inside each HittableList
    AABB ObjBox(Obj o) const { return o.kind == 0 ? tris[o.idx].bbox : nodes[o.idx].bbox; }
    double ObjArea(Obj o) const { return o.kind == 0 ? tris[o.idx].area : nodes[o.idx].area; }
};

// Triangle::Triangle, Source/Triangle.cpp:11-53
void make_triangle(Tri& T, const double* pv, const double* pn, const double* puv) {
    for (int k = 0; k < 3; ++k) T.v[k] = v3(pv[3 * k], pv[3 * k + 1], pv[3 * k + 2]);
    for (int k = 0; k < 3; ++k) T.uv[k] = V2{puv[2 * k], puv[2 * k + 1]};
    T.e0 = T.v[1] - T.v[0];
    T.e1 = T.v[2] - T.v[0];
    V3 n = cross(T.e0, T.e1);
    T.normal = normalize(n);
    if (anynan(T.normal)) {
        V3 n0 = pn ? v3(pn[0], pn[1], pn[2]) : v3(0, 0, 0);
        V3 n1 = pn ? v3(pn[3], pn[4], pn[5]) : v3(0, 0, 0);
        V3 n2 = pn ? v3(pn[6], pn[7], pn[8]) : v3(0, 0, 0);
        T.normal = normalize(n0 + n1 + n2);
        if (anynan(T.normal)) T.normal = v3(0.0, 0.0, 1.0);
    }
    V2 d0 = T.uv[1] - T.uv[0], d1 = T.uv[2] - T.uv[0];
    double f = 1.0 / (d0.x * d1.y - d1.x * d0.y);
    T.tangent.x = f * (d1.y * T.e0.x - d0.y * T.e1.x);
    T.tangent.y = f * (d1.y * T.e0.y - d0.y * T.e1.y);
    T.tangent.z = f * (d1.y * T.e0.z - d0.y * T.e1.z);
    T.tangent = normalize(T.tangent);
    if (anynan(T.tangent)) {  // Triangle.cpp:38-45; `abs(v.x) < 0.9f` compares against (double)0.9f
        V3 v = T.normal;
        V3 helper = (std::fabs(v.x) < (double)0.9f) ? v3(1, 0, 0) : v3(0, 1, 0);
        T.tangent = normalize(cross(v, helper));
    }
    T.area = length(n) * 0.5;
    T.D = dot(T.normal, T.v[0]);
    T.w = n / dot(n, n);
    T.bbox = AABB(AABB(T.v[0], T.v[1]), AABB(T.v[0], T.v[2]));  // Triangle.cpp:94-99
}

// BVHNode::BVHNode(objects, start, end), Source/BVH.cpp:7-48. `objects` is sorted IN PLACE.
int64_t build_bvh(Scene& s, std::vector<Obj>& objects, size_t start, size_t end) {
    Node node;
    node.bbox = AABB();
    for (size_t i = start; i < end; ++i) node.bbox = AABB(node.bbox, s.ObjBox(objects[i]));
    int axis = node.bbox.LongestAxis();
    auto comparator = [&s, axis](const Obj& a, const Obj& b) {  // BVH.cpp:68-73
        return s.ObjBox(a).axis(axis).min < s.ObjBox(b).axis(axis).min;
    };
    size_t span = end - start;
    if (span == 1) {
        node.left = node.right = objects[start];
        node.area = s.ObjArea(objects[start]);
    } else if (span == 2) {
        node.left = objects[start];
        node.right = objects[start + 1];
        node.area = s.ObjArea(objects[start]) + s.ObjArea(objects[start + 1]);
    } else {
        std::sort(objects.begin() + start, objects.begin() + end, comparator);
        size_t mid = start + span / 2;
        int64_t l = build_bvh(s, objects, start, mid);
        int64_t r = build_bvh(s, objects, mid, end);
        node.left = Obj{1, l};
        node.right = Obj{1, r};
        node.area = s.nodes[l].area + s.nodes[r].area;  // both children are BVHNodes when span >= 3
    }
    s.nodes.push_back(node);
    return (int64_t)s.nodes.size() - 1;
}

// Triangle::Hit + IsInterior, Source/Triangle.cpp:54-83,100-113
bool tri_hit(const Scene& s, const Tri& T, V3 o, V3 d, Interval dom, HitRec& rec) {
    double denom = dot(T.normal, d);
    if (std::fabs(denom) < 1e-8) return false;
    double t = (T.D - dot(T.normal, o)) / denom;
    if (!(dom.min <= t && t <= dom.max)) return false;  // Interval::Contains, inclusive
    V3 p = o + d * t;                                   // Ray.h:13
    V3 v0p = p - T.v[0];
    double alpha = dot(T.w, cross(v0p, T.e1));
    double beta = dot(T.w, cross(T.e0, v0p));
    if (alpha != alpha || beta != beta) return false;
    if ((alpha < 0) || (beta < 0) || (alpha + beta > 1)) return false;
    rec.uv = (1. - alpha - beta) * T.uv[0] + alpha * T.uv[1] + beta * T.uv[2];
    rec.position = p;
    rec.set_position = true;
    rec.time = t;
    rec.mat = T.mat;
    rec.tangent = T.tangent;
    rec.SetFaceNormal(d, T.normal);
    rec.prim = T.prim;
    rec.alpha = alpha;
    rec.beta = beta;
    (void)s;
    return true;
}

bool obj_hit(const Scene& s, Obj ob, V3 o, V3 d, Interval dom, HitRec& rec);

// BVHNode::Hit, Source/BVH.cpp:51-61
bool node_hit(const Scene& s, const Node& n, V3 o, V3 d, Interval dom, HitRec& rec) {
    if (!n.bbox.Hit(o, d, dom)) return false;
    bool hitL = obj_hit(s, n.left, o, d, dom, rec);
    bool hitR = obj_hit(s, n.right, o, d, Interval(dom.min, hitL ? rec.time : dom.max), rec);
    return hitL || hitR;
}
bool obj_hit(const Scene& s, Obj ob, V3 o, V3 d, Interval dom, HitRec& rec) {
    return ob.kind == 0 ? tri_hit(s, s.tris[ob.idx], o, d, dom, rec) : node_hit(s, s.nodes[ob.idx], o, d, dom, rec);
}
// HittableList::Hit over its single BVHNode, Source/HittableList.h:26-39
bool world_hit(const Scene& s, V3 o, V3 d, Interval dom, HitRec& rec) {
    if (s.world_root < 0) return false;
    HitRec tmp;
    if (node_hit(s, s.nodes[s.world_root], o, d, Interval(dom.min, dom.max), tmp)) {
        rec = tmp;
        return true;
    }
    return false;
}

// Triangle::Sample, Source/Triangle.cpp:84-93
void tri_sample(const Tri& T, V3 origin, HitRec& rec, double& pdf, Rng& rng) {
    double x = std::sqrt(rng.RandomDouble()), y = rng.RandomDouble();
    V3 p = T.v[0] * (1.0 - x) + T.v[1] * (x * (1.0 - y)) + T.v[2] * (x * y);
    rec.position = p;
    V3 direction = p - origin;
    rec.SetFaceNormal(direction, T.normal);
    rec.mat = T.mat;
    rec.prim = T.prim;
    pdf = 1.0 / T.area;
}
// BVHNode::TraverseSample, Source/BVH.cpp:86-100 — `float p`
void traverse_sample(const Scene& s, V3 origin, Obj node, float p, HitRec& rec, double& pdf, Rng& rng) {
    if (node.kind == 0) {
        tri_sample(s.tris[node.idx], origin, rec, pdf, rng);
        pdf *= s.tris[node.idx].area;
        return;
    }
    const Node& n = s.nodes[node.idx];
    if (p < s.ObjArea(n.left)) traverse_sample(s, origin, n.left, p, rec, pdf, rng);
    else traverse_sample(s, origin, n.right, (float)(p - s.ObjArea(n.left)), rec, pdf, rng);
}
// The opt-in "corrected" estimators of a render (prt_render_params.flags: PRT_RENDER_UNBIASED_LIGHT_PICK,
// PRT_RENDER_LOBE_PROBABILITY), set per worker thread by orc_render / orc_pixel_samples. 0 = the reference.
static thread_local uint32_t tl_render_flags = 0;

// HittableList::Sample -> BVHNode::Sample, Source/HittableList.h:44-59, Source/BVH.cpp:62-67
void lights_sample(const Scene& s, V3 origin, HitRec& rec, double& pdf, Rng& rng) {
    const Node& top = s.nodes[s.lights_root];
    double areaSum = 0.0;
    areaSum += top.area;
    double p0 = rng.RandomDouble() * areaSum;
    areaSum = 0.0;
    areaSum += top.area;
    if (p0 <= areaSum) {
        const double u = rng.RandomDouble();
        double p = ((tl_render_flags & PRT_RENDER_UNBIASED_LIGHT_PICK) ? u : std::sqrt(u)) * top.area;
        traverse_sample(s, origin, Obj{1, s.lights_root}, (float)p, rec, pdf, rng);
        pdf /= top.area;
    }
}

/* ---------------------------------------------------------------------------------------------
 * samplers (Source/RandomNumberGenerator.h:39-64)
 * ------------------------------------------------------------------------------------------- */
V2 SampleUniformDiskConcentric(V2 u) {
    V2 uo{2. * u.x - 1., 2. * u.y - 1.};
    if (uo.x == 0. && uo.y == 0.) return {0., 0.};
    double theta, r;
    if (std::fabs(uo.x) > std::fabs(uo.y)) {
        r = uo.x;
        theta = PiOver4 * (uo.y / uo.x);
    } else {
        r = uo.y;
        theta = PiOver2 - PiOver4 * (uo.x / uo.y);
    }
    return {r * std::cos(theta), r * std::sin(theta)};
}
V3 SampleCosineHemisphere(Rng& rng) {
    // `glm::dvec2(RandomDouble(), RandomDouble())`: g++ evaluates the constructor arguments right
    // to left, so the SECOND draw of the stream lands in u.x (pinned against oracle/_ref by
    // tests/test_oracle_pinning.py::test_material_scatter).
    double second_arg = rng.RandomDouble();
    double first_arg = rng.RandomDouble();
    V2 u{first_arg, second_arg};
    V2 d = SampleUniformDiskConcentric(u);
    double z = std::sqrt(std::max(0.0, 1. - d.x * d.x - d.y * d.y));
    return v3(d.x, d.y, z);
}

/* ---------------------------------------------------------------------------------------------
 * materials (Source/Material.h)
 * ------------------------------------------------------------------------------------------- */
struct EvalCtx { V3 p; V2 uv; V3 wo, n, dpdus; };

V3 LocalToWorld(V3 local, const EvalCtx& c) {  // Material.h:76-83
    V3 bitangent = cross(c.dpdus, c.n);
    return normalize(local.x * c.dpdus + local.y * bitangent + local.z * c.n);
}
V3 WorldToLocal(V3 world, const HitRec& r) {  // Material.h:84-95
    V3 bitangent = cross(r.tangent, r.normal);
    return v3(dot(world, r.tangent), dot(world, bitangent), dot(world, r.normal));
}
V3 Reflect(V3 wo, V3 n) { return -wo + 2. * dot(wo, n) * n; }  // Material.h:96-98

V3 tex_or_color(const Scene& s, int tex, V3 c, V2 uv) {
    if (tex >= 0) return s.texs[tex].Value(uv.x, uv.y);
    return c;
}


/* ---------------------------------------------------------------------------------------------
 * CookTorrance (Source/Material.h:368-521) and its helpers (Source/MaterialUtils.h)
 * ------------------------------------------------------------------------------------------- */
struct Cx { double re, im; };  // MaterialUtils.h:7-43 (Complex<double>)
inline Cx cx(double re) { return {re, 0}; }
inline Cx operator+(Cx a, Cx b) { return {a.re + b.re, a.im + b.im}; }
inline Cx operator-(Cx a, Cx b) { return {a.re - b.re, a.im - b.im}; }
inline Cx operator*(Cx a, Cx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
inline Cx operator/(Cx a, Cx z) {
    double scale = 1 / (z.re * z.re + z.im * z.im);
    return {scale * (a.re * z.re + a.im * z.im), scale * (a.im * z.re - a.re * z.im)};
}
inline double Norm(Cx z) { return z.re * z.re + z.im * z.im; }  // MaterialUtils.h:45-48
inline Cx CxSqrt(Cx z) {                                        // MaterialUtils.h:53-65
    double n = std::sqrt(Norm(z)), t1 = std::sqrt(.5 * (n + std::fabs(z.re))), t2 = .5 * z.im / t1;
    if (n == 0) return cx(0);
    if (z.re >= 0) return {t1, t2};
    return {std::fabs(t2), std::copysign(t1, z.im)};
}
inline double Clamp(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }  // :67-71
inline double Sqr(double v) { return v * v; }
inline double Cos2Theta(V3 w) { return Sqr(w.z); }                               // :75-91
inline double Sin2Theta(V3 w) { return std::max<double>(0., 1 - Cos2Theta(w)); }
inline double Tan2Theta(V3 w) { return Sin2Theta(w) / Cos2Theta(w); }
inline double SinTheta(V3 w) { return std::sqrt(Sin2Theta(w)); }
inline double AbsCosTheta(V3 w) { return std::abs(w.z); }
inline double CosPhi(V3 w) { double st = SinTheta(w); return (st == 0) ? 1 : Clamp(w.x / st, -1, 1); }
inline double SinPhi(V3 w) { double st = SinTheta(w); return (st == 0) ? 0 : Clamp(w.y / st, -1, 1); }
inline double AbsDot(V3 a, V3 b) { return std::fabs(dot(a, b)); }
inline double Lerp(double x, double a, double b) { return (1 - x) * a + x * b; }
inline double LengthSquared3(V3 v) { return Sqr(v.x) + Sqr(v.y) + Sqr(v.z); }
inline double LengthSquared2(V2 v) { return Sqr(v.x) + Sqr(v.y); }
inline bool SameHemisphere(V3 w, V3 wp) { return w.z * wp.z > 0; }                // :112-114
inline double FrComplex(double cosTheta_i, Cx eta) {                             // :98-110
    cosTheta_i = Clamp(cosTheta_i, 0, 1);
    double sin2Theta_i = 1 - Sqr(cosTheta_i);
    Cx sin2Theta_t = cx(sin2Theta_i) / (eta * eta);
    Cx cosTheta_t = CxSqrt(cx(1) - sin2Theta_t);
    Cx r_parl = (eta * cx(cosTheta_i) - cosTheta_t) / (eta * cx(cosTheta_i) + cosTheta_t);
    Cx r_perp = (cx(cosTheta_i) - eta * cosTheta_t) / (cx(cosTheta_i) + eta * cosTheta_t);
    return (Norm(r_parl) + Norm(r_perp)) / 2;
}
struct CookTorrance {
    double alphaX, alphaY;
    V3 eta, k;
    double D(V3 wm) const {  // Material.h:374-381
        double tan2Theta = Tan2Theta(wm);
        if (std::isinf(tan2Theta)) return 0;
        double cos4Theta = Sqr(Cos2Theta(wm));
        double e = tan2Theta * (Sqr(CosPhi(wm) / alphaX) + Sqr(SinPhi(wm) / alphaY));
        return 1 / (Pi * alphaX * alphaY * cos4Theta * Sqr(1 + e));
    }
    double Lambda(V3 w) const {  // :382-387
        double tan2Theta = Tan2Theta(w);
        if (std::isinf(tan2Theta)) return 0;
        double alpha2 = Sqr(CosPhi(w) * alphaX) + Sqr(SinPhi(w) * alphaY);
        return (std::sqrt(1 + alpha2 * tan2Theta) - 1) / 2;
    }
    double G1(V3 w) const { return 1 / (1 + Lambda(w)); }                       // :388
    double G(V3 wo, V3 wi) const { return 1 / (1 + Lambda(wo) + Lambda(wi)); }  // :389-391
    double Dw(V3 w, V3 wm) const { return G1(w) / AbsCosTheta(w) * D(wm) * AbsDot(w, wm); }  // :392-398
    V3 Fresnel(V3 wo, V3 wm) const {  // :452,477
        double c = AbsDot(wo, wm);
        return v3(FrComplex(c, Cx{eta.x, k.x}), FrComplex(AbsDot(wo, wm), Cx{eta.y, k.y}), FrComplex(AbsDot(wo, wm), Cx{eta.z, k.z}));
    }
    V3 SampleWm(V3 w, V2 u) const {  // :410-430
        V3 wh = normalize(v3(alphaX * w.x, alphaY * w.y, w.z));
        if (wh.z < 0) wh = -wh;
        V3 T1 = (wh.z < 0.99999) ? normalize(cross(v3(0., 0., 1.), wh)) : v3(1, 0, 0);
        V3 T2 = cross(wh, T1);
        // SampleUniformDiskPolar, RandomNumberGenerator.h:69-73
        double r = std::sqrt(u.x);
        double theta = 2 * Pi * u.y;
        V2 p{r * std::cos(theta), r * std::sin(theta)};
        double h = std::sqrt(1 - Sqr(p.x));
        p.y = Lerp((1 + wh.z) / 2, h, p.y);
        double pz = std::sqrt(std::max<double>(0., 1. - LengthSquared2(p)));
        V3 nh = p.x * T1 + p.y * T2 + pz * wh;
        return normalize(v3(alphaX * nh.x, alphaY * nh.y, std::max<double>(1e-6, nh.z)));
    }
    V3 Eval(V3 wi, V3 wo) const {  // :464-484
        if (!SameHemisphere(wo, wi)) return v3(0, 0, 0);
        double cosTheta_o = AbsCosTheta(wo), cosTheta_i = AbsCosTheta(wi);
        if (cosTheta_i == 0 || cosTheta_o == 0) return v3(0, 0, 0);
        V3 wm = wi + wo;
        if (LengthSquared3(wm) == 0) return v3(0, 0, 0);
        wm = normalize(wm);
        V3 F = Fresnel(wo, wm);
        return D(wm) * F * G(wo, wi) / (4 * cosTheta_i * cosTheta_o);
    }
    // Sample, :432-462. false = SampleFlags::Unset
    bool Sample(V3 wo, Rng& rng, V3& wi, V3& f, double& pdf) const {
        if (wo.z == 0) return false;
        // vec2(RandomDouble(), RandomDouble()): arguments evaluated right to left by g++
        double second_arg = rng.RandomDouble();
        double first_arg = rng.RandomDouble();
        V3 wm = SampleWm(wo, V2{first_arg, second_arg});
        wi = Reflect(wo, wm);
        if (!SameHemisphere(wo, wi)) return false;
        pdf = Dw(wo, wm) / (4. * AbsDot(wo, wm));
        double cosTheta_o = AbsCosTheta(wo), cosTheta_i = AbsCosTheta(wi);
        if (cosTheta_i == 0 || cosTheta_o == 0) return false;
        V3 F = Fresnel(wo, wm);
        f = D(wm) * F * G(wo, wi) / (4. * cosTheta_i * cosTheta_o);
        return true;
    }
};
inline CookTorrance cook(const Mat& m) { return CookTorrance{m.alphaX, m.alphaY, m.eta, m.k}; }

// Material::Eval for the NEE estimator (Camera.cpp:168)
V3 mat_eval(const Scene& s, const Mat& m, V3 wi, const EvalCtx& c, Rng& rng) {
    switch (m.type) {
        case PRT_MAT_LAMBERTIAN:  // Material.h:128-130
            return tex_or_color(s, m.kd_tex, m.kd, c.uv) * InvPi;
        case PRT_MAT_PHONG: {  // Material.h:227-248: draws a number (Material.h:228); the stream is REWOUND after it, as in
                               // ref_driver.cpp's PhongDefined (the advance would depend on the shadow ray's outcome)
            uint32_t saved = rng.dim;
            double u = rng.RandomDouble();
            if (rng.mode == ORC_RNG_COUNTER && !rng.script) rng.dim = saved;
            if (u < m.pkd) {
                if (wi.z <= 0) return v3(0, 0, 0);
                return tex_or_color(s, m.kd_tex, m.kd, c.uv) * InvPi;
            } else if (m.pkd <= u && u < m.pkd + m.pks) {
                if (wi.z <= 0) return v3(0, 0, 0);
                V3 localReflect = normalize(Reflect(c.wo, v3(0., 0., 1.)));
                double ca = std::max(0., dot(wi, localReflect));
                if (ca <= 0.) return v3(0, 0, 0);
                // textured Phong aliases Ks to the Kd texture (Material.h:178-181)
                return tex_or_color(s, m.kd_tex, m.ks, c.uv) * (m.ns + 2.) * Inv2Pi * std::pow(ca, m.ns);
            }
            return v3(0, 0, 0);
        }
        case PRT_MAT_COOK_TORRANCE:  // Material.h:464-484
            return cook(m).Eval(wi, c.wo);
        default:
            return v3(0, 0, 0);  // Material.h:65-67 (Mirror, Empty, lights never reach Eval anyway)
    }
}

// Material::Scatter. Returns false when the material does not scatter. `att` is pre-set to 0 by the
// caller (defined meaning of Camera.cpp:177).
bool mat_scatter(const Scene& s, const Mat& m, V3 rayDir, const HitRec& rec, V3& att, V3& outDir, Rng& rng, double* cosinePdf = nullptr) {
    if (cosinePdf) *cosinePdf = 0.0; /* PRT_RENDER_MIS: pdf of the sampled direction where the BSDF is the cosine lobe */
    EvalCtx c;
    c.p = rec.position; c.uv = rec.uv; c.n = rec.normal; c.dpdus = rec.tangent;
    c.wo = WorldToLocal(-rayDir, rec);
    switch (m.type) {
        case PRT_MAT_LAMBERTIAN: {  // Material.h:106-151
            V3 wi = SampleCosineHemisphere(rng);
            while (wi.z <= 0.) wi = SampleCosineHemisphere(rng);
            double pdf = wi.z * InvPi;
            V3 f = tex_or_color(s, m.kd_tex, m.kd, c.uv) * InvPi;
            outDir = LocalToWorld(wi, c);
            att = f * wi.z / pdf;
            if (cosinePdf) *cosinePdf = pdf;
            return true;
        }
        case PRT_MAT_PHONG: {  // Material.h:183-226,262-285
            V3 wi = v3(0, 0, 0), f = v3(0, 0, 0);
            double pdf = 0.0;
            double u = rng.RandomDouble();
            if (u < m.pkd) {
                wi = SampleCosineHemisphere(rng);
                while (wi.z <= 0.) wi = SampleCosineHemisphere(rng);
                pdf = wi.z * InvPi;
                f = tex_or_color(s, m.kd_tex, m.kd, c.uv) * InvPi;
            } else if (m.pkd <= u && u < m.pkd + m.pks) {
                double u1 = rng.RandomDouble(), u2 = rng.RandomDouble();
                double alpha = std::acos(std::pow(u1, 1.0 / (m.ns + 1.0)));
                double phi = 2.0 * Pi * u2;
                double sa = std::sin(alpha), ca = std::cos(alpha), sp = std::sin(phi), cp = std::cos(phi);
                V3 rw = v3(sa * cp, sa * sp, ca);
                // ReflectiveSpaceToLocal, Material.h:302-312
                V3 localR = normalize(Reflect(c.wo, v3(0., 0., 1.)));
                V3 V = (std::fabs(localR.x) > 0.9 ? v3(0., 1., 0.) : v3(1., 0., 0.));
                V3 T = normalize(cross(V, localR));
                V3 B = cross(localR, T);
                wi = rw.x * T + rw.y * B + rw.z * localR;
                // SpecularPDF, Material.h:255-261
                if (wi.z <= 0.) pdf = 0.0;
                else pdf = (m.ns + 1.0) * Inv2Pi * std::pow(dot(wi, localR), m.ns);
                double lca = std::max(0.0, dot(wi, localR));
                if (wi.z > 0. && lca > 0.)
                    f = tex_or_color(s, m.kd_tex, m.ks, c.uv) * (m.ns + 2.) * Inv2Pi * std::pow(lca, m.ns);
            }
            outDir = LocalToWorld(wi, c);
            if (pdf > 0. && wi.z > 0) att = f * wi.z / pdf;
            if (tl_render_flags & PRT_RENDER_LOBE_PROBABILITY) att = att / (u < m.pkd ? m.pkd : m.pks);
            if (cosinePdf && u < m.pkd && m.pkd == 1.0) *cosinePdf = pdf;
            return true;
        }
        case PRT_MAT_MIRROR: {  // Material.h:334-363
            V3 wi = Reflect(c.wo, v3(0., 0., 1.));
            double cosTheta = wi.z;
            V3 f = v3(1.0, 1.0, 1.0) / cosTheta;
            outDir = LocalToWorld(wi, c);
            att = f * cosTheta / 1.0;
            return true;
        }
        case PRT_MAT_COOK_TORRANCE: {  // Material.h:485-509; wo is normalised here and only here
            V3 wo = normalize(c.wo), wi, f;
            double pdf;
            if (!cook(m).Sample(wo, rng, wi, f, pdf)) return false;
            att = f * wi.z / pdf;
            outDir = LocalToWorld(wi, c);
            return true;
        }
        default:
            return false;  // Material.h:59-61 (Empty, lights)
    }
}

/* ---------------------------------------------------------------------------------------------
 * camera + integrator (Source/Camera.cpp)
 * ------------------------------------------------------------------------------------------- */
struct Cam {
    int W, H;
    V3 center, pixel00, dU, dV;
    void Init(const prt_camera* c) {  // Camera.cpp:75-106
        W = c->image_width < 1 ? 1 : c->image_width;
        H = c->image_height < 1 ? 1 : c->image_height;
        double aspect = double(W) / double(H);
        V3 eye = v3(c->eye[0], c->eye[1], c->eye[2]), look = v3(c->look_at[0], c->look_at[1], c->look_at[2]);
        V3 up = v3(c->up[0], c->up[1], c->up[2]);
        center = eye;
        double focal = length(eye - look);
        double theta = c->fovy * 0.01745329251994329576923690768489;  // glm::radians
        double h = std::tan(theta / 2.0);
        double vh = 2. * h * focal;
        double vw = vh * aspect;
        V3 w = normalize(eye - look);
        V3 u = normalize(cross(up, w));
        V3 v = cross(w, u);
        V3 vu = vw * u;
        V3 vv = vh * (-v);
        dU = vu / (double)W;
        dV = vv / (double)H;
        V3 upperLeft = center - (focal * w) - vu / 2. - vv / 2.;
        pixel00 = upperLeft + 0.5 * (dU + dV);
    }
    void GetRay(int i, int j, V3& o, V3& d) const {  // Camera.cpp:108-117
        V3 ps = pixel00 + ((double)i) * dU + ((double)j) * dV;
        o = center;
        d = ps - o;
    }
    // PRT_RENDER_PIXEL_JITTER: the commented-out lines Camera.cpp:110-111 with SampleSquare()
    // (RandomNumberGenerator.h:65-68); NOT what the shipped reference renders
    void GetRayJittered(int i, int j, Rng& rng, V3& o, V3& d) const {
        const double ox = rng.RandomDouble() - 0.5;
        const double oy = rng.RandomDouble() - 0.5;
        V3 ps = pixel00 + ((double)i + ox) * dU + ((double)j + oy) * dV;
        o = center;
        d = ps - o;
    }
};

struct Counters { uint64_t hits = 0, shadows = 0; };

struct Integrator {
    const Scene& s;
    int maxDepth;
    bool sampleLights;
    double rr;
    V3 background;
    Counters cnt;

    bool Hit(V3 o, V3 d, Interval dom, HitRec& rec) {
        cnt.hits++;
        if (dom.min == 0.001) cnt.shadows++;
        return world_hit(s, o, d, dom, rec);
    }

    static bool CosineLobe(const Mat& m) { return m.type == PRT_MAT_LAMBERTIAN || (m.type == PRT_MAT_PHONG && m.pkd == 1.0); }

    // Camera::RayColor, Source/Camera.cpp:119-204
    V3 RayColor(V3 ro, V3 rd, int depth, Rng& rng) {
        if (depth < 0.) return v3(0, 0, 0);
        HitRec rec;
        if (!Hit(ro, rd, Interval(0.0001, INF), rec)) return background;
        const Mat& m = s.mats[rec.mat];
        if (m.HasEmission()) return m.type == PRT_MAT_DEBUG ? m.kd : m.emission;
        V3 ps = rec.position;
        V3 direct = v3(0, 0, 0), scatter = v3(0, 0, 0);
        if (sampleLights && !m.SkipLightSampling()) {
            double pdfLights = 0.0;
            HitRec ls;
            lights_sample(s, ps, ls, pdfLights, rng);
            V3 pl = ls.position;
            V3 lightDirection = normalize(pl - ps);
            V3 lightNormal = ls.normal;
            HitRec lr;
            double distance = length(pl - ps);
            bool anyHit = Hit(ps, lightDirection, Interval(0.001, std::numeric_limits<double>::max()), lr);
            if (dot(rec.normal, lightDirection) > 0.0 && ls.front && anyHit &&
                distance - length(ps - lr.position) < 0.001) {
                const Mat& lm = s.mats[ls.mat];
                V3 emission = lm.type == PRT_MAT_DEBUG ? lm.kd : lm.emission;
                EvalCtx c;
                c.p = rec.position; c.uv = rec.uv; c.n = rec.normal; c.dpdus = rec.tangent;
                c.wo = WorldToLocal(-rd, rec);
                V3 localWi = WorldToLocal(lightDirection, rec);
                V3 localLightNormal = WorldToLocal(lightNormal, rec);
                V3 fr = mat_eval(s, m, localWi, c, rng);
                double cosTheta = localWi.z;
                double cosThetaBar = dot(localLightNormal, -localWi);
                direct = emission * fr * cosTheta * cosThetaBar / (distance * distance) / pdfLights;
                if ((tl_render_flags & PRT_RENDER_MIS) && CosineLobe(m)) {
                    const double pl = pdfLights * (distance * distance) / cosThetaBar;
                    const double pb = cosTheta * InvPi;
                    direct = direct * (pl / (pl + pb));
                }
            }
        }
        V3 att = v3(0, 0, 0);
        V3 sd;
        if (rng.RandomDouble() < rr) {
            double cosinePdf = 0.0;
            if (mat_scatter(s, m, rd, rec, att, sd, rng, &cosinePdf)) {
                if (sampleLights) {
                    HitRec peek;
                    if (Hit(ps, sd, Interval(0.0001, INF), peek)) {
                        if (!s.mats[peek.mat].HasEmission()) {
                            scatter = att * RayColor(ps, sd, depth - 1, rng) / rr;
                        } else if (m.SkipLightSampling()) {
                            scatter = att * RayColor(ps, sd, depth - 1, rng) / rr;
                        } else if ((tl_render_flags & PRT_RENDER_MIS) && cosinePdf > 0.0 && depth - 1 >= 0) {
                            /* PRT_RENDER_MIS: the emitter the reference drops here, weighted against the
                             * light sampler's density for the same point (0 for a back-facing emitter) */
                            const Mat& em = s.mats[peek.mat];
                            const V3 le = em.type == PRT_MAT_DEBUG ? em.kd : em.emission;
                            const double pb = (double)(float)cosinePdf; /* carried as a float by the CUDA path record */
                            const double d2 = dot(sd, sd);
                            const double cosl = std::fabs(dot(peek.normal, sd)) / std::sqrt(d2);
                            double pl = 0.0;
                            if (peek.front && cosl > 0.0) pl = (1.0 / s.nodes[s.lights_root].area) * (peek.time * peek.time * d2) / cosl;
                            scatter = att * le * (pb / (pb + pl)) / rr;
                        }
                    }
                } else {
                    scatter = att * RayColor(ps, sd, depth - 1, rng) / rr;
                }
            }
        }
        return direct + scatter;
    }
};

void parallel_for(int64_t n, int nthreads, const std::function<void(int64_t, int64_t, int)>& body) {
    if (nthreads <= 1 || n < 2) { body(0, n, 0); return; }
    std::atomic<int64_t> next{0};
    const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(4096, n / (nthreads * 8) + 1));
    std::vector<std::thread> th;
    for (int k = 0; k < nthreads; ++k)
        th.emplace_back([&, k]() {
            for (;;) {
                int64_t b = next.fetch_add(chunk);
                if (b >= n) break;
                body(b, std::min(n, b + chunk), k);
            }
        });
    for (auto& t : th) t.join();
}

void walk_leaves(const Scene& s, Obj o, std::vector<int64_t>& out) {
    if (o.kind == 0) { out.push_back(s.tris[o.idx].prim); return; }
    const Node& n = s.nodes[o.idx];
    walk_leaves(s, n.left, out);
    if (!(n.right.kind == n.left.kind && n.right.idx == n.left.idx)) walk_leaves(s, n.right, out);
}

}  // namespace

/* =============================================================================================
 * C interface (oracle_api.h)
 * ============================================================================================= */
extern "C" {

int orc_scene_create(const prt_scene_desc* d, const char* const* texture_paths, int flags, void** out) {
    (void)texture_paths;  // the port takes decoded texels from the description
    if (!d || !out || d->n_triangles < 0) return PRT_ERR_INVALID_ARGUMENT;
    auto s = std::make_unique<Scene>();
    s->flags = flags;
    for (int i = 0; i < d->n_textures; ++i) {
        Tex t;
        const prt_texture& pt = d->textures[i];
        if (pt.texels && pt.width > 0 && pt.height > 0 && pt.channels > 0) {
            t.w = pt.width; t.h = pt.height; t.c = pt.channels;
            t.data.assign(pt.texels, pt.texels + (size_t)pt.width * pt.height * pt.channels);
        }
        s->texs.push_back(std::move(t));
    }
    for (int i = 0; i < d->n_materials; ++i) {
        const prt_material& pm = d->materials[i];
        Mat m;
        m.type = pm.type;
        m.kd = v3(pm.kd[0], pm.kd[1], pm.kd[2]);
        m.ks = v3(pm.ks[0], pm.ks[1], pm.ks[2]);
        m.emission = v3(pm.emission[0], pm.emission[1], pm.emission[2]);
        m.ns = pm.ns;
        m.alphaX = pm.alpha_x; m.alphaY = pm.alpha_y;
        m.eta = v3(pm.eta[0], pm.eta[1], pm.eta[2]);
        m.k = v3(pm.k[0], pm.k[1], pm.k[2]);
        m.kd_tex = (pm.kd_tex >= 0 && pm.kd_tex < d->n_textures) ? pm.kd_tex : -1;
        if (m.ns <= 9.) { m.pkd = 1.0; m.pks = 0.0; } else { m.pkd = 0.6; m.pks = 0.4; }  // Material.h:318-327
        s->mats.push_back(m);
    }
    s->tris.resize((size_t)d->n_triangles);
    const double uv_default[6] = {0, 0, 1, 0, 1, 1};  // Model.cpp:170-175
    const double uv_tag[6] = {0, 0, 1, 0, 0, 1};
    for (int m = 0; m < d->n_meshes; ++m)
        for (int64_t f = d->mesh_first[m]; f < d->mesh_first[m + 1]; ++f) {
            Tri& T = s->tris[(size_t)f];
            const double* puv = (flags & ORC_UV_TAG) ? uv_tag : (d->texcoords ? d->texcoords + f * 6 : uv_default);
            make_triangle(T, d->vertices + f * 9, d->normals ? d->normals + f * 9 : nullptr, puv);
            T.mat = d->mesh_material[m];
            T.prim = f;
        }
    // main.cpp:36-45
    std::vector<Obj> world, lights;
    for (int m = 0; m < d->n_meshes; ++m) {
        if (d->mesh_first[m + 1] <= d->mesh_first[m]) continue;
        std::vector<Obj> objs;  // mesh->objects, sorted in place by BOTH builds (BVH.cpp:49)
        for (int64_t f = d->mesh_first[m]; f < d->mesh_first[m + 1]; ++f) objs.push_back(Obj{0, f});
        world.push_back(Obj{1, build_bvh(*s, objs, 0, objs.size())});
        if (s->mats[d->mesh_material[m]].HasEmission()) lights.push_back(Obj{1, build_bvh(*s, objs, 0, objs.size())});
    }
    if (!world.empty()) s->world_root = build_bvh(*s, world, 0, world.size());
    if (!lights.empty()) s->lights_root = build_bvh(*s, lights, 0, lights.size());
    *out = s.release();
    return 0;
}

void orc_scene_destroy(void* p) { delete (Scene*)p; }

int orc_trace_closest(void* p, int64_t n, const double* o, const double* d, const double* tmin,
                      const double* tmax, int32_t* prim, double* t, double* alpha, double* beta, int nthreads) {
    Scene* s = (Scene*)p;
    if (!s) return PRT_ERR_INVALID_ARGUMENT;
    parallel_for(n, nthreads, [&](int64_t b, int64_t e, int) {
        for (int64_t i = b; i < e; ++i) {
            prim[i] = -1;
            if (t) t[i] = 0.;
            if (alpha) alpha[i] = 0.;
            if (beta) beta[i] = 0.;
            HitRec rec;
            if (world_hit(*s, v3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), v3(d[3 * i], d[3 * i + 1], d[3 * i + 2]),
                          Interval(tmin[i], tmax[i]), rec)) {
                prim[i] = (int32_t)rec.prim;
                if (t) t[i] = rec.time;
                if (alpha) alpha[i] = (s->flags & ORC_UV_TAG) ? rec.uv.x : rec.alpha;
                if (beta) beta[i] = (s->flags & ORC_UV_TAG) ? rec.uv.y : rec.beta;
            }
        }
    });
    return 0;
}

int orc_reference_order(void* p, int64_t* world_rank, int64_t* light_order, int64_t* n_light) {
    Scene* s = (Scene*)p;
    if (!s) return PRT_ERR_INVALID_ARGUMENT;
    if (world_rank) {
        std::vector<int64_t> leaves;
        if (s->world_root >= 0) walk_leaves(*s, Obj{1, s->world_root}, leaves);
        for (size_t i = 0; i < s->tris.size(); ++i) world_rank[i] = -1;
        for (size_t r = 0; r < leaves.size(); ++r) world_rank[leaves[r]] = (int64_t)r;
    }
    std::vector<int64_t> ll;
    if (s->lights_root >= 0) walk_leaves(*s, Obj{1, s->lights_root}, ll);
    if (n_light) *n_light = (int64_t)ll.size();
    if (light_order) std::copy(ll.begin(), ll.end(), light_order);
    return 0;
}

int orc_sample_lights(void* p, int64_t n, const double* origins, const double* u, double* position,
                      double* normal, uint8_t* front, double* pdf, int32_t* prim) {
    Scene* s = (Scene*)p;
    if (!s || s->lights_root < 0) return PRT_ERR_INVALID_ARGUMENT;
    for (int64_t i = 0; i < n; ++i) {
        int32_t script[4];
        for (int k = 0; k < 4; ++k) script[k] = (int32_t)(u[4 * i + k] * 2147483648.0);
        Rng rng;
        rng.script = script; rng.script_n = 4;
        HitRec rec;
        double q = 0.;
        lights_sample(*s, v3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), rec, q, rng);
        position[3 * i] = rec.position.x; position[3 * i + 1] = rec.position.y; position[3 * i + 2] = rec.position.z;
        normal[3 * i] = rec.normal.x; normal[3 * i + 1] = rec.normal.y; normal[3 * i + 2] = rec.normal.z;
        front[i] = rec.front ? 1 : 0;
        pdf[i] = q;
        prim[i] = (int32_t)rec.prim;
    }
    return 0;
}

int orc_render(void* p, const prt_camera* c, const prt_render_params* prm, int rng_mode, int nthreads,
               double* rgb, orc_render_info* info) {
    Scene* s = (Scene*)p;
    if (!s || !c || !prm || !rgb) return PRT_ERR_INVALID_ARGUMENT;
    if (prm->sample_lights && s->lights_root < 0) return PRT_ERR_INVALID_ARGUMENT;
    Cam cam;
    cam.Init(c);
    const int W = cam.W, H = cam.H;
    int sb = prm->spp_begin, se = prm->spp_end;
    if (sb == 0 && se == 0) se = prm->samples_per_pixel;
    if (nthreads < 1) nthreads = 1;
    const double scale = 1.0 / prm->samples_per_pixel;  // Camera.cpp:83
    std::vector<Counters> cnts((size_t)nthreads);
    auto t0 = std::chrono::steady_clock::now();
    parallel_for(H, nthreads, [&](int64_t jb, int64_t je, int tid) {
        Integrator I{*s, prm->max_depth, prm->sample_lights != 0, prm->russian_roulette,
                     v3(prm->background[0], prm->background[1], prm->background[2]), Counters{}};
        tl_render_flags = prm->flags;
        Rng rng;
        rng.mode = rng_mode == ORC_RNG_COUNTER ? ORC_RNG_COUNTER : ORC_RNG_THREAD_LOCAL;
        rng.xs = orc_mix64(prm->rng_key + ORC_GOLDEN * (uint64_t)(jb + 1)) | 1ull;
        for (int64_t j = jb; j < je; ++j)
            for (int i = 0; i < W; ++i) {
                V3 ro, rd;
                cam.GetRay(i, (int)j, ro, rd);
                const uint32_t pix = (uint32_t)(j * W + i);
                V3 acc = v3(0, 0, 0);
                for (int smp = sb; smp < se; ++smp) {  // Camera.cpp:54-57
                    rng.seed = orc_stream_seed(prm->rng_key, pix, (uint32_t)smp);
                    rng.dim = 0;
                    if (prm->flags & PRT_RENDER_PIXEL_JITTER) cam.GetRayJittered(i, (int)j, rng, ro, rd);
                    acc = acc + I.RayColor(ro, rd, prm->max_depth, rng) * scale;
                }
                rgb[3 * (size_t)pix] = acc.x; rgb[3 * (size_t)pix + 1] = acc.y; rgb[3 * (size_t)pix + 2] = acc.z;
            }
        cnts[(size_t)tid].hits += I.cnt.hits;
        cnts[(size_t)tid].shadows += I.cnt.shadows;
        tl_render_flags = 0;
    });
    auto t1 = std::chrono::steady_clock::now();
    if (info) {
        info->world_hit_calls = 0; info->shadow_calls = 0;
        for (auto& k : cnts) { info->world_hit_calls += k.hits; info->shadow_calls += k.shadows; }
        info->paths = (uint64_t)W * H * (uint64_t)(se - sb);
        info->seconds = std::chrono::duration<double>(t1 - t0).count();
        info->threads = nthreads;
        info->reserved = 0;
    }
    return 0;
}

int orc_pixel_samples(void* p, const prt_camera* c, const prt_render_params* prm, int32_t px, int32_t py, double* out) {
    Scene* s = (Scene*)p;
    if (!s || !c || !prm || !out) return PRT_ERR_INVALID_ARGUMENT;
    Cam cam;
    cam.Init(c);
    int sb = prm->spp_begin, se = prm->spp_end;
    if (sb == 0 && se == 0) se = prm->samples_per_pixel;
    Integrator I{*s, prm->max_depth, prm->sample_lights != 0, prm->russian_roulette,
                 v3(prm->background[0], prm->background[1], prm->background[2]), Counters{}};
    V3 ro, rd;
    cam.GetRay(px, py, ro, rd);
    tl_render_flags = prm->flags;
    Rng rng;
    const uint32_t pix = (uint32_t)(py * cam.W + px);
    for (int smp = sb; smp < se; ++smp) {
        rng.seed = orc_stream_seed(prm->rng_key, pix, (uint32_t)smp);
        rng.dim = 0;
        if (prm->flags & PRT_RENDER_PIXEL_JITTER) cam.GetRayJittered(px, py, rng, ro, rd);
        V3 L = I.RayColor(ro, rd, prm->max_depth, rng);
        out[3 * (smp - sb)] = L.x; out[3 * (smp - sb) + 1] = L.y; out[3 * (smp - sb) + 2] = L.z;
    }
    tl_render_flags = 0;
    return 0;
}

int orc_triangle_record(void* p, int64_t prim, double* o) {
    Scene* s = (Scene*)p;
    if (!s || prim < 0 || (size_t)prim >= s->tris.size()) return PRT_ERR_INVALID_ARGUMENT;
    const Tri& T = s->tris[(size_t)prim];
    int k = 0;
    for (int i = 0; i < 3; ++i) o[k++] = T.normal[i];
    o[k++] = T.D;
    for (int i = 0; i < 3; ++i) o[k++] = T.w[i];
    for (int i = 0; i < 3; ++i) o[k++] = T.v[0][i];
    for (int i = 0; i < 3; ++i) o[k++] = T.e0[i];
    for (int i = 0; i < 3; ++i) o[k++] = T.e1[i];
    for (int i = 0; i < 3; ++i) o[k++] = T.tangent[i];
    o[k++] = T.area;
    return 0;
}

int orc_material_scatter(void* p, int32_t material, const double* ray_dir, const double* normal,
                         const double* tangent, const double* uv, const double* u, int32_t n_u,
                         double* attenuation, double* out_dir, int32_t* n_draws, int32_t* scattered) {
    Scene* s = (Scene*)p;
    if (!s || material < 0 || (size_t)material >= s->mats.size()) return PRT_ERR_INVALID_ARGUMENT;
    std::vector<int32_t> script((size_t)n_u);
    for (int i = 0; i < n_u; ++i) script[(size_t)i] = (int32_t)(u[i] * 2147483648.0);
    Rng rng;
    rng.script = script.data(); rng.script_n = n_u;
    HitRec rec;
    rec.position = v3(0, 0, 0);
    rec.time = 1.0;
    rec.tangent = v3(tangent[0], tangent[1], tangent[2]);
    rec.uv = V2{uv[0], uv[1]};
    rec.mat = material;
    V3 rd = v3(ray_dir[0], ray_dir[1], ray_dir[2]);
    rec.SetFaceNormal(rd, v3(normal[0], normal[1], normal[2]));
    const double nan = std::numeric_limits<double>::quiet_NaN();
    V3 att = v3(nan, nan, nan), od = v3(0, 0, 0);
    if (s->mats[(size_t)material].type == PRT_MAT_PHONG && (s->flags & ORC_DEFINE_UB)) att = v3(0, 0, 0);
    bool ok = mat_scatter(*s, s->mats[(size_t)material], rd, rec, att, od, rng);
    attenuation[0] = att.x; attenuation[1] = att.y; attenuation[2] = att.z;
    out_dir[0] = od.x; out_dir[1] = od.y; out_dir[2] = od.z;
    *n_draws = rng.script_i;
    *scattered = ok ? 1 : 0;
    return 0;
}

int orc_material_eval(void* p, int32_t material, const double* ray_dir, const double* normal,
                      const double* tangent, const double* uv, const double* wi_world, const double* u, int32_t n_u,
                      double* f) {
    Scene* s = (Scene*)p;
    if (!s || material < 0 || (size_t)material >= s->mats.size()) return PRT_ERR_INVALID_ARGUMENT;
    std::vector<int32_t> script((size_t)n_u);
    for (int i = 0; i < n_u; ++i) script[(size_t)i] = (int32_t)(u[i] * 2147483648.0);
    Rng rng;
    rng.script = script.data(); rng.script_n = n_u;
    HitRec rec;
    rec.position = v3(0, 0, 0);
    rec.time = 1.0;
    rec.tangent = v3(tangent[0], tangent[1], tangent[2]);
    rec.uv = V2{uv[0], uv[1]};
    rec.mat = material;
    V3 rd = v3(ray_dir[0], ray_dir[1], ray_dir[2]);
    rec.SetFaceNormal(rd, v3(normal[0], normal[1], normal[2]));
    EvalCtx c;  // Camera.cpp:160-165
    c.p = rec.position; c.uv = rec.uv; c.n = rec.normal; c.dpdus = rec.tangent;
    c.wo = WorldToLocal(-rd, rec);
    V3 localWi = WorldToLocal(v3(wi_world[0], wi_world[1], wi_world[2]), rec);
    V3 fr = mat_eval(*s, s->mats[(size_t)material], localWi, c, rng);
    f[0] = fr.x; f[1] = fr.y; f[2] = fr.z;
    return 0;
}

int orc_texture_value(void* p, int32_t texture, int64_t n, const double* uv, double* rgb) {
    Scene* s = (Scene*)p;
    if (!s || texture < 0 || (size_t)texture >= s->texs.size()) return PRT_ERR_INVALID_ARGUMENT;
    for (int64_t i = 0; i < n; ++i) {
        V3 c = s->texs[(size_t)texture].Value(uv[2 * i], uv[2 * i + 1]);
        rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    }
    return 0;
}

int orc_camera_ray(const prt_camera* c, int32_t i, int32_t j, double* origin, double* dir) {
    Cam cam;
    cam.Init(c);
    V3 o, d;
    cam.GetRay(i, j, o, d);
    origin[0] = o.x; origin[1] = o.y; origin[2] = o.z;
    dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
    return 0;
}

}  // extern "C"

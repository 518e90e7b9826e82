/*
 * scene_host.cpp — host flattener: triangle records, the reference's median-split BVHs (needed for
 * the tie rule and for light sampling whatever acceleration structure is traversed), the SAH BVH2
 * the extend / shadow kernels walk, and the tables for materials, textures and lights.
 *
 * Reference behaviour restated here (paths relative to /root/reference/):
 *   Triangle::Triangle            Source/Triangle.cpp:11-53      make_triangle()
 *   Triangle::SetBoundingBox      Source/Triangle.cpp:94-99      make_triangle()
 *   AABB ctors / PadToMinimus     Source/AABB.cpp:16-29,76-82    box_of_points(), box_union()
 *   BVHNode::BVHNode              Source/BVH.cpp:7-48,68-73      RefBuilder::build()
 *   world / lights assembly       main.cpp:36-45                 build_reference_bvhs()
 *   PhoneReflectance probabilities Source/Material.h:318-327     ingest()
 *   ImageTexture::SRGBToLinear    Source/Texture.cpp:66-71       ingest() (256-entry table)
 */
#include "scene_host.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <limits>
#include <memory>
#include <thread>

namespace prt {

namespace {

const double INF = std::numeric_limits<double>::infinity();

/* ---- Interval / AABB in the reference's own construction order --------------------------------- */
inline void interval_union(double amin, double amax, double bmin, double bmax, double& omin, double& omax) {
    omin = amin <= bmin ? amin : bmin; /* Interval.h:13-16 */
    omax = amax >= bmax ? amax : bmax;
}
inline Box6 box_empty() {
    Box6 b;
    for (int a = 0; a < 3; ++a) { b.v[2 * a] = +INF; b.v[2 * a + 1] = -INF; }
    return b;
}
inline Box6 box_union(const Box6& a, const Box6& b) { /* AABB.cpp:24-29: no re-padding */
    Box6 r;
    for (int k = 0; k < 3; ++k) interval_union(a.v[2 * k], a.v[2 * k + 1], b.v[2 * k], b.v[2 * k + 1], r.v[2 * k], r.v[2 * k + 1]);
    return r;
}
inline Box6 box_of_points(d3 a, d3 b) { /* AABB.cpp:16-22 + 76-82 */
    Box6 r;
    const double pa[3] = {a.x, a.y, a.z}, pb[3] = {b.x, b.y, b.z};
    for (int k = 0; k < 3; ++k) {
        double lo = pa[k] <= pb[k] ? pa[k] : pb[k];
        double hi = pa[k] <= pb[k] ? pb[k] : pa[k];
        const double delta = 0.0001;
        if (hi - lo < delta) { /* Interval::Expand, Interval.h:33-36 */
            const double padding = delta / 2.;
            lo = lo - padding;
            hi = hi + padding;
        }
        r.v[2 * k] = lo;
        r.v[2 * k + 1] = hi;
    }
    return r;
}
inline int longest_axis(const Box6& b) { /* AABB.cpp:66-74 */
    const double lx = b.v[1] - b.v[0], ly = b.v[3] - b.v[2], lz = b.v[5] - b.v[4];
    if (lx > ly) return lx > lz ? 0 : 2;
    return ly > lz ? 1 : 2;
}

/* Triangle::Triangle */
void make_triangle(const double* pv, const double* pn, const double* puv, TriIsect& I, TriShade& S, double& area, Box6& box) {
    d3 v0 = mk3(pv[0], pv[1], pv[2]), v1 = mk3(pv[3], pv[4], pv[5]), v2 = mk3(pv[6], pv[7], pv[8]);
    d2 uv0 = mk2(puv[0], puv[1]), uv1 = mk2(puv[2], puv[3]), uv2 = mk2(puv[4], puv[5]);
    d3 e0 = v1 - v0, e1 = v2 - v0;
    d3 n = cross(e0, e1);
    d3 normal = normalize(n);
    if (anynan(normal)) {
        d3 n0 = pn ? mk3(pn[0], pn[1], pn[2]) : mk3(0, 0, 0);
        d3 n1 = pn ? mk3(pn[3], pn[4], pn[5]) : mk3(0, 0, 0);
        d3 n2 = pn ? mk3(pn[6], pn[7], pn[8]) : mk3(0, 0, 0);
        normal = normalize(n0 + n1 + n2);
        if (anynan(normal)) normal = mk3(0.0, 0.0, 1.0);
    }
    d2 du0 = uv1 - uv0, du1 = uv2 - uv0;
    double f = 1.0 / (du0.x * du1.y - du1.x * du0.y);
    d3 tangent;
    tangent.x = f * (du1.y * e0.x - du0.y * e1.x);
    tangent.y = f * (du1.y * e0.y - du0.y * e1.y);
    tangent.z = f * (du1.y * e0.z - du0.y * e1.z);
    tangent = normalize(tangent);
    if (anynan(tangent)) { /* `abs(v.x) < 0.9f`: the literal is a float promoted to double */
        d3 helper = (std::fabs(normal.x) < (double)0.9f) ? mk3(1, 0, 0) : mk3(0, 1, 0);
        tangent = normalize(cross(normal, helper));
    }
    area = length(n) * 0.5;
    double D = dot(normal, v0);
    d3 w = n / dot(n, n);
    box = box_union(box_of_points(v0, v1), box_of_points(v0, v2));

    I.n[0] = normal.x; I.n[1] = normal.y; I.n[2] = normal.z;
    I.D = D;
    I.w[0] = w.x; I.w[1] = w.y; I.w[2] = w.z;
    I.v0[0] = v0.x; I.v0[1] = v0.y; I.v0[2] = v0.z;
    I.e0[0] = e0.x; I.e0[1] = e0.y; I.e0[2] = e0.z;
    I.e1[0] = e1.x; I.e1[1] = e1.y; I.e1[2] = e1.z;
    for (int k = 0; k < 6; ++k) S.uv[k] = puv[k];
    S.tangent[0] = tangent.x; S.tangent[1] = tangent.y; S.tangent[2] = tangent.z;
    S.normal[0] = normal.x; S.normal[1] = normal.y; S.normal[2] = normal.z;
    S.material = 0;
    for (int k = 0; k < 7; ++k) S.pad[k] = 0;
}

void parallel_ranges(int64_t n, int n_threads, const std::function<void(int64_t, int64_t)>& body);

}  // namespace

int default_host_threads() {
    if (const char* e = std::getenv("PRT_HOST_THREADS")) { /* every host-side build gives the same result for any count */
        const int v = std::atoi(e);
        if (v >= 1) return v > 32 ? 32 : v;
    }
    unsigned h = std::thread::hardware_concurrency();
    if (h == 0) h = 4;
    if (h > 32) h = 32;
    return (int)h;
}

namespace {
void parallel_ranges(int64_t n, int n_threads, const std::function<void(int64_t, int64_t)>& body) {
    if (n_threads <= 1 || n < 65536) { body(0, n); return; }
    std::vector<std::thread> th;
    const int64_t chunk = (n + n_threads - 1) / n_threads;
    for (int k = 0; k < n_threads; ++k) {
        const int64_t b = k * chunk, e = std::min(n, b + chunk);
        if (b >= e) break;
        th.emplace_back([=, &body]() { body(b, e); });
    }
    for (auto& t : th) t.join();
}
}  // namespace

/* ------------------------------------------------------------------------------------------------
 * ingest: copy the description, derive per-triangle records, material / texture tables
 * ---------------------------------------------------------------------------------------------- */
bool HostScene::ingest(const prt_scene_desc* d) {
    if (!d || d->n_triangles < 0 || d->n_meshes < 0 || d->n_materials < 0 || d->n_textures < 0) {
        error = "prt_scene_create: negative count in prt_scene_desc";
        return false;
    }
    if (d->n_triangles >= (int64_t(1) << 27)) {
        error = "prt_scene_create: more than 2^27 triangles";
        return false;
    }
    if (d->n_triangles > 0 && (!d->vertices || !d->mesh_first || !d->mesh_material || d->n_meshes == 0)) {
        error = "prt_scene_create: triangles without vertices / mesh tables";
        return false;
    }
    n_tris = d->n_triangles;
    mesh_first.assign(d->mesh_first, d->mesh_first + (d->n_meshes ? d->n_meshes + 1 : 0));
    mesh_material.assign(d->mesh_material, d->mesh_material + d->n_meshes);
    if (d->n_meshes) {
        if (mesh_first.front() != 0 || mesh_first.back() != n_tris) {
            error = "prt_scene_create: mesh_first must run from 0 to n_triangles";
            return false;
        }
        for (int m = 0; m < d->n_meshes; ++m) {
            if (mesh_first[m + 1] < mesh_first[m]) { error = "prt_scene_create: mesh_first not monotone"; return false; }
            if (mesh_material[m] < 0 || mesh_material[m] >= d->n_materials) { error = "prt_scene_create: mesh_material out of range"; return false; }
        }
    }
    materials.assign(d->materials, d->materials + d->n_materials);
    vertices.assign(d->vertices, d->vertices + 9 * n_tris);
    texcoords.resize((size_t)6 * n_tris);
    const double uv_default[6] = {0, 0, 1, 0, 1, 1}; /* Model.cpp:170-175 */
    for (int64_t f = 0; f < n_tris; ++f)
        std::memcpy(&texcoords[6 * f], d->texcoords ? d->texcoords + 6 * f : uv_default, 6 * sizeof(double));

    /* textures: texel pool + the sRGB table. GetPixel (Texture.cpp:50-65) feeds SRGBToLinear with
     * colorScale * byte; the 256 possible results are computed here with the host libm `pow`,
     * which is the very function the reference calls, so the table holds the reference's bits. */
    const double colorScale = 1.0 / 255.0;
    for (int i = 0; i < 256; ++i) {
        const double v = colorScale * (double)i;
        srgb_lut[i] = (v <= 0.04045) ? v * (1. / 12.92) : std::pow((v + 0.055) * (1. / 1.055), 2.4);
    }
    textures.clear();
    dev_textures.clear();
    texel_pool.clear();
    for (int i = 0; i < d->n_textures; ++i) {
        const prt_texture& pt = d->textures[i];
        HostTexture t;
        DevTexture dt;
        dt.offset = (int64_t)texel_pool.size();
        dt.width = dt.height = dt.channels = dt.valid = 0;
        if (pt.texels && pt.width > 0 && pt.height > 0 && pt.channels > 0) {
            t.width = pt.width; t.height = pt.height; t.channels = pt.channels;
            const size_t nb = (size_t)pt.width * pt.height * pt.channels;
            t.texels.assign(pt.texels, pt.texels + nb);
            texel_pool.insert(texel_pool.end(), t.texels.begin(), t.texels.end());
            while (texel_pool.size() % 16) texel_pool.push_back(0);
            dt.width = pt.width; dt.height = pt.height; dt.channels = pt.channels; dt.valid = 1;
        }
        textures.push_back(std::move(t));
        dev_textures.push_back(dt);
    }

    dev_materials.clear();
    for (int i = 0; i < d->n_materials; ++i) {
        const prt_material& pm = materials[i];
        DevMaterial m;
        std::memset(&m, 0, sizeof(m));
        for (int k = 0; k < 3; ++k) { m.kd[k] = pm.kd[k]; m.ks[k] = pm.ks[k]; }
        m.type = pm.type;
        m.kd_tex = (pm.kd_tex >= 0 && pm.kd_tex < d->n_textures) ? pm.kd_tex : -1;
        m.ns = pm.ns;
        m.alpha_x = pm.alpha_x; m.alpha_y = pm.alpha_y;
        for (int k = 0; k < 3; ++k) { m.eta[k] = pm.eta[k]; m.k[k] = pm.k[k]; }
        if (pm.ns <= 9.) { m.pkd = 1.0; m.pks = 0.0; } else { m.pkd = 0.6; m.pks = 0.4; } /* Material.h:318-327 */
        m.has_emission = (pm.type == PRT_MAT_DIFFUSE_LIGHT || pm.type == PRT_MAT_DEBUG) ? 1 : 0; /* :166,528 */
        for (int k = 0; k < 3; ++k) m.emission[k] = pm.type == PRT_MAT_DEBUG ? pm.kd[k] : (pm.type == PRT_MAT_DIFFUSE_LIGHT ? pm.emission[k] : 0.0);
        if (pm.type == PRT_MAT_PHONG) m.skip_light_sampling = pm.ns > 1. ? 1 : 0; /* Material.h:328 */
        else m.skip_light_sampling = (pm.type == PRT_MAT_MIRROR || pm.type == PRT_MAT_EMPTY) ? 1 : 0; /* :365,539 */
        dev_materials.push_back(m);
    }

    isect.resize((size_t)n_tris);
    shade.resize((size_t)n_tris);
    area.resize((size_t)n_tris);
    tri_box.resize((size_t)n_tris);
    const double* pn = d->normals;
    parallel_ranges(n_tris, default_host_threads(), [&](int64_t b, int64_t e) {
        for (int64_t f = b; f < e; ++f)
            make_triangle(&vertices[9 * f], pn ? pn + 9 * f : nullptr, &texcoords[6 * f], isect[f], shade[f], area[f], tri_box[f]);
    });
    for (size_t m = 0; m + 1 < mesh_first.size(); ++m)
        for (int64_t f = mesh_first[m]; f < mesh_first[m + 1]; ++f) shade[f].material = mesh_material[m];
    return true;
}

/* ------------------------------------------------------------------------------------------------
 * the reference's BVH (median split on the longest axis by bbox.min, std::sort) — BVH.cpp:7-48
 * ---------------------------------------------------------------------------------------------- */
namespace {

struct KeyRef {
    double key;
    ObjRef ref;
};

int64_t ref_node_count(int64_t span) { /* nodes a build over `span` objects creates */
    if (span <= 2) return 1;
    return 1 + ref_node_count(span / 2) + ref_node_count(span - span / 2);
}

struct RefBuilder {
    HostScene& s;
    std::vector<KeyRef>& items;
    RefBuilder(HostScene& s, std::vector<KeyRef>& items) : s(s), items(items) {}

    const Box6& box_of(ObjRef o) const { return o.kind == 0 ? s.tri_box[o.idx] : s.ref_nodes[o.idx].box; }
    double area_of(ObjRef o) const { return o.kind == 0 ? s.area[o.idx] : s.ref_nodes[o.idx].area; }

    /* builds the node `at` over items[start,end); nodes of the subtree occupy [at, at+count) */
    void build(int64_t start, int64_t end, int32_t at, int par_depth) {
        HostRefNode node;
        node.box = box_empty();
        for (int64_t i = start; i < end; ++i) node.box = box_union(node.box, box_of(items[i].ref));
        const int axis = longest_axis(node.box);
        const int64_t span = end - start;
        if (span == 1) {
            node.left = node.right = items[start].ref;
            node.area = area_of(items[start].ref);
        } else if (span == 2) {
            node.left = items[start].ref;
            node.right = items[start + 1].ref;
            node.area = area_of(items[start].ref) + area_of(items[start + 1].ref);
        } else {
            for (int64_t i = start; i < end; ++i) items[i].key = box_of(items[i].ref).v[2 * axis];
            /* the same libstdc++ std::sort, driven by the same comparison outcomes as
             * BVH.cpp:31,68-73, leaves the same permutation */
            std::sort(items.begin() + start, items.begin() + end, [](const KeyRef& a, const KeyRef& b) { return a.key < b.key; });
            const int64_t mid = start + span / 2;
            const int32_t l = at + 1;
            const int32_t r = (int32_t)(at + 1 + ref_node_count(mid - start));
            if (par_depth > 0 && span > 32768) {
                std::thread t([&, l]() { build(start, mid, l, par_depth - 1); });
                build(mid, end, r, par_depth - 1);
                t.join();
            } else {
                build(start, mid, l, 0);
                build(mid, end, r, 0);
            }
            node.left = ObjRef{1, l};
            node.right = ObjRef{1, r};
            node.area = s.ref_nodes[l].area + s.ref_nodes[r].area;
        }
        s.ref_nodes[at] = node;
    }
};

/* builds a tree over `items` into ref_nodes[at, at + ref_node_count(size)), which the caller has already made room for */
void build_ref_tree_at(HostScene& s, std::vector<KeyRef>& items, int32_t at, int n_threads) {
    int par = 0;
    while ((1 << par) < n_threads) ++par;
    RefBuilder b(s, items);
    b.build(0, (int64_t)items.size(), at, par + 1);
}
int32_t build_ref_tree(HostScene& s, std::vector<KeyRef>& items, int n_threads) {
    const int32_t at = (int32_t)s.ref_nodes.size();
    s.ref_nodes.resize(s.ref_nodes.size() + (size_t)ref_node_count((int64_t)items.size()));
    build_ref_tree_at(s, items, at, n_threads);
    return at;
}

void walk_ref_leaves(const HostScene& s, ObjRef o, std::vector<int32_t>& out) {
    /* iterative DFS, left before right; a span-1 node (left == right) contributes its leaf once */
    std::vector<ObjRef> stack{o};
    while (!stack.empty()) {
        ObjRef c = stack.back();
        stack.pop_back();
        if (c.kind == 0) { out.push_back(c.idx); continue; }
        const HostRefNode& n = s.ref_nodes[c.idx];
        if (!(n.right.kind == n.left.kind && n.right.idx == n.left.idx)) stack.push_back(n.right);
        stack.push_back(n.left);
    }
}

}  // namespace

void HostScene::build_reference_bvhs(int n_threads) {
    ref_nodes.clear();
    world_root = lights_root = -1;
    std::vector<ObjRef> world, lights;
    const int n_meshes = (int)mesh_material.size();
    /* one tree per mesh (main.cpp:37-43), a second one for an emissive mesh. The trees are independent, so they are built
     * side by side on the host threads; every tree's place in ref_nodes is fixed beforehand, in mesh order, so the
     * node numbering does not depend on which thread finishes first */
    struct MeshJob { int m; int32_t at_world, at_light; };
    std::vector<MeshJob> jobs;
    size_t total = 0;
    for (int m = 0; m < n_meshes; ++m) {
        if (mesh_first[m + 1] <= mesh_first[m]) continue;
        const int64_t cnt = ref_node_count(mesh_first[m + 1] - mesh_first[m]);
        MeshJob j{m, (int32_t)total, -1};
        total += (size_t)cnt;
        if (dev_materials[mesh_material[m]].has_emission) { j.at_light = (int32_t)total; total += (size_t)cnt; }
        jobs.push_back(j);
    }
    ref_nodes.resize(total);
    {
        std::vector<size_t> by_size(jobs.size());
        for (size_t i = 0; i < jobs.size(); ++i) by_size[i] = i;
        std::sort(by_size.begin(), by_size.end(), [&](size_t a, size_t b) { /* largest first: the long jobs start early */
            const int64_t sa = mesh_first[jobs[a].m + 1] - mesh_first[jobs[a].m], sb = mesh_first[jobs[b].m + 1] - mesh_first[jobs[b].m];
            return sa != sb ? sa > sb : a < b;
        });
        std::atomic<size_t> next_job{0};
        auto worker = [&]() {
            for (;;) {
                const size_t k = next_job.fetch_add(1);
                if (k >= jobs.size()) return;
                const MeshJob& j = jobs[by_size[k]];
                std::vector<KeyRef> items;
                items.reserve((size_t)(mesh_first[j.m + 1] - mesh_first[j.m]));
                for (int64_t f = mesh_first[j.m]; f < mesh_first[j.m + 1]; ++f) items.push_back(KeyRef{0.0, ObjRef{0, (int32_t)f}});
                /* BVHNode(shared_ptr<Mesh>) sorts mesh->objects IN PLACE (BVH.cpp:49): the second build
                 * for an emissive mesh starts from the order the first one left behind */
                build_ref_tree_at(*this, items, j.at_world, n_threads);
                if (j.at_light >= 0) build_ref_tree_at(*this, items, j.at_light, n_threads);
            }
        };
        const int n_workers = (int)std::min<size_t>((size_t)std::max(1, n_threads), jobs.size());
        std::vector<std::thread> pool;
        for (int t = 1; t < n_workers; ++t) pool.emplace_back(worker);
        worker();
        for (auto& t : pool) t.join();
    }
    for (const MeshJob& j : jobs) {
        world.push_back(ObjRef{1, j.at_world});
        if (j.at_light >= 0) lights.push_back(ObjRef{1, j.at_light});
    }
    auto top = [&](std::vector<ObjRef>& list) -> int32_t { /* main.cpp:44-45 */
        if (list.empty()) return -1;
        std::vector<KeyRef> items;
        for (ObjRef o : list) items.push_back(KeyRef{0.0, o});
        return build_ref_tree(*this, items, 1);
    };
    world_root = top(world);
    lights_root = top(lights);

    world_rank.assign((size_t)n_tris, -1);
    if (world_root >= 0) {
        std::vector<int32_t> leaves;
        leaves.reserve((size_t)n_tris);
        walk_ref_leaves(*this, ObjRef{1, world_root}, leaves);
        for (size_t r = 0; r < leaves.size(); ++r) world_rank[leaves[r]] = (int64_t)r;
    }

    /* lights: flatten the lights BVH for TraverseSample (BVH.cpp:86-100) */
    light_order.clear();
    light_tris.clear();
    light_nodes.clear();
    light_root_flat = -1;
    lights_area = 0.0;
    if (lights_root >= 0) {
        std::vector<int32_t> leaves;
        walk_ref_leaves(*this, ObjRef{1, lights_root}, leaves);
        std::vector<int32_t> light_index_of_prim((size_t)n_tris, -1);
        for (size_t i = 0; i < leaves.size(); ++i) {
            const int32_t p = leaves[i];
            light_order.push_back(p);
            light_index_of_prim[p] = (int32_t)i;
            LightTri L;
            std::memset(&L, 0, sizeof(L));
            for (int k = 0; k < 3; ++k) {
                L.v0[k] = vertices[9 * (size_t)p + k];
                L.v1[k] = vertices[9 * (size_t)p + 3 + k];
                L.v2[k] = vertices[9 * (size_t)p + 6 + k];
                L.normal[k] = shade[p].normal[k];
            }
            L.area = area[p];
            L.material = shade[p].material;
            L.prim = p;
            light_tris.push_back(L);
        }
        /* nodes reachable from lights_root, renumbered in DFS order */
        std::vector<int32_t> flat_of((size_t)ref_nodes.size(), -1);
        std::vector<int32_t> order;
        std::vector<int32_t> stack{lights_root};
        while (!stack.empty()) {
            int32_t c = stack.back();
            stack.pop_back();
            if (flat_of[c] >= 0) continue;
            flat_of[c] = (int32_t)order.size();
            order.push_back(c);
            const HostRefNode& n = ref_nodes[c];
            if (n.right.kind == 1) stack.push_back(n.right.idx);
            if (n.left.kind == 1) stack.push_back(n.left.idx);
        }
        light_nodes.resize(order.size());
        auto ref_of = [&](ObjRef o) -> int32_t { return o.kind == 1 ? flat_of[o.idx] : ~light_index_of_prim[o.idx]; };
        for (size_t i = 0; i < order.size(); ++i) {
            const HostRefNode& n = ref_nodes[order[i]];
            LightNode ln;
            ln.left_area = n.left.kind == 0 ? area[n.left.idx] : ref_nodes[n.left.idx].area;
            ln.left = ref_of(n.left);
            ln.right = ref_of(n.right);
            light_nodes[i] = ln;
        }
        light_root_flat = 0;
        lights_area = ref_nodes[lights_root].area;
    }
}

/* ------------------------------------------------------------------------------------------------
 * SAH BVH2 (binned, top-down), FP32 child boxes rounded outward
 * ---------------------------------------------------------------------------------------------- */
namespace {

inline float round_down(double x) {
    float f = (float)x;
    if ((double)f > x) f = std::nextafterf(f, -std::numeric_limits<float>::infinity());
    return f;
}
inline float round_up(double x) {
    float f = (float)x;
    if ((double)f < x) f = std::nextafterf(f, std::numeric_limits<float>::infinity());
    return f;
}
inline double half_area(const Box6& b) {
    const double dx = b.v[1] - b.v[0], dy = b.v[3] - b.v[2], dz = b.v[5] - b.v[4];
    return dx * dy + dy * dz + dz * dx;
}
inline void box_grow(Box6& a, const Box6& b) {
    for (int k = 0; k < 3; ++k) {
        if (b.v[2 * k] < a.v[2 * k]) a.v[2 * k] = b.v[2 * k];
        if (b.v[2 * k + 1] > a.v[2 * k + 1]) a.v[2 * k + 1] = b.v[2 * k + 1];
    }
}

struct TmpNode {
    Box6 box;
    int32_t left = -1, right = -1; /* TmpNode indices; -1/-1 = leaf */
    int32_t first = 0, count = 0;  /* leaf: range in the index array */
};

/* what the builder permutes: a triangle's box travels with its index, so that every pass over a node's triangles is a
 * sequential read (an index array alone made each pass a gather from the box array: one cache miss per triangle and level) */
struct BoxRef {
    Box6 box;
    int32_t prim;
};

struct SahBuilder {
    const HostScene& s;
    std::vector<BoxRef>& work;
    std::vector<TmpNode>& nodes;
    std::atomic<int32_t>& next;
    int max_leaf;
    static constexpr int NB_MAX = 64;
    static constexpr int32_t PAR_BINS = 131072; /* nodes of at least this many triangles bin on several host threads */
    int n_threads = 1;
    int NB = 16;           /* bins per axis, at most NB_MAX (build_sah_tree picks; PRT_SAH_BINS overrides) */
    int sweep_below = 0;   /* PRT_SAH_SWEEP_BELOW (experiments): subtrees of fewer triangles try EVERY split position */
    double C_TRI = 1.25, C_NODE = 1.0; /* PRT_SAH_CTRI overrides C_TRI (experiments) */

    int32_t alloc() { return next.fetch_add(1); }

    /* pre_box / pre_cbox: the node's box and the box of its centroids where the parent already knows them (the unions of
     * the bins on either side of its split: min / max are exact, so they are the very values the pass below would find) */
    void build(int32_t at, int32_t first, int32_t count, int par_depth, const Box6* pre_box = nullptr, const Box6* pre_cbox = nullptr) {
        TmpNode& N = nodes[at];
        Box6 box = box_empty(), cbox = box_empty();
        if (pre_box && pre_cbox) {
            box = *pre_box;
            cbox = *pre_cbox;
        } else {
            for (int32_t i = first; i < first + count; ++i) {
                const Box6& b = work[i].box;
                box_grow(box, b);
                for (int k = 0; k < 3; ++k) {
                    const double c = 0.5 * (b.v[2 * k] + b.v[2 * k + 1]);
                    if (c < cbox.v[2 * k]) cbox.v[2 * k] = c;
                    if (c > cbox.v[2 * k + 1]) cbox.v[2 * k + 1] = c;
                }
            }
        }
        N.box = box;
        N.first = first;
        N.count = count;
        if (count == 1) return;

        int best_axis = -1, best_bin = -1;
        double best_cost = std::numeric_limits<double>::infinity();
        const double parent_area = half_area(box);
        int32_t sweep_mid = -1; /* >= 0: idx[first, first + count) is already ordered for the best sweep split */
        Box6 child_box[2], child_cbox[2];
        bool child_known = false;
        if (count < sweep_below) {
            /* full sweep: order by centroid on each axis, every one of the count - 1 split positions is a candidate */
            std::vector<int32_t> order((size_t)count), best_order;
            for (int32_t i = 0; i < count; ++i) order[i] = work[first + i].prim;
            std::vector<double> right((size_t)count);
            for (int axis = 0; axis < 3; ++axis) {
                std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
                    const Box6 &A = s.tri_box[a], &B = s.tri_box[b];
                    return A.v[2 * axis] + A.v[2 * axis + 1] < B.v[2 * axis] + B.v[2 * axis + 1];
                });
                Box6 acc = box_empty();
                for (int32_t i = count - 1; i > 0; --i) { box_grow(acc, s.tri_box[order[i]]); right[i] = half_area(acc); }
                acc = box_empty();
                for (int32_t i = 0; i < count - 1; ++i) {
                    box_grow(acc, s.tri_box[order[i]]);
                    const double cost = half_area(acc) * (i + 1) + right[i + 1] * (count - 1 - i);
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; sweep_mid = first + i + 1; }
                }
                if (best_axis == axis) best_order = order;
            }
            if (sweep_mid >= 0)
                for (int32_t i = 0; i < count; ++i) work[first + i] = BoxRef{s.tri_box[best_order[i]], best_order[i]};
        } else {
            /* ONE pass over the triangles fills the bins of all three axes (tri_box[idx[i]] is a random access: three
             * passes were three cache misses per triangle and level); above PAR_BINS triangles the pass is cut into
             * ranges for the host threads, whose bins are merged in range order — counts add, boxes take min / max, so
             * the merged bins, and with them the tree, do not depend on the number of threads */
            double lo3[3], scale3[3];
            bool use[3];
            for (int axis = 0; axis < 3; ++axis) {
                lo3[axis] = cbox.v[2 * axis];
                const double ext = cbox.v[2 * axis + 1] - lo3[axis];
                use[axis] = ext > 0.0;
                scale3[axis] = use[axis] ? NB / ext : 0.0;
            }
            struct Bins {
                int cnt[3][NB_MAX];
                Box6 bb[3][NB_MAX]; /* union of the triangle boxes of the bin */
                Box6 cb[3][NB_MAX]; /* box of their centroids */
            };
            /* only the counts are cleared (768 bytes): a bin's boxes are initialised when its first triangle arrives, and
             * the scans below skip empty bins. Most nodes of a tree hold a handful of triangles; clearing and scanning
             * 3 x 64 boxes for each of them was most of the build time (5.5 GB of stores for 1.1 M triangles). */
            auto clear_bins = [&](Bins& B) {
                for (int axis = 0; axis < 3; ++axis)
                    for (int b = 0; b < NB; ++b) B.cnt[axis][b] = 0;
            };
            auto fill_bins = [&](int32_t from, int32_t to, Bins& B) {
                for (int32_t i = from; i < to; ++i) {
                    const Box6& b = work[i].box;
                    const double c3[3] = {0.5 * (b.v[0] + b.v[1]), 0.5 * (b.v[2] + b.v[3]), 0.5 * (b.v[4] + b.v[5])};
                    for (int axis = 0; axis < 3; ++axis) {
                        if (!use[axis]) continue;
                        int bin = (int)((c3[axis] - lo3[axis]) * scale3[axis]);
                        if (bin < 0) bin = 0;
                        if (bin >= NB) bin = NB - 1;
                        if (B.cnt[axis][bin]++ == 0) { B.bb[axis][bin] = box_empty(); B.cb[axis][bin] = box_empty(); }
                        box_grow(B.bb[axis][bin], b);
                        Box6& cbn = B.cb[axis][bin];
                        for (int k = 0; k < 3; ++k) {
                            if (c3[k] < cbn.v[2 * k]) cbn.v[2 * k] = c3[k];
                            if (c3[k] > cbn.v[2 * k + 1]) cbn.v[2 * k + 1] = c3[k];
                        }
                    }
                }
            };
            static thread_local Bins tl_bins; /* done with before this call recurses */
            static thread_local bool tl_clean = false; /* every count of tl_bins is zero */
            Bins* const bins = &tl_bins;
            if (!tl_clean) { clear_bins(*bins); tl_clean = true; }
            /* non-empty bins per axis, ascending. A node of a few triangles (most nodes) keeps the list as it bins and
             * touches nothing else: no clearing and no scan of 3 x 64 bins for a handful of triangles */
            constexpr int SPARSE_MAX = 16;
            const bool sparse = count <= SPARSE_MAX;
            int ids[3][NB_MAX], n_ids[3] = {0, 0, 0};
            const int helpers = count >= PAR_BINS ? std::min(n_threads, 16) : 1;
            if (helpers > 1 && std::getenv("PRT_DEBUG_BUILD")) std::fprintf(stderr, "[prt] big node: %d triangles\n", count);
            if (sparse) {
                for (int32_t i = first; i < first + count; ++i) {
                    const Box6& bx = work[i].box;
                    const double c3[3] = {0.5 * (bx.v[0] + bx.v[1]), 0.5 * (bx.v[2] + bx.v[3]), 0.5 * (bx.v[4] + bx.v[5])};
                    for (int axis = 0; axis < 3; ++axis) {
                        if (!use[axis]) continue;
                        int bin = (int)((c3[axis] - lo3[axis]) * scale3[axis]);
                        if (bin < 0) bin = 0;
                        if (bin >= NB) bin = NB - 1;
                        if (bins->cnt[axis][bin]++ == 0) {
                            bins->bb[axis][bin] = box_empty();
                            bins->cb[axis][bin] = box_empty();
                            int at_ = n_ids[axis]++;
                            while (at_ > 0 && ids[axis][at_ - 1] > bin) { ids[axis][at_] = ids[axis][at_ - 1]; --at_; }
                            ids[axis][at_] = bin;
                        }
                        box_grow(bins->bb[axis][bin], bx);
                        Box6& cbn = bins->cb[axis][bin];
                        for (int k = 0; k < 3; ++k) {
                            if (c3[k] < cbn.v[2 * k]) cbn.v[2 * k] = c3[k];
                            if (c3[k] > cbn.v[2 * k + 1]) cbn.v[2 * k + 1] = c3[k];
                        }
                    }
                }
            } else if (helpers > 1) {
                std::vector<std::unique_ptr<Bins>> part((size_t)helpers);
                std::vector<std::thread> th;
                for (int t = 0; t < helpers; ++t) {
                    part[t].reset(new Bins);
                    clear_bins(*part[t]);
                    const int32_t from = first + (int32_t)((int64_t)count * t / helpers), to = first + (int32_t)((int64_t)count * (t + 1) / helpers);
                    th.emplace_back([&, t, from, to] { fill_bins(from, to, *part[t]); });
                }
                for (auto& t : th) t.join();
                for (int t = 0; t < helpers; ++t)
                    for (int axis = 0; axis < 3; ++axis)
                        for (int b = 0; b < NB; ++b) {
                            if (part[t]->cnt[axis][b] == 0) continue;
                            if (bins->cnt[axis][b] == 0) { bins->bb[axis][b] = box_empty(); bins->cb[axis][b] = box_empty(); }
                            bins->cnt[axis][b] += part[t]->cnt[axis][b];
                            box_grow(bins->bb[axis][b], part[t]->bb[axis][b]);
                            box_grow(bins->cb[axis][b], part[t]->cb[axis][b]);
                        }
            } else {
                fill_bins(first, first + count, *bins);
            }
            if (!sparse)
                for (int axis = 0; axis < 3; ++axis)
                    for (int b = 0; b < NB; ++b)
                        if (use[axis] && bins->cnt[axis][b]) ids[axis][n_ids[axis]++] = b;
            /* candidates: a split after every non-empty bin but the last. (Between two non-empty bins the split is the same
             * two sets at the same cost; the strict comparison keeps the first, which is the one after the non-empty bin.) */
            for (int axis = 0; axis < 3; ++axis) {
                if (!use[axis]) continue;
                const int* cnt = bins->cnt[axis];
                const Box6* bb = bins->bb[axis];
                const int m = n_ids[axis];
                double right_area[NB_MAX];
                int right_cnt[NB_MAX];
                Box6 acc = box_empty();
                int c = 0;
                for (int j = m - 1; j > 0; --j) {
                    box_grow(acc, bb[ids[axis][j]]);
                    c += cnt[ids[axis][j]];
                    right_area[j] = half_area(acc);
                    right_cnt[j] = c;
                }
                acc = box_empty();
                c = 0;
                for (int j = 0; j + 1 < m; ++j) {
                    const int bnow = ids[axis][j];
                    box_grow(acc, bb[bnow]);
                    c += cnt[bnow];
                    const double cost = half_area(acc) * c + right_area[j + 1] * right_cnt[j + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = bnow; }
                }
            }
            if (best_axis >= 0) { /* what the two children will hold, known without another pass over them */
                for (int side = 0; side < 2; ++side) { child_box[side] = box_empty(); child_cbox[side] = box_empty(); }
                for (int j = 0; j < n_ids[best_axis]; ++j) {
                    const int bnow = ids[best_axis][j];
                    const int side = bnow <= best_bin ? 0 : 1;
                    box_grow(child_box[side], bins->bb[best_axis][bnow]);
                    box_grow(child_cbox[side], bins->cb[best_axis][bnow]);
                }
                child_known = true;
            }
            if (sparse) { /* leave the counts as they were found: zero */
                for (int axis = 0; axis < 3; ++axis)
                    for (int j = 0; j < n_ids[axis]; ++j) bins->cnt[axis][ids[axis][j]] = 0;
            } else {
                tl_clean = false;
            }
        }
        if (count <= max_leaf) {
            /* SAH termination: a leaf of `count` triangles against one more node + the split */
            const double leaf_cost = C_TRI * count;
            const double split_cost = best_axis >= 0 && parent_area > 0.0 ? C_NODE + C_TRI * best_cost / parent_area
                                                                          : std::numeric_limits<double>::infinity();
            if (leaf_cost <= split_cost) return;
        }
        int32_t mid;
        if (sweep_mid >= 0) {
            mid = sweep_mid;
        } else if (best_axis >= 0) {
            const double lo = cbox.v[2 * best_axis], ext = cbox.v[2 * best_axis + 1] - lo;
            const double scale = NB / ext;
            auto it = std::partition(work.begin() + first, work.begin() + first + count, [&](const BoxRef& p) {
                const Box6& b = p.box;
                const double c = 0.5 * (b.v[2 * best_axis] + b.v[2 * best_axis + 1]);
                int bin = (int)((c - lo) * scale);
                if (bin < 0) bin = 0;
                if (bin >= NB) bin = NB - 1;
                return bin <= best_bin;
            });
            mid = (int32_t)(it - work.begin());
        } else {
            mid = first + count / 2; /* coincident centroids: split by position */
        }
        if (mid == first || mid == first + count) { mid = first + count / 2; child_known = false; }
        const int32_t l = alloc(), r = alloc();
        nodes[at].left = l;
        nodes[at].right = r;
        const Box6 *lb = child_known ? &child_box[0] : nullptr, *lc = child_known ? &child_cbox[0] : nullptr;
        const Box6 *rb = child_known ? &child_box[1] : nullptr, *rcb = child_known ? &child_cbox[1] : nullptr;
        const int32_t nl = mid - first, nr = first + count - mid;
        /* a thread per split only where BOTH sides are worth one: SAH peels small pieces off a large rest (walls off a
         * room), and a level of the thread budget spent on a 12-triangle side is a level the large rest does not get */
        if (par_depth > 0 && std::min(nl, nr) > 32768) {
            std::thread t([&, l]() { build(l, first, nl, par_depth - 1, lb, lc); });
            build(r, mid, nr, par_depth - 1, rb, rcb);
            t.join();
        } else {
            build(l, first, nl, nl > 65536 ? par_depth : 0, lb, lc);
            build(r, mid, nr, nr > 65536 ? par_depth : 0, rb, rcb);
        }
    }
};

inline void set_child_box(float* lo, float* hi, const Box6& b) {
    for (int k = 0; k < 3; ++k) { lo[k] = round_down(b.v[2 * k]); hi[k] = round_up(b.v[2 * k + 1]); }
}

}  // namespace

/* Tree rotations (Kensler 2008) on the finished binary tree: at every inner node, exchanging one child with a grandchild
 * on the other side leaves the node's own box as it is and changes ONE box, the other child's; the exchange that shrinks
 * that box most is applied if it shrinks it at all. Subtrees move whole, so every box stays the exact union of its
 * leaves. Post-order, a few passes. PRT_SAH_ROTATE = number of passes (experiments; 0 = off). */
static int rotate_tree(std::vector<TmpNode>& t, int passes) {
    std::vector<int32_t> order, stack;
    stack.push_back(0);
    while (!stack.empty()) { /* pre-order, reversed below: children before parents */
        const int32_t n = stack.back();
        stack.pop_back();
        if (t[n].left < 0) continue;
        order.push_back(n);
        stack.push_back(t[n].left);
        stack.push_back(t[n].right);
    }
    int applied = 0;
    auto merged_area = [&](int32_t a, int32_t b) {
        Box6 u = t[a].box;
        box_grow(u, t[b].box);
        return half_area(u);
    };
    for (int pass = 0; pass < passes; ++pass) {
        int changed = 0;
        for (size_t k = order.size(); k-- > 0;) {
            TmpNode& N = t[order[k]];
            int best_side = -1, best_opt = 0;
            double best_gain = 0.0;
            for (int side = 0; side < 2; ++side) {
                const int32_t c = side ? N.right : N.left, other = side ? N.left : N.right;
                if (t[c].left < 0) continue;
                const double a_c = half_area(t[c].box);
                const double g0 = a_c - merged_area(other, t[c].right); /* other <-> c.left  */
                const double g1 = a_c - merged_area(t[c].left, other);  /* other <-> c.right */
                if (g0 > best_gain) { best_gain = g0; best_side = side; best_opt = 0; }
                if (g1 > best_gain) { best_gain = g1; best_side = side; best_opt = 1; }
            }
            if (best_side < 0) continue;
            const int32_t c = best_side ? N.right : N.left;
            int32_t& other_slot = best_side ? N.left : N.right;
            int32_t& grand_slot = best_opt ? t[c].right : t[c].left;
            if (!(best_gain > 1e-12 * half_area(t[c].box))) continue;
            std::swap(other_slot, grand_slot);
            t[c].box = t[t[c].left].box;
            box_grow(t[c].box, t[t[c].right].box);
            ++changed;
        }
        applied += changed;
        if (!changed) break;
    }
    return applied;
}

/* binned-SAH binary tree over all triangles (TmpNode[0] = root) + the permuted index array */
static void build_sah_tree(HostScene& s, int n_threads, int max_leaf, std::vector<int32_t>& idx, std::vector<TmpNode>& tmp) {
    std::vector<BoxRef> work((size_t)s.n_tris);
    for (int64_t i = 0; i < s.n_tris; ++i) work[i] = BoxRef{s.tri_box[i], (int32_t)i};
    tmp.assign((size_t)(2 * s.n_tris + 1), TmpNode());
    std::atomic<int32_t> next{1};
    int par = 0;
    while ((1 << par) < n_threads) ++par;
    SahBuilder b{s, work, tmp, next, max_leaf};
    b.n_threads = n_threads;
    /* 64 bins per axis on large meshes: -1.2 % nodes per ray and +2 % spp/s on the 1.1 M-triangle room for 0.1 s more build
     * time; no effect on the small scenes, whose trees stay as they were (profiles/r2_s19_builder_ab.txt) */
    b.NB = s.n_tris >= 262144 ? 64 : 16;
    if (const char* e = std::getenv("PRT_SAH_CTRI")) b.C_TRI = std::atof(e);
    if (const char* e = std::getenv("PRT_SAH_BINS")) b.NB = std::max(2, std::min((int)SahBuilder::NB_MAX, std::atoi(e)));
    if (const char* e = std::getenv("PRT_SAH_SWEEP_BELOW")) b.sweep_below = std::max(0, std::atoi(e));
    b.build(0, 0, (int32_t)s.n_tris, par + 1);
    idx.resize((size_t)s.n_tris);
    for (int64_t i = 0; i < s.n_tris; ++i) idx[i] = work[i].prim;
    if (const char* e = std::getenv("PRT_SAH_ROTATE")) {
        const int applied = rotate_tree(tmp, std::max(0, std::atoi(e)));
        if (std::getenv("PRT_DEBUG_BUILD")) std::fprintf(stderr, "[prt] tree rotations applied: %d\n", applied);
    }
}

/* PRT_DEBUG_BUILD=1: FNV-1a of the emitted nodes and of the slot order on stderr (two builds made the same tree iff equal) */
static void debug_tree_checksum(const HostScene& s) {
    if (!std::getenv("PRT_DEBUG_BUILD")) return;
    auto fnv = [](const void* p, size_t n) {
        uint64_t h = 1469598103934665603ull;
        const unsigned char* b = (const unsigned char*)p;
        for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
        return h;
    };
    std::fprintf(stderr, "[prt] tree checksum: nodes2 %016llx nodes4 %016llx slots %016llx (%zu / %zu nodes)\n",
                 (unsigned long long)fnv(s.nodes2.data(), s.nodes2.size() * sizeof(Node2)), (unsigned long long)fnv(s.nodes4.data(), s.nodes4.size() * sizeof(Node4)),
                 (unsigned long long)fnv(s.slot_prim.data(), s.slot_prim.size() * sizeof(int32_t)), s.nodes2.size(), s.nodes4.size());
}

static void finish_slots(HostScene& s) {
    debug_tree_checksum(s);
    s.isect_slots.resize((size_t)s.n_tris);
    parallel_ranges(s.n_tris, default_host_threads(), [&](int64_t b, int64_t e) { /* a permutation: every entry written once */
        for (int64_t sidx = b; sidx < e; ++sidx) {
            s.prim_slot[s.slot_prim[sidx]] = (int32_t)sidx;
            s.isect_slots[sidx] = s.isect[s.slot_prim[sidx]];
        }
    });
}

void HostScene::build_sah_bvh2(int n_threads, int max_leaf) {
    nodes2.clear();
    nodes4.clear();
    nodes4q.clear();
    slot_prim.clear();
    prim_slot.assign((size_t)n_tris, -1);
    if (max_leaf < 1) max_leaf = 1;
    if (max_leaf > 8) max_leaf = 8;
    sah_max_leaf = max_leaf;
    if (n_tris == 0) return;
    std::vector<int32_t> idx;
    std::vector<TmpNode> tmp;
    build_sah_tree(*this, n_threads, max_leaf, idx, tmp);

    /* flatten in DFS preorder; leaves define the slot order */
    slot_prim.reserve((size_t)n_tris);
    auto leaf_ref = [&](const TmpNode& n) -> int32_t {
        const int32_t first_slot = (int32_t)slot_prim.size();
        for (int32_t i = n.first; i < n.first + n.count; ++i) slot_prim.push_back(idx[i]);
        return ~((first_slot << 3) | (n.count - 1));
    };
    const int32_t NO_CHILD = INT32_MIN;
    if (tmp[0].left < 0) { /* the whole scene is one leaf: a root with one real child */
        Node2 root;
        std::memset(&root, 0, sizeof(root));
        set_child_box(root.lo0, root.hi0, tmp[0].box);
        for (int k = 0; k < 3; ++k) { root.lo1[k] = std::numeric_limits<float>::infinity(); root.hi1[k] = -std::numeric_limits<float>::infinity(); }
        root.child0 = leaf_ref(tmp[0]);
        root.child1 = NO_CHILD;
        nodes2.push_back(root);
    } else {
        struct Work { int32_t tmp_idx, flat_idx; };
        std::vector<Work> stack;
        nodes2.emplace_back();
        stack.push_back({0, 0});
        while (!stack.empty()) {
            Work w = stack.back();
            stack.pop_back();
            const TmpNode& n = tmp[w.tmp_idx];
            const TmpNode& L = tmp[n.left];
            const TmpNode& R = tmp[n.right];
            Node2 out;
            std::memset(&out, 0, sizeof(out));
            set_child_box(out.lo0, out.hi0, L.box);
            set_child_box(out.lo1, out.hi1, R.box);
            /* left subtree first so that its nodes and slots directly follow the parent */
            int32_t c0, c1;
            if (L.left < 0) c0 = leaf_ref(L);
            else { c0 = (int32_t)nodes2.size(); nodes2.emplace_back(); }
            if (R.left < 0) c1 = leaf_ref(R);
            else { c1 = (int32_t)nodes2.size(); nodes2.emplace_back(); }
            out.child0 = c0;
            out.child1 = c1;
            nodes2[w.flat_idx] = out;
            if (R.left >= 0) stack.push_back({n.right, c1});
            if (L.left >= 0) stack.push_back({n.left, c0});
        }
    }
    finish_slots(*this);
}

/* BVH4 = the binned-SAH binary tree with, at every node, the inner child of largest surface area
 * replaced by its own two children until four children are held (or only leaves remain). The boxes
 * are the binary tree's, so everything said about Node2 (outward FP32 rounding, containment of the
 * reference's boxes) holds unchanged. */
void HostScene::build_sah_bvh4(int n_threads, int max_leaf) {
    nodes2.clear();
    nodes4.clear();
    nodes4q.clear();
    slot_prim.clear();
    prim_slot.assign((size_t)n_tris, -1);
    if (max_leaf < 1) max_leaf = 1;
    if (max_leaf > 8) max_leaf = 8;
    sah_max_leaf = max_leaf;
    if (n_tris == 0) return;
    std::vector<int32_t> idx;
    std::vector<TmpNode> tmp;
    const bool timing = std::getenv("PRT_BUILD_TIMING") != nullptr;
    auto t_lap = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[prt] host   %-26s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_lap).count());
        t_lap = now;
    };
    build_sah_tree(*this, n_threads, max_leaf, idx, tmp);
    lap("binned SAH binary tree");

    slot_prim.reserve((size_t)n_tris);
    auto leaf_ref = [&](const TmpNode& n) -> int32_t {
        const int32_t first_slot = (int32_t)slot_prim.size();
        for (int32_t i = n.first; i < n.first + n.count; ++i) slot_prim.push_back(idx[i]);
        return ~((first_slot << 3) | (n.count - 1));
    };
    auto area_of = [](const Box6& b) {
        const double dx = b.v[1] - b.v[0], dy = b.v[3] - b.v[2], dz = b.v[5] - b.v[4];
        return dx * dy + dy * dz + dz * dx;
    };
    const int32_t NO_CHILD = INT32_MIN;
    const float inf = std::numeric_limits<float>::infinity();
    struct Work { int32_t tmp_idx, flat_idx; };
    std::vector<Work> stack;
    nodes4.emplace_back();
    stack.push_back({0, 0});
    while (!stack.empty()) {
        Work w = stack.back();
        stack.pop_back();
        int32_t kids[4];
        int nk = 0;
        if (tmp[w.tmp_idx].left < 0) {
            kids[nk++] = w.tmp_idx; /* the whole scene is one leaf */
        } else {
            kids[nk++] = tmp[w.tmp_idx].left;
            kids[nk++] = tmp[w.tmp_idx].right;
            while (nk < 4) {
                int best = -1;
                double best_area = -1.0;
                for (int i = 0; i < nk; ++i)
                    if (tmp[kids[i]].left >= 0) {
                        const double a = area_of(tmp[kids[i]].box);
                        if (a > best_area) { best_area = a; best = i; }
                    }
                if (best < 0) break;
                const int32_t t = kids[best];
                for (int i = nk; i > best + 1; --i) kids[i] = kids[i - 1]; /* keep the left-to-right order */
                kids[best] = tmp[t].left;
                kids[best + 1] = tmp[t].right;
                ++nk;
            }
        }
        Node4 out;
        std::memset(&out, 0, sizeof(out));
        int32_t refs[4];
        for (int i = 0; i < 4; ++i) {
            if (i < nk) {
                float lo[3], hi[3];
                set_child_box(lo, hi, tmp[kids[i]].box);
                for (int k = 0; k < 3; ++k) { out.lo[k][i] = lo[k]; out.hi[k][i] = hi[k]; }
                if (tmp[kids[i]].left < 0) refs[i] = leaf_ref(tmp[kids[i]]);
                else { refs[i] = (int32_t)nodes4.size(); nodes4.emplace_back(); }
            } else {
                for (int k = 0; k < 3; ++k) { out.lo[k][i] = inf; out.hi[k][i] = -inf; }
                refs[i] = NO_CHILD;
            }
            out.child[i] = refs[i];
        }
        nodes4[w.flat_idx] = out;
        for (int i = nk - 1; i >= 0; --i)
            if (refs[i] >= 0) stack.push_back({kids[i], refs[i]});
    }
    lap("collapse to BVH4 + emit");
    finish_slots(*this);
    lap("slot tables + triangle gather");
}

/* Node4 -> Node4Q. Per node and axis: origin = the smallest child low (already an outward-rounded
 * float), step = the smallest power of two with 255 steps covering the node's extent, lows floored,
 * highs ceiled; every decoded plane is then checked in double against the plane it replaces and moved
 * one more step outward if rounding got in the way. */
void HostScene::quantise_bvh4() {
    nodes4q.assign(nodes4.size(), Node4Q());
    parallel_ranges((int64_t)nodes4.size(), default_host_threads(), [&](int64_t n_begin, int64_t n_end) { /* node by node, independent */
    for (int64_t n = n_begin; n < n_end; ++n) {
        const Node4& N = nodes4[n];
        Node4Q Q;
        std::memset(&Q, 0, sizeof(Q));
        for (int c = 0; c < 4; ++c) Q.child[c] = N.child[c];
        const int32_t NO_CHILD = INT32_MIN;
        for (int k = 0; k < 3; ++k) {
            double lo = std::numeric_limits<double>::infinity(), hi = -std::numeric_limits<double>::infinity();
            for (int c = 0; c < 4; ++c)
                if (N.child[c] != NO_CHILD) {
                    lo = std::min(lo, (double)N.lo[k][c]);
                    hi = std::max(hi, (double)N.hi[k][c]);
                }
            if (!(lo <= hi)) { lo = 0.0; hi = 0.0; }
            const float origin = (float)lo; /* exact: a float */
            Q.origin[k] = origin;
            const double extent = hi - (double)origin;
            int e = 0; /* biased exponent byte; 0 = step 0 */
            double step = 0.0;
            if (extent > 0.0 && std::isfinite(extent)) {
                int ex;
                std::frexp(extent / 255.0, &ex); /* extent / 255 = m 2^ex, m in [0.5, 1) -> 2^ex >= extent / 255 */
                e = ex + 127;
                if (e < 1) e = 1;
                if (e > 254) e = 254;
                step = std::ldexp(1.0, e - 127);
                while (255.0 * step < extent && e < 254) { ++e; step = std::ldexp(1.0, e - 127); }
            } else if (!std::isfinite(extent)) {
                e = 254;
                step = std::ldexp(1.0, e - 127);
            }
            Q.exps |= (uint32_t)e << (8 * k);
            for (int c = 0; c < 4; ++c) {
                uint32_t ql = 255u, qh = 0u; /* unused child (the traversal masks it by its reference) */
                if (N.child[c] != NO_CHILD && step > 0.0) {
                    double a = std::floor(((double)N.lo[k][c] - (double)origin) / step);
                    double b = std::ceil(((double)N.hi[k][c] - (double)origin) / step);
                    if (a < 0.0) a = 0.0;
                    if (a > 255.0) a = 255.0;
                    if (b < 0.0) b = 0.0;
                    if (b > 255.0) b = 255.0;
                    while (a > 0.0 && (double)origin + a * step > (double)N.lo[k][c]) a -= 1.0;
                    while (b < 255.0 && (double)origin + b * step < (double)N.hi[k][c]) b += 1.0;
                    ql = (uint32_t)a;
                    qh = (uint32_t)b;
                } else if (N.child[c] != NO_CHILD) {
                    ql = 0u; qh = 0u; /* flat axis: every plane is the origin */
                }
                Q.qlo[k] |= ql << (8 * c);
                Q.qhi[k] |= qh << (8 * c);
            }
        }
        nodes4q[n] = Q;
    }
    });
    if (std::getenv("PRT_DEBUG_BUILD")) {
        uint64_t h = 1469598103934665603ull;
        const unsigned char* b = (const unsigned char*)nodes4q.data();
        for (size_t i = 0; i < nodes4q.size() * sizeof(Node4Q); ++i) { h ^= b[i]; h *= 1099511628211ull; }
        std::fprintf(stderr, "[prt] tree checksum: nodes4q %016llx\n", (unsigned long long)h);
    }
}

void HostScene::use_reference_slots() {
    nodes2.clear();
    nodes4.clear();
    nodes4q.clear();
    slot_prim.assign((size_t)n_tris, 0);
    prim_slot.assign((size_t)n_tris, -1);
    for (int64_t p = 0; p < n_tris; ++p) {
        const int64_t r = world_rank[p];
        slot_prim[(size_t)r] = (int32_t)p;
        prim_slot[(size_t)p] = (int32_t)r;
    }
    isect_slots.resize((size_t)n_tris);
    for (int64_t sidx = 0; sidx < n_tris; ++sidx) isect_slots[sidx] = isect[slot_prim[sidx]];
}

/* reference world BVH -> RefNode[], nodes renumbered in DFS order, triangle leaves = ~slot */
void HostScene::flatten_reference_world() {
    ref_flat.clear();
    ref_root_flat = -1;
    if (world_root < 0) return;
    std::vector<int32_t> flat_of((size_t)ref_nodes.size(), -1);
    std::vector<int32_t> order;
    std::vector<int32_t> stack{world_root};
    while (!stack.empty()) {
        int32_t c = stack.back();
        stack.pop_back();
        if (flat_of[c] >= 0) continue;
        flat_of[c] = (int32_t)order.size();
        order.push_back(c);
        const HostRefNode& n = ref_nodes[c];
        if (n.right.kind == 1) stack.push_back(n.right.idx);
        if (n.left.kind == 1) stack.push_back(n.left.idx);
    }
    ref_flat.resize(order.size());
    auto ref_of = [&](ObjRef o) -> int32_t { return o.kind == 1 ? flat_of[o.idx] : ~prim_slot[o.idx]; };
    for (size_t i = 0; i < order.size(); ++i) {
        const HostRefNode& n = ref_nodes[order[i]];
        RefNode out;
        std::memset(&out, 0, sizeof(out));
        for (int k = 0; k < 6; ++k) out.box[k] = n.box.v[k];
        out.left = ref_of(n.left);
        out.right = ref_of(n.right);
        ref_flat[i] = out;
    }
    ref_root_flat = 0;
}

}  // namespace prt

// rtb_api.cu -- the extern "C" launch layer of librtb.so (include/rtb.h): context, scene
// flattening (Object[] -> SoA primitive records + material table, creation order preserved),
// GPU LBVH build, persistent device buffers, render entry points.  Replaces src/dispatch.cu of the
// reference (per-frame cudaMalloc/cudaMallocManaged/cudaFree, element-wise managed-memory copy out,
// three __constant__ symbols).  No CPU fallback: every entry point needs a working CUDA device.
#include "rtb_ctx.h"
#include "rtb_lbvh.h"
#include "rtb_refbvh.h"
#include "rtb_meshprep.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <chrono>
#include <string>
#include <thread>
#include <utility>
#include <vector>

std::string &rtb_create_error() {
    static std::string e;
    return e;
}

namespace {

void free_scene(rtb_ctx *c) {
    for (void *p : c->scene_allocs) cudaFree(p);
    c->scene_allocs.clear();
    c->d = DScene{};
    c->committed = false;
    c->scene_bytes = 0;
}

template <typename T> cudaError_t upload(rtb_ctx *c, const std::vector<T> &v, const T **out) {
    *out = nullptr;
    size_t bytes = sizeof(T) * (v.empty() ? 1 : v.size());
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return e;
    c->scene_allocs.push_back(p);
    c->scene_bytes += bytes;
    if (!v.empty()) e = cudaMemcpyAsync(p, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice, c->stream);
    *out = (const T *)p;
    return e;
}

// device array of `total` elements (at least one) whose first v.size() come from the host
template <typename T> cudaError_t upload_front(rtb_ctx *c, const std::vector<T> &v, size_t total, T **out) {
    *out = nullptr;
    size_t bytes = sizeof(T) * (total == 0 ? 1 : total);
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return e;
    c->scene_allocs.push_back(p);
    c->scene_bytes += bytes;
    if (!v.empty()) e = cudaMemcpyAsync(p, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice, c->stream);
    *out = (T *)p;
    return e;
}

// ---- host float helpers: the expressions of Triangle::precompute (objects.cu:175-186) ----------
struct H3 {
    float x, y, z;
};
inline H3 hsub(H3 a, H3 b) { return H3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline H3 hadd(H3 a, H3 b) { return H3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline H3 hcross(H3 a, H3 b) { return H3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline H3 hnorm(H3 a) {
    float mag_sq = a.x * a.x + a.y * a.y + a.z * a.z;
    float inv = 1 / sqrtf(mag_sq);
    return H3{a.x * inv, a.y * inv, a.z * inv};
}
inline H3 h3(RtbVec3 v) { return H3{v.x, v.y, v.z}; }

struct TriRec {
    float4 a, b, c;
    H3 normal;
};
TriRec make_tri(H3 p0, H3 p1, H3 p2) {
    H3 s1 = hsub(p1, p0), s2 = hsub(p2, p0);
    H3 n = hnorm(hcross(s1, s2));
    TriRec t;
    t.a = make_float4(p0.x, p0.y, p0.z, s1.x);
    t.b = make_float4(s1.y, s1.z, s2.x, s2.y);
    t.c = make_float4(s2.z, n.x, n.y, n.z);
    t.normal = n;
    return t;
}

// a mesh whose partition and records are made on the device (rtb_meshprep.cu): its triangles come after all host-built
// ones in the scene arrays
struct DevMesh {
    int obj = 0, n = 0, node_base = -1;
    float *d_pts = nullptr, *d_uv = nullptr;
    MeshPrep prep;
    size_t tri_base = 0, leaf_base = 0;
};

// the vertices of the triangles as their records give them back, NaN-skipping min / max (fminf / fmaxf written out:
// libm calls otherwise, 18 per triangle)
void rec_box(const float4 *rec, int count, float *blo, float *bhi) {
    for (int c = 0; c < 3; c++) { blo[c] = 3e38f; bhi[c] = -3e38f; }
    for (int k = 0; k < count; k++) {
        float4 a = rec[3 * k], b = rec[3 * k + 1], cc = rec[3 * k + 2];
        float p[3][3] = {{a.x, a.y, a.z}, {a.x + a.w, a.y + b.x, a.z + b.y}, {a.x + b.z, a.y + b.w, a.z + cc.x}};
        for (int v = 0; v < 3; v++)
            for (int c = 0; c < 3; c++) {
                blo[c] = (blo[c] < p[v][c] || p[v][c] != p[v][c]) ? blo[c] : p[v][c];
                bhi[c] = (bhi[c] > p[v][c] || p[v][c] != p[v][c]) ? bhi[c] : p[v][c];
            }
    }
}

struct Flat {
    std::vector<DevMesh> dev_meshes;
    ~Flat() {
        for (DevMesh &m : dev_meshes) { cudaFree(m.d_pts); cudaFree(m.d_uv); m.prep.release(); }
    }
    std::vector<float4> tris, spheres, ref_nodes;
    std::vector<int2> tri_ids;
    std::vector<float2> tri_uv;
    std::vector<int> tri_refleaf, sphere_ids, obj_material, leaf_refs;
    std::vector<DMaterial> materials;
    std::vector<float> image_rgb;
    std::vector<float> leaf_area;  // surface area of each leaf's bounding box (parallel to leaf_refs)
    std::vector<float> leaf_box;   // lo.xyz hi.xyz of each leaf (parallel to leaf_refs)
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};  // scene bounds
    int culled = 0;

    void note_box(const float *blo, const float *bhi) {
        float e[3];
        for (int c = 0; c < 3; c++) {
            e[c] = bhi[c] - blo[c];
            if (blo[c] < lo[c]) lo[c] = blo[c];
            if (bhi[c] > hi[c]) hi[c] = bhi[c];
        }
        leaf_area.push_back(2.f * (e[0] * e[1] + e[1] * e[2] + e[2] * e[0]));
        for (int c = 0; c < 3; c++) leaf_box.push_back(blo[c]);
        for (int c = 0; c < 3; c++) leaf_box.push_back(bhi[c]);
    }
    void tri_box(int idx, int count, float *blo, float *bhi) const { rec_box(&tris[3 * (size_t)idx], count, blo, bhi); }

    int push_tri(TriRec t, int obj, int sub, const float *uv, int refleaf) {
        int idx = (int)tri_ids.size();
        tris.push_back(t.a); tris.push_back(t.b); tris.push_back(t.c);
        tri_ids.push_back(make_int2(obj, sub));
        for (int k = 0; k < 3; k++) tri_uv.push_back(uv ? make_float2(uv[2 * k], uv[2 * k + 1]) : make_float2(0, 0));
        tri_refleaf.push_back(refleaf);
        return idx;
    }
    // the same records as add_tri, written at known positions of pre-sized arrays (triangle `idx`, leaf
    // `leaf_pos`), so that the triangles of a big mesh can be flattened by several host threads; the box goes
    // into blo / bhi for the caller's own bounds
    void set_tri(size_t idx, size_t leaf_pos, TriRec t, int obj, int sub, const float *uv, int refleaf, float *blo, float *bhi) {
        tris[3 * idx] = t.a; tris[3 * idx + 1] = t.b; tris[3 * idx + 2] = t.c;
        tri_ids[idx] = make_int2(obj, sub);
        for (int k = 0; k < 3; k++) tri_uv[3 * idx + k] = uv ? make_float2(uv[2 * k], uv[2 * k + 1]) : make_float2(0, 0);
        tri_refleaf[idx] = refleaf;
        leaf_refs[leaf_pos] = ~(((int)idx << RTB_KIND_BITS) | KIND_TRI);
        tri_box((int)idx, 1, blo, bhi);
        float e[3] = {bhi[0] - blo[0], bhi[1] - blo[1], bhi[2] - blo[2]};
        leaf_area[leaf_pos] = 2.f * (e[0] * e[1] + e[1] * e[2] + e[2] * e[0]);
        for (int c = 0; c < 3; c++) { leaf_box[6 * leaf_pos + c] = blo[c]; leaf_box[6 * leaf_pos + 3 + c] = bhi[c]; }
    }
    void add_tri(TriRec t, int obj, int sub, const float *uv, int refleaf) {
        int idx = push_tri(t, obj, sub, uv, refleaf);
        leaf_refs.push_back(~((idx << RTB_KIND_BITS) | KIND_TRI));
        float blo[3], bhi[3];
        tri_box(idx, 1, blo, bhi);
        note_box(blo, bhi);
    }
    // Quad::create_triangles (objects.cu:244-253): t1 = (p1,p2,p3) uv (0,0)(1,0)(1,1), t2 = (p1,p4,p3) uv (0,0)(0,1)(1,1).
    // OneWayQuad's normal is +-t1.normal (objects.cu:285-289): carried by the leaf kind.
    void add_quad(H3 p1, H3 p2, H3 p3, H3 p4, int obj, int sub, bool one_way, bool invert) {
        static const float uv1[6] = {0, 0, 1, 0, 1, 1}, uv2[6] = {0, 0, 0, 1, 1, 1};
        int idx = push_tri(make_tri(p1, p2, p3), obj, sub, uv1, -1);
        push_tri(make_tri(p1, p4, p3), obj, sub + 1, uv2, -1);
        int kind = one_way ? (invert ? KIND_QUAD_ONEWAY_INV : KIND_QUAD_ONEWAY) : KIND_QUAD;
        leaf_refs.push_back(~((idx << RTB_KIND_BITS) | kind));
        float blo[3], bhi[3];
        tri_box(idx, 2, blo, bhi);
        note_box(blo, bhi);
    }
};

// ---- EXPERIMENT: full-sweep SAH, top-down, one primitive per leaf, on the host -------------------------------
struct SahBox {
    float lo[3], hi[3];
    void grow(const float *b) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b[k]); hi[k] = std::max(hi[k], b[3 + k]); } }
    void grow(const SahBox &o) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], o.lo[k]); hi[k] = std::max(hi[k], o.hi[k]); } }
    float area() const { float e0 = hi[0] - lo[0], e1 = hi[1] - lo[1], e2 = hi[2] - lo[2]; return e0 * e1 + e1 * e2 + e2 * e0; }
    static SahBox empty() { return SahBox{{3e38f, 3e38f, 3e38f}, {-3e38f, -3e38f, -3e38f}}; }
};

void sah_pad(const SahBox &b, int ref, float4 &olo, float4 &ohi) {  // the padding of rtb_lbvh.cu padded()
    float p[3];
    for (int k = 0; k < 3; k++) p[k] = 4e-7f * fmaxf(fmaxf(fabsf(b.lo[k]), fabsf(b.hi[k])), b.hi[k] - b.lo[k]) + 1e-12f;
    olo = make_float4(b.lo[0] - p[0], b.lo[1] - p[1], b.lo[2] - p[2], 0.f);
    std::memcpy(&olo.w, &ref, sizeof(int));
    ohi = make_float4(b.hi[0] + p[0], b.hi[1] + p[1], b.hi[2] + p[2], 0.f);
}

// builds the subtree over prims[first, first + n) (n >= 2) as node `self`; returns its box
SahBox sah_build_rec(const std::vector<float> &leaf_box, const std::vector<int> &idx, const std::vector<int> &refs, std::vector<int> &prims,
                     int first, int n, std::vector<float4> &nodes, int self, int depth, int &height) {
    if (depth > height) height = depth;
    auto box_of = [&](int prim) { const float *b = &leaf_box[6 * (size_t)idx[prim]]; SahBox r = SahBox::empty(); r.grow(b); return r; };
    int best_axis = 0, best_split = n / 2;
    float best_cost = 3e38f;
    std::vector<float> right_area(n);
    std::vector<int> order(prims.begin() + first, prims.begin() + first + n), best_order;
    for (int axis = 0; axis < 3; axis++) {
        std::sort(order.begin(), order.end(), [&](int a, int b) {
            const float *ba = &leaf_box[6 * (size_t)idx[a]], *bb = &leaf_box[6 * (size_t)idx[b]];
            float ca = ba[axis] + ba[3 + axis], cb = bb[axis] + bb[3 + axis];
            return ca < cb || (ca == cb && a < b);
        });
        SahBox acc = SahBox::empty();
        for (int i = n - 1; i > 0; i--) { acc.grow(box_of(order[i])); right_area[i] = acc.area(); }
        acc = SahBox::empty();
        for (int i = 1; i < n; i++) {
            acc.grow(box_of(order[i - 1]));
            float cost = acc.area() * i + right_area[i] * (n - i);
            if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = i; best_order = order; }
        }
    }
    (void)best_axis;
    if (best_order.empty()) best_order = order;
    std::copy(best_order.begin(), best_order.end(), prims.begin() + first);
    int nl = best_split, nr = n - best_split;
    SahBox bl, br;
    int refl, refr;
    if (nl == 1) { bl = box_of(prims[first]); refl = refs[prims[first]]; }
    else { refl = (int)(nodes.size() / 4); nodes.resize(nodes.size() + 4); bl = sah_build_rec(leaf_box, idx, refs, prims, first, nl, nodes, refl, depth + 1, height); }
    if (nr == 1) { br = box_of(prims[first + nl]); refr = refs[prims[first + nl]]; }
    else { refr = (int)(nodes.size() / 4); nodes.resize(nodes.size() + 4); br = sah_build_rec(leaf_box, idx, refs, prims, first + nl, nr, nodes, refr, depth + 1, height); }
    float4 a, b2, c2, d2;
    sah_pad(bl, refl, a, b2);
    sah_pad(br, refr, c2, d2);
    nodes[4 * (size_t)self] = a; nodes[4 * (size_t)self + 1] = b2; nodes[4 * (size_t)self + 2] = c2; nodes[4 * (size_t)self + 3] = d2;
    SahBox all = bl;
    all.grow(br);
    return all;
}

int sah_build_host(const std::vector<float> &leaf_box, const std::vector<int> &idx, const std::vector<int> &refs, std::vector<float4> &nodes) {
    int n = (int)refs.size(), height = 0;
    std::vector<int> prims(n);
    for (int i = 0; i < n; i++) prims[i] = i;
    nodes.assign(4, make_float4(0, 0, 0, 0));
    sah_build_rec(leaf_box, idx, refs, prims, 0, n, nodes, 0, 1, height);
    return height;
}

int flatten(rtb_ctx *c, Flat &f) {
    // image store -> one float array
    std::vector<int> img_offset;
    for (const RtbImage &img : c->images) {
        img_offset.push_back((int)f.image_rgb.size());
        f.image_rgb.insert(f.image_rgb.end(), img.rgb, img.rgb + (size_t)img.width * img.height * 3);
    }
    for (const RtbMaterial &m : c->materials) {
        DMaterial d;
        std::memset(&d, 0, sizeof d);
        d.type = m.type;
        d.tex_type = m.texture.type;
        d.smoothness = m.smoothness;
        d.refractive_index = m.refractive_index;
        d.need_uv = m.need_uv;
        d.num_squares = m.texture.num_squares;
        if (m.texture.type == RTB_TEX_CHECKERBOARD) {
            d.a[0] = m.texture.light.x; d.a[1] = m.texture.light.y; d.a[2] = m.texture.light.z;
            d.b[0] = m.texture.dark.x; d.b[1] = m.texture.dark.y; d.b[2] = m.texture.dark.z;
        } else {
            d.a[0] = m.texture.colour.x; d.a[1] = m.texture.colour.y; d.a[2] = m.texture.colour.z;
        }
        if (m.type == RTB_MAT_EMISSIVE) { d.b[0] = m.emitted_light.x; d.b[1] = m.emitted_light.y; d.b[2] = m.emitted_light.z; }
        if (m.texture.type == RTB_TEX_IMAGE) {
            if (m.texture.image < 0 || m.texture.image >= (int)c->images.size()) return fail(c, "flattening scene", "material refers to an unknown image");
            d.img_w = c->images[m.texture.image].width;
            d.img_h = c->images[m.texture.image].height;
            d.img_offset = img_offset[m.texture.image];
        }
        f.materials.push_back(d);
    }
    for (int i = 0; i < (int)c->objects.size(); i++) {
        const RtbObject &o = c->objects[i];
        if (o.material < 0 || o.material >= (int)c->materials.size()) return fail(c, "flattening scene", "object refers to an unknown material");
        f.obj_material.push_back(o.material);
        switch (o.type) {
            case RTB_OBJ_SPHERE: {
                int idx = (int)f.sphere_ids.size();
                f.spheres.push_back(make_float4(o.p[0].x, o.p[0].y, o.p[0].z, o.f[0]));
                f.sphere_ids.push_back(i);
                f.leaf_refs.push_back(~((idx << RTB_KIND_BITS) | KIND_SPHERE));
                float r = fabsf(o.f[0]);
                float blo[3] = {o.p[0].x - r, o.p[0].y - r, o.p[0].z - r}, bhi[3] = {o.p[0].x + r, o.p[0].y + r, o.p[0].z + r};
                f.note_box(blo, bhi);
                break;
            }
            case RTB_OBJ_TRIANGLE: f.add_tri(make_tri(h3(o.p[0]), h3(o.p[1]), h3(o.p[2])), i, 0, o.uv, -1); break;
            case RTB_OBJ_QUAD: f.add_quad(h3(o.p[0]), h3(o.p[1]), h3(o.p[2]), h3(o.p[3]), i, 0, false, false); break;
            case RTB_OBJ_ONE_WAY_QUAD: f.add_quad(h3(o.p[0]), h3(o.p[1]), h3(o.p[2]), h3(o.p[3]), i, 0, true, o.invert_normal != 0); break;
            case RTB_OBJ_CUBOID: {  // Cuboid::create_faces, objects.cu:327-349
                H3 tl_near = h3(o.p[0]), w{o.f[0], 0, 0}, h{0, o.f[1], 0}, d{0, 0, o.f[2]};
                H3 tr_near = hadd(tl_near, w), br_near = hsub(tr_near, h), bl_near = hsub(tl_near, h);
                H3 tl_far = hadd(tl_near, d), tr_far = hadd(tl_far, w), br_far = hsub(tr_far, h), bl_far = hsub(tl_far, h);
                f.add_quad(tl_near, tr_near, br_near, bl_near, i, 0, false, false);
                f.add_quad(tl_far, tr_far, br_far, bl_far, i, 2, false, false);
                f.add_quad(tl_near, bl_near, bl_far, tl_far, i, 4, false, false);
                f.add_quad(tr_near, br_near, br_far, tr_far, i, 6, false, false);
                f.add_quad(bl_near, br_near, br_far, bl_far, i, 8, false, false);
                f.add_quad(tl_near, tr_near, tr_far, tl_far, i, 10, false, false);
                break;
            }
            case RTB_OBJ_MESH: {
                if (o.tri_first < 0 || o.tri_count < 0 || (size_t)(o.tri_first + o.tri_count) * 9 > c->tri_points.size())
                    return fail(c, "flattening scene", "mesh triangle range out of bounds");
                const float *pts = c->tri_points.data() + (size_t)o.tri_first * 9;
                int node_base = (int)(f.ref_nodes.size() / 2);
                // a big mesh is partitioned and flattened on the device (rtb_meshprep.cu; option "device_flatten":
                // 0 never, 1 every mesh, 2 = from 32768 triangles up); non-finite coordinates take the host path
                if (o.tri_count > 0 && c->opt_builder != 2 && (c->opt_device_flatten == 1 || (c->opt_device_flatten == 2 && o.tri_count >= 32768))) {
                    auto tp0 = std::chrono::steady_clock::now();
                    f.dev_meshes.emplace_back();
                    DevMesh &m = f.dev_meshes.back();
                    m.obj = i; m.n = o.tri_count;
                    size_t pts_bytes = sizeof(float) * 9 * (size_t)o.tri_count;
                    RTB_CUDA(c, cudaMalloc((void **)&m.d_pts, pts_bytes), "allocating mesh vertices");
                    RTB_CUDA(c, cudaMemcpyAsync(m.d_pts, pts, pts_bytes, cudaMemcpyHostToDevice, c->stream), "uploading mesh vertices");
                    if (c->mesh_uv_given && !c->tri_uv.empty()) {
                        size_t uv_bytes = sizeof(float) * 6 * (size_t)o.tri_count;
                        RTB_CUDA(c, cudaMalloc((void **)&m.d_uv, uv_bytes), "allocating mesh texture points");
                        RTB_CUDA(c, cudaMemcpyAsync(m.d_uv, c->tri_uv.data() + (size_t)o.tri_first * 6, uv_bytes, cudaMemcpyHostToDevice, c->stream), "uploading mesh texture points");
                    }
                    int nonfinite = 0;
                    m.prep.n = o.tri_count; m.prep.kept = o.tri_count;
                    if (c->opt_ref_quirks) RTB_CUDA(c, mesh_partition_device(m.d_pts, o.tri_count, c->stream, m.prep, &nonfinite), "partitioning the mesh");
                    c->commit_partition_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tp0).count();
                    if (!nonfinite) {
                        if (c->opt_ref_quirks) {
                            m.node_base = node_base;
                            for (size_t q = 0; q < m.prep.nodes.size(); q += 2) {
                                float4 lo = m.prep.nodes[q];
                                int parent;
                                std::memcpy(&parent, &lo.w, sizeof(int));
                                parent = parent < 0 ? -1 : node_base + parent;
                                std::memcpy(&lo.w, &parent, sizeof(int));
                                f.ref_nodes.push_back(lo);
                                f.ref_nodes.push_back(m.prep.nodes[q + 1]);
                            }
                        }
                        f.culled += o.tri_count - m.prep.kept;
                        break;
                    }
                    cudaFree(m.d_pts); cudaFree(m.d_uv); m.prep.release();
                    f.dev_meshes.pop_back();
                }
                RefBvh rb;
                if (c->opt_ref_quirks) {
                    auto tp0 = std::chrono::steady_clock::now();
                    ref_bvh_partition(pts, o.tri_count, rb);
                    c->commit_partition_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tp0).count();
                    for (const RefBvhNode &n : rb.nodes) {
                        int parent = n.parent < 0 ? -1 : node_base + n.parent;
                        float4 lo = make_float4(n.lo[0], n.lo[1], n.lo[2], 0.f);
                        std::memcpy(&lo.w, &parent, sizeof(int));
                        f.ref_nodes.push_back(lo);
                        f.ref_nodes.push_back(make_float4(n.hi[0], n.hi[1], n.hi[2], 0.f));
                    }
                }
                // kept triangles get consecutive positions (degenerate leaves are invisible in the reference and
                // dropped), then the records are written in place, by several host threads for a big mesh
                std::vector<int> slot((size_t)o.tri_count);
                int kept = 0;
                for (int k = 0; k < o.tri_count; k++) slot[k] = (c->opt_ref_quirks && rb.degenerate[k]) ? -1 : kept++;
                f.culled += o.tri_count - kept;
                const size_t tri_base = f.tri_ids.size(), leaf_base = f.leaf_refs.size();
                f.tris.resize(3 * (tri_base + kept)); f.tri_ids.resize(tri_base + kept); f.tri_uv.resize(3 * (tri_base + kept));
                f.tri_refleaf.resize(tri_base + kept);
                f.leaf_refs.resize(leaf_base + kept); f.leaf_area.resize(leaf_base + kept); f.leaf_box.resize(6 * (leaf_base + kept));
                int workers = o.tri_count >= 32768 ? (int)std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1;
                std::vector<float> wlo((size_t)workers * 3, 3e38f), whi((size_t)workers * 3, -3e38f);
                auto fill = [&](int w) {
                    int k0 = (int)((long long)o.tri_count * w / workers), k1 = (int)((long long)o.tri_count * (w + 1) / workers);
                    for (int k = k0; k < k1; k++) {
                        if (slot[k] < 0) continue;
                        const float *p = pts + (size_t)k * 9;
                        const float *uv = c->tri_uv.empty() ? nullptr : c->tri_uv.data() + (size_t)(o.tri_first + k) * 6;
                        float blo[3], bhi[3];
                        f.set_tri(tri_base + slot[k], leaf_base + slot[k], make_tri(H3{p[0], p[1], p[2]}, H3{p[3], p[4], p[5]}, H3{p[6], p[7], p[8]}), i, k, uv,
                                  c->opt_ref_quirks ? node_base + rb.leaf_of[k] : -1, blo, bhi);
                        for (int a = 0; a < 3; a++) {
                            if (blo[a] < wlo[3 * w + a]) wlo[3 * w + a] = blo[a];
                            if (bhi[a] > whi[3 * w + a]) whi[3 * w + a] = bhi[a];
                        }
                    }
                };
                if (workers == 1) {
                    fill(0);
                } else {
                    std::vector<std::thread> pool;
                    for (int w = 0; w < workers; w++) pool.emplace_back(fill, w);
                    for (std::thread &t : pool) t.join();
                }
                for (int w = 0; w < workers; w++)
                    for (int a = 0; a < 3; a++) {
                        if (wlo[3 * w + a] < f.lo[a]) f.lo[a] = wlo[3 * w + a];
                        if (whi[3 * w + a] > f.hi[a]) f.hi[a] = whi[3 * w + a];
                    }
                break;
            }
            default: return fail(c, "flattening scene", "unknown object type");
        }
    }
    return 0;
}

}  // namespace

// ---- shard helpers ---------------------------------------------------------------------------------
namespace {

int shard_local_rows(const rtb_ctx *c, int rank) { return rtb_shard_row_count(c->height, rank, c->shard_world, c->band_rows); }

__global__ void shard_pack_kernel(const float *image, float *dense, int width, int local_rows, int rank, int world, int band_rows, int height) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int row_floats = width * 3;
    if (i >= local_rows * row_floats) return;
    int lr = i / row_floats, col = i - lr * row_floats;
    int band = lr / band_rows, within = lr - band * band_rows;
    int y = (band * world + rank) * band_rows + within;
    if (y < height) dense[i] = image[(size_t)y * row_floats + col];
}

__global__ void shard_unpack_kernel(float *image, const float *gathered, int width, int rows_padded, int world, int band_rows, int height) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    int row_floats = width * 3;
    size_t per_rank = (size_t)rows_padded * row_floats;
    if (i >= per_rank * world) return;
    int rank = (int)(i / per_rank);
    size_t j = i - (size_t)rank * per_rank;
    int lr = (int)(j / row_floats), col = (int)(j - (size_t)lr * row_floats);
    int band = lr / band_rows, within = lr - band * band_rows;
    int y = (band * world + rank) * band_rows + within;
    if (y < height) image[(size_t)y * row_floats + col] = gathered[i];
}

// parse_pixel_colours, src/main.cu:343-371.  The reference converts on the host with
// `int colour = value * 255` (x86 cvttss2si: NaN / out of range -> INT_MIN -> clamped to 0).
__global__ void display_rgba8_kernel(const float *image, uchar4 *rgba, size_t pixels) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    unsigned char out[3];
    for (int k = 0; k < 3; k++) {
        float v = image[3 * i + k] * 255;
        int colour = (v >= 2147483648.0f || v < -2147483648.0f || v != v) ? (-2147483647 - 1) : (int)v;
        colour = colour > 255 ? 255 : (colour < 0 ? 0 : colour);
        out[k] = (unsigned char)colour;
    }
    rgba[i] = make_uchar4(out[0], out[1], out[2], 255);
}

int ensure_image(rtb_ctx *c) {
    size_t need = (size_t)c->width * c->height * 3;
    if (need == 0) return fail(c, "allocating image", "camera/resolution not set");
    if (c->image_floats != need) {
        if (c->d_image) cudaFree(c->d_image);
        c->d_image = nullptr;
        RTB_CUDA(c, cudaMalloc((void **)&c->d_image, need * sizeof(float)), "allocating image");
        RTB_CUDA(c, cudaMemsetAsync(c->d_image, 0, need * sizeof(float), c->stream), "clearing image");
        c->image_floats = need;
        c->frame_num = 0;
    }
    return 0;
}

int launch_frame_wavefront(rtb_ctx *c, int time_ms, int frame_num);

int launch_frame_megakernel(rtb_ctx *c, int time_ms, int frame_num);

int launch_frame_pool(rtb_ctx *c, int time_ms, int frame_num);

int launch_frame_stream(rtb_ctx *c, int time_ms, int frame_num);

int check_stream_watchdog(rtb_ctx *c);

int launch_kernel_variant(rtb_ctx *c, int kern, int time_ms, int frame_num) {
    return kern == 1 ? launch_frame_wavefront(c, time_ms, frame_num) : kern == 3 ? launch_frame_pool(c, time_ms, frame_num)
         : kern == 4 ? launch_frame_stream(c, time_ms, frame_num) : launch_frame_megakernel(c, time_ms, frame_num);
}

// option "kernel" = 2: a short frame (RTB_TUNE_SPP samples per pixel of the real scene, camera, shard
// and bounce limit) through each variant into a scratch image; the winner renders until something
// that shapes the workload changes.  Which variant wins depends on the workload (long
// paths in a closed room favour the wavefront kernels, short paths or few pixels per GPU the
// megakernel); the images are bit-identical either way.
#define RTB_TUNE_SPP 8
// the wavefront kernels pack the bounce index in 8 bits and the sample index in 24 (rtb_wavefront.cuh);
// the reference has no such limits, so beyond them the automatic choice is the megakernel
bool wavefront_fits(const rtb_ctx *c) { return c->rd.reflection_limit <= 255 && (long long)c->rd.rays_per_pixel < (1 << 24); }

int autotune_kernel(rtb_ctx *c) {
    c->tuned_prims = c->d.num_prims;
    if (!wavefront_fits(c)) { c->auto_choice = 0; return 0; }
    c->auto_choice = 1;
    if (c->opt_counters || c->image_floats == 0) return 0;  // counted builds are not timed
    if (c->tune_scratch_floats != c->image_floats) {
        if (c->d_tune_scratch) cudaFree(c->d_tune_scratch);
        c->d_tune_scratch = nullptr; c->tune_scratch_floats = 0;
        if (cudaMalloc((void **)&c->d_tune_scratch, c->image_floats * sizeof(float)) != cudaSuccess) {
            cudaGetLastError();
            return 0;  // no room for a scratch image: keep the wavefront kernels
        }
        c->tune_scratch_floats = c->image_floats;
    }
    float *scratch = c->d_tune_scratch;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float *image = c->d_image;
    const RtbRenderData rd = c->rd;
    c->d_image = scratch;
    int rc = 0;
    // candidates: each kernel with its traversal policies (walk_min, refill); the megakernel has one
    struct Cand { int kern, walk_min, refill; float ms; };
    std::vector<Cand> cand;
    if (c->opt_tune_set & 16) { cand.push_back({4, 16, 8, 0.f}); cand.push_back({4, 12, 16, 0.f}); }
    if (c->opt_tune_set & 8) { cand.push_back({3, 16, 8, 0.f}); cand.push_back({3, 12, 16, 0.f}); }
    if (c->opt_tune_set & 2) { cand.push_back({1, 16, 8, 0.f}); cand.push_back({1, 12, 16, 0.f}); }
    if (c->opt_tune_set & 1) {   // megakernel: `refill` = its early-pop variant (both, unless option "mega_early_pop" pins one)
        if (c->opt_mega_early_pop != 1) cand.push_back({0, 0, 0, 0.f});
        if (c->opt_mega_early_pop != 0) cand.push_back({0, 0, 1, 0.f});
    }
    const bool pinned = c->opt_walk_min > 0 && c->opt_wf_refill > 0;  // both knobs pinned by the caller: one policy per kernel
    int warmed = 0;  // kernels whose first (1-spp) frame has run
    for (size_t i = 0; i < cand.size() && rc == 0; i++) {
        const bool second_policy = i > 0 && cand[i - 1].kern == cand[i].kern;
        if (second_policy && pinned && cand[i].kern != 0) { cand[i].ms = 1e30f; continue; }
        if (cand[i].kern == 0) c->tuned_mega_early_pop = cand[i].refill;
        else { c->tuned_walk_min = cand[i].walk_min; c->tuned_refill = cand[i].refill; }
        // pass plan: a 1-spp frame warms a kernel up the first time it runs, then the timed frame.  The streaming
        // kernel balances its SM split from the utilisation the previous frame measured: two more short frames
        // let it settle before the timed one.
        std::vector<int> plan;
        if (!((warmed >> cand[i].kern) & 1)) {
            plan.push_back(1);
            if (cand[i].kern == 4) { plan.push_back(4); plan.push_back(4); }
        }
        plan.push_back(RTB_TUNE_SPP);
        for (size_t pass = 0; pass < plan.size() && rc == 0; pass++) {
            c->rd.rays_per_pixel = std::min(rd.rays_per_pixel, plan[pass]);
            cudaMemsetAsync(scratch, 0, c->image_floats * sizeof(float), c->stream);
            cudaEventRecord(e0, c->stream);
            rc = launch_kernel_variant(c, cand[i].kern, 12345, 0);
            cudaEventRecord(e1, c->stream);
            if (rc == 0 && cudaStreamSynchronize(c->stream) != cudaSuccess) rc = fail(c, "tuning the kernel choice", cudaGetErrorString(cudaGetLastError()));
            float ms = 0;
            if (rc == 0) cudaEventElapsedTime(&ms, e0, e1);
            cand[i].ms = ms;
        }
        warmed |= 1 << cand[i].kern;
    }
    c->rd = rd;
    c->d_image = image;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rc != 0 || cand.empty()) { c->auto_choice = rc != 0 ? -1 : 1; return rc; }
    for (int k = 0; k < 5; k++) c->tune_ms[k] = 0;
    size_t best = 0;
    for (size_t i = 0; i < cand.size(); i++) {
        if (cand[i].ms < cand[best].ms) best = i;
        double &slot = c->tune_ms[cand[i].kern];
        if (slot == 0 || cand[i].ms < slot) slot = cand[i].ms;
    }
    c->auto_choice = cand[best].kern;
    // the megakernel keeps its better variant even when another kernel wins (pinning kernel = 0 later uses it)
    {
        float best_mega = 1e30f;
        for (const Cand &cd : cand)
            if (cd.kern == 0 && cd.ms > 0 && cd.ms < best_mega) { best_mega = cd.ms; c->tuned_mega_early_pop = cd.refill; }
    }
    if (cand[best].kern != 0) { c->tuned_walk_min = cand[best].walk_min; c->tuned_refill = cand[best].refill; }
    else { c->tuned_walk_min = 16; c->tuned_refill = 8; }
    return 0;
}

int effective_kernel(const rtb_ctx *c) { return c->opt_kernel == 2 ? (c->auto_choice < 0 ? (wavefront_fits(c) ? 1 : 0) : c->auto_choice) : c->opt_kernel; }

int launch_frame(rtb_ctx *c, int time_ms, int frame_num) {
    if (c->opt_kernel == 2 && c->auto_choice < 0 && autotune_kernel(c) != 0) return 1;
    return launch_kernel_variant(c, effective_kernel(c), time_ms, frame_num);
}

// One frame through the streaming kernel (rtb_stream.cuh): one launch; the SM split follows the utilisation the
// previous frame measured (shade work ~ shade SMs x busy fraction, likewise traversal).
int launch_frame_stream(rtb_ctx *c, int time_ms, int frame_num) {
    if (!wavefront_fits(c) || c->width > 65535 || c->height > 65535)
        return fail(c, "render", "the streaming kernel packs the bounce index in 8 bits, the sample index in 24 and a pixel coordinate in 16 "
                                 "(reflection_limit <= 255, rays_per_pixel < 2^24, width and height < 65536): use option kernel = 0 or 2");
    const bool strict_fp = c->opt_strict != 0;
    const int local_rows = shard_local_rows(c, c->shard_rank);
    const size_t pixels = (size_t)c->width * local_rows;
    if (pixels == 0) return 0;
    if (c->stream_per_sm == 0) {
        cudaError_t e = strict_fp ? strict::stream_occupancy(&c->stream_per_sm) : fast::stream_occupancy(&c->stream_per_sm);
        if (e != cudaSuccess || c->stream_per_sm <= 0) return fail(c, "querying occupancy", cudaGetErrorString(e));
    }
    const int threads = strict_fp ? strict::stream_block_threads() : fast::stream_block_threads();
    int per_sm = c->opt_blocks_per_sm > 0 ? std::min(c->opt_blocks_per_sm, c->stream_per_sm) : c->stream_per_sm;
    const int blocks = c->num_sms * per_sm;
    // slots in flight: enough to keep every lane of both roles busy with queue slack, few enough to stay in the L2
    size_t want_slots = std::min(pixels, (size_t)c->num_sms * 1024 * (size_t)std::max(1, c->opt_stream_slots));
    want_slots = std::max<size_t>(want_slots, 32);
    // ring entries per sub-queue: a power of two, at least the slot count (a slot sits in at most one queue) and large
    // enough that the tickets drawn ahead of the tail (at most one per lane of the grid) stay within one lap
    size_t cap = (size_t)1 << 18;
    while (cap < want_slots) cap <<= 1;
    const int NQ = RTB_STREAM_SHARDS;
    const size_t ctl_words = 8 + 2 * 2 * NQ * 32 + 1024 + 8;   // ctl, head/tail of two queues (NQ sub-queues, 32 words apart), sm_role, stats (4 x u64)
    if (c->stream_slots < want_slots || c->stream_cap < cap) {
        if (c->d_stream_state) cudaFree(c->d_stream_state);
        if (c->d_stream_pix) cudaFree(c->d_stream_pix);
        if (c->d_stream_rings) cudaFree(c->d_stream_rings);
        c->d_stream_state = nullptr; c->d_stream_pix = nullptr; c->d_stream_rings = nullptr; c->stream_slots = 0; c->stream_cap = 0;
        RTB_CUDA(c, cudaMalloc((void **)&c->d_stream_state, want_slots * 6 * sizeof(float4)), "allocating the slot records");
        RTB_CUDA(c, cudaMalloc((void **)&c->d_stream_pix, want_slots * sizeof(unsigned int)), "allocating the slot records");
        RTB_CUDA(c, cudaMalloc((void **)&c->d_stream_rings, 2 * NQ * cap * sizeof(int)), "allocating the stream queues");
        RTB_CUDA(c, cudaMemsetAsync(c->d_stream_rings, 0, 2 * NQ * cap * sizeof(int), c->stream), "clearing the stream queues");
        c->stream_slots = want_slots; c->stream_cap = cap;
    }
    if (!c->d_stream_ctl) RTB_CUDA(c, cudaMalloc((void **)&c->d_stream_ctl, ctl_words * sizeof(unsigned int)), "allocating the stream control block");
    // the split: the caller's, else the one balanced from the previous frame's utilisation, else 40 % shading
    int shade_sms = c->opt_stream_shade_sms > 0 ? c->opt_stream_shade_sms : (c->stream_shade_sms > 0 ? c->stream_shade_sms : (c->num_sms * 2 + 2) / 5);
    shade_sms = std::max(1, std::min(c->num_sms - 1, shade_sms));
    c->stream_shade_sms = shade_sms;
    StreamParams p;
    p.scene = c->d;
    std::memcpy(&p.cam, &c->cam, sizeof(DCamera));
    p.rd.rays_per_pixel = c->rd.rays_per_pixel;
    p.rd.reflection_limit = c->rd.reflection_limit;
    p.rd.antialias = c->rd.antialias;
    p.rd.sky[0] = c->rd.sky_colour.x; p.rd.sky[1] = c->rd.sky_colour.y; p.rd.sky[2] = c->rd.sky_colour.z;
    p.width = c->width; p.height = c->height;
    p.time_ms = time_ms; p.frame_num = frame_num;
    p.prev = c->d_image; p.out = c->d_image;
    p.peers.n = 0;
    for (int r = 0; r < RTB_MAX_PEERS; r++) p.peers.image[r] = nullptr;
    if (c->opt_exchange == 2 && !c->opt_slices && c->n_peers == c->shard_world && c->shard_world > 1 && c->d_image == c->peer_image[c->shard_rank]) {
        p.peers.n = c->n_peers;  // fused exchange: finished pixels go to every GPU's image
        for (int r = 0; r < c->n_peers; r++) p.peers.image[r] = c->peer_image[r];
    }
    p.shard_rank = c->shard_rank; p.shard_world = c->shard_world; p.band_rows = c->band_rows;
    p.local_rows = local_rows;
    p.refill = c->opt_wf_refill < 1 ? c->tuned_refill : (c->opt_wf_refill > 32 ? 32 : c->opt_wf_refill);
    p.walk_min = c->opt_walk_min < 1 ? c->tuned_walk_min : (c->opt_walk_min > 32 ? 32 : c->opt_walk_min);
    const size_t slots = c->stream_slots;
    float4 *st = c->d_stream_state;
    p.st.A = st; p.st.B = st + slots; p.st.C = st + 2 * slots; p.st.O = st + 3 * slots; p.st.D = st + 4 * slots; p.st.H = st + 5 * slots;
    p.pix = c->d_stream_pix;
    unsigned int *ctl = c->d_stream_ctl;
    p.ctl = ctl;
    p.rq.ring = c->d_stream_rings; p.rq.head = ctl + 8; p.rq.tail = ctl + 8 + NQ * 32; p.rq.mask = (unsigned int)(c->stream_cap - 1);
    p.hq.ring = c->d_stream_rings + NQ * c->stream_cap; p.hq.head = ctl + 8 + 2 * NQ * 32; p.hq.tail = ctl + 8 + 3 * NQ * 32; p.hq.mask = (unsigned int)(c->stream_cap - 1);
    p.sm_role = (int *)(ctl + 8 + 4 * NQ * 32);
    p.stats = (unsigned long long *)(ctl + 8 + 4 * NQ * 32 + 1024);
    p.num_slots = (int)std::min(want_slots, slots);
    p.total_pixels = (unsigned int)pixels;
    p.shade_sms = shade_sms; p.total_sms = c->num_sms;
    p.counters = c->opt_counters ? c->d_counters : nullptr;
    // the previous frame's utilisation -> the next frame's split (read before the control block is cleared)
    // (a host round trip: every frame while the split settles, then now and then)
    if (c->stream_frames_launched > 0 && c->opt_stream_shade_sms <= 0 && (c->stream_frames_launched < 8 || c->stream_frames_launched % 64 == 0)) {
        unsigned long long h[4];
        RTB_CUDA(c, cudaMemcpyAsync(h, p.stats, sizeof h, cudaMemcpyDeviceToHost, c->stream), "reading the stream statistics");
        RTB_CUDA(c, cudaStreamSynchronize(c->stream), "reading the stream statistics");
        if (h[1] > 0 && h[3] > 0) {
            double us = (double)h[0] / (double)h[1], ut = (double)h[2] / (double)h[3];
            c->stream_util[0] = us; c->stream_util[1] = ut;
            double ws = c->stream_last_shade_sms * us, wt = (c->num_sms - c->stream_last_shade_sms) * ut;
            if (ws + wt > 0) {
                int next = (int)std::lround(c->num_sms * ws / (ws + wt));
                next = std::max(4, std::min(c->num_sms - 4, next));
                // move towards it (half the way damps the noise of a short frame), never by less than one SM
                int cur = c->stream_last_shade_sms, step = (next - cur) / 2;
                if (step == 0 && next != cur) step = next > cur ? 1 : -1;
                shade_sms = cur + step;
                c->stream_shade_sms = shade_sms;
                p.shade_sms = shade_sms;
            }
        }
    }
    RTB_CUDA(c, cudaMemsetAsync(ctl, 0, ctl_words * sizeof(unsigned int), c->stream), "clearing the stream control block");
    RTB_CUDA(c, cudaMemsetAsync(c->d_stream_rings, 0, 2 * NQ * c->stream_cap * sizeof(int), c->stream), "clearing the stream queues");
    cudaError_t e = strict_fp ? strict::launch_stream(p, c->opt_traversal == 1, c->opt_counters != 0, blocks, c->stream)
                              : fast::launch_stream(p, c->opt_traversal == 1, c->opt_counters != 0, blocks, c->stream);
    if (e != cudaSuccess) return fail(c, "launching render kernel", cudaGetErrorString(e));
    c->stream_last_shade_sms = p.shade_sms;
    c->stream_frames_launched++;
    // the watchdog word is copied out right behind the kernel (pinned host word, no synchronisation here)
    if (!c->h_stream_error) RTB_CUDA(c, cudaMallocHost((void **)&c->h_stream_error, sizeof(unsigned int)), "allocating the watchdog word");
    RTB_CUDA(c, cudaMemcpyAsync(c->h_stream_error, ctl + 5, sizeof(unsigned int), cudaMemcpyDeviceToHost, c->stream), "reading the watchdog word");
    c->stream_error_check_pending = true;
    return 0;
}

// after a synchronisation: did the streaming kernel's watchdog fire?
int check_stream_watchdog(rtb_ctx *c) {
    if (!c->stream_error_check_pending || !c->h_stream_error) return 0;
    c->stream_error_check_pending = false;
    if (*c->h_stream_error != 0) return fail(c, "render", "the streaming kernel stalled (its warps waited for each other for seconds): please report; option kernel = 1 renders the same image");
    return 0;
}

// One frame through the pooled persistent kernel (rtb_pool.cuh): one launch.
int launch_frame_pool(rtb_ctx *c, int time_ms, int frame_num) {
    if (!wavefront_fits(c) || c->width > 65535 || c->height > 65535)
        return fail(c, "render", "the pooled kernel packs the bounce index in 8 bits, the sample index in 24 and a pixel coordinate in 16 "
                                 "(reflection_limit <= 255, rays_per_pixel < 2^24, width and height < 65536): use option kernel = 0 or 2");
    const bool strict_fp = c->opt_strict != 0;
    if (c->pool_per_sm == 0) {
        cudaError_t e = strict_fp ? strict::pool_occupancy(&c->pool_per_sm) : fast::pool_occupancy(&c->pool_per_sm);
        if (e != cudaSuccess || c->pool_per_sm <= 0) return fail(c, "querying occupancy", cudaGetErrorString(e));
    }
    const int threads = strict_fp ? strict::pool_block_threads() : fast::pool_block_threads();
    int per_sm = c->opt_blocks_per_sm > 0 ? std::min(c->opt_blocks_per_sm, c->pool_per_sm) : c->pool_per_sm;
    const int blocks = c->num_sms * per_sm;
    const size_t warps = (size_t)blocks * (threads / 32);
    if (c->pool_warps < warps) {
        if (c->d_pool) cudaFree(c->d_pool);
        c->d_pool = nullptr; c->pool_warps = 0;
        RTB_CUDA(c, cudaMalloc((void **)&c->d_pool, warps * 6 * 64 * sizeof(float4)), "allocating the slot pool");
        c->pool_warps = warps;
    }
    PoolParams p;
    p.scene = c->d;
    std::memcpy(&p.cam, &c->cam, sizeof(DCamera));
    p.rd.rays_per_pixel = c->rd.rays_per_pixel;
    p.rd.reflection_limit = c->rd.reflection_limit;
    p.rd.antialias = c->rd.antialias;
    p.rd.sky[0] = c->rd.sky_colour.x; p.rd.sky[1] = c->rd.sky_colour.y; p.rd.sky[2] = c->rd.sky_colour.z;
    p.width = c->width; p.height = c->height;
    p.time_ms = time_ms; p.frame_num = frame_num;
    p.prev = c->d_image; p.out = c->d_image;
    p.peers.n = 0;
    for (int r = 0; r < RTB_MAX_PEERS; r++) p.peers.image[r] = nullptr;
    if (c->opt_exchange == 2 && !c->opt_slices && c->n_peers == c->shard_world && c->shard_world > 1 && c->d_image == c->peer_image[c->shard_rank]) {
        p.peers.n = c->n_peers;  // fused exchange: finished pixels go to every GPU's image
        for (int r = 0; r < c->n_peers; r++) p.peers.image[r] = c->peer_image[r];
    }
    p.shard_rank = c->shard_rank; p.shard_world = c->shard_world; p.band_rows = c->band_rows;
    p.local_rows = shard_local_rows(c, c->shard_rank);
    p.refill = c->opt_wf_refill < 1 ? c->tuned_refill : (c->opt_wf_refill > 32 ? 32 : c->opt_wf_refill);
    p.walk_min = c->opt_walk_min < 1 ? c->tuned_walk_min : (c->opt_walk_min > 32 ? 32 : c->opt_walk_min);
    p.pool = c->d_pool;
    p.work_counter = c->d_work;
    p.counters = c->opt_counters ? c->d_counters : nullptr;
    RTB_CUDA(c, cudaMemsetAsync(c->d_work, 0, sizeof(unsigned int), c->stream), "resetting work counter");
    cudaError_t e = strict_fp ? strict::launch_pool(p, c->opt_traversal == 1, c->opt_counters != 0, blocks, c->stream)
                              : fast::launch_pool(p, c->opt_traversal == 1, c->opt_counters != 0, blocks, c->stream);
    if (e != cudaSuccess) return fail(c, "launching render kernel", cudaGetErrorString(e));
    return 0;
}

int launch_frame_megakernel(rtb_ctx *c, int time_ms, int frame_num) {
    // the traversal-loop variant (trav_run, EARLY_POP): pinned by option "mega_early_pop", else what the timed choice kept
    const bool early_pop = c->opt_mega_early_pop >= 0 ? c->opt_mega_early_pop != 0 : c->tuned_mega_early_pop != 0;
    RenderParams p;
    p.scene = c->d;
    std::memcpy(&p.cam, &c->cam, sizeof(DCamera));
    p.rd.rays_per_pixel = c->rd.rays_per_pixel;
    p.rd.reflection_limit = c->rd.reflection_limit;
    p.rd.antialias = c->rd.antialias;
    p.rd.sky[0] = c->rd.sky_colour.x; p.rd.sky[1] = c->rd.sky_colour.y; p.rd.sky[2] = c->rd.sky_colour.z;
    p.width = c->width; p.height = c->height;
    p.time_ms = time_ms; p.frame_num = frame_num;
    p.prev = c->d_image; p.out = c->d_image;
    p.peers.n = 0;
    for (int r = 0; r < RTB_MAX_PEERS; r++) p.peers.image[r] = nullptr;
    if (c->opt_exchange == 2 && !c->opt_slices && c->n_peers == c->shard_world && c->shard_world > 1 && c->d_image == c->peer_image[c->shard_rank]) {
        p.peers.n = c->n_peers;  // fused exchange: finished pixels go to every GPU's image
        for (int r = 0; r < c->n_peers; r++) p.peers.image[r] = c->peer_image[r];
    }
    p.shard_rank = c->shard_rank; p.shard_world = c->shard_world; p.band_rows = c->band_rows;
    p.local_rows = shard_local_rows(c, c->shard_rank);
    p.refill = c->opt_refill < 1 ? 1 : (c->opt_refill > 32 ? 32 : c->opt_refill);
    p.walk_min = c->opt_walk_min < 1 ? 8 : (c->opt_walk_min > 32 ? 32 : c->opt_walk_min);
    p.work_counter = c->d_work;
    p.counters = c->opt_counters ? c->d_counters : nullptr;
    RTB_CUDA(c, cudaMemsetAsync(c->d_work, 0, sizeof(unsigned int), c->stream), "resetting work counter");
    LaunchConfig cfg;
    cfg.threads = c->opt_threads;
    int per_sm = c->opt_blocks_per_sm;
    if (per_sm <= 0) {
        cudaError_t e = c->opt_strict ? strict::render_occupancy(cfg.threads, early_pop, &per_sm) : fast::render_occupancy(cfg.threads, early_pop, &per_sm);
        if (e != cudaSuccess || per_sm <= 0) return fail(c, "querying occupancy", cudaGetErrorString(e));
    }
    cfg.blocks = c->num_sms * per_sm;
    cudaError_t e = c->opt_strict ? strict::launch_render(p, c->opt_traversal == 1, c->opt_counters != 0, early_pop, cfg, c->stream)
                                  : fast::launch_render(p, c->opt_traversal == 1, c->opt_counters != 0, early_pop, cfg, c->stream);
    if (e != cudaSuccess) return fail(c, "launching render kernel", cudaGetErrorString(e));
    return 0;
}

// One frame through the wavefront kernels (rtb_wavefront.cuh): first wf_shade starts every pixel,
// then spp * reflection_limit iterations of {wf_traverse, wf_shade}; an iteration advances every
// live pixel by one path segment, so after the last one every pixel has written its result.
int launch_frame_wavefront(rtb_ctx *c, int time_ms, int frame_num) {
    int local_rows = shard_local_rows(c, c->shard_rank);
    size_t slots = (size_t)c->width * local_rows;
    if (slots == 0) return 0;
    if (!wavefront_fits(c))
        return fail(c, "render", "the wavefront kernels pack the bounce index in 8 bits and the sample index in 24 (reflection_limit <= 255, rays_per_pixel < 2^24): "
                                 "use option kernel = 0 or 2");
    if (c->wf_slots < slots) {
        if (c->d_wf_state) cudaFree(c->d_wf_state);
        if (c->d_wf_queue) cudaFree(c->d_wf_queue);
        if (c->d_wf_alive) cudaFree(c->d_wf_alive);
        if (c->d_wf_block_count) cudaFree(c->d_wf_block_count);
        c->d_wf_state = nullptr; c->d_wf_queue = nullptr; c->d_wf_alive = nullptr; c->d_wf_block_count = nullptr; c->wf_slots = 0;
        RTB_CUDA(c, cudaMalloc((void **)&c->d_wf_state, slots * 6 * sizeof(float4)), "allocating wavefront state");
        RTB_CUDA(c, cudaMalloc((void **)&c->d_wf_queue, slots * 3 * sizeof(int)), "allocating wavefront queues");
        RTB_CUDA(c, cudaMalloc((void **)&c->d_wf_alive, slots), "allocating wavefront flags");
        RTB_CUDA(c, cudaMalloc((void **)&c->d_wf_block_count, RTB_WF_MAX_LANES * ((slots + 255) / 256 + 2) * sizeof(int)), "allocating wavefront scan buffers");
        c->wf_slots = slots;
    }
    if (!c->d_wf_counts) RTB_CUDA(c, cudaMalloc((void **)&c->d_wf_counts, 8 * RTB_WF_MAX_LANES * sizeof(unsigned int)), "allocating wavefront counters");
    if (c->wf_shade_per_sm == 0) {
        cudaError_t e = c->opt_strict ? strict::wf_occupancy(&c->wf_shade_per_sm, &c->wf_trav_per_sm) : fast::wf_occupancy(&c->wf_shade_per_sm, &c->wf_trav_per_sm);
        if (e != cudaSuccess || c->wf_shade_per_sm <= 0 || c->wf_trav_per_sm <= 0) return fail(c, "querying occupancy", cudaGetErrorString(e));
    }
    WfParams base;
    base.scene = c->d;
    std::memcpy(&base.cam, &c->cam, sizeof(DCamera));
    base.rd.rays_per_pixel = c->rd.rays_per_pixel;
    base.rd.reflection_limit = c->rd.reflection_limit;
    base.rd.antialias = c->rd.antialias;
    base.rd.sky[0] = c->rd.sky_colour.x; base.rd.sky[1] = c->rd.sky_colour.y; base.rd.sky[2] = c->rd.sky_colour.z;
    base.width = c->width; base.height = c->height;
    base.time_ms = time_ms; base.frame_num = frame_num;
    base.prev = c->d_image; base.out = c->d_image;
    base.peers.n = 0;
    for (int r = 0; r < RTB_MAX_PEERS; r++) base.peers.image[r] = nullptr;
    if (c->opt_exchange == 2 && !c->opt_slices && c->n_peers == c->shard_world && c->shard_world > 1 && c->d_image == c->peer_image[c->shard_rank]) {
        base.peers.n = c->n_peers;  // fused exchange: finished pixels go to every GPU's image
        for (int r = 0; r < c->n_peers; r++) base.peers.image[r] = c->peer_image[r];
    }
    base.shard_rank = c->shard_rank; base.shard_world = c->shard_world; base.band_rows = c->band_rows;
    base.local_rows = local_rows;
    float4 *st = c->d_wf_state;
    base.st.A = st; base.st.B = st + slots; base.st.C = st + 2 * slots; base.st.O = st + 3 * slots; base.st.D = st + 4 * slots; base.st.H = st + 5 * slots;
    base.alive = c->d_wf_alive;
    base.refill = c->opt_wf_refill < 1 ? c->tuned_refill : (c->opt_wf_refill > 32 ? 32 : c->opt_wf_refill);
    base.walk_min = c->opt_walk_min < 1 ? c->tuned_walk_min : (c->opt_walk_min > 32 ? 32 : c->opt_walk_min);
    base.counters = c->opt_counters ? c->d_counters : nullptr;
    const bool brute = c->opt_traversal == 1, count_ops = c->opt_counters != 0, strict_fp = c->opt_strict != 0;
    int trav_per_sm = c->opt_blocks_per_sm > 0 ? std::min(c->opt_blocks_per_sm, c->wf_trav_per_sm) : c->wf_trav_per_sm;
    const long long iterations = (long long)std::max(0, c->rd.rays_per_pixel) * std::max(0, c->rd.reflection_limit);
    const long long rebuild_every = c->opt_wf_rebuild;
    constexpr int WF_PACE_STEP = 8, WF_PACE_RING = 64;
    const int pace = std::min(c->opt_wf_pace, WF_PACE_STEP * (WF_PACE_RING - 2)) / WF_PACE_STEP * WF_PACE_STEP;   // iterations the host may run ahead (0 = unbounded)

    // The pixels are processed in `groups` consecutive row groups, each taken through ALL its
    // iterations before the next (a group's records then stay L2-resident), two groups at a time on
    // two streams (the launch gaps and the tails of one group's kernels are filled by the other's).
    int groups = std::max(1, std::min(c->opt_wf_groups, local_rows));
    // `lanes` row groups are in flight at once, each on its own stream with its own queues and counters: a group's
    // launch gaps, ramps and tails -- and the longest ray of its iteration -- are filled by the other groups' kernels.
    // Beyond two lanes every group's traversal grid shrinks, so that the groups' persistent blocks fit side by side.
    const int lanes = std::max(1, std::min(std::min(c->opt_wf_streams, groups), RTB_WF_MAX_LANES));
    if (lanes > 2 && c->opt_blocks_per_sm <= 0) trav_per_sm = std::max(1, 2 * trav_per_sm / lanes);
    const int trav_blocks = c->num_sms * trav_per_sm;
    cudaStream_t streams[RTB_WF_MAX_LANES];
    streams[0] = c->stream;
    for (int l = 1; l < lanes; l++) {
        if (!c->lane_stream[l]) {
            RTB_CUDA(c, cudaStreamCreateWithFlags(&c->lane_stream[l], cudaStreamNonBlocking), "creating a render stream");
            RTB_CUDA(c, cudaEventCreateWithFlags(&c->ev_join[l], cudaEventDisableTiming), "creating events");
        }
        streams[l] = c->lane_stream[l];
    }
    if (lanes > 1) {
        if (!c->ev_fork) RTB_CUDA(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming), "creating events");
        RTB_CUDA(c, cudaEventRecord(c->ev_fork, c->stream), "forking the render streams");
        for (int l = 1; l < lanes; l++) RTB_CUDA(c, cudaStreamWaitEvent(streams[l], c->ev_fork, 0), "forking the render streams");
    }
    const bool timing = c->opt_wf_timing != 0 && groups == 1;
    c->wf_events_used = 0;
    if (timing) {
        size_t need = (size_t)(2 * iterations + 2);
        while (c->wf_events.size() < need) {
            cudaEvent_t ev;
            RTB_CUDA(c, cudaEventCreate(&ev), "creating timing events");
            c->wf_events.push_back(ev);
        }
    }
    auto mark = [&]() { if (timing) cudaEventRecord(c->wf_events[c->wf_events_used++], c->stream); };
    mark();

    struct Group {
        WfParams p;
        int *queue[2];
        unsigned int *count[2], *tcount[2];
        int *block_count;
        size_t slot_base, slot_count;
        int shade_blocks;
        cudaStream_t st;
    };
    auto launch_shade = [&](Group &g, int first, int in, int out) {
        g.p.first = first;
        g.p.queue_in = g.queue[in]; g.p.n_in = g.count[in];
        g.p.queue_out = g.queue[out]; g.p.n_out = g.count[out];
        g.p.tn = g.tcount[first ? 0 : out];  // the traversal queue of iteration k is counted in tcount[k & 1]
        return strict_fp ? strict::launch_wf_shade(g.p, brute, count_ops, g.shade_blocks, g.st) : fast::launch_wf_shade(g.p, brute, count_ops, g.shade_blocks, g.st);
    };
    cudaError_t e = cudaSuccess;
    const int rows_per_group = (local_rows + groups - 1) / groups;
    for (int g0 = 0; g0 < groups && e == cudaSuccess; g0 += lanes) {
        Group gr[RTB_WF_MAX_LANES];
        int active = 0;
        for (int l = 0; l < lanes && g0 + l < groups; l++) {
            Group &g = gr[active];
            int r0 = (g0 + l) * rows_per_group, r1 = std::min(local_rows, r0 + rows_per_group);
            if (r1 <= r0) continue;
            g.slot_base = (size_t)r0 * c->width;
            g.slot_count = (size_t)(r1 - r0) * c->width;
            g.st = streams[l];
            g.p = base;
            g.p.slot_base = (int)g.slot_base;
            g.p.slot_count = (int)g.slot_count;
            g.queue[0] = c->d_wf_queue + g.slot_base; g.queue[1] = c->d_wf_queue + slots + g.slot_base;
            g.p.tqueue = c->d_wf_queue + 2 * slots + g.slot_base;
            unsigned int *cnt = c->d_wf_counts + 8 * l;
            g.count[0] = cnt; g.count[1] = cnt + 1; g.p.fetch = cnt + 2; g.tcount[0] = cnt + 4; g.tcount[1] = cnt + 5;
            g.block_count = c->d_wf_block_count + l * (((slots + 255) / 256) + 2);
            g.shade_blocks = (int)std::min<size_t>((g.slot_count + RTB_SHADE_THREADS - 1) / RTB_SHADE_THREADS, (size_t)c->num_sms * c->wf_shade_per_sm);
            e = cudaMemsetAsync(cnt, 0, 8 * sizeof(unsigned int), g.st);
            if (e == cudaSuccess) e = launch_shade(g, 1, 1, 0);
            active++;
        }
        mark();
        for (long long k = 0; k < iterations && e == cudaSuccess; k++) {
            int in = (int)(k & 1), out = in ^ 1;
            // Optional pacing of the host (option "wf_pace", off by default): every WF_PACE_STEP iterations each stream
            // records an event and the host waits for the one recorded wf_pace iterations earlier, so the launch
            // queues stay shallow.  Measured on C3 and C4: no effect on the frame time either way.
            if (pace > 0 && k % WF_PACE_STEP == 0) {
                const long long mark_idx = k / WF_PACE_STEP;
                for (int l = 0; l < active && e == cudaSuccess; l++) {
                    cudaEvent_t &ev = c->pace_events[l][mark_idx % WF_PACE_RING];
                    if (!ev) e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
                    if (e == cudaSuccess) e = cudaEventRecord(ev, gr[l].st);
                    long long back = mark_idx - pace / WF_PACE_STEP;
                    if (e == cudaSuccess && back >= 0) e = cudaEventSynchronize(c->pace_events[l][back % WF_PACE_RING]);
                }
            }
            for (int l = 0; l < active && e == cudaSuccess; l++) {
                Group &g = gr[l];
                if (rebuild_every > 0 && k > 0 && k % rebuild_every == 0) {  // queue[in] back into slot order
                    e = strict_fp ? strict::launch_wf_rebuild((int)g.slot_base, (int)g.slot_count, c->d_wf_alive, g.block_count, g.queue[in], g.count[in], g.st)
                                  : fast::launch_wf_rebuild((int)g.slot_base, (int)g.slot_count, c->d_wf_alive, g.block_count, g.queue[in], g.count[in], g.st);
                    if (e != cudaSuccess) break;
                }
                g.p.first = 0;
                g.p.queue_in = g.queue[in]; g.p.n_in = g.count[in];
                g.p.queue_out = g.queue[out]; g.p.n_out = g.count[out];
                g.p.tn = g.tcount[in]; g.p.tn_next = g.tcount[out];
                e = strict_fp ? strict::launch_wf_traverse(g.p, brute, count_ops, trav_blocks, g.st) : fast::launch_wf_traverse(g.p, brute, count_ops, trav_blocks, g.st);
                mark();
                if (e == cudaSuccess) e = launch_shade(g, 0, in, out);
                mark();
            }
        }
    }
    for (int l = 1; l < lanes && e == cudaSuccess; l++) {
        e = cudaEventRecord(c->ev_join[l], streams[l]);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, c->ev_join[l], 0);
    }
    if (e != cudaSuccess) return fail(c, "launching wavefront kernels", cudaGetErrorString(e));
    if (timing) {  // events: [start] shade0 [1] {traverse [2k+2] shade [2k+3]}
        RTB_CUDA(c, cudaStreamSynchronize(c->stream), "timing the wavefront kernels");
        c->wf_shade_ms = c->wf_traverse_ms = 0;
        for (size_t i = 1; i < c->wf_events_used; i++) {
            float ms = 0;
            cudaEventElapsedTime(&ms, c->wf_events[i - 1], c->wf_events[i]);
            ((i == 1 || (i & 1)) ? c->wf_shade_ms : c->wf_traverse_ms) += ms;
        }
        float total = 0;
        cudaEventElapsedTime(&total, c->wf_events[0], c->wf_events[c->wf_events_used - 1]);
        c->wf_total_ms = total;
    }
    return 0;
}

// seed of frame k of a render call.  Sample slices (SURVEY.md 8(e) partitioning B): the ranks render disjoint
// frames of one progressive sequence, frame f = k * world + rank, seeded time0 + f * step.
int frame_time(const rtb_ctx *c, int time0_ms, int k, int time_step_ms) {
    if (c->opt_slices && c->comm && c->comm_world > 1) return time0_ms + (k * c->comm_world + c->comm_rank) * time_step_ms;
    return time0_ms + k * time_step_ms;
}

int check_ready(rtb_ctx *c) {
    if (!c) return 1;
    if (!c->have_cam) return fail(c, "render", "camera not set (rtb_set_camera)");
    if (!c->have_rd) return fail(c, "render", "render data not set (rtb_set_render_data)");
    if (!c->committed) return fail(c, "render", "scene not committed (rtb_scene_commit)");
    return ensure_image(c);
}

}  // namespace

namespace {
// 8 independent FFMA chains per thread; 2 flop per FFMA
__global__ void fma_peak_kernel(float *out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == 12345.678f) out[0] = s;
}
}  // namespace

// ======================================= extern "C" ==================================================
extern "C" {

int rtb_abi_version(void) { return RTB_ABI_VERSION; }

int rtb_build_has_checks(void) {
#ifdef RTB_CHECKS
    return 1;
#else
    return 0;
#endif
}

int rtb_shard_row_of(int local_row, int rank, int world, int band_rows) {
    if (world <= 1 || band_rows <= 0) return local_row;
    int band = local_row / band_rows, within = local_row - band * band_rows;
    return (band * world + rank) * band_rows + within;
}

int rtb_shard_row_count(int height, int rank, int world, int band_rows) {
    if (world <= 1 || band_rows <= 0) return height;
    if (rank < 0 || rank >= world) return 0;
    int period = world * band_rows, full = height / period, rest = height - full * period;
    int extra = rest - rank * band_rows;
    extra = extra < 0 ? 0 : (extra > band_rows ? band_rows : extra);
    return full * band_rows + extra;
}

const char *rtb_last_error(rtb_ctx *ctx) { return ctx ? ctx->err.c_str() : rtb_create_error().c_str(); }

int rtb_create(rtb_ctx **out, int device) {
    if (!out) return fail(nullptr, "rtb_create", "null output pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess) return fail(nullptr, "rtb_create", cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(nullptr, "rtb_create", "no such CUDA device");
    int caller_device = -1;
    if (cudaGetDevice(&caller_device) != cudaSuccess) caller_device = -1;
    RTB_CUDA(nullptr, cudaSetDevice(device), "selecting device");
    struct Restore {  // the caller's current device is the caller's business
        int prev, dev;
        ~Restore() { if (prev >= 0 && prev != dev) cudaSetDevice(prev); }
    } restore{caller_device, device};
    rtb_ctx *c = new rtb_ctx();
    c->device = device;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev1);
    if (e == cudaSuccess) e = cudaMalloc((void **)&c->d_work, sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMalloc((void **)&c->d_counters, 16 * sizeof(unsigned long long));
    if (e != cudaSuccess) {
        fail(nullptr, "rtb_create", cudaGetErrorString(e));
        rtb_destroy(c);
        return 1;
    }
    c->num_sms = prop.multiProcessorCount;
    // RTB_DEVICE_FLATTEN = 0 / 1 / 2 presets option "device_flatten" (the GPU test suite runs once with 1: every mesh
    // of every parity scene then goes through the device partition + flatten)
    if (const char *env = std::getenv("RTB_DEVICE_FLATTEN")) {
        int v = std::atoi(env);
        if (v >= 0 && v <= 2) c->opt_device_flatten = v;
    }
    *out = c;
    return 0;
}

void rtb_destroy(rtb_ctx *c) {
    if (!c) return;
    Bind b(c);
    if (c->stream) cudaStreamSynchronize(c->stream);
    free_scene(c);
    rtb_comm_release(c);
    if (c->d_image) cudaFree(c->d_image);
    if (c->d_rgba) cudaFree(c->d_rgba);
    if (c->d_work) cudaFree(c->d_work);
    if (c->d_counters) cudaFree(c->d_counters);
    if (c->d_wf_state) cudaFree(c->d_wf_state);
    if (c->d_wf_queue) cudaFree(c->d_wf_queue);
    if (c->d_wf_counts) cudaFree(c->d_wf_counts);
    if (c->d_wf_alive) cudaFree(c->d_wf_alive);
    if (c->d_wf_block_count) cudaFree(c->d_wf_block_count);
    if (c->d_pool) cudaFree(c->d_pool);
    if (c->d_stream_state) cudaFree(c->d_stream_state);
    if (c->d_stream_pix) cudaFree(c->d_stream_pix);
    if (c->d_stream_rings) cudaFree(c->d_stream_rings);
    for (cudaEvent_t &ev : c->frame_done) if (ev) cudaEventDestroy(ev);
    for (auto &row : c->pace_events) for (cudaEvent_t &ev : row) if (ev) cudaEventDestroy(ev);
    if (c->d_stream_ctl) cudaFree(c->d_stream_ctl);
    if (c->h_stream_error) cudaFreeHost(c->h_stream_error);
    if (c->d_tune_scratch) cudaFree(c->d_tune_scratch);
    for (cudaEvent_t ev : c->wf_events) cudaEventDestroy(ev);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (int l = 1; l < RTB_WF_MAX_LANES; l++) {
        if (c->ev_join[l]) cudaEventDestroy(c->ev_join[l]);
        if (c->lane_stream[l]) { cudaStreamSynchronize(c->lane_stream[l]); cudaStreamDestroy(c->lane_stream[l]); }
    }
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int rtb_set_camera(rtb_ctx *c, const RtbCamData *cam, int width, int height) {
    if (!c || !cam) return 1;
    if (width <= 0 || height <= 0 || (long long)width * height * 3 > 0x7fffffffLL) return fail(c, "rtb_set_camera", "bad resolution");
    if (width != c->width || height != c->height) c->auto_choice = -1;  // a new pose keeps the tuned kernel choice, a new size does not
    c->cam = *cam;
    c->width = width; c->height = height;
    c->have_cam = true;
    return 0;
}

int rtb_set_render_data(rtb_ctx *c, const RtbRenderData *rd) {
    if (!c || !rd) return 1;
    if (rd->rays_per_pixel != c->rd.rays_per_pixel || rd->reflection_limit != c->rd.reflection_limit) c->auto_choice = -1;
    c->rd = *rd;
    c->have_rd = true;
    return 0;
}

int rtb_scene_begin(rtb_ctx *c) {
    if (!c) return 1;
    Bind b(c);
    cudaStreamSynchronize(c->stream);
    free_scene(c);
    c->objects.clear(); c->materials.clear(); c->tri_points.clear(); c->tri_uv.clear();
    c->image_data.clear(); c->images.clear();
    c->mesh_uv_given = false;
    return 0;
}

int rtb_scene_add_material(rtb_ctx *c, const RtbMaterial *m) {
    if (!c || !m) return -1;
    c->materials.push_back(*m);
    return (int)c->materials.size() - 1;
}

int rtb_scene_add_image(rtb_ctx *c, const float *rgb, int width, int height) {
    if (!c || !rgb || width <= 0 || height <= 0) return -1;
    c->image_data.emplace_back(rgb, rgb + (size_t)width * height * 3);
    c->images.push_back(RtbImage{width, height, nullptr});
    for (size_t i = 0; i < c->images.size(); i++) c->images[i].rgb = c->image_data[i].data();
    return (int)c->images.size() - 1;
}

static int add_object(rtb_ctx *c, const RtbObject &o) {
    if (!c) return -1;
    if (o.material < 0 || o.material >= (int)c->materials.size()) { fail(c, "adding object", "unknown material index"); return -1; }
    c->objects.push_back(o);
    return (int)c->objects.size() - 1;
}

int rtb_scene_add_sphere(rtb_ctx *c, const float center[3], float radius, int material) {
    RtbObject o; std::memset(&o, 0, sizeof o);
    o.type = RTB_OBJ_SPHERE; o.material = material;
    o.p[0] = RtbVec3{center[0], center[1], center[2]}; o.f[0] = radius;
    return add_object(c, o);
}

int rtb_scene_add_triangle(rtb_ctx *c, const float pts[9], const float uv[6], int material) {
    RtbObject o; std::memset(&o, 0, sizeof o);
    o.type = RTB_OBJ_TRIANGLE; o.material = material;
    for (int k = 0; k < 3; k++) o.p[k] = RtbVec3{pts[3 * k], pts[3 * k + 1], pts[3 * k + 2]};
    if (uv) std::memcpy(o.uv, uv, sizeof o.uv);
    return add_object(c, o);
}

int rtb_scene_add_quad(rtb_ctx *c, const float pts[12], int material) {
    RtbObject o; std::memset(&o, 0, sizeof o);
    o.type = RTB_OBJ_QUAD; o.material = material;
    for (int k = 0; k < 4; k++) o.p[k] = RtbVec3{pts[3 * k], pts[3 * k + 1], pts[3 * k + 2]};
    return add_object(c, o);
}

int rtb_scene_add_one_way_quad(rtb_ctx *c, const float pts[12], int invert_normal, int material) {
    RtbObject o; std::memset(&o, 0, sizeof o);
    o.type = RTB_OBJ_ONE_WAY_QUAD; o.material = material; o.invert_normal = invert_normal;
    for (int k = 0; k < 4; k++) o.p[k] = RtbVec3{pts[3 * k], pts[3 * k + 1], pts[3 * k + 2]};
    return add_object(c, o);
}

int rtb_scene_add_cuboid(rtb_ctx *c, const float tl_near[3], float width, float height, float depth, int material) {
    RtbObject o; std::memset(&o, 0, sizeof o);
    o.type = RTB_OBJ_CUBOID; o.material = material;
    o.p[0] = RtbVec3{tl_near[0], tl_near[1], tl_near[2]};
    o.f[0] = width; o.f[1] = height; o.f[2] = depth;
    return add_object(c, o);
}

int rtb_scene_add_mesh(rtb_ctx *c, const float *tri_points, const float *tri_uv, int num_tris, int material) {
    if (!c || num_tris < 0 || (num_tris > 0 && !tri_points)) return -1;
    RtbObject o; std::memset(&o, 0, sizeof o);
    o.type = RTB_OBJ_MESH; o.material = material;
    o.tri_first = (int)(c->tri_points.size() / 9); o.tri_count = num_tris;
    int idx = add_object(c, o);
    if (idx < 0) return idx;
    c->tri_points.insert(c->tri_points.end(), tri_points, tri_points + (size_t)num_tris * 9);
    if (tri_uv) c->mesh_uv_given = true;
    if (tri_uv) c->tri_uv.insert(c->tri_uv.end(), tri_uv, tri_uv + (size_t)num_tris * 6);
    else c->tri_uv.insert(c->tri_uv.end(), (size_t)num_tris * 6, 0.f);
    return idx;
}

// Dynamic scenes: the host description stays in the context after a commit,
// so a frame loop edits objects in place and commits again (flatten + upload + GPU tree rebuild).
int rtb_scene_update_object(rtb_ctx *c, int index, const RtbObject *o) {
    if (!c || !o) return 1;
    if (index < 0 || index >= (int)c->objects.size()) return fail(c, "rtb_scene_update_object", "object index out of range");
    RtbObject &cur = c->objects[index];
    if (o->material < 0 || o->material >= (int)c->materials.size()) return fail(c, "rtb_scene_update_object", "unknown material index");
    if (cur.type == RTB_OBJ_MESH || o->type == RTB_OBJ_MESH) {
        if (cur.type != o->type) return fail(c, "rtb_scene_update_object", "a mesh cannot change type (use rtb_scene_update_mesh for its triangles)");
        cur.material = o->material;  // the triangle range of a mesh is owned by the context
        return 0;
    }
    if (o->type < RTB_OBJ_SPHERE || o->type > RTB_OBJ_CUBOID) return fail(c, "rtb_scene_update_object", "unknown object type");
    cur = *o;
    return 0;
}

int rtb_scene_update_mesh(rtb_ctx *c, int index, const float *tri_points, const float *tri_uv) {
    if (!c || !tri_points) return 1;
    if (index < 0 || index >= (int)c->objects.size() || c->objects[index].type != RTB_OBJ_MESH) return fail(c, "rtb_scene_update_mesh", "not a mesh object");
    const RtbObject &o = c->objects[index];
    std::memcpy(&c->tri_points[(size_t)o.tri_first * 9], tri_points, (size_t)o.tri_count * 9 * sizeof(float));
    if (tri_uv) c->mesh_uv_given = true;
    if (tri_uv) std::memcpy(&c->tri_uv[(size_t)o.tri_first * 6], tri_uv, (size_t)o.tri_count * 6 * sizeof(float));
    return 0;
}

static int scene_commit_timed(rtb_ctx *c);

int rtb_scene_commit(rtb_ctx *c) {
    if (!c) return 1;
    auto t0 = std::chrono::steady_clock::now();
    int rc = scene_commit_timed(c);
    c->commit_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return rc;
}

static int scene_commit_timed(rtb_ctx *c) {
    Bind b(c);
    cudaStreamSynchronize(c->stream);
    free_scene(c);
    Flat f;
    auto tf0 = std::chrono::steady_clock::now();
    c->commit_partition_ms = 0;
    if (flatten(c, f) != 0) return 1;
    auto tf1 = std::chrono::steady_clock::now();
    c->commit_flatten_ms = std::chrono::duration<double, std::milli>(tf1 - tf0).count();
    DScene d{};
    // device-flattened meshes come after all host-built records
    const size_t host_tris = f.tri_ids.size(), host_leaves = f.leaf_refs.size();
    size_t dev_kept = 0;
    for (DevMesh &m : f.dev_meshes) { m.tri_base = host_tris + dev_kept; m.leaf_base = host_leaves + dev_kept; dev_kept += (size_t)m.prep.kept; }
    const size_t total_tris = host_tris + dev_kept, total_leaves = host_leaves + dev_kept;
    if (total_tris > (size_t)(0x7fffffff >> RTB_KIND_BITS)) return fail(c, "flattening scene", "too many triangles for the leaf encoding");
    float4 *w_tris = nullptr;
    int2 *w_tri_ids = nullptr;
    float2 *w_tri_uv = nullptr;
    int *w_tri_refleaf = nullptr, *w_all_refs = nullptr;
    RTB_CUDA(c, upload_front(c, f.tris, 3 * total_tris, &w_tris), "uploading triangles");
    RTB_CUDA(c, upload_front(c, f.tri_ids, total_tris, &w_tri_ids), "uploading triangle ids");
    RTB_CUDA(c, upload_front(c, f.tri_uv, 3 * total_tris, &w_tri_uv), "uploading triangle uvs");
    RTB_CUDA(c, upload_front(c, f.tri_refleaf, total_tris, &w_tri_refleaf), "uploading reference leaves");
    RTB_CUDA(c, upload_front(c, f.leaf_refs, total_leaves, &w_all_refs), "uploading primitive list");
    d.tris = w_tris; d.tri_ids = w_tri_ids; d.tri_uv = w_tri_uv; d.tri_refleaf = w_tri_refleaf; d.all_refs = w_all_refs;
    RTB_CUDA(c, upload(c, f.spheres, &d.spheres), "uploading spheres");
    RTB_CUDA(c, upload(c, f.sphere_ids, &d.sphere_ids), "uploading sphere ids");
    RTB_CUDA(c, upload(c, f.obj_material, &d.obj_material), "uploading object materials");
    RTB_CUDA(c, upload(c, f.materials, &d.materials), "uploading materials");
    RTB_CUDA(c, upload(c, f.image_rgb, &d.image_rgb), "uploading texture rgb values");
    RTB_CUDA(c, upload(c, f.ref_nodes, &d.ref_nodes), "uploading reference boxes");
    int *d_scratch = nullptr;  // [0..5] bounds of the device meshes, [6] candidate count, [8..] candidates (int2)
    constexpr int CAND_CAP = 256;
    struct ScratchGuard { int *&p; ~ScratchGuard() { cudaFree(p); } } scratch_guard{d_scratch};
    if (!f.dev_meshes.empty()) {
        RTB_CUDA(c, cudaMalloc((void **)&d_scratch, sizeof(int) * (8 + 2 * CAND_CAP)), "allocating flatten scratch");
        mesh_bounds_init(d_scratch, c->stream);
        for (DevMesh &m : f.dev_meshes) {
            MeshFill mf;
            mf.d_pts = m.d_pts; mf.d_uv = m.d_uv; mf.n = m.n;
            mf.d_slot = m.prep.d_slot; mf.d_leaf_of = m.prep.d_leaf_of;
            mf.obj = m.obj; mf.tri_base = (int)m.tri_base; mf.leaf_base = (int)m.leaf_base; mf.node_base = m.node_base;
            mf.tris = w_tris; mf.tri_ids = w_tri_ids; mf.tri_uv = w_tri_uv; mf.tri_refleaf = w_tri_refleaf; mf.leaf_refs = w_all_refs;
            mf.bounds = d_scratch;
            RTB_CUDA(c, mesh_fill_device(mf, c->stream), "flattening the mesh");
        }
        int hb[6];
        RTB_CUDA(c, cudaMemcpyAsync(hb, d_scratch, sizeof hb, cudaMemcpyDeviceToHost, c->stream), "reading mesh bounds");
        RTB_CUDA(c, cudaStreamSynchronize(c->stream), "flattening the mesh");
        if (dev_kept > 0) {
            float blo[3], bhi[3];
            mesh_bounds_decode(hb, blo, bhi);
            for (int a = 0; a < 3; a++) {
                if (blo[a] < f.lo[a]) f.lo[a] = blo[a];
                if (bhi[a] > f.hi[a]) f.hi[a] = bhi[a];
            }
        }
        for (DevMesh &m : f.dev_meshes) {  // the inputs are no longer needed
            cudaFree(m.d_pts); cudaFree(m.d_uv); m.prep.release();
            m.d_pts = nullptr; m.d_uv = nullptr;
        }
    }
    c->commit_upload_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tf1).count();
    d.num_prims = (int)total_leaves;
    // large-primitive split: a leaf whose box area exceeds 1/20 of the scene box's is tested in the
    // warp-uniform list instead of living in the tree (at most RTB_MAX_LARGE, the largest first)
    struct LargeCand {
        float area;
        long long order;   // position of the leaf in the primitive list (creation order; device meshes last)
        int ref;
        int mesh;          // device mesh it belongs to, -1 = a host-built leaf
        float box[6];
        float4 rec[6];     // records of its one or two triangles
    };
    std::vector<LargeCand> cands;
    if (c->opt_large_split && total_leaves > 0) {
        float e[3] = {f.hi[0] - f.lo[0], f.hi[1] - f.lo[1], f.hi[2] - f.lo[2]};
        float scene_area = 2.f * (e[0] * e[1] + e[1] * e[2] + e[2] * e[0]);
        const float threshold = 0.05f * scene_area;
        for (size_t i = 0; i < host_leaves; i++)
            if (f.leaf_area[i] > threshold) {
                LargeCand lc{};
                lc.area = f.leaf_area[i]; lc.order = (long long)i; lc.ref = f.leaf_refs[i]; lc.mesh = -1;
                for (int k = 0; k < 6; k++) lc.box[k] = f.leaf_box[6 * i + k];
                int enc = ~lc.ref, kind = enc & 7, idx = enc >> RTB_KIND_BITS;
                if (kind != KIND_SPHERE)
                    for (int t = 0; t < 3 * (kind == KIND_TRI ? 1 : 2); t++) lc.rec[t] = f.tris[3 * (size_t)idx + t];
                cands.push_back(lc);
            }
        for (size_t mi = 0; mi < f.dev_meshes.size(); mi++) {
            const DevMesh &m = f.dev_meshes[mi];
            if (m.prep.kept <= 0) continue;
            int2 *d_list = (int2 *)(d_scratch + 8);
            RTB_CUDA(c, cudaMemsetAsync(d_scratch + 6, 0, sizeof(int), c->stream), "large primitives of the mesh");
            RTB_CUDA(c, mesh_large_candidates(w_tris, (int)m.tri_base, m.prep.kept, threshold, d_scratch + 6, d_list, CAND_CAP, c->stream), "large primitives of the mesh");
            int count = 0;
            std::vector<int2> list(CAND_CAP);
            RTB_CUDA(c, cudaMemcpyAsync(&count, d_scratch + 6, sizeof(int), cudaMemcpyDeviceToHost, c->stream), "large primitives of the mesh");
            RTB_CUDA(c, cudaMemcpyAsync(list.data(), d_list, sizeof(int2) * CAND_CAP, cudaMemcpyDeviceToHost, c->stream), "large primitives of the mesh");
            RTB_CUDA(c, cudaStreamSynchronize(c->stream), "large primitives of the mesh");
            for (int q = 0; q < std::min(count, CAND_CAP); q++) {
                LargeCand lc{};
                std::memcpy(&lc.area, &list[q].y, sizeof(float));
                lc.order = (long long)m.leaf_base + list[q].x;
                lc.ref = ~(((int)(m.tri_base + list[q].x) << RTB_KIND_BITS) | KIND_TRI);
                lc.mesh = (int)mi;
                cands.push_back(lc);
            }
        }
        std::sort(cands.begin(), cands.end(), [](const LargeCand &x, const LargeCand &y) { return x.area != y.area ? x.area > y.area : x.order < y.order; });
        size_t limit = (size_t)std::max(0, std::min(c->opt_max_large, 32));  // bit masks of 32
        if (cands.size() > limit) cands.resize(limit);
        std::sort(cands.begin(), cands.end(), [](const LargeCand &x, const LargeCand &y) { return x.order < y.order; });
        for (LargeCand &lc : cands)
            if (lc.mesh >= 0) {  // a triangle of a device mesh: fetch its record
                int idx = (~lc.ref) >> RTB_KIND_BITS;
                RTB_CUDA(c, cudaMemcpyAsync(lc.rec, w_tris + 3 * (size_t)idx, 3 * sizeof(float4), cudaMemcpyDeviceToHost, c->stream), "large primitives of the mesh");
                RTB_CUDA(c, cudaStreamSynchronize(c->stream), "large primitives of the mesh");
                rec_box(lc.rec, 1, lc.box, lc.box + 3);
            }
    }
    std::vector<int> large, small, small_idx;  // small_idx: position of each (host-built) tree primitive in f.leaf_refs / f.leaf_box
    {
        std::vector<char> is_large(host_leaves, 0);
        for (const LargeCand &lc : cands) {
            large.push_back(lc.ref);
            if (lc.mesh < 0) is_large[(size_t)lc.order] = 1;
        }
        for (size_t i = 0; i < host_leaves; i++)
            if (!is_large[i]) { small.push_back(f.leaf_refs[i]); small_idx.push_back((int)i); }
    }
    // padded boxes of the large primitives (conservative pre-test before the exact one)
    std::vector<float4> large_boxes;
    for (const LargeCand &lc : cands) {
        const float *b = lc.box;
        float lo[3], hi[3];
        for (int k = 0; k < 3; k++) {
            float pad = 4e-7f * fmaxf(fmaxf(fabsf(b[k]), fabsf(b[3 + k])), b[3 + k] - b[k]) + 1e-12f;
            lo[k] = b[k] - pad; hi[k] = b[3 + k] + pad;
        }
        // A primitive that is flat on a coordinate axis (walls, floors: every side vector has an exactly zero
        // component there) gets that axis + 1 in lo.w: large_prims_best then proves, from the slab of that axis
        // alone, that Triangle::hit would return a distance below its 1e-6 threshold (the ray STARTS on this
        // wall) and skips the two exact tests.  See the derivation there.
        int flat_axis = 0;
        int kind = (~lc.ref) & 7;
        if (kind != KIND_SPHERE) {
            int count = kind == KIND_TRI ? 1 : 2;
            bool flat[3] = {true, true, true};
            for (int t = 0; t < count; t++) {
                float4 ta = lc.rec[3 * t], tb = lc.rec[3 * t + 1], tc = lc.rec[3 * t + 2];
                float s1[3] = {ta.w, tb.x, tb.y}, s2[3] = {tb.z, tb.w, tc.x};
                for (int k2 = 0; k2 < 3; k2++) flat[k2] = flat[k2] && s1[k2] == 0.f && s2[k2] == 0.f;
                if (t == 1) {  // both triangles of a quad start at the same point (Quad::create_triangles, objects.cu:244-253)
                    float4 t0 = lc.rec[0];
                    if (t0.x != ta.x || t0.y != ta.y || t0.z != ta.z) flat[0] = flat[1] = flat[2] = false;
                }
            }
            for (int k2 = 0; k2 < 3; k2++)
                if (flat[k2] && flat_axis == 0) flat_axis = k2 + 1;
        }
        float4 lo4 = make_float4(lo[0], lo[1], lo[2], 0.f);
        std::memcpy(&lo4.w, &flat_axis, sizeof(int));
        large_boxes.push_back(lo4);
        large_boxes.push_back(make_float4(hi[0], hi[1], hi[2], 0.f));
    }
    RTB_CUDA(c, upload(c, large_boxes, &d.large_boxes), "uploading large-primitive boxes");
    for (size_t i = 0; i < 2 * RTB_MAX_LARGE; i++) d.large_box_c[i] = i < large_boxes.size() ? large_boxes[i] : make_float4(0, 0, 0, 0);
    RTB_CUDA(c, upload(c, large, &d.large_refs), "uploading large-primitive list");
    // the tree's primitive list: the host-built leaves that are not large, then the device meshes' (a large one is
    // replaced by the last of its mesh: the order of the list does not matter to the builder)
    size_t dev_large = 0;
    for (const LargeCand &lc : cands) dev_large += lc.mesh >= 0;
    const size_t small_total = small.size() + dev_kept - dev_large;
    int *w_small = nullptr;
    RTB_CUDA(c, upload_front(c, small, small.size() + dev_kept, &w_small), "uploading tree primitive list");  // room for every mesh's whole list: the large ones leave after the copy
    {
        size_t cursor = small.size();
        for (size_t mi = 0; mi < f.dev_meshes.size(); mi++) {
            const DevMesh &m = f.dev_meshes[mi];
            size_t span = (size_t)m.prep.kept;
            if (span == 0) continue;
            {
                cudaError_t ce = cudaMemcpyAsync(w_small + cursor, w_all_refs + m.leaf_base, sizeof(int) * span, cudaMemcpyDeviceToDevice, c->stream);
                if (ce != cudaSuccess) {
                    char msg[256];
                    std::snprintf(msg, sizeof msg, "%s (mesh %zu: span %zu at %zu of %zu, from %zu of %zu; last error %s)", cudaGetErrorString(ce), mi, span, cursor, small_total,
                                  m.leaf_base, total_leaves, cudaGetErrorString(cudaPeekAtLastError()));
                    return fail(c, "copying the mesh's primitive list", msg);
                }
            }
            std::vector<size_t> removed;
            for (const LargeCand &lc : cands)
                if (lc.mesh == (int)mi) removed.push_back((size_t)(lc.order - (long long)m.leaf_base));
            std::sort(removed.rbegin(), removed.rend());
            for (size_t pos : removed) {  // descending: the tail never holds a removed entry
                if (pos != span - 1)
                    RTB_CUDA(c, cudaMemcpyAsync(w_small + cursor + pos, w_small + cursor + span - 1, sizeof(int), cudaMemcpyDeviceToDevice, c->stream), "filling the gap of a large primitive in the mesh's list");
                span--;
            }
            cursor += span;
        }
    }
    const int *d_small = w_small;
    d.num_large = (int)large.size();
    d.large_sphere_mask = 0;
    for (size_t i = 0; i < large.size() && i < 32; i++)
        if (((~large[i]) & 7) == KIND_SPHERE) d.large_sphere_mask |= 1u << i;
    d.num_bvh_prims = (int)small_total;
    d.num_tris = (int)total_tris;
    d.num_spheres = (int)f.sphere_ids.size();
    d.num_objects = (int)c->objects.size();
    c->culled_tris = f.culled;

    LbvhInput in;
    in.num_prims = d.num_bvh_prims;
    in.leaf_refs = d_small;
    in.tris = d.tris; in.spheres = d.spheres;
    float4 *nodes = nullptr;
    int root = 0, height = 0;
    RTB_CUDA(c, cudaEventRecord(c->ev0, c->stream), "timing LBVH build");
    if (c->opt_builder == 2 && in.num_prims >= 2) {
        // EXPERIMENT: full-sweep SAH on the host (tree-quality yardstick for the GPU builders)
        std::vector<float4> hn;
        height = sah_build_host(f.leaf_box, small_idx, small, hn);
        RTB_CUDA(c, cudaMalloc((void **)&nodes, hn.size() * sizeof(float4)), "allocating SAH nodes");
        RTB_CUDA(c, cudaMemcpyAsync(nodes, hn.data(), hn.size() * sizeof(float4), cudaMemcpyHostToDevice, c->stream), "uploading SAH nodes");
        root = 0;
    } else {
        cudaError_t be = lbvh_build(in, &nodes, &root, &height, c->stream, c->opt_builder == 2 ? 1 : c->opt_builder, c->opt_ploc_radius);
        if (be == cudaErrorNotSupported)  // a PLOC round merged nothing: the Karras hierarchy always exists
            be = lbvh_build(in, &nodes, &root, &height, c->stream, 0, c->opt_ploc_radius);
        RTB_CUDA(c, be, "building LBVH");
    }
    RTB_CUDA(c, cudaEventRecord(c->ev1, c->stream), "timing LBVH build");
    RTB_CUDA(c, cudaStreamSynchronize(c->stream), "building LBVH");
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->build_ms = ms;
    if (nodes) {
        c->scene_allocs.push_back(nodes);
        c->scene_bytes += sizeof(float4) * 4 * (size_t)(d.num_bvh_prims - 1);
    }
    c->tree_height = height;
    // the traversal keeps one stack entry per level (rtb_trace.cuh): a deeper tree would drop far children
    if (height > RTB_STACK - 2 && c->opt_builder != 0) {  // PLOC has no depth bound (nested / sorted input): try the Karras hierarchy
        float4 *nodes2 = nullptr;
        int root2 = 0, height2 = 0;
        RTB_CUDA(c, lbvh_build(in, &nodes2, &root2, &height2, c->stream, 0, c->opt_ploc_radius), "building LBVH");
        if (nodes2 && height2 < height) {
            c->scene_allocs.back() = nodes2;
            cudaFree(nodes);
            nodes = nodes2; root = root2; height = height2;
            c->tree_height = height;
        } else if (nodes2) {
            cudaFree(nodes2);
        }
    }
    if (height > RTB_STACK - 2) {
        char msg[160];
        std::snprintf(msg, sizeof msg, "the tree over this scene is %d levels high; the traversal stack holds %d (duplicate or nested primitives?)", height, RTB_STACK - 2);
        free_scene(c);
        return fail(c, "building LBVH", msg);
    }
    d.nodes = nodes;
    d.root_ref = root;
    for (int k = 0; k < 4; k++) d.root_node[k] = make_float4(0, 0, 0, 0);
    if (nodes && root >= 0)
        RTB_CUDA(c, cudaMemcpy(d.root_node, nodes + 4 * (size_t)root, 4 * sizeof(float4), cudaMemcpyDeviceToHost), "reading the root node");
    c->d = d;
    c->committed = true;
    // the timed kernel choice survives a commit that leaves the scene about the same size (dynamic scenes
    // commit every frame); more than 2x more or fewer primitives tune again
    if (c->tuned_prims < 0 || d.num_prims > 2 * (long long)c->tuned_prims || 2 * (long long)d.num_prims < c->tuned_prims) c->auto_choice = -1;
    return 0;
}

int rtb_scene_set(rtb_ctx *c, const RtbSceneDesc *s) {
    if (!c || !s) return 1;
    if (rtb_scene_begin(c) != 0) return 1;
    for (int i = 0; i < s->num_images; i++)
        if (rtb_scene_add_image(c, s->images[i].rgb, s->images[i].width, s->images[i].height) < 0) return fail(c, "rtb_scene_set", "bad image");
    for (int i = 0; i < s->num_materials; i++) rtb_scene_add_material(c, &s->materials[i]);
    for (int i = 0; i < s->num_objects; i++) {
        RtbObject o = s->objects[i];
        if (o.type == RTB_OBJ_MESH) {
            if (o.tri_first < 0 || o.tri_count < 0 || o.tri_first + o.tri_count > s->num_mesh_tris) return fail(c, "rtb_scene_set", "mesh triangle range out of bounds");
            const float *uv = s->mesh_tri_uv ? s->mesh_tri_uv + (size_t)o.tri_first * 6 : nullptr;
            if (rtb_scene_add_mesh(c, s->mesh_tri_points + (size_t)o.tri_first * 9, uv, o.tri_count, o.material) < 0) return 1;
        } else if (add_object(c, o) < 0) {
            return 1;
        }
    }
    return rtb_scene_commit(c);
}

int rtb_scene_desc(rtb_ctx *c, RtbSceneDesc *out) {
    if (!c || !out) return 1;
    out->num_objects = (int)c->objects.size(); out->objects = c->objects.data();
    out->num_materials = (int)c->materials.size(); out->materials = c->materials.data();
    out->num_mesh_tris = (int)(c->tri_points.size() / 9); out->mesh_tri_points = c->tri_points.data();
    out->mesh_tri_uv = c->tri_uv.data();
    out->num_images = (int)c->images.size(); out->images = c->images.data();
    return 0;
}

static int fetch_counters(rtb_ctx *c) {
    if (!c->opt_counters) return 0;
    unsigned long long h[16];
    RTB_CUDA(c, cudaMemcpyAsync(h, c->d_counters, sizeof h, cudaMemcpyDeviceToHost, c->stream), "reading counters");
    RTB_CUDA(c, cudaStreamSynchronize(c->stream), "reading counters");
    c->last_counters = RtbCounters{h[0], h[1], h[2], h[3], h[4], h[5]};
    for (int k = 0; k < 8; k++) c->last_lane_stats[k] = h[8 + k];
    return 0;
}

int rtb_render(rtb_ctx *c, float *rgb_inout, int *frame_num_inout, int current_time_ms) {
    if (!c || !rgb_inout || !frame_num_inout) return 1;
    Bind b(c);
    if (check_ready(c) != 0) return 1;
    size_t bytes = c->image_floats * sizeof(float);
    RTB_CUDA(c, cudaMemcpyAsync(c->d_image, rgb_inout, bytes, cudaMemcpyHostToDevice, c->stream), "copying previous render");
    if (c->opt_counters) RTB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 16 * sizeof(unsigned long long), c->stream), "clearing counters");
    if (rtb_exchange_after_local_write(c) != 0) return 1;
    RTB_CUDA(c, cudaEventRecord(c->ev0, c->stream), "render");
    if (launch_frame(c, frame_time(c, current_time_ms, 0, 1), *frame_num_inout) != 0) return 1;
    if (rtb_exchange_frame(c) != 0) return 1;
    RTB_CUDA(c, cudaEventRecord(c->ev1, c->stream), "render");
    RTB_CUDA(c, cudaMemcpyAsync(rgb_inout, c->d_image, bytes, cudaMemcpyDeviceToHost, c->stream), "copying image to host");
    RTB_CUDA(c, cudaStreamSynchronize(c->stream), "final check after render");
    if (check_stream_watchdog(c) != 0) return 1;
    c->timing_pending = true;
    (*frame_num_inout)++;
    c->frame_num = *frame_num_inout;
    return fetch_counters(c);
}

static int render_frames(rtb_ctx *c, int n_frames, int time0_ms, int time_step_ms, bool exchange) {
    if (!c || n_frames < 0) return 1;
    Bind b(c);
    if (check_ready(c) != 0) return 1;
    if (c->opt_counters) RTB_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 16 * sizeof(unsigned long long), c->stream), "clearing counters");
    RTB_CUDA(c, cudaEventRecord(c->ev0, c->stream), "render");
    for (int k = 0; k < n_frames; k++) {
        // Optional bounded run-ahead (option "run_ahead", off by default): at most that many frames are enqueued beyond
        // the one the device works on; for callers that want rtb_render_device to return close to the device's pace.
        if (c->opt_run_ahead > 0) {
            if (c->frames_enqueued >= c->opt_run_ahead) {
                cudaEvent_t &old = c->frame_done[(c->frames_enqueued - c->opt_run_ahead) % 4];
                if (old) RTB_CUDA(c, cudaEventSynchronize(old), "waiting for an earlier frame");
            }
        }
        if (launch_frame(c, frame_time(c, time0_ms, k, time_step_ms), c->frame_num) != 0) return 1;
        {
            cudaEvent_t &done = c->frame_done[c->frames_enqueued % 4];
            if (!done) RTB_CUDA(c, cudaEventCreateWithFlags(&done, cudaEventDisableTiming), "creating frame events");
            RTB_CUDA(c, cudaEventRecord(done, c->stream), "marking the end of a frame");
            c->frames_enqueued++;
        }
        // tiles: the image is complete on every rank after each frame (the next frame's running mean reads it);
        // sample slices: the mean over all ranks' frames so far
        if (exchange && rtb_exchange_frame(c) != 0) return 1;
        c->frame_num++;
    }
    RTB_CUDA(c, cudaEventRecord(c->ev1, c->stream), "render");
    c->timing_pending = true;
    return 0;
}

int rtb_render_device(rtb_ctx *c, int n_frames, int time0_ms, int time_step_ms) { return render_frames(c, n_frames, time0_ms, time_step_ms, true); }

int rtb_sync(rtb_ctx *c) {
    if (!c) return 1;
    Bind b(c);
    RTB_CUDA(c, cudaStreamSynchronize(c->stream), "final check after render");
    if (check_stream_watchdog(c) != 0) return 1;
    return fetch_counters(c);
}

int rtb_reset_accumulation(rtb_ctx *c) {
    if (!c) return 1;
    Bind b(c);
    if (ensure_image(c) != 0) return 1;
    RTB_CUDA(c, cudaMemsetAsync(c->d_image, 0, c->image_floats * sizeof(float), c->stream), "clearing image");
    c->frame_num = 0;
    return rtb_exchange_after_local_write(c);
}

int rtb_set_accumulation(rtb_ctx *c, const float *rgb_host, int frame_num) {
    if (!c || !rgb_host) return 1;
    Bind b(c);
    if (ensure_image(c) != 0) return 1;
    RTB_CUDA(c, cudaMemcpyAsync(c->d_image, rgb_host, c->image_floats * sizeof(float), cudaMemcpyHostToDevice, c->stream), "copying previous render");
    c->frame_num = frame_num;
    if (rtb_exchange_after_local_write(c) != 0) return 1;
    RTB_CUDA(c, cudaStreamSynchronize(c->stream), "copying previous render");
    return 0;
}

int rtb_download(rtb_ctx *c, float *rgb_host, int *frame_num_out) {
    if (!c || !rgb_host) return 1;
    Bind b(c);
    if (ensure_image(c) != 0) return 1;
    RTB_CUDA(c, cudaMemcpyAsync(rgb_host, c->d_image, c->image_floats * sizeof(float), cudaMemcpyDeviceToHost, c->stream), "copying image to host");
    RTB_CUDA(c, cudaStreamSynchronize(c->stream), "copying image to host");
    if (check_stream_watchdog(c) != 0) return 1;
    if (frame_num_out) *frame_num_out = c->frame_num;
    return 0;
}

int rtb_display_rgba8(rtb_ctx *c, uint8_t *rgba_host) {
    if (!c || !rgba_host) return 1;
    Bind b(c);
    if (ensure_image(c) != 0) return 1;
    size_t pixels = c->image_floats / 3;
    if (c->rgba_pixels != pixels) {
        if (c->d_rgba) cudaFree(c->d_rgba);
        c->d_rgba = nullptr; c->rgba_pixels = 0;
        RTB_CUDA(c, cudaMalloc((void **)&c->d_rgba, pixels * 4), "allocating display buffer");
        c->rgba_pixels = pixels;
    }
    display_rgba8_kernel<<<(unsigned int)((pixels + 255) / 256), 256, 0, c->stream>>>(c->d_image, (uchar4 *)c->d_rgba, pixels);
    RTB_CUDA(c, cudaGetLastError(), "launching display conversion");
    RTB_CUDA(c, cudaMemcpyAsync(rgba_host, c->d_rgba, pixels * 4, cudaMemcpyDeviceToHost, c->stream), "copying display buffer to host");
    RTB_CUDA(c, cudaStreamSynchronize(c->stream), "copying display buffer to host");
    return 0;
}

int rtb_write_ppm(rtb_ctx *c, const char *path) {
    if (!c || !path) return 1;
    std::vector<uint8_t> rgba((size_t)c->width * c->height * 4);
    if (rgba.empty()) return fail(c, "rtb_write_ppm", "camera/resolution not set");
    if (rtb_display_rgba8(c, rgba.data()) != 0) return 1;
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(c, "rtb_write_ppm", "cannot open the output file");
    std::fprintf(f, "P6\n%d %d\n255\n", c->width, c->height);
    std::vector<uint8_t> rgb((size_t)c->width * c->height * 3);
    for (size_t i = 0; i < (size_t)c->width * c->height; i++) { rgb[3 * i] = rgba[4 * i]; rgb[3 * i + 1] = rgba[4 * i + 1]; rgb[3 * i + 2] = rgba[4 * i + 2]; }
    size_t w = std::fwrite(rgb.data(), 1, rgb.size(), f);
    std::fclose(f);
    if (w != rgb.size()) return fail(c, "rtb_write_ppm", "short write");
    return 0;
}

void *rtb_device_image(rtb_ctx *c) {
    if (!c) return nullptr;
    Bind b(c);
    if (ensure_image(c) != 0) return nullptr;
    return c->d_image;
}

void *rtb_stream(rtb_ctx *c) { return c ? (void *)c->stream : nullptr; }

int rtb_last_device_ms(rtb_ctx *c, float *ms_out) {
    if (!c || !ms_out) return 1;
    Bind b(c);
    if (!c->timing_pending) return fail(c, "rtb_last_device_ms", "no timed render yet");
    RTB_CUDA(c, cudaEventSynchronize(c->ev1), "timing");
    RTB_CUDA(c, cudaEventElapsedTime(ms_out, c->ev0, c->ev1), "timing");
    return 0;
}

int rtb_set_shard(rtb_ctx *c, int rank, int world, int band_rows) {
    if (!c) return 1;
    if (world < 1 || rank < 0 || rank >= world || band_rows < 1) return fail(c, "rtb_set_shard", "bad shard arguments");
    if (world != c->shard_world || band_rows != c->band_rows) c->auto_choice = -1;
    c->shard_rank = rank; c->shard_world = world; c->band_rows = band_rows;
    return 0;
}

int rtb_shard_rows(rtb_ctx *c, int *rows_this_rank, int *rows_padded) {
    if (!c) return 1;
    int mx = 0;
    for (int r = 0; r < c->shard_world; r++) { int n = shard_local_rows(c, r); if (n > mx) mx = n; }
    if (rows_this_rank) *rows_this_rank = shard_local_rows(c, c->shard_rank);
    if (rows_padded) *rows_padded = mx;
    return 0;
}

int rtb_shard_pack(rtb_ctx *c, void *dense_dev) {
    if (!c || !dense_dev) return 1;
    Bind b(c);
    if (ensure_image(c) != 0) return 1;
    int rows = shard_local_rows(c, c->shard_rank);
    int n = rows * c->width * 3;
    if (n > 0) shard_pack_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(c->d_image, (float *)dense_dev, c->width, rows, c->shard_rank, c->shard_world, c->band_rows, c->height);
    RTB_CUDA(c, cudaGetLastError(), "packing shard");
    return 0;
}

int rtb_shard_unpack(rtb_ctx *c, const void *gathered_dev) {
    if (!c || !gathered_dev) return 1;
    Bind b(c);
    if (ensure_image(c) != 0) return 1;
    int padded = 0;
    rtb_shard_rows(c, nullptr, &padded);
    size_t n = (size_t)padded * c->width * 3 * c->shard_world;
    if (n > 0) shard_unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d_image, (const float *)gathered_dev, c->width, padded, c->shard_world, c->band_rows, c->height);
    RTB_CUDA(c, cudaGetLastError(), "unpacking shards");
    return 0;
}

int rtb_primary_ids(rtb_ctx *c, int32_t *obj_id, int32_t *sub_id, float *t) {
    if (!c || !obj_id || !sub_id) return 1;
    Bind b(c);
    if (!c->have_cam) return fail(c, "rtb_primary_ids", "camera not set");
    if (!c->committed) return fail(c, "rtb_primary_ids", "scene not committed");
    size_t n = (size_t)c->width * c->height;
    int32_t *d_obj = nullptr, *d_sub = nullptr;
    float *d_t = nullptr;
    RTB_CUDA(c, cudaMalloc((void **)&d_obj, n * 4), "allocating id buffers");
    cudaError_t e = cudaMalloc((void **)&d_sub, n * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_t, n * 4);
    if (e == cudaSuccess) {
        IdsParams p;
        p.scene = c->d;
        std::memcpy(&p.cam, &c->cam, sizeof(DCamera));
        p.width = c->width; p.height = c->height;
        p.obj_id = d_obj; p.sub_id = d_sub; p.t = d_t;
        e = c->opt_strict ? strict::launch_primary_ids(p, c->opt_traversal == 1, c->stream) : fast::launch_primary_ids(p, c->opt_traversal == 1, c->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(obj_id, d_obj, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sub_id, d_sub, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && t) e = cudaMemcpyAsync(t, d_t, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_obj); cudaFree(d_sub); cudaFree(d_t);
    if (e != cudaSuccess) return fail(c, "rtb_primary_ids", cudaGetErrorString(e));
    return 0;
}

// Test hook: the reference-BVH partition of one mesh by the host restatement (rtb_refbvh.cpp) or by the device version
// (rtb_meshprep.cu), so that a test can compare the two array by array.
int rtb_debug_mesh_partition(int device, const float *tri_points, int n, int on_device, int32_t *leaf_of, uint8_t *degenerate, float *nodes,
                             int nodes_cap, int *num_nodes) {
    if (!tri_points || n <= 0 || !leaf_of || !degenerate || !nodes || !num_nodes) return 1;
    if (!on_device) {
        RefBvh rb;
        ref_bvh_partition(tri_points, n, rb);
        *num_nodes = (int)rb.nodes.size();
        if (*num_nodes > nodes_cap) return 1;
        for (int i = 0; i < n; i++) { leaf_of[i] = rb.leaf_of[i]; degenerate[i] = (uint8_t)rb.degenerate[i]; }
        for (int q = 0; q < *num_nodes; q++) {
            const RefBvhNode &nd = rb.nodes[q];
            float *o = nodes + 8 * (size_t)q;
            o[0] = nd.lo[0]; o[1] = nd.lo[1]; o[2] = nd.lo[2]; o[3] = (float)nd.parent;
            o[4] = nd.hi[0]; o[5] = nd.hi[1]; o[6] = nd.hi[2]; o[7] = 0.f;
        }
        return 0;
    }
    int prev = -1;
    cudaGetDevice(&prev);
    if (cudaSetDevice(device) != cudaSuccess) return 1;
    float *d_pts = nullptr;
    MeshPrep prep;
    int nonfinite = 0, rc = 1;
    std::vector<int> slot((size_t)n);
    cudaError_t e = cudaMalloc((void **)&d_pts, sizeof(float) * 9 * (size_t)n);
    if (e == cudaSuccess) e = cudaMemcpy(d_pts, tri_points, sizeof(float) * 9 * (size_t)n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = mesh_partition_device(d_pts, n, nullptr, prep, &nonfinite);
    if (e == cudaSuccess) e = cudaMemcpy(leaf_of, prep.d_leaf_of, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(slot.data(), prep.d_slot, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && !nonfinite && (int)(prep.nodes.size() / 2) <= nodes_cap) {
        *num_nodes = (int)(prep.nodes.size() / 2);
        int kept = 0;
        bool ordered = true;
        for (int i = 0; i < n; i++) {
            degenerate[i] = slot[i] < 0;
            if (slot[i] >= 0) ordered = ordered && slot[i] == kept++;
        }
        for (int q = 0; q < *num_nodes; q++) {
            float4 lo = prep.nodes[2 * (size_t)q], hi = prep.nodes[2 * (size_t)q + 1];
            int parent;
            std::memcpy(&parent, &lo.w, sizeof(int));
            float *o = nodes + 8 * (size_t)q;
            o[0] = lo.x; o[1] = lo.y; o[2] = lo.z; o[3] = (float)parent;
            o[4] = hi.x; o[5] = hi.y; o[6] = hi.z; o[7] = 0.f;
        }
        rc = (ordered && kept == prep.kept) ? 0 : 1;
    }
    cudaFree(d_pts);
    prep.release();
    if (prev >= 0) cudaSetDevice(prev);
    return rc;
}

int rtb_counters(rtb_ctx *c, RtbCounters *out) {
    if (!c || !out) return 1;
    *out = c->last_counters;
    return 0;
}

int rtb_set_option(rtb_ctx *c, const char *key, int value) {
    if (!c || !key) return 1;
    std::string k(key);
    c->auto_choice = -1;  // every option shapes the kernels: tune again at the next frame
    if (k == "counters") c->opt_counters = value;
    else if (k == "traversal") c->opt_traversal = value;
    else if (k == "strict_fp") { c->opt_strict = value; c->wf_shade_per_sm = 0; c->pool_per_sm = 0; c->stream_per_sm = 0; }
    else if (k == "exchange") {
        if (value < 0 || value > 2) return fail(c, "rtb_set_option", "exchange: 0 = caller's, 1 = NCCL all-gather / all-reduce, 2 = fused peer stores");
        c->opt_exchange = value;
    }
    else if (k == "slices") c->opt_slices = value != 0;
    else if (k == "wf_pace") c->opt_wf_pace = value < 0 ? 0 : value;   // iterations the host runs ahead of each wavefront stream (0 = unbounded)
    else if (k == "run_ahead") c->opt_run_ahead = value < 0 ? 0 : (value > 3 ? 3 : value);   // frames enqueued ahead of the device (0 = unbounded)
    else if (k == "tune_set") c->opt_tune_set = value & 27;   // kernels the timed choice tries: 1 megakernel, 2 wavefront, 8 pooled, 16 streaming
    else if (k == "stream_slots") c->opt_stream_slots = value < 1 ? 1 : value;
    else if (k == "stream_shade_sms") { c->opt_stream_shade_sms = value; c->stream_shade_sms = 0; }
    else if (k == "reference_bvh_quirks") c->opt_ref_quirks = value;  // takes effect at the next commit
    else if (k == "mega_early_pop") c->opt_mega_early_pop = value < 0 ? -1 : (value != 0);   // megakernel traversal variant: -1 = timed choice
    else if (k == "builder") c->opt_builder = value;
    else if (k == "device_flatten") {   // takes effect at the next commit
        if (value < 0 || value > 2) return fail(c, "rtb_set_option", "device_flatten: 0 = meshes flattened on the host, 1 = on the device, 2 = on the device from 32768 triangles up");
        c->opt_device_flatten = value;
    }
    else if (k == "ploc_radius") c->opt_ploc_radius = value;
    else if (k == "large_split") c->opt_large_split = value;          // takes effect at the next commit
    else if (k == "max_large") c->opt_max_large = value;
    else if (k == "refill") c->opt_refill = value;  // lanes waiting before a warp leaves the tree to shade
    else if (k == "walk_min") c->opt_walk_min = value;
    else if (k == "wf_timing") c->opt_wf_timing = value;
    else if (k == "wf_rebuild") c->opt_wf_rebuild = value < 0 ? 0 : value;
    else if (k == "wf_groups") c->opt_wf_groups = value < 1 ? 1 : value;
    else if (k == "wf_streams") c->opt_wf_streams = value < 1 ? 1 : value;
    else if (k == "wf_refill") c->opt_wf_refill = value;  // free lanes before a traversal warp fetches new rays
    else if (k == "kernel") {
        if (value < 0 || value > 4) return fail(c, "rtb_set_option", "kernel: 0 = megakernel, 1 = wavefront, 2 = auto (timed choice), 3 = pooled persistent kernel, 4 = streaming kernel");
        c->opt_kernel = value;
    }
    else if (k == "block_threads") {
        if (value != RTB_BLOCK_THREADS) return fail(c, "rtb_set_option", "block_threads is fixed at build time");
    } else if (k == "blocks_per_sm") c->opt_blocks_per_sm = value;
    else return fail(c, "rtb_set_option", "unknown option");
    return 0;
}

int rtb_measure_fp32_peak(rtb_ctx *c, double ms_target, double *tflops_out) {
    if (!c || !tflops_out) return 1;
    Bind b(c);
    float *d = nullptr;
    RTB_CUDA(c, cudaMalloc((void **)&d, 4), "fp32 peak");
    const int threads = 256, blocks = c->num_sms * 8;
    int iters = 2000;
    float ms = 0;
    double best = 0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(c->ev0, c->stream);
        fma_peak_kernel<<<blocks, threads, 0, c->stream>>>(d, iters, 0.999f, 0.001f);
        cudaEventRecord(c->ev1, c->stream);
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { cudaFree(d); return fail(c, "fp32 peak", cudaGetErrorString(e)); }
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        double flops = 2.0 * 8 * 16 * (double)iters * threads * blocks;
        double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
        if (ms < ms_target && iters < (1 << 24)) iters = (int)(iters * (ms_target / (ms > 0.01f ? ms : 0.01f)) * 1.1) + 1;
    }
    cudaFree(d);
    c->timing_pending = false;
    *tflops_out = best;
    return 0;
}

int rtb_get_info(rtb_ctx *c, const char *key, double *value) {
    if (!c || !key || !value) return 1;
    std::string k(key);
    if (k == "num_prims") *value = c->d.num_prims;
    else if (k == "num_tris") *value = c->d.num_tris;
    else if (k == "num_spheres") *value = c->d.num_spheres;
    else if (k == "num_nodes") *value = c->d.num_bvh_prims > 1 ? c->d.num_bvh_prims - 1 : 0;
    else if (k == "num_large") *value = c->d.num_large;
    else if (k == "tree_height") *value = c->tree_height;
    else if (k == "exchange") *value = c->opt_exchange;
    else if (k == "exchange_fused") *value = (c->opt_exchange == 2 && !c->opt_slices && c->n_peers == c->shard_world && c->shard_world > 1) ? 1 : 0;
    else if (k == "comm_world") *value = c->comm ? c->comm_world : 1;
    else if (k == "build_ms") *value = c->build_ms;
    else if (k == "commit_ms") *value = c->commit_ms;
    else if (k == "commit_partition_ms") *value = c->commit_partition_ms;
    else if (k == "commit_flatten_ms") *value = c->commit_flatten_ms;
    else if (k == "commit_upload_ms") *value = c->commit_upload_ms;
    else if (k == "kernel") *value = effective_kernel(c);      // the variant frames are rendered with
    else if (k == "kernel_option") *value = c->opt_kernel;     // 0, 1 or 2 = auto
    else if (k == "kernel_tuned") *value = c->opt_kernel == 2 && c->auto_choice >= 0 ? 1 : 0;
    else if (k == "tune_megakernel_ms") *value = c->tune_ms[0];
    else if (k == "tune_wavefront_ms") *value = c->tune_ms[1];
    else if (k == "tune_pool_ms") *value = c->tune_ms[3];
    else if (k == "tune_stream_ms") *value = c->tune_ms[4];
    else if (k == "stream_shade_sms") *value = c->stream_shade_sms;
    else if (k == "stream_util_shade") *value = c->stream_util[0];
    else if (k == "stream_util_traverse") *value = c->stream_util[1];
    else if (k == "stream_slots") *value = (double)c->stream_slots;
    else if (k == "walk_min") *value = c->opt_walk_min > 0 ? c->opt_walk_min : (effective_kernel(c) != 0 ? c->tuned_walk_min : 8);
    else if (k == "wf_refill") *value = c->opt_wf_refill > 0 ? c->opt_wf_refill : c->tuned_refill;
    else if (k == "wf_shade_ms") *value = c->wf_shade_ms;
    else if (k == "wf_traverse_ms") *value = c->wf_traverse_ms;
    else if (k == "wf_total_ms") *value = c->wf_total_ms;
    else if (k == "device_flatten") *value = c->opt_device_flatten;
    else if (k == "mega_early_pop") *value = c->opt_mega_early_pop >= 0 ? c->opt_mega_early_pop : c->tuned_mega_early_pop;
    else if (k == "builder") *value = c->opt_builder;
    else if (k == "launches_per_frame") {  // kernels of ONE frame: megakernel 1; wavefront per row group 1 + 2 * spp * bounces + 3 per queue rebuild
        double it = (double)std::max(0, c->rd.rays_per_pixel) * std::max(0, c->rd.reflection_limit);
        double rebuilds = c->opt_wf_rebuild > 0 && it > 0 ? std::floor((it - 1) / c->opt_wf_rebuild) : 0;
        int groups = std::max(1, std::min(c->opt_wf_groups, shard_local_rows(c, c->shard_rank)));
        *value = effective_kernel(c) == 1 ? groups * (1.0 + 2.0 * it + 3.0 * rebuilds) : 1.0;  // megakernel and pooled kernel: one launch
    }
    else if (k.rfind("lane_stat", 0) == 0 && k.size() == 10 && k[9] >= '0' && k[9] <= '7') *value = (double)c->last_lane_stats[k[9] - '0'];
    else if (k == "scene_bytes") *value = (double)c->scene_bytes;
    else if (k == "culled_tris") *value = c->culled_tris;
    else if (k == "num_sms") *value = c->num_sms;
    else if (k == "block_threads") *value = c->opt_threads;
    else if (k == "blocks_per_sm") {
        int per_sm = c->opt_blocks_per_sm;
        if (per_sm <= 0) { Bind b(c); (c->opt_strict ? strict::render_occupancy : fast::render_occupancy)(c->opt_threads, false, &per_sm); }
        *value = per_sm;
    } else if (k == "frame_num") *value = c->frame_num;
    else return fail(c, "rtb_get_info", "unknown key");
    return 0;
}

}  // extern "C"

// ---- internals used by rtb_comm.cu ---------------------------------------------------------------------------------
int rtb_set_accumulation_async(rtb_ctx *c, const float *rgb_host, int frame_num) {
    Bind b(c);
    if (ensure_image(c) != 0) return 1;
    RTB_CUDA(c, cudaMemcpyAsync(c->d_image, rgb_host, c->image_floats * sizeof(float), cudaMemcpyHostToDevice, c->stream), "copying previous render");
    c->frame_num = frame_num;
    return 0;
}

int rtb_render_frames_no_exchange(rtb_ctx *c, int n_frames, int time0_ms, int time_step_ms) { return render_frames(c, n_frames, time0_ms, time_step_ms, false); }

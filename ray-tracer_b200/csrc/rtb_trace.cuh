// rtb_trace.cuh -- closest-hit search: primitive tests with the reference's exact expressions,
// an LBVH traversal that is conservative with respect to them, and the reference's tie rules made
// traversal-order independent.
#pragma once
#include "rtb_device.cuh"

namespace rtb {

struct Ray {
    V3 o, d, inv;  // origin, direction (normalised), 1/direction (ray.cu:198-202)
};

// best candidate so far.  code = (index << 3) | (which << 2) | kind, or -1 for none.
struct Best {
    float t, u, v;  // distance and the barycentrics Triangle::hit computes (objects.cu:143-146)
    int code;
};

#ifdef RTB_COUNTERS
struct Counters {
    unsigned long long box, tri, sph;
};
#define RTB_COUNT(c, f) ((c).f++)
#else
struct Counters {};
#define RTB_COUNT(c, f)
#endif

// ---- Triangle::hit, objects.cu:135-163 ---------------------------------------------------------
// T -> {p0.xyz, s1.x} {s1.yz, s2.xy} {s2.z, n.xyz}.  Early exits only skip work whose result the
// reference discards (hit is a conjunction); t, u, v of an accepted hit are the reference's.
__device__ __forceinline__ bool tri_test(const float4 *__restrict__ T, const Ray &r, float &t, float &u, float &v) {
    float4 a = __ldg(T), b = __ldg(T + 1), c = __ldg(T + 2);
    V3 p0 = v3(a.x, a.y, a.z), side1 = v3(a.w, b.x, b.y), side2 = v3(b.z, b.w, c.x);
    V3 p_vec = cross(r.d, side2);
    float det = dot(side1, p_vec);
    float inv_det = 1 / det;
    V3 t_vec = r.o - p0;
    u = dot(t_vec, p_vec) * inv_det;
    if (!(u >= 0)) return false;
    V3 q_vec = cross(t_vec, side1);
    v = dot(r.d, q_vec) * inv_det;
    float w = 1 - u - v;
    if (!(v >= 0) || !(w >= 0)) return false;
    t = dot(side2, q_vec) * inv_det;
    return t > RTB_EPS;
}

// ---- Sphere::hit, objects.cu:40-79 (near root only, no inside hits) ----------------------------
__device__ __forceinline__ bool sphere_test(float4 s, const Ray &r, float &t) {
    V3 c_min_q = v3(s.x, s.y, s.z) - r.o;
    float a = dot(r.d, r.d);
    float b = dot(r.d, c_min_q) * (-2);
    float c = dot(c_min_q, c_min_q) - s.w * s.w;
    float disc = b * b - 4 * a * c;
    if (!(disc >= 0)) return false;
    t = (-b - sqrtf(disc)) / (2 * a);
    return t > RTB_EPS;
}

// ---- ids and the tie rule --------------------------------------------------------------------
__device__ __forceinline__ int2 code_ids(const DScene &s, int code) {
    int kind = code & 3, idx = code >> 3;
    if (kind == KIND_TRI) return __ldg(s.tri_ids + idx);
    if (kind == KIND_QUAD) {
        int2 q = __ldg(s.quad_ids + idx);
        q.y += (code >> 2) & 1;
        return q;
    }
    return make_int2(__ldg(s.sphere_ids + idx), 0);
}

// Exact tie (bit-equal t).  Across objects the reference keeps the LATER object (`<=`,
// raytracer.cu:36); inside a cuboid or mesh the FIRST tested primitive (`<`, objects.cu:315,596).
static __device__ __noinline__ bool wins_tie(const DScene &s, int cand, int best) {
    if (best < 0) return true;
    int2 c = code_ids(s, cand), b = code_ids(s, best);
    return c.x > b.x || (c.x == b.x && c.y < b.y);
}

// ---- BoundingBox::ray_hits on the reference's own tree (objects.cu:404-434) ---------------------
// Only consulted when a direction component is exactly zero: then 0 * inf = NaN can make the
// reference miss a box that contains the hit (its strict `tmin < tmax` and NaN-dropping min/max).
static __device__ __noinline__ bool ref_path_hit(const DScene &s, int leaf, const Ray &r) {
    for (int n = leaf; n >= 0;) {
        float4 lo = __ldg(s.ref_nodes + 2 * n), hi = __ldg(s.ref_nodes + 2 * n + 1);
        float tmin = 0, tmax = RTB_INF_F;
        float t1 = (lo.x - r.o.x) * r.inv.x, t2 = (hi.x - r.o.x) * r.inv.x;
        tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
        t1 = (lo.y - r.o.y) * r.inv.y; t2 = (hi.y - r.o.y) * r.inv.y;
        tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
        t1 = (lo.z - r.o.z) * r.inv.z; t2 = (hi.z - r.o.z) * r.inv.z;
        tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
        if (!(tmin < tmax && tmax > 0)) return false;
        n = __float_as_int(lo.w);
    }
    return true;
}

__device__ __forceinline__ void consider(const DScene &s, Best &best, float t, float u, float v, int code) {
    if (t < best.t || (t == best.t && wins_tie(s, code, best.code))) {
        best.t = t; best.u = u; best.v = v; best.code = code;
    }
}

// test one leaf (ref < 0) and fold it into `best`
__device__ __forceinline__ void test_leaf(const DScene &s, const Ray &r, int ref, Best &best, Counters &cnt) {
    int enc = ~ref, kind = enc & 3, idx = enc >> 2;
    float t, u = 0, v = 0;
    if (kind == KIND_TRI) {
        RTB_COUNT(cnt, tri);
        if (tri_test(s.tris + 3 * idx, r, t, u, v)) {
            if (r.d.x == 0 || r.d.y == 0 || r.d.z == 0) {
                int leaf = __ldg(s.tri_refleaf + idx);
                if (leaf >= 0 && !ref_path_hit(s, leaf, r)) return;
            }
            consider(s, best, t, u, v, (idx << 3) | KIND_TRI);
        }
    } else if (kind == KIND_QUAD) {
        const float4 *Q = s.quads + 7 * idx;
        float4 ow = __ldg(Q + 6);
        // OneWayQuad::hit, objects.cu:273-280
        if (__float_as_int(ow.w) != 0 && dot(r.d, v3(ow.x, ow.y, ow.z)) < 0) return;
        // Quad::hit, objects.cu:223-236: triangle 1 if it hits, else whatever triangle 2 says
        RTB_COUNT(cnt, tri);
        if (tri_test(Q, r, t, u, v)) {
            consider(s, best, t, u, v, (idx << 3) | KIND_QUAD);
        } else {
            RTB_COUNT(cnt, tri);
            if (tri_test(Q + 3, r, t, u, v)) consider(s, best, t, u, v, (idx << 3) | 4 | KIND_QUAD);
        }
    } else {
        RTB_COUNT(cnt, sph);
        if (sphere_test(__ldg(s.spheres + idx), r, t)) consider(s, best, t, 0.f, 0.f, (idx << 3) | KIND_SPHERE);
    }
}

// Slab test of one child box against [0, limit].  Not parity surface: it only has to be
// conservative, i.e. never reject a box that holds a primitive the exact tests above would accept
// closer than `best`.  NaNs (0 * inf) are dropped by fminf/fmaxf, which widens the interval; the
// final comparison carries a relative slack for the rounding of the six products.
__device__ __forceinline__ bool box_test(float4 lo, float4 hi, const Ray &r, float limit, float &tnear) {
    float t1 = (lo.x - r.o.x) * r.inv.x, t2 = (hi.x - r.o.x) * r.inv.x;
    float tmin = fminf(t1, t2), tmax = fmaxf(t1, t2);
    t1 = (lo.y - r.o.y) * r.inv.y; t2 = (hi.y - r.o.y) * r.inv.y;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = (lo.z - r.o.z) * r.inv.z; t2 = (hi.z - r.o.z) * r.inv.z;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    tmin = fmaxf(tmin, 0.f);
    tmax = fminf(tmax, limit);
    tnear = tmin;
    return tmin <= tmax * 1.000004f;
}

#ifndef RTB_STACK
#define RTB_STACK 64
#endif

// closest hit over the whole scene through the LBVH
__device__ __forceinline__ Best trace_bvh(const DScene &s, const Ray &r, Counters &cnt) {
    Best best;
    best.t = RTB_INF_F; best.u = 0; best.v = 0; best.code = -1;
    if (s.num_prims == 0) return best;
    int stack[RTB_STACK];
    int sp = 0;
    int cur = s.root_ref;
    for (;;) {
        if (cur >= 0) {
            const float4 *N = s.nodes + 4 * cur;
            float4 lo0 = __ldg(N), hi0 = __ldg(N + 1), lo1 = __ldg(N + 2), hi1 = __ldg(N + 3);
            // relative slack on the distance cull: a primitive's own t carries more rounding than
            // the box planes do (cancellation in Moller-Trumbore), so keep boxes that start a
            // hair beyond the current best
            float limit = best.t * 1.0001f;
            float n0, n1;
            RTB_COUNT(cnt, box); RTB_COUNT(cnt, box);
            bool h0 = box_test(lo0, hi0, r, limit, n0);
            bool h1 = box_test(lo1, hi1, r, limit, n1);
            int r0 = __float_as_int(lo0.w), r1 = __float_as_int(lo1.w);
            if (h0 && h1) {
                bool first0 = n0 <= n1;
                if (sp < RTB_STACK) stack[sp++] = first0 ? r1 : r0;
                cur = first0 ? r0 : r1;
                continue;
            }
            if (h0) { cur = r0; continue; }
            if (h1) { cur = r1; continue; }
        } else {
            test_leaf(s, r, cur, best, cnt);
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return best;
}

// brute force over every primitive in creation order (option "traversal" = 1): the LBVH's checker
__device__ __forceinline__ Best trace_brute(const DScene &s, const Ray &r, Counters &cnt) {
    Best best;
    best.t = RTB_INF_F; best.u = 0; best.v = 0; best.code = -1;
    for (int i = 0; i < s.num_prims; i++) test_leaf(s, r, __ldg(s.all_refs + i), best, cnt);
    return best;
}

}  // namespace rtb

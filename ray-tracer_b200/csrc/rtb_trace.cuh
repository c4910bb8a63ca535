// rtb_trace.cuh -- closest-hit search: primitive tests with the reference's exact expressions,
// an LBVH traversal that is conservative with respect to them, and the reference's tie rules made
// traversal-order independent.
#pragma once
#include "rtb_device.cuh"

namespace rtb {

struct Ray {
    V3 o, d, inv;  // origin, direction (normalised), 1/direction (ray.cu:198-202)
};

// best candidate so far.  code = (index << 1) | is_sphere with index into tris / spheres, -1 = none.
struct Best {
    float t, u, v;  // distance and the barycentrics Triangle::hit computes (objects.cu:143-146)
    int code;
};

#ifdef RTB_COUNTERS
struct Counters {
    unsigned long long box, tri, sph;
    // lane-occupancy statistics of the traversal loops (counted by lane 0 of each warp):
    // node_iters / leaf_iters = warp-level iterations, *_lanes = lanes doing useful work in them,
    // blocked = lanes holding two leaves, done = lanes whose ray is finished, in node iterations
    unsigned long long node_iters, node_lanes, node_blocked, node_done, leaf_iters, leaf_lanes, mask_lanes;
};
#define RTB_COUNT(c, f) ((c).f++)
#define RTB_COUNT_N(c, f, n) ((c).f += (n))
#else
struct Counters {};
#define RTB_COUNT(c, f)
#define RTB_COUNT_N(c, f, n)
#endif

// ---- Triangle::hit, objects.cu:135-163 ---------------------------------------------------------
// T -> {p0.xyz, s1.x} {s1.yz, s2.xy} {s2.z, n.xyz}.  Early exits only skip work whose result the
// reference discards (hit is a conjunction); t, u, v of an accepted hit are the reference's.
// (Measured and rejected: fetching the 48-byte record with two requests -- a 256-bit + a 128-bit load
// picked by the record's parity, or a 64-byte padded record with two 256-bit loads -- is 1-5 % slower
// than three 128-bit loads: more registers in flight / more sectors, and the third word is needed last.)
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
    if (code & 1) return make_int2(__ldg(s.sphere_ids + (code >> 1)), 0);
    return __ldg(s.tri_ids + (code >> 1));
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

// test one leaf (ref < 0) and fold it into `best`.  One code path serves single triangles and
// quads (two consecutive triangles) so that lanes holding either kind stay converged.
__device__ __forceinline__ void test_leaf(const DScene &s, const Ray &r, int ref, Best &best, Counters &cnt) {
    int enc = ~ref, kind = enc & 7, idx = enc >> RTB_KIND_BITS;
    RTB_ASSERT(ref < 0 && idx >= 0 && (kind == KIND_SPHERE ? idx < s.num_spheres : idx + (kind == KIND_TRI ? 0 : 1) < s.num_tris));
    float t, u = 0, v = 0;
    if (kind == KIND_SPHERE) {
        RTB_COUNT(cnt, sph);
        if (sphere_test(__ldg(s.spheres + idx), r, t)) consider(s, best, t, 0.f, 0.f, (idx << 1) | 1);
        return;
    }
    if (kind >= KIND_QUAD_ONEWAY) {
        // OneWayQuad::hit, objects.cu:273-280; normal_vec = t1.normal * (1 - 2*invert), :285-289
        float4 c = __ldg(s.tris + 3 * idx + 2);
        float m = (float)(1 - 2 * (kind == KIND_QUAD_ONEWAY_INV));
        if (dot(r.d, v3(c.y * m, c.z * m, c.w * m)) < 0) return;
    }
    // Quad::hit, objects.cu:223-236: triangle 1 if it hits, else whatever triangle 2 says
    int count = kind == KIND_TRI ? 1 : 2;
    for (int k = 0; k < count; k++) {
        RTB_COUNT(cnt, tri);
        if (tri_test(s.tris + 3 * (idx + k), r, t, u, v)) {
            if (kind == KIND_TRI && (r.d.x == 0 || r.d.y == 0 || r.d.z == 0)) {
                int leaf = __ldg(s.tri_refleaf + idx);
                if (leaf >= 0 && !ref_path_hit(s, leaf, r)) return;
            }
            consider(s, best, t, u, v, (idx + k) << 1);
            return;
        }
    }
}

// Slab test of one child box against [0, limit].  Not parity surface: it only has to be
// conservative, i.e. never reject a box that holds a primitive the exact tests above would accept
// closer than `best`.  NaNs (0 * inf) are dropped by fminf/fmaxf, which widens the interval; the
// final comparison carries a relative slack for the rounding of the six products.
// (Measured and rejected: fma(plane, inv, -(o * inv)) halves the slab arithmetic, but the rounding
// of o * inv is absolute, and covering it -- or switching flat axes off -- makes the rare rays that
// are nearly parallel to an axis walk most of a large tree: 1M-triangle scene 1.4 G -> 6 M rays/s.)
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

// box_test against [0, inf) that also hands out, per axis, the larger of the two slab distances (what the flat-wall
// skip of large_prims_best needs: the same products, not recomputed)
__device__ __forceinline__ bool box_test_far(float4 lo, float4 hi, const Ray &r, float &fx, float &fy, float &fz) {
    float t1 = (lo.x - r.o.x) * r.inv.x, t2 = (hi.x - r.o.x) * r.inv.x;
    float tmin = fminf(t1, t2), tmax = fx = fmaxf(t1, t2);
    t1 = (lo.y - r.o.y) * r.inv.y; t2 = (hi.y - r.o.y) * r.inv.y;
    fy = fmaxf(t1, t2);
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fy);
    t1 = (lo.z - r.o.z) * r.inv.z; t2 = (hi.z - r.o.z) * r.inv.z;
    fz = fmaxf(t1, t2);
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fz);
    tmin = fmaxf(tmin, 0.f);
    tmax = fminf(tmax, RTB_INF_F);
    return tmin <= tmax * 1.000004f;
}

#define RTB_DONE 0x7fffffff

// Resumable traversal state of one ray: lives in the lane across iterations of the megakernel's
// scheduling loop, so that lanes whose rays end early can go and shade / fetch new rays while the
// long rays of the warp keep their place in the tree (dynamic fetch, Aila & Laine 2009).
struct Trav {
    Best best;
    int cur;   // >= 0 internal node, < 0 leaf, RTB_DONE finished
    int leaf;  // postponed leaf (< 0) or 0
    int top;   // the top of the stack, kept in a register: a pop then has no load on its critical path
               // (the load that refills `top` is not needed before the NEXT pop); RTB_DONE = nothing left
    int sp;
    int stack[RTB_STACK];
};

__device__ __forceinline__ bool trav_done(const Trav &T) { return T.cur == RTB_DONE && T.leaf == 0; }

// The large primitives (walls, floors) are tested before the tree: the hit they give culls most of
// it.  `mask` = the lanes of the warp that call this together.  The box tests run converged (same
// address in every lane: broadcast loads) and only record which large primitives the ray may
// touch; the primitive tests then run in a warp-synchronous loop in which every lane pops ITS next
// candidate, so the long test is issued for many lanes at once instead of once per large primitive
// for the few lanes that touch that one.
__device__ __forceinline__ void large_prims_best(const DScene &s, const Ray &r, Best &best, unsigned mask, Counters &cnt) {
    best.t = RTB_INF_F; best.u = 0; best.v = 0; best.code = -1;
    unsigned cand = 0;
    for (int i = 0; i < s.num_large; i++) {
        RTB_COUNT(cnt, box);
        const float4 lo = s.large_box_c[2 * i], hi = s.large_box_c[2 * i + 1];
        float fx, fy, fz;
        bool c = box_test_far(lo, hi, r, fx, fy, fz);
        // A ray that STARTS on a wall meets that wall's box at t ~ 0, and Triangle::hit (objects.cu:135-163) then
        // runs twice only to fail `dist > 1e-6`.  For a primitive that is flat on axis k (every side vector has an
        // exactly zero k component, checked at commit) the products of Triangle::hit that multiply those zeros vanish
        // exactly, and what remains is dist = -(o_k - p0_k) / d_k up to the rounding of a 2-D cross product that is
        // formed the same way in numerator and denominator (relative error a few 1e-7 / sin(angle between the sides)).
        // o_k - p0_k is exact (Sterbenz) and lies inside the padded slab, so max((lo_k - o_k) / d_k, (hi_k - o_k) / d_k)
        // -- the far slab distance of axis k, which the box test has just formed -- bounds dist from above: below
        // 0.99e-6 the reference's test rejects both triangles, whatever u and v say.
        // NaN / inf (d_k == 0) compare false and take the exact tests.  flat_axis is uniform (kernel parameter).
        int flat_axis = __float_as_int(lo.w);
        if (flat_axis) {
            float tk = flat_axis == 1 ? fx : (flat_axis == 2 ? fy : fz);
            if (tk < 0.99e-6f) c = false;
        }
        if (c) cand |= 1u << i;
    }
    // Triangles and quads: one Triangle::hit per lane and step, whatever the lane's candidate is (a
    // single triangle, or triangle 1 / triangle 2 of a quad), so lanes with different candidate lists
    // stay converged and the loop runs max-over-lanes(total triangle tests) steps.
    unsigned part = cand & ~s.large_sphere_mask;
    int tri = 0, left = 0;   // next triangle of the candidate in progress, triangles it still has
    bool single = false;
    while (__any_sync(mask, (part != 0) | (left > 0))) {
        if (left == 0 && part) {
            int i = __ffs(part) - 1;
            part &= part - 1;
            int enc = ~__ldg(s.large_refs + i), kind = enc & 7;
            tri = enc >> RTB_KIND_BITS;
            single = kind == KIND_TRI;
            left = single ? 1 : 2;
            if (kind >= KIND_QUAD_ONEWAY) {  // OneWayQuad::hit, objects.cu:273-280 (see test_leaf)
                float4 c = __ldg(s.tris + 3 * tri + 2);
                float m = (float)(1 - 2 * (kind == KIND_QUAD_ONEWAY_INV));
                if (dot(r.d, v3(c.y * m, c.z * m, c.w * m)) < 0) left = 0;
            }
        }
        if (left > 0) {
            float t, u = 0, v = 0;
            RTB_COUNT(cnt, tri);
            bool hit = tri_test(s.tris + 3 * tri, r, t, u, v);
            if (hit && single && (r.d.x == 0 || r.d.y == 0 || r.d.z == 0)) {
                int leaf = __ldg(s.tri_refleaf + tri);
                if (leaf >= 0 && !ref_path_hit(s, leaf, r)) hit = false;
            }
            if (hit) consider(s, best, t, u, v, tri << 1);
            // Quad::hit, objects.cu:223-236: triangle 1 if it hits, else whatever triangle 2 says
            left = hit ? 0 : left - 1;
            tri++;
        }
    }
    part = cand & s.large_sphere_mask;
    while (__any_sync(mask, part != 0)) {
        if (part) {
            int i = __ffs(part) - 1;
            part &= part - 1;
            test_leaf(s, r, __ldg(s.large_refs + i), best, cnt);
        }
    }
}

// A ray with a NaN in its origin or direction hits nothing in the reference: Triangle::hit gets a NaN determinant and
// fails `u >= 0` (objects.cu:141-152), Sphere::hit a NaN discriminant and fails `disc >= 0` (:52-55); quads and cuboids
// are triangles.  Such rays exist: Box-Muller with u2 == 0 gives rho = inf (utils.cu:236-238, one draw in 2^32), the
// normalisation turns it into NaN, and the path keeps that NaN for its remaining bounces.  The conservative box test
// drops NaNs, i.e. ACCEPTS every box, so without this check one such ray walks the whole tree: a 1 M-triangle scene
// showed one 1.2 s frame in forty (measured: tools/c4_probe.py).  The tree is skipped; the result (a miss) is the same.
__device__ __forceinline__ bool ray_has_nan(const Ray &r) {
    // x * 0 is NaN for a NaN or infinite x (an infinite coordinate misses everything as well: the barycentrics or the
    // discriminant become inf - inf), 0 otherwise
    float s = r.o.x * 0.f + r.o.y * 0.f + r.o.z * 0.f + r.d.x * 0.f + r.d.y * 0.f + r.d.z * 0.f;
    return s != s;
}

// The root's node step, taken with the node in the kernel parameters (uniform loads, no L1
// traffic) by the kernel that creates the ray, where the whole warp is converged.  Returns the
// child to start at (`first`; RTB_DONE = the ray cannot touch the tree closer than `best`) and the
// child to visit afterwards (`second`, 0 = none).
__device__ __forceinline__ void root_step(const DScene &s, const Ray &r, const Best &best, int &first, int &second, Counters &cnt) {
    first = RTB_DONE; second = 0;
    if (s.num_bvh_prims <= 0 || ray_has_nan(r)) return;
    if (s.root_ref < 0) { first = s.root_ref; return; }  // a one-primitive tree: the root is a leaf
    float limit = best.t * 1.0001f, n0, n1;
    RTB_COUNT(cnt, box); RTB_COUNT(cnt, box);
    bool h0 = box_test(s.root_node[0], s.root_node[1], r, limit, n0);
    bool h1 = box_test(s.root_node[2], s.root_node[3], r, limit, n1);
    int r0 = __float_as_int(s.root_node[0].w), r1 = __float_as_int(s.root_node[2].w);
    bool first0 = h0 && (!h1 || n0 <= n1);
    if (h0 || h1) first = first0 ? r0 : r1;
    if (h0 && h1) second = first0 ? r1 : r0;
}

// enter the tree below the root with the outcome of root_step
__device__ __forceinline__ void trav_start_below_root(Trav &T, int first, int second) {
    T.stack[0] = RTB_DONE;  // sentinel: popping an empty stack yields "finished"
    T.sp = 1;
    T.top = second != 0 ? second : RTB_DONE;
    T.leaf = 0;
    T.cur = first;
    if (T.cur < 0) {  // a leaf: postpone it, go on with the other child
        T.leaf = T.cur;
        T.cur = T.top;
        T.top = RTB_DONE;
    }
}

// point the cursor at the root (T.best already holds the best large-primitive hit)
__device__ __forceinline__ void trav_start(const DScene &s, Trav &T) {
    T.stack[0] = RTB_DONE;  // sentinel: popping an empty stack yields "finished"
    T.sp = 1;
    T.top = RTB_DONE;
    T.leaf = 0;
    T.cur = s.num_bvh_prims > 0 ? s.root_ref : RTB_DONE;
    if (T.cur < 0) {  // a one-primitive tree: the root is a leaf
        T.leaf = T.cur;
        T.cur = RTB_DONE;
    }
}

__device__ __forceinline__ void trav_begin(const DScene &s, const Ray &r, Trav &T, unsigned mask, Counters &cnt) {
    large_prims_best(s, r, T.best, mask, cnt);
    trav_start(s, T);
    if (ray_has_nan(r)) { T.cur = RTB_DONE; T.leaf = 0; }  // hits nothing (see ray_has_nan): do not walk the tree
}

// Advance the rays of the warp through the LBVH: warp-synchronous "while-while" traversal with
// postponed leaves.  All lanes of `mask` move through the two phases together (the ballots are the
// reconvergence points, so independent thread scheduling cannot let early lanes run ahead into a
// leaf test on their own):
//   NODE phase  lanes whose cursor is an internal node take one step per iteration (one 64-byte
//               fetch, two slab tests, branch-free cursor update); a lane that reaches a leaf
//               postpones it and keeps walking, and stalls only when it reaches a second one;
//   LEAF phase  every lane that holds a leaf tests it (the long primitive test is issued once for
//               all of them).
// The NODE phase goes on while at least `walk_min` lanes are walking; below that it ends as soon
// as the other lanes have something better to do: a leaf to test, or -- when `quota` lanes have
// finished their ray -- new rays to fetch (the call returns; quota <= 0 = run to completion).
// One 64-byte node.  sm_100a has 256-bit global loads (LDG.E.256): two requests per node visit instead
// of four, which is what the L1 data pipe -- the limiter of the traversal kernel (ncu: 80 % busy,
// 4.3 wavefronts per divergent request) -- is paid in.
__device__ __forceinline__ void load_node(const float4 *N, float4 &lo0, float4 &hi0, float4 &lo1, float4 &hi1) {
#ifdef RTB_NODE_LDG128
    lo0 = __ldg(N); hi0 = __ldg(N + 1); lo1 = __ldg(N + 2); hi1 = __ldg(N + 3);
#else
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo0.x), "=f"(lo0.y), "=f"(lo0.z), "=f"(lo0.w), "=f"(hi0.x), "=f"(hi0.y), "=f"(hi0.z), "=f"(hi0.w) : "l"(N));
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo1.x), "=f"(lo1.y), "=f"(lo1.z), "=f"(lo1.w), "=f"(hi1.x), "=f"(hi1.y), "=f"(hi1.z), "=f"(hi1.w) : "l"(N + 2));
#endif
}

#ifdef RTB_DYNAMIC_MASK
#define RTB_TRAV_MASK mask
#else
#define RTB_TRAV_MASK 0xffffffffu  // all 32 lanes call trav_run together: idle lanes hold a finished state (cur = RTB_DONE, leaf = 0)
#endif
// quota: leave once this many lanes of the WARP hold a finished (or no) ray; <= 0 = run to completion.
// NODE_STEPS: node steps between two ballots of the NODE phase (measured: 2 is +1-2 % in the wavefront
// traversal kernel and, since the megakernel runs 7 blocks per SM, +2-3.5 % there as well: rtb_render.cuh).
// max_rounds > 0: come back after that many rounds whatever the lanes did (a caller that polls a queue).
template <int NODE_STEPS = 1, bool EARLY_POP = false>
__device__ __forceinline__ void trav_run(const DScene &s, const Ray &r, Trav &T, unsigned mask, int quota, int walk_min,
                                         Counters &cnt, int max_rounds = 0) {
    (void)mask;
    int cur = T.cur, leaf = T.leaf, sp = T.sp, top = T.top;
    Best best = T.best;
    // every round retires at least one node step or one leaf of some lane; the cap only turns a
    // logic error into a wrong pixel (caught by the parity tests) instead of a hung GPU
    const int round_cap = max_rounds > 0 ? max_rounds : (1 << 22);
    for (int round = 0; round < round_cap; round++) {
        unsigned finished;
        for (;;) {
            bool walk = (unsigned)cur < (unsigned)RTB_DONE;  // an internal node index
            unsigned walking = __ballot_sync(RTB_TRAV_MASK, walk);
            unsigned empty = __ballot_sync(RTB_TRAV_MASK, leaf == 0);
            // invariant: a leaf at the cursor implies a postponed leaf, so !walk && leaf == 0 means done
            finished = RTB_TRAV_MASK & ~walking & empty;
            if (!walking) break;
            if (__popc(walking) < walk_min && ((RTB_TRAV_MASK & ~empty) || (quota > 0 && __popc(finished) >= quota))) break;
            RTB_COUNT(cnt, node_iters); RTB_COUNT_N(cnt, node_lanes, __popc(walking)); RTB_COUNT_N(cnt, node_blocked, __popc(RTB_TRAV_MASK & ~walking & ~empty));
            RTB_COUNT_N(cnt, node_done, __popc(finished)); RTB_COUNT_N(cnt, mask_lanes, __popc(RTB_TRAV_MASK));
#pragma unroll
            for (int step = 0; step < NODE_STEPS; step++) {
                if ((unsigned)cur < (unsigned)RTB_DONE) {
                    RTB_ASSERT(cur < s.num_bvh_prims - 1);
                    const float4 *N = s.nodes + 4 * cur;
                    float4 lo0, hi0, lo1, hi1;
                    load_node(N, lo0, hi0, lo1, hi1);
                    // EARLY_POP: the two entries a pop (or two) of this step would bring up are fetched WITH the node, so
                    // that their latency hides behind the node's and the cursor update below is register-only.  (Fetched
                    // inside the pops, the loads sit on the critical path of the step: even predicated off, the second pop
                    // waits for the first one's load -- 11 % of wf_traverse's stall samples.)  Measured
                    // (profiles/r2_pop_ab.jsonl): the megakernel on the 1 M-triangle scene -5.5 % (nodes miss L1, the
                    // step is long), every L1-resident scene and the wavefront traversal kernel +1 ... +4 % (two more
                    // unconditional local loads on an L1 pipe that is 75 % busy): a variant of the megakernel only, taken
                    // by the timed kernel choice where it wins.  stack[0] is the sentinel RTB_DONE and sp >= 1.
                    int below1 = 0, below2 = 0;
                    if (EARLY_POP) { below1 = T.stack[sp - 1]; below2 = T.stack[max(sp - 2, 0)]; }
                    // relative slack on the distance cull: a primitive's own t carries more rounding
                    // than the box planes do (cancellation in Moller-Trumbore), so keep boxes that
                    // start a hair beyond the current best
                    float limit = best.t * 1.0001f;
                    float n0, n1;
                    RTB_COUNT(cnt, box); RTB_COUNT(cnt, box);
                    bool h0 = box_test(lo0, hi0, r, limit, n0);
                    bool h1 = box_test(lo1, hi1, r, limit, n1);
                    int r0 = __float_as_int(lo0.w), r1 = __float_as_int(lo1.w);
                    // branch-free cursor update (selects, a predicated push, predicated pops): the
                    // four outcomes would otherwise run as four divergent paths
                    bool first0 = h0 && (!h1 || n0 <= n1);
                    int near = first0 ? r0 : r1, far = first0 ? r1 : r0;
                    if (!EARLY_POP) {
                    // the stack cannot overflow: rtb_scene_commit refuses a tree higher than RTB_STACK - 2
                    RTB_ASSERT(!(h0 && h1) || sp < RTB_STACK);
                    if (h0 && h1) { T.stack[sp++] = top; top = far; }
                    cur = near;
                    // a pop takes the register-held top; stack[0] holds RTB_DONE as a sentinel (sp >= 1), so
                    // the refill of `top` is an unconditional load
                    if (!(h0 || h1)) { cur = top; sp = max(sp - 1, 0); top = T.stack[sp]; sp = max(sp, 1); }
                    if (cur < 0 && leaf == 0) {  // first leaf: postpone it and keep walking
                        leaf = cur;
                        cur = top; sp = max(sp - 1, 0); top = T.stack[sp]; sp = max(sp, 1);
                    }
                    } else {
                    const bool both = h0 && h1, any = h0 || h1;
                    const bool take = any && near < 0 && leaf == 0;   // the first leaf: postpone it and keep walking
                    if (take) leaf = near;
                    const bool go_near = any && !take;   // the near child, not postponed
                    const bool go_far = both && take;    // near was postponed: a push and a pop that cancel
                    const bool push = both && !take;     // go near, keep far for later
                    const bool pop1 = !go_near && !go_far;   // nothing to go to here: the register-held top
                    // the stack cannot overflow: rtb_scene_commit refuses a tree higher than RTB_STACK - 2
                    RTB_ASSERT(!push || sp < RTB_STACK);
                    if (push) T.stack[sp] = top;
                    int ncur = go_near ? near : (go_far ? far : top);
                    int ntop = push ? far : (pop1 ? below1 : top);
                    sp = push ? sp + 1 : (pop1 ? max(sp - 1, 1) : sp);
                    const bool pop2 = pop1 && ncur < 0 && leaf == 0;   // the popped entry is a first leaf: postpone it, pop again
                    if (pop2) leaf = ncur;
                    cur = pop2 ? ntop : ncur;
                    top = pop2 ? below2 : ntop;
                    sp = pop2 ? max(sp - 1, 1) : sp;
                    }
                }
            }
        }
        RTB_COUNT(cnt, leaf_iters); RTB_COUNT_N(cnt, leaf_lanes, __popc(__ballot_sync(RTB_TRAV_MASK, leaf < 0)));
        if (leaf < 0) {
            test_leaf(s, r, leaf, best, cnt);
            leaf = 0;
            if (cur < 0) {  // a second leaf was reached while the first was pending
                leaf = cur;
                cur = top; sp = max(sp - 1, 0); top = T.stack[sp]; sp = max(sp, 1);
            }
        }
        unsigned alive = __ballot_sync(RTB_TRAV_MASK, cur != RTB_DONE || leaf < 0);
        if (!alive) break;
        if (quota > 0 && __popc(RTB_TRAV_MASK & ~alive) >= quota) break;
    }
    T.cur = cur; T.leaf = leaf; T.sp = sp; T.top = top; T.best = best;
}

// closest hit over the whole scene through the LBVH, in one go; `mask` = the calling lanes
__device__ __forceinline__ Best trace_bvh(const DScene &s, const Ray &r, unsigned mask, Counters &cnt) {
    Trav T;
    trav_begin(s, r, T, mask, cnt);
    trav_run(s, r, T, mask, 0, 16, cnt);
    return T.best;
}

// brute force over every primitive in creation order (option "traversal" = 1): the LBVH's checker
__device__ __forceinline__ Best trace_brute(const DScene &s, const Ray &r, Counters &cnt) {
    Best best;
    best.t = RTB_INF_F; best.u = 0; best.v = 0; best.code = -1;
    for (int i = 0; i < s.num_prims; i++) test_leaf(s, r, __ldg(s.all_refs + i), best, cnt);
    return best;
}

}  // namespace rtb

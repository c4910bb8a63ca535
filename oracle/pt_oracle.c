/* TEST INFRASTRUCTURE (oracle). Not part of the product path; the product never links this.
 *
 * pt_oracle.c -- plain-C, single-precision CPU restatement of the reference's per-pixel
 * path-tracing function.  Each function cites the reference file:line it follows (paths are
 * relative to /root/reference).  Closest-hit inside a MESH is a brute-force loop over the mesh
 * triangles (lowest index wins an exact tie); the reference's own host-built BVH
 * (src/objects.cu:448-771) is NOT restated -- its tree shape is not parity surface, only the
 * closest hit is.
 *
 * Parity status: PINNED against oracle/_ref/libref_host.so (the reference's own source compiled
 * for the host) by tests/test_oracle_ref.py, and against tests/golden/ generated from it.
 * Compile with -ffp-contract=off: every + - * / sqrt below must round exactly as written.
 */
#define _GNU_SOURCE
#include "pt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define PTO_INF 1073741824 /* `1 << 31 - 1` == 1 << 30, src/objects.cu:6 */
#define PTO_EPS 0.000001f  /* FLOAT_PRECISION_ERROR, src/objects.cu:7 */
static const float PTO_PI = 3.141592653589793; /* src/camera.cu:9 (float) */
static const float PTO_AA_RANGE = 0.001;       /* ANTIALIAS_OFFSET_RANGE, src/ray.cu:4 (float) */

typedef struct { float x, y, z; } V3;

/* ---- Vec3, src/utils.cu:13-163 ---------------------------------------------------------- */
static inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
static inline V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 vmuls(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline V3 vmul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 vdivs(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
/* utils.cu:132-138: products first, then left-to-right sum */
static inline float vdot(V3 a, V3 b) { float nx = a.x * b.x, ny = a.y * b.y, nz = a.z * b.z; return nx + ny + nz; }
/* utils.cu:148-155 */
static inline V3 vcross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
/* utils.cu:120-123 */
static inline float vmag(V3 a) { float m = a.x * a.x + a.y * a.y + a.z * a.z; return sqrtf(m); }
/* utils.cu:125-130: v * (1 / |v|) */
static inline V3 vnorm(V3 a) { float inv = 1 / vmag(a); return v3(a.x * inv, a.y * inv, a.z * inv); }
static inline V3 from_rtb(RtbVec3 a) { return v3(a.x, a.y, a.z); }

/* ---- RNG, src/utils.cu:220-239 ----------------------------------------------------------- */
unsigned int pto_seed(int array_index, int time_ms) {
    /* raytracer.cu:127: int products that wrap, converted to uint */
    return (unsigned int)array_index * 3145739u + (unsigned int)time_ms * 6291469u;
}

float pto_rng(unsigned int *state) {
    unsigned int ns = *state * 747796405u + 2891336453u;
    *state = ns;
    unsigned int r = ((ns >> ((ns >> 28) + 4)) ^ ns) * 277803737u;
    r = (r >> 22) ^ r;
    return (float)((double)r / 4294967295.0); /* utils.cu:228: uint / double, rounded to float */
}

float pto_normal(unsigned int *state) {
    float theta = (float)(2 * 3.14159 * (double)pto_rng(state)); /* utils.cu:236: double product */
    float rho = sqrtf(-2 * logf(pto_rng(state)));                /* utils.cu:237: float log/sqrt */
    return rho * cosf(theta);
}

/* ---- flattened scene ----------------------------------------------------------------------- */
typedef struct {
    V3 p0, side1, side2, normal; /* Triangle::precompute, objects.cu:175-186 */
    float uv[6];                 /* texture_points[3] */
    V3 p1, p2;                   /* kept for the reference-BVH boxes below */
} Tri;

typedef struct {
    int type, material;
    V3 center; float radius;     /* SPHERE */
    int tri_first, tri_count;    /* into PtoScene.tris: TRIANGLE 1, QUAD/ONE_WAY 2, CUBOID 12, MESH n */
    V3 one_way_normal;
    struct RefBvhTag *ref_bvh;   /* MESH: the reference's boxes (see ref_bvh_build) */
} Obj;

struct PtoScene {
    int num_objects; Obj *objects;
    int num_materials; RtbMaterial *materials;
    int num_tris; Tri *tris;
    int num_images; RtbImage *images; /* deep copies */
};

static Tri make_tri(V3 a, V3 b, V3 c, const float *uv) {
    Tri t;
    t.p0 = a;
    t.side1 = vsub(b, a);
    t.side2 = vsub(c, a);
    t.normal = vnorm(vcross(t.side1, t.side2));
    for (int i = 0; i < 6; i++) t.uv[i] = uv ? uv[i] : 0.f;
    t.p1 = b; t.p2 = c;
    return t;
}

/* Quad::create_triangles, objects.cu:244-253 */
static void make_quad(Tri *out, V3 p1, V3 p2, V3 p3, V3 p4) {
    static const float uv1[6] = {0, 0, 1, 0, 1, 1};
    static const float uv2[6] = {0, 0, 0, 1, 1, 1};
    out[0] = make_tri(p1, p2, p3, uv1);
    out[1] = make_tri(p1, p4, p3, uv2);
}

/* Cuboid::create_faces, objects.cu:327-349 */
static void make_cuboid(Tri *out, V3 tl_near, float width, float height, float depth) {
    V3 w = v3(width, 0, 0), h = v3(0, height, 0), d = v3(0, 0, depth);
    V3 tr_near = vadd(tl_near, w), br_near = vsub(tr_near, h), bl_near = vsub(tl_near, h);
    V3 tl_far = vadd(tl_near, d), tr_far = vadd(tl_far, w), br_far = vsub(tr_far, h), bl_far = vsub(tl_far, h);
    make_quad(out + 0, tl_near, tr_near, br_near, bl_near);
    make_quad(out + 2, tl_far, tr_far, br_far, bl_far);
    make_quad(out + 4, tl_near, bl_near, bl_far, tl_far);
    make_quad(out + 6, tr_near, br_near, br_far, tr_far);
    make_quad(out + 8, bl_near, br_near, br_far, bl_far);
    make_quad(out + 10, tl_near, tr_near, tr_far, tl_far);
}

/* ---- the reference's BVH boxes (culling is parity surface in degenerate cases) -----------------
 * The reference reaches a mesh triangle only through the boxes on the path root -> leaf of its
 * host-built BVH, each tested with BoundingBox::ray_hits (objects.cu:404-434): `tmin < tmax &&
 * tmax > 0`, strict, with CUDA's NaN-dropping min/max.  For a generic ray a triangle hit implies
 * that all those boxes are hit, so the tree is invisible.  Two degenerate cases are not:
 *   (1) a box with zero extent on an axis (axis-aligned faces, e.g. the top/bottom of
 *       models/cube.obj) gives t1 == t2, so tmin >= tmax and EVERY ray misses it;
 *   (2) a zero direction component with the origin exactly on a box plane gives 0 * inf = NaN,
 *       which min/max drop, and the box can be missed (row H/2 of an antialias-off render).
 * To be exact the oracle rebuilds the reference's tree -- BVH::build :602-624, split_triangles
 * :626-653 (first n/2+1 by distance of points[0] to the reference point go left), sort_triangles
 * :655-706 (top-down merge sort taking the RIGHT element on ties), get_ref_point :708-719,
 * BoundingBox::grow :369-393, fixed depth 10 (:786) -- and accepts a triangle hit only if every
 * box on its path passes the reference's slab test.  (The traversal's distance culls, :501,:517,
 * cannot reject a hit that is closer than the best so far and are not restated.)
 */
typedef struct { int idx; float dist; } RefPair;
typedef struct { V3 lo, hi; int parent; } RefNode;
typedef struct RefBvhTag { RefNode *nodes; int num_nodes, cap; int *leaf_of; /* per mesh triangle */ } RefBvh;

static void ref_merge_sort(RefPair *a, int n, RefPair *tmp) {
    if (n <= 1) return;
    int mid = n / 2, nr = n - mid;
    ref_merge_sort(a, mid, tmp);
    ref_merge_sort(a + mid, nr, tmp);
    int l = 0, r = 0;
    for (int k = 0; k < n; k++) {
        int add_left;
        if (l >= mid) add_left = 0;
        else if (r >= nr) add_left = 1;
        else add_left = a[l].dist < a[mid + r].dist;
        tmp[k] = add_left ? a[l++] : a[mid + r++];
    }
    memcpy(a, tmp, sizeof(RefPair) * n);
}

/* returns the node index, -1 for an empty subtree (the reference creates empty nodes whose boxes
 * are (0,0,0)-(0,0,0): degenerate, never hit, and they hold no triangles) */
static int ref_bvh_build(RefBvh *bvh, const Tri *tris, RefPair *items, int n, int depth, RefPair *tmp) {
    if (n == 0) return -1;
    V3 lo = tris[items[0].idx].p0, hi = lo;
    for (int i = 0; i < n; i++) {
        const Tri *t = &tris[items[i].idx];
        V3 pts[3] = {t->p0, t->p1, t->p2};
        for (int k = 0; k < 3; k++) {
            lo.x = fminf(lo.x, pts[k].x); lo.y = fminf(lo.y, pts[k].y); lo.z = fminf(lo.z, pts[k].z);
            hi.x = fmaxf(hi.x, pts[k].x); hi.y = fmaxf(hi.y, pts[k].y); hi.z = fmaxf(hi.z, pts[k].z);
        }
    }
    if (bvh->num_nodes == bvh->cap) {
        bvh->cap = bvh->cap ? bvh->cap * 2 : 256;
        bvh->nodes = (RefNode *)realloc(bvh->nodes, sizeof(RefNode) * bvh->cap);
    }
    int me = bvh->num_nodes++;
    bvh->nodes[me].lo = lo; bvh->nodes[me].hi = hi; bvh->nodes[me].parent = -1;
    if (depth <= 0) {
        for (int i = 0; i < n; i++) bvh->leaf_of[items[i].idx] = me;
        return me;
    }
    float w = hi.x - lo.x, h = hi.y - lo.y, d = hi.z - lo.z; /* update_dimensions :395-402 */
    V3 ref;
    if (w >= h && w >= d) ref = v3(lo.x + w / 2, lo.y, lo.z + d / 2);
    else if (h >= w && h >= d) ref = v3(lo.x, lo.y + h / 2, lo.z + d / 2);
    else ref = v3(lo.x + w / 2, lo.y + h / 2, lo.z);
    for (int i = 0; i < n; i++) items[i].dist = vmag(vsub(tris[items[i].idx].p0, ref));
    ref_merge_sort(items, n, tmp);
    int nleft = n / 2 + 1; /* i <= mid */
    if (nleft > n) nleft = n;
    int l = ref_bvh_build(bvh, tris, items, nleft, depth - 1, tmp);
    int r = ref_bvh_build(bvh, tris, items + nleft, n - nleft, depth - 1, tmp);
    if (l >= 0) bvh->nodes[l].parent = me;
    if (r >= 0) bvh->nodes[r].parent = me;
    return me;
}

static RefBvh *ref_bvh_create(const Tri *tris, int n) {
    RefBvh *bvh = (RefBvh *)calloc(1, sizeof *bvh);
    bvh->leaf_of = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    RefPair *items = (RefPair *)malloc(sizeof(RefPair) * (size_t)(n > 0 ? n : 1)), *tmp = (RefPair *)malloc(sizeof(RefPair) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) { items[i].idx = i; items[i].dist = 0; bvh->leaf_of[i] = -1; }
    ref_bvh_build(bvh, tris, items, n, 10, tmp);
    free(items); free(tmp);
    return bvh;
}

PtoScene *pto_scene_create(const RtbSceneDesc *d) {
    PtoScene *s = (PtoScene *)calloc(1, sizeof *s);
    s->num_objects = d->num_objects;
    s->objects = (Obj *)calloc(d->num_objects ? d->num_objects : 1, sizeof(Obj));
    s->num_materials = d->num_materials;
    s->materials = (RtbMaterial *)malloc(sizeof(RtbMaterial) * (d->num_materials ? d->num_materials : 1));
    memcpy(s->materials, d->materials, sizeof(RtbMaterial) * d->num_materials);
    s->num_images = d->num_images;
    s->images = (RtbImage *)calloc(d->num_images ? d->num_images : 1, sizeof(RtbImage));
    for (int i = 0; i < d->num_images; i++) {
        size_t n = (size_t)d->images[i].width * d->images[i].height * 3;
        float *px = (float *)malloc(n * sizeof(float));
        memcpy(px, d->images[i].rgb, n * sizeof(float));
        s->images[i].width = d->images[i].width;
        s->images[i].height = d->images[i].height;
        s->images[i].rgb = px;
    }
    size_t nt = 0;
    for (int i = 0; i < d->num_objects; i++) {
        const RtbObject *o = &d->objects[i];
        nt += o->type == RTB_OBJ_TRIANGLE ? 1 : (o->type == RTB_OBJ_QUAD || o->type == RTB_OBJ_ONE_WAY_QUAD) ? 2
              : o->type == RTB_OBJ_CUBOID ? 12 : o->type == RTB_OBJ_MESH ? (size_t)o->tri_count : 0;
    }
    s->tris = (Tri *)malloc(sizeof(Tri) * (nt ? nt : 1));
    size_t k = 0;
    for (int i = 0; i < d->num_objects; i++) {
        const RtbObject *o = &d->objects[i];
        Obj *q = &s->objects[i];
        q->type = o->type;
        q->material = o->material;
        q->tri_first = (int)k;
        switch (o->type) {
            case RTB_OBJ_SPHERE: q->center = from_rtb(o->p[0]); q->radius = o->f[0]; q->tri_count = 0; break;
            case RTB_OBJ_TRIANGLE:
                s->tris[k] = make_tri(from_rtb(o->p[0]), from_rtb(o->p[1]), from_rtb(o->p[2]), o->uv);
                q->tri_count = 1; break;
            case RTB_OBJ_QUAD:
            case RTB_OBJ_ONE_WAY_QUAD:
                make_quad(&s->tris[k], from_rtb(o->p[0]), from_rtb(o->p[1]), from_rtb(o->p[2]), from_rtb(o->p[3]));
                q->tri_count = 2;
                /* OneWayQuad::get_normal_vec, objects.cu:285-289 */
                q->one_way_normal = vmuls(s->tris[k].normal, (float)(1 - 2 * (o->invert_normal ? 1 : 0)));
                break;
            case RTB_OBJ_CUBOID:
                make_cuboid(&s->tris[k], from_rtb(o->p[0]), o->f[0], o->f[1], o->f[2]);
                q->tri_count = 12; break;
            case RTB_OBJ_MESH:
                for (int t = 0; t < o->tri_count; t++) {
                    const float *p = d->mesh_tri_points + (size_t)(o->tri_first + t) * 9;
                    const float *uv = d->mesh_tri_uv ? d->mesh_tri_uv + (size_t)(o->tri_first + t) * 6 : NULL;
                    s->tris[k + t] = make_tri(v3(p[0], p[1], p[2]), v3(p[3], p[4], p[5]), v3(p[6], p[7], p[8]), uv);
                }
                q->tri_count = o->tri_count;
                q->ref_bvh = ref_bvh_create(&s->tris[k], o->tri_count);
                break;
            default: q->tri_count = 0;
        }
        k += q->tri_count;
    }
    s->num_tris = (int)k;
    return s;
}

void pto_scene_destroy(PtoScene *s) {
    if (!s) return;
    for (int i = 0; i < s->num_images; i++) free((void *)s->images[i].rgb);
    for (int i = 0; i < s->num_objects; i++)
        if (s->objects[i].ref_bvh) { free(s->objects[i].ref_bvh->nodes); free(s->objects[i].ref_bvh->leaf_of); free(s->objects[i].ref_bvh); }
    free(s->images); free(s->tris); free(s->materials); free(s->objects); free(s);
}

/* ---- ray + hit data, src/ray.cu:22-61 ------------------------------------------------------- */
typedef struct {
    V3 origin, dir, dir_inv;
    unsigned int *rng;
    int antialias;
    float cur_n; /* current_refractive_index, ray.cu:56,144 */
} Ray;

typedef struct {
    int hits;
    float dist;
    V3 point, normal;
    float u, v;
    int sub; /* winning triangle inside the object (harness id) */
} Hit;

static inline void change_direction(Ray *r, V3 d) { /* ray.cu:198-202 */
    r->dir = d;
    r->dir_inv = v3(1 / d.x, 1 / d.y, 1 / d.z);
}

/* cam_pixel_to_world (camera.cu:24-29) + Ray::set_direction_origin (ray.cu:147-155) */
static void ray_init(Ray *r, const RtbCamData *cam, int x, int y, unsigned int *rng, int aa) {
    V3 du = from_rtb(cam->delta_u), dv = from_rtb(cam->delta_v);
    V3 plane_point = vadd(vmuls(du, (float)x), vmuls(dv, (float)y));
    V3 view_pos = vadd(from_rtb(cam->tl_pixel_pos), plane_point);
    V3 o = from_rtb(cam->cam_pos);
    r->origin = o;
    r->rng = rng;
    r->antialias = aa;
    r->cur_n = 1;
    change_direction(r, vnorm(vsub(view_pos, o)));
}

/* Ray::apply_antialias, ray.cu:130-142 -- (u - 0.5) * 2 * RANGE in double, RANGE a float const */
static void apply_antialias(Ray *r) {
    if (!r->antialias) return;
    V3 off;
    off.x = (float)(((double)pto_rng(r->rng) - 0.5) * 2 * (double)PTO_AA_RANGE);
    off.y = (float)(((double)pto_rng(r->rng) - 0.5) * 2 * (double)PTO_AA_RANGE);
    off.z = (float)(((double)pto_rng(r->rng) - 0.5) * 2 * (double)PTO_AA_RANGE);
    change_direction(r, vnorm(vadd(r->dir, off)));
}

/* Triangle::hit, objects.cu:135-163 */
static Hit tri_hit(const Tri *t, const Ray *r, int need_uv) {
    V3 p_vec = vcross(r->dir, t->side2);
    float det = vdot(t->side1, p_vec);
    float inv_det = 1 / det;
    V3 t_vec = vsub(r->origin, t->p0);
    float u = vdot(t_vec, p_vec) * inv_det;
    V3 q_vec = vcross(t_vec, t->side1);
    float v = vdot(r->dir, q_vec) * inv_det;
    float w = 1 - u - v;
    float dist = vdot(t->side2, q_vec) * inv_det;
    int hits = dist > PTO_EPS && u >= 0 && v >= 0 && w >= 0;
    Hit h;
    h.hits = hits;
    h.dist = hits ? dist : (float)PTO_INF;
    h.point = vadd(vmuls(r->dir, dist), r->origin);
    h.normal = vmuls(t->normal, (float)(1 - 2 * (vdot(t->normal, r->dir) > 0)));
    h.u = h.v = 0;
    h.sub = 0;
    if (need_uv) { /* assign_texture_coords(hit, w, u, v): tp0*w + tp1*u + tp2*v, objects.cu:160,196-199 */
        h.u = (t->uv[0] * w + t->uv[2] * u) + t->uv[4] * v;
        h.v = (t->uv[1] * w + t->uv[3] * u) + t->uv[5] * v;
    }
    return h;
}

/* Quad::hit, objects.cu:223-236: t1 if it hits, else t2 (whatever t2 says) */
static Hit quad_hit(const Tri *t, const Ray *r, int need_uv) {
    Hit a = tri_hit(&t[0], r, need_uv);
    Hit b = tri_hit(&t[1], r, need_uv);
    if (a.hits) { a.sub = 0; return a; }
    b.sub = 1;
    return b;
}

/* Sphere::hit + assign_texture_coords, objects.cu:40-97 */
static Hit sphere_hit(const Obj *o, const Ray *r, int need_uv) {
    Hit h;
    h.hits = 0; h.dist = (float)PTO_INF; h.u = h.v = 0; h.sub = 0;
    h.point = v3(0, 0, 0); h.normal = v3(0, 0, 0);
    V3 c_min_q = vsub(o->center, r->origin);
    float a = vdot(r->dir, r->dir);
    float b = vdot(r->dir, c_min_q) * (-2);
    float c = vdot(c_min_q, c_min_q) - o->radius * o->radius;
    float disc = b * b - 4 * a * c;
    if (disc >= 0) {
        float d = (-b - sqrtf(disc)) / (2 * a);
        if (d > PTO_EPS) {
            V3 p = vadd(vmuls(r->dir, d), r->origin);
            h.hits = 1; h.dist = d; h.point = p;
            h.normal = vnorm(vsub(p, o->center));
            if (need_uv) {
                float theta = asinf((p.y - o->center.y) / o->radius);
                float phi = acosf((p.x - o->center.x) / o->radius);
                float u = (theta + PTO_PI / 2) / PTO_PI;
                float v_ratio = (1 - phi / PTO_PI) / 2;
                int behind = p.z > o->center.z;
                int mult = 1 - 2 * behind;
                h.u = u;
                h.v = 1 * behind + mult * v_ratio;
            }
        }
    }
    return h;
}

/* BoundingBox::ray_hits, objects.cu:404-434 (fminf/fmaxf == CUDA's NaN-dropping min/max) */
static int ref_box_hit(const RefNode *b, const Ray *r) {
    float tmin = 0, tmax = (float)PTO_INF;
    float t1 = (b->lo.x - r->origin.x) * r->dir_inv.x, t2 = (b->hi.x - r->origin.x) * r->dir_inv.x;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = (b->lo.y - r->origin.y) * r->dir_inv.y; t2 = (b->hi.y - r->origin.y) * r->dir_inv.y;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = (b->lo.z - r->origin.z) * r->dir_inv.z; t2 = (b->hi.z - r->origin.z) * r->dir_inv.z;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    return tmin < tmax && tmax > 0;
}

static int ref_path_hit(const struct RefBvhTag *bvh, int tri, const Ray *r) {
    for (int n = bvh->leaf_of[tri]; n >= 0; n = bvh->nodes[n].parent)
        if (!ref_box_hit(&bvh->nodes[n], r)) return 0;
    return 1;
}

/* Object::hit dispatch, objects.cu:827-842 */
static Hit object_hit(const PtoScene *s, const Obj *o, const Ray *r) {
    int need_uv = s->materials[o->material].need_uv;
    const Tri *t = s->tris + o->tri_first;
    Hit miss;
    miss.hits = 0; miss.dist = (float)PTO_INF; miss.u = miss.v = 0; miss.sub = 0;
    miss.point = v3(0, 0, 0); miss.normal = v3(0, 0, 0);
    switch (o->type) {
        case RTB_OBJ_SPHERE: return sphere_hit(o, r, need_uv);
        case RTB_OBJ_TRIANGLE: return tri_hit(t, r, need_uv);
        case RTB_OBJ_QUAD: return quad_hit(t, r, need_uv);
        case RTB_OBJ_ONE_WAY_QUAD: /* objects.cu:273-280 */
            if (vdot(r->dir, o->one_way_normal) < 0) return miss;
            return quad_hit(t, r, need_uv);
        case RTB_OBJ_CUBOID: { /* objects.cu:305-322: strict <, first face wins ties */
            Hit best = miss;
            for (int f = 0; f < 6; f++) {
                Hit fh = quad_hit(t + 2 * f, r, need_uv);
                if (fh.hits && fh.dist < best.dist) { best = fh; best.sub = f * 2 + fh.sub; }
            }
            return best;
        }
        case RTB_OBJ_MESH: { /* closest triangle, strict < (check_leaf_node objects.cu:596, traverse :509) */
            Hit best = miss;
            for (int k = 0; k < o->tri_count; k++) {
                Hit th = tri_hit(t + k, r, need_uv);
                if (th.hits && th.dist < best.dist && ref_path_hit(o->ref_bvh, k, r)) { best = th; best.sub = k; }
            }
            return best;
        }
    }
    return miss;
}

/* get_ray_collision, raytracer.cu:24-46: `<=` -> the later object wins an exact tie; the
 * precision_error chained comparison (:37) is always false for real hits and is omitted */
static Hit closest_hit(const PtoScene *s, const Ray *r, int *obj_out, RtbCounters *c) {
    Hit best;
    best.hits = 0; best.dist = (float)PTO_INF; best.u = best.v = 0; best.sub = 0;
    best.point = v3(0, 0, 0); best.normal = v3(0, 0, 0);
    int bi = -1;
    for (int i = 0; i < s->num_objects; i++) {
        Hit h = object_hit(s, &s->objects[i], r);
        if (!h.hits) continue;
        if (h.dist <= best.dist) { best = h; bi = i; }
    }
    if (c) { c->segments++; if (bi >= 0) c->hit_segments++; }
    *obj_out = bi;
    return best;
}

/* ---- scatter, src/ray.cu:67-128,157-196 ------------------------------------------------------ */
static void ray_reflect(Ray *r, const Hit *h, float smoothness) {
    /* diffuse_reflect, ray.cu:157-170: three Box-Muller normals (6 draws) */
    float dx = pto_normal(r->rng);
    float dy = pto_normal(r->rng);
    float dz = pto_normal(r->rng);
    V3 rv = v3(dx, dy, dz);
    if (vdot(rv, h->normal) < 0) rv = vmuls(rv, -1);
    rv = vnorm(rv);
    V3 diffuse = vnorm(vadd(h->normal, rv)); /* true_lambertian_reflect, ray.cu:172-178 */
    /* perfect_reflect, ray.cu:180-186: d - (n*2)*dot */
    float dot = vdot(r->dir, h->normal);
    V3 spec = vnorm(vsub(r->dir, vmuls(vmuls(h->normal, 2), dot)));
    r->origin = h->point;
    /* lerp(a, b, t) = a + (b - a) * t, ray.cu:32-34 */
    change_direction(r, vnorm(vadd(diffuse, vmuls(vsub(spec, diffuse), smoothness))));
}

static void ray_refract(Ray *r, const Hit *h, const RtbMaterial *m) {
    float n1, n2;
    V3 refn;
    if (vdot(h->normal, r->dir) > 0) { /* leaving, ray.cu:84-89 */
        n1 = m->refractive_index; n2 = r->cur_n; refn = h->normal;
    } else {
        n1 = r->cur_n; n2 = m->refractive_index; refn = vmuls(h->normal, -1);
    }
    r->cur_n = n2; /* ray.cu:98: before the reflect decision */
    /* ray.cu:101-102: min(float, 1.0) is the double overload -> acos/asin in FP64 */
    float theta1 = (float)acos(fmin((double)vdot(r->dir, refn), 1.0));
    float theta2 = (float)asin(fmin((double)(n1 * sinf(theta1) / n2), 1.0));
    float critical = asinf(n2 / n1);
    /* get_reflection_coeff, ray.cu:188-196 (pow(float, int) -> float) */
    float sqrt_r0 = (n1 - n2) / (n1 + n2);
    float r0 = sqrt_r0 * sqrt_r0;
    float cos_theta = cosf(theta1);
    float coeff = r0 + (1 - r0) * powf(1 - cos_theta, 5.0f);
    /* ray.cu:108: short-circuit -- the draw happens only when theta1 <= critical (or NaN) */
    if (theta1 > critical || coeff > pto_rng(r->rng)) {
        ray_reflect(r, h, m->smoothness);
        return;
    }
    V3 perp;
    if (theta1 != 0) perp = vdivs(vsub(r->dir, vmuls(refn, cosf(theta1))), sinf(theta1));
    else perp = v3(0, 0, 0);
    V3 res = vadd(vmuls(refn, cosf(theta2)), vmuls(perp, sinf(theta2)));
    r->origin = h->point;
    change_direction(r, vnorm(res));
}

/* Texture::get_texture_colour, material.cu:53-124 */
static V3 texture_colour(const PtoScene *s, const RtbTexture *t, float u, float v) {
    switch (t->type) {
        case RTB_TEX_COLOUR: return from_rtb(t->colour);
        case RTB_TEX_GRADIENT: return v3(u, v, 0);
        case RTB_TEX_CHECKERBOARD: {
            int uc = (int)(u * t->num_squares);
            int vc = (int)(v * t->num_squares);
            return ((uc + vc) % 2 == 0) ? from_rtb(t->light) : from_rtb(t->dark);
        }
        case RTB_TEX_IMAGE: {
            int uc = (int)((t->img_width - 1) * u);
            int vc = (int)((t->img_height - 1) * v);
            long idx = (long)vc * t->img_width + uc;
            long n = (long)t->img_width * t->img_height;
            if (idx < 0) idx = 0;           /* the reference reads out of bounds here; clamp */
            if (idx >= n) idx = n - 1;
            const float *px = s->images[t->image].rgb + idx * 3;
            return v3(px[0], px[1], px[2]);
        }
    }
    return v3(0, 0, 0);
}

/* trace_ray, raytracer.cu:67-94 (ray by value) */
static V3 trace_ray(const PtoScene *s, Ray ray, const RtbRenderData *rd, RtbCounters *c) {
    V3 final = v3(0, 0, 0), cur = v3(1, 1, 1);
    V3 sky = from_rtb(rd->sky_colour);
    if (c) c->samples++;
    for (int b = 0; b < rd->reflection_limit; b++) {
        apply_antialias(&ray);
        int oi;
        Hit h = closest_hit(s, &ray, &oi, c);
        if (!h.hits) {
            final = vadd(final, vmul(sky, cur));
            break;
        }
        const RtbMaterial *m = &s->materials[s->objects[oi].material];
        switch (m->type) { /* update_ray, raytracer.cu:49-64 */
            case RTB_MAT_STANDARD:
            case RTB_MAT_EMISSIVE: ray_reflect(&ray, &h, m->smoothness); break;
            case RTB_MAT_REFRACTIVE: ray_refract(&ray, &h, m); break;
        }
        if (m->type == RTB_MAT_EMISSIVE) final = vadd(final, vmul(from_rtb(m->emitted_light), cur));
        else cur = vmul(cur, texture_colour(s, &m->texture, h.u, h.v));
    }
    return final;
}

/* get_pixel_colour + get_ray_colour, raytracer.cu:97-136 */
static void render_pixel(const PtoScene *s, const RtbCamData *cam, int width, const RtbRenderData *rd, int time_ms,
                         int frame_num, const float *prev, float *out, int x, int y, RtbCounters *c) {
    int array_index = (y * width + x) * 3;
    V3 previous = prev ? v3(prev[array_index], prev[array_index + 1], prev[array_index + 2]) : v3(0, 0, 0);
    unsigned int rng = pto_seed(array_index, time_ms);
    Ray ray;
    ray_init(&ray, cam, x, y, &rng, rd->antialias);
    V3 colour = v3(0, 0, 0);
    for (int i = 0; i < rd->rays_per_pixel; i++) colour = vadd(colour, trace_ray(s, ray, rd, c));
    colour = vdivs(colour, (float)rd->rays_per_pixel);
    V3 previous_sum = vmuls(previous, (float)frame_num);
    V3 res = vdivs(vadd(colour, previous_sum), (float)(frame_num + 1));
    out[array_index] = res.x; out[array_index + 1] = res.y; out[array_index + 2] = res.z;
}

/* ---- row-parallel drivers -------------------------------------------------------------------- */
typedef struct {
    const PtoScene *s; const RtbCamData *cam; int width, height; const RtbRenderData *rd;
    int time_ms, frame_num; const float *prev; float *out; int y0, y1;
    int32_t *obj_id, *sub_id; float *t;
    int *next_row; pthread_mutex_t *mu; RtbCounters counters; int want_counters;
} Job;

static int next_row(Job *j) {
    pthread_mutex_lock(j->mu);
    int y = (*j->next_row)++;
    pthread_mutex_unlock(j->mu);
    return y;
}

static void *render_worker(void *arg) {
    Job *j = (Job *)arg;
    for (;;) {
        int y = next_row(j);
        if (y >= j->y1) break;
        for (int x = 0; x < j->width; x++)
            render_pixel(j->s, j->cam, j->width, j->rd, j->time_ms, j->frame_num, j->prev, j->out, x, y,
                         j->want_counters ? &j->counters : NULL);
    }
    return NULL;
}

static void *ids_worker(void *arg) {
    Job *j = (Job *)arg;
    for (;;) {
        int y = next_row(j);
        if (y >= j->y1) break;
        for (int x = 0; x < j->width; x++) {
            unsigned int rng = 0;
            Ray ray;
            ray_init(&ray, j->cam, x, y, &rng, 0);
            int oi;
            Hit h = closest_hit(j->s, &ray, &oi, NULL);
            size_t p = (size_t)y * j->width + x;
            j->obj_id[p] = oi;
            j->sub_id[p] = oi >= 0 ? h.sub : -1;
            if (j->t) j->t[p] = oi >= 0 ? h.dist : (float)PTO_INF;
        }
    }
    return NULL;
}

static void run_jobs(Job *proto, int nthreads, void *(*fn)(void *)) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    int row = proto->y0;
    Job *jobs = (Job *)malloc(sizeof(Job) * nthreads);
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * nthreads);
    for (int i = 0; i < nthreads; i++) {
        jobs[i] = *proto;
        jobs[i].next_row = &row;
        jobs[i].mu = &mu;
        memset(&jobs[i].counters, 0, sizeof(RtbCounters));
    }
    if (nthreads == 1) fn(&jobs[0]);
    else {
        for (int i = 0; i < nthreads; i++) pthread_create(&th[i], NULL, fn, &jobs[i]);
        for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
    }
    memset(&proto->counters, 0, sizeof(RtbCounters));
    for (int i = 0; i < nthreads; i++) {
        proto->counters.segments += jobs[i].counters.segments;
        proto->counters.hit_segments += jobs[i].counters.hit_segments;
        proto->counters.samples += jobs[i].counters.samples;
    }
    free(jobs); free(th);
}

int pto_render(const PtoScene *s, const RtbCamData *cam, int width, int height, const RtbRenderData *rd, int time_ms,
               int frame_num, const float *prev, float *out, int y0, int y1, int nthreads, RtbCounters *counters) {
    Job j;
    memset(&j, 0, sizeof j);
    j.s = s; j.cam = cam; j.width = width; j.height = height; j.rd = rd; j.time_ms = time_ms; j.frame_num = frame_num;
    j.prev = prev; j.out = out; j.y0 = y0; j.y1 = y1; j.want_counters = counters != NULL;
    run_jobs(&j, nthreads, render_worker);
    if (counters) *counters = j.counters;
    return 0;
}

int pto_primary_ids(const PtoScene *s, const RtbCamData *cam, int width, int height, int32_t *obj_id, int32_t *sub_id,
                    float *t, int nthreads) {
    Job j;
    memset(&j, 0, sizeof j);
    j.s = s; j.cam = cam; j.width = width; j.height = height; j.y0 = 0; j.y1 = height;
    j.obj_id = obj_id; j.sub_id = sub_id; j.t = t;
    run_jobs(&j, nthreads, ids_worker);
    return 0;
}

void pto_primary_dir(const RtbCamData *cam, int x, int y, float *dir3) {
    unsigned int rng = 0;
    Ray r;
    ray_init(&r, cam, x, y, &rng, 0);
    dir3[0] = r.dir.x; dir3[1] = r.dir.y; dir3[2] = r.dir.z;
}

/* Camera::assign_constant_mem at the default pose (all rotations 0), camera.cu:46-108.
 * volatile keeps tanf a run-time libm call as in the reference (no compile-time folding). */
void pto_default_camera(int width, int height, RtbCamData *out) {
    volatile float fov_v = 60 * (PTO_PI / 180);
    volatile float focal_v = 0.1f;
    float fov = fov_v, focal = focal_v;
    float aspect = (float)width / (float)height;
    float vw = 2 * focal * tanf(fov / 2);
    float vh = vw / aspect;
    V3 u = v3(1, 0, 0), v = v3(0, -1, 0);
    float su = (vw / (float)width) / vmag(u);   /* set_mag, utils.cu:157-163 */
    u = vmuls(u, su);
    float sv = (vh / (float)height) / vmag(v);
    v = vmuls(v, sv);
    V3 n = vnorm(vcross(v, u));
    V3 u_step = vdivs(vmuls(u, (float)(-width)), 2);  /* delta_u * -W / 2, camera.cu:101 */
    V3 v_step = vdivs(vmuls(v, (float)(-height)), 2);
    V3 cam_pos = v3(0, 0, 0);
    V3 focal_off = vadd(vmuls(n, focal), cam_pos);
    V3 tl = vadd(vadd(u_step, v_step), focal_off);
    out->cam_pos.x = 0; out->cam_pos.y = 0; out->cam_pos.z = 0;
    out->tl_pixel_pos.x = tl.x; out->tl_pixel_pos.y = tl.y; out->tl_pixel_pos.z = tl.z;
    out->delta_u.x = u.x; out->delta_u.y = u.y; out->delta_u.z = u.z;
    out->delta_v.x = v.x; out->delta_v.y = v.y; out->delta_v.z = v.z;
}

/* ---- display conversion, src/main.cu:343-371 -----------------------------------------------------
 * `int colour = pixel * 255` is a float -> int conversion in host code: on x86-64 (cvttss2si) a NaN
 * or an out-of-range product gives INT_MIN, which the clamp turns into 0.  Restated explicitly so
 * that the result does not depend on the compiler's treatment of the undefined conversion. */
void pto_display_rgba8(const float *rgb, int width, int height, unsigned char *rgba) {
    for (size_t i = 0; i < (size_t)width * height; i++) {
        for (int k = 0; k < 3; k++) {
            float v = rgb[3 * i + k] * 255;
            int colour = (v >= 2147483648.0f || v < -2147483648.0f || v != v) ? (-2147483647 - 1) : (int)v;
            if (colour > 255) colour = 255;
            else if (colour < 0) colour = 0;
            rgba[4 * i + k] = (unsigned char)colour;
        }
        rgba[4 * i + 3] = 255;
    }
}

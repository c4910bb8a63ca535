// rtb_refbvh.cpp -- which boxes the reference puts around the triangles of a mesh.
//
// The reference reaches a mesh triangle only through the boxes of its host-built BVH
// (src/objects.cu:448-771), each tested with `tmin < tmax && tmax > 0` (strict, :430).  For a
// generic ray a triangle hit implies all enclosing boxes are hit and the tree is invisible, with
// two exceptions that do change pixels:
//   * a leaf box with zero extent on an axis (axis-aligned faces such as the top and bottom of
//     models/cube.obj) can never satisfy tmin < tmax: those triangles are INVISIBLE in the
//     reference.  librtb drops them from the LBVH (`degenerate`).
//   * a ray with an exactly-zero direction component whose origin lies on a box plane gives
//     0 * inf = NaN in the slab test; the reference's NaN-dropping min/max can then miss the box.
//     The device code re-runs the reference's slab test along leaf_of -> parent for such rays.
// So the partition itself (not the traversal order) has to be reproduced: fixed depth 10 (:786);
// at each level the triangles are ordered by the distance of points[0] to a reference point on the
// box (:708-719) with a top-down merge sort that takes from the RIGHT run on ties (:655-706) and
// the first n/2+1 go left (:641-650).  Everything is float arithmetic on the host: compile with
// -ffp-contract=off.
#include "rtb_refbvh.h"
#include <cmath>
#include <cstring>

namespace rtb {
namespace {

struct Item {
    int tri;
    float dist;
};

void merge_sort(Item *a, int n, Item *tmp) {
    if (n <= 1) return;
    int nl = n / 2, nr = n - nl;
    merge_sort(a, nl, tmp);
    merge_sort(a + nl, nr, tmp);
    int l = 0, r = 0;
    for (int k = 0; k < n; k++) {
        bool take_left;
        if (l >= nl) take_left = false;
        else if (r >= nr) take_left = true;
        else take_left = a[l].dist < a[nl + r].dist;
        tmp[k] = take_left ? a[l++] : a[nl + r++];
    }
    std::memcpy(a, tmp, sizeof(Item) * n);
}

int build(const float *pts, Item *items, int n, int depth, Item *tmp, RefBvh &out) {
    if (n == 0) return -1;  // the reference makes an empty (0,0,0) box here: degenerate, never hit, holds nothing
    RefBvhNode node;
    for (int c = 0; c < 3; c++) node.lo[c] = node.hi[c] = pts[9 * (size_t)items[0].tri + c];
    for (int i = 0; i < n; i++) {
        const float *p = pts + 9 * (size_t)items[i].tri;
        for (int v = 0; v < 3; v++)
            for (int c = 0; c < 3; c++) {
                node.lo[c] = fminf(node.lo[c], p[3 * v + c]);
                node.hi[c] = fmaxf(node.hi[c], p[3 * v + c]);
            }
    }
    node.parent = -1;
    int me = (int)out.nodes.size();
    out.nodes.push_back(node);
    if (depth <= 0) {
        bool flat = node.lo[0] == node.hi[0] || node.lo[1] == node.hi[1] || node.lo[2] == node.hi[2];
        for (int i = 0; i < n; i++) {
            out.leaf_of[items[i].tri] = me;
            out.degenerate[items[i].tri] = flat ? 1 : 0;
        }
        return me;
    }
    float w = node.hi[0] - node.lo[0], h = node.hi[1] - node.lo[1], d = node.hi[2] - node.lo[2];
    float ref[3];
    if (w >= h && w >= d) { ref[0] = node.lo[0] + w / 2; ref[1] = node.lo[1]; ref[2] = node.lo[2] + d / 2; }
    else if (h >= w && h >= d) { ref[0] = node.lo[0]; ref[1] = node.lo[1] + h / 2; ref[2] = node.lo[2] + d / 2; }
    else { ref[0] = node.lo[0] + w / 2; ref[1] = node.lo[1] + h / 2; ref[2] = node.lo[2]; }
    for (int i = 0; i < n; i++) {
        const float *p = pts + 9 * (size_t)items[i].tri;
        float x = p[0] - ref[0], y = p[1] - ref[1], z = p[2] - ref[2];
        float mag_sq = x * x + y * y + z * z;
        items[i].dist = sqrtf(mag_sq);
    }
    merge_sort(items, n, tmp);
    int n_left = n / 2 + 1;
    if (n_left > n) n_left = n;
    int l = build(pts, items, n_left, depth - 1, tmp, out);
    int r = build(pts, items + n_left, n - n_left, depth - 1, tmp, out);
    if (l >= 0) out.nodes[l].parent = me;
    if (r >= 0) out.nodes[r].parent = me;
    return me;
}

}  // namespace

void ref_bvh_partition(const float *points, int n, RefBvh &out) {
    out.nodes.clear();
    out.leaf_of.assign(n > 0 ? n : 0, -1);
    out.degenerate.assign(n > 0 ? n : 0, 0);
    if (n <= 0) return;
    std::vector<Item> items(n), tmp(n);
    for (int i = 0; i < n; i++) items[i] = Item{i, 0.f};
    build(points, items.data(), n, 10, tmp.data(), out);
}

}  // namespace rtb

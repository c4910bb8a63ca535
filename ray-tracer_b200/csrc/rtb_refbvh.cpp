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
// The preorder node numbers of the reference's recursion follow from the triangle counts alone
// (count_nodes), so the two subtrees of a node -- and the two halves of a merge sort -- are built by
// separate threads near the root: the result is the serial one, a million-triangle mesh takes a
// fraction of the time (the commit of a dynamic scene waits for this).
#include "rtb_refbvh.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>

namespace rtb {
namespace {

struct Item {
    int tri;
    float dist;
};

// glibc's fminf / fmaxf written out (the second argument on ties, the other one if an argument is NaN):
// g++ calls libm for them unless -ffinite-math-only, and the box loops below run 18 of them per triangle
inline float min_f(float x, float y) { return (x < y || y != y) ? x : y; }
inline float max_f(float x, float y) { return (x > y || y != y) ? x : y; }

// `par` = levels of the recursion that still fork a thread (0 = serial).  The two halves are disjoint
// ranges of `a` and `tmp`, so the result does not depend on it.
void merge_sort(Item *a, int n, Item *tmp, int par) {
    if (n <= 1) return;
    int nl = n / 2, nr = n - nl;
    if (par > 0 && n >= 4096) {
        std::thread left([=] { merge_sort(a, nl, tmp, par - 1); });
        merge_sort(a + nl, nr, tmp + nl, par - 1);
        left.join();
    } else {
        merge_sort(a, nl, tmp, 0);
        merge_sort(a + nl, nr, tmp + nl, 0);
    }
    // ties take from the RIGHT run (src/objects.cu:655-706).  The body has no data-dependent branch (the
    // outcome of the comparison is a coin flip on real data): select, then advance one of the cursors.
    const Item *b = a + nl;
    int l = 0, r = 0, k = 0;
    while (l < nl && r < nr) {
        bool take_left = a[l].dist < b[r].dist;
        tmp[k++] = take_left ? a[l] : b[r];
        l += take_left;
        r += !take_left;
    }
    if (l < nl) std::memcpy(tmp + k, a + l, sizeof(Item) * (nl - l));
    else std::memcpy(tmp + k, b + r, sizeof(Item) * (nr - r));
    std::memcpy(a, tmp, sizeof(Item) * n);
}

// The merge sort above orders by (dist ascending, position in the input DESCENDING): a merge that takes
// from the right run on ties puts equal keys in reverse input order, at every level of the recursion.
// That is a total order when no dist is NaN, so any sort that realises it gives the same array: here a
// stable LSD radix sort (3 passes of 11 bits over the float's bit pattern -- dist is a square root, never
// negative, so the patterns order like the values) of the REVERSED input.  Linear time instead of
// n log n unpredictable branches: the sorts were three quarters of the partition's time.
void sort_items(Item *a, int n, Item *tmp, int par) {
    bool ordered_bits = true;
    for (int i = 0; i < n; i++) ordered_bits &= a[i].dist >= 0.f;  // false for NaN (and for a negative value, which sqrtf never returns)
    if (n < 2048 || !ordered_bits) {
        merge_sort(a, n, tmp, par);
        return;
    }
    std::reverse(a, a + n);
    std::vector<unsigned int> hist(3 * 2048, 0u);
    auto bits = [](const Item &it) { unsigned int u; std::memcpy(&u, &it.dist, 4); return u; };
    for (int i = 0; i < n; i++) {
        unsigned int u = bits(a[i]);
        hist[u & 2047u]++; hist[2048 + ((u >> 11) & 2047u)]++; hist[4096 + (u >> 22)]++;
    }
    for (int pass = 0; pass < 3; pass++) {
        unsigned int *h = hist.data() + 2048 * pass, sum = 0;
        for (int d = 0; d < 2048; d++) { unsigned int c = h[d]; h[d] = sum; sum += c; }
    }
    Item *src = a, *dst = tmp;
    for (int pass = 0; pass < 3; pass++) {
        unsigned int *h = hist.data() + 2048 * pass;
        const int shift = 11 * pass;
        for (int i = 0; i < n; i++) dst[h[(bits(src[i]) >> shift) & 2047u]++] = src[i];
        Item *t = src; src = dst; dst = t;
    }
    std::memcpy(a, tmp, sizeof(Item) * n);  // three passes: the result is in tmp
}

int left_count(int n) { return n / 2 + 1 > n ? n : n / 2 + 1; }  // src/objects.cu:641-650

// nodes of the subtree over n triangles with `depth` levels below it: the preorder numbering of the
// reference's recursive build, known before the build runs, so subtrees can be built concurrently
int count_nodes(int n, int depth) {
    if (n == 0) return 0;
    if (depth <= 0) return 1;
    int nl = left_count(n);
    return 1 + count_nodes(nl, depth - 1) + count_nodes(n - nl, depth - 1);
}

// builds the subtree whose root is node `me` (nodes are pre-sized); returns me, or -1 for an empty range
int build(const float *pts, Item *items, int n, int depth, Item *tmp, RefBvh &out, int me, int par) {
    if (n == 0) return -1;  // the reference makes an empty (0,0,0) box here: degenerate, never hit, holds nothing
    RefBvhNode node;
    for (int c = 0; c < 3; c++) node.lo[c] = node.hi[c] = pts[9 * (size_t)items[0].tri + c];
    for (int i = 0; i < n; i++) {
        const float *p = pts + 9 * (size_t)items[i].tri;
        for (int v = 0; v < 3; v++)
            for (int c = 0; c < 3; c++) {
                node.lo[c] = min_f(node.lo[c], p[3 * v + c]);
                node.hi[c] = max_f(node.hi[c], p[3 * v + c]);
            }
    }
    node.parent = -1;
    out.nodes[me] = node;
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
    sort_items(items, n, tmp, par);
    int n_left = left_count(n);
    int left_root = me + 1, right_root = me + 1 + count_nodes(n_left, depth - 1);
    int l, r;
    if (par > 0 && n >= 4096) {
        std::thread left([&] { l = build(pts, items, n_left, depth - 1, tmp, out, left_root, par - 1); });
        r = build(pts, items + n_left, n - n_left, depth - 1, tmp + n_left, out, right_root, par - 1);
        left.join();
    } else {
        l = build(pts, items, n_left, depth - 1, tmp, out, left_root, 0);
        r = build(pts, items + n_left, n - n_left, depth - 1, tmp + n_left, out, right_root, 0);
    }
    if (l >= 0) out.nodes[l].parent = me;
    if (r >= 0) out.nodes[r].parent = me;
    return me;
}

}  // namespace

void ref_bvh_partition(const float *points, int n, RefBvh &out, int threads) {
    out.nodes.clear();
    out.leaf_of.assign(n > 0 ? n : 0, -1);
    out.degenerate.assign(n > 0 ? n : 0, 0);
    if (n <= 0) return;
    std::vector<Item> items(n), tmp(n);
    for (int i = 0; i < n; i++) items[i] = Item{i, 0.f};
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    int par = 0;  // fork levels: 2^par concurrent subtrees / sort runs
    while (par < 5 && (2 << par) <= threads) par++;
    out.nodes.resize(count_nodes(n, 10));
    build(points, items.data(), n, 10, tmp.data(), out, 0, par);
}

}  // namespace rtb

// Host logic check (CPU suite): ray-tracer_b200/csrc/rtb_refbvh.cpp -- radix sort, inlined min / max, threads
// near the root -- against a literal single-threaded restatement of the reference's BVH partition
// (src/objects.cu:602-770: recursive build to depth 10, top-down merge sort that takes from the right run on
// ties, first n/2+1 triangles go left).  Inputs are seeded and full of exact ties, duplicate triangles,
// axis-aligned (flat) triangles and signed zeros.  Prints "mismatches 0" on success.
#include "rtb_refbvh.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace lit {
struct Item { int tri; float dist; };
struct Node { float lo[3], hi[3]; int parent; };
struct Out { std::vector<Node> nodes; std::vector<int> leaf_of; std::vector<char> degenerate; };

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

int build(const float *pts, Item *items, int n, int depth, Item *tmp, Out &out) {
    if (n == 0) return -1;
    Node node;
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
        for (int i = 0; i < n; i++) { out.leaf_of[items[i].tri] = me; out.degenerate[items[i].tri] = flat ? 1 : 0; }
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

void partition(const float *pts, int n, Out &out) {
    out.nodes.clear();
    out.leaf_of.assign(n, -1);
    out.degenerate.assign(n, 0);
    if (n <= 0) return;
    std::vector<Item> items(n), tmp(n);
    for (int i = 0; i < n; i++) items[i] = Item{i, 0.f};
    build(pts, items.data(), n, 10, tmp.data(), out);
}
}  // namespace lit

int main() {
    long mismatches = 0;
    const int sizes[] = {0, 1, 2, 3, 7, 12, 723, 2047, 2048, 2049, 5000, 40000, 150000};
    for (int n : sizes) {
        std::vector<float> pts((size_t)n * 9 + 9);
        srand(1234 + n);
        for (auto &p : pts) p = (float)(rand() % 1000) / 37.f - 5.f;
        for (int i = 0; i < n / 10; i++) std::memcpy(&pts[(size_t)(rand() % n) * 9], &pts[(size_t)(rand() % n) * 9], 36);
        for (int i = 0; i < n / 50; i++) { size_t k = (size_t)(rand() % n) * 9; pts[k + 5] = pts[k + 8] = pts[k + 2]; }
        for (int i = 0; i < n / 100; i++) pts[(size_t)(rand() % n) * 9 + rand() % 9] = (rand() & 1) ? 0.f : -0.f;
        lit::Out want;
        lit::partition(pts.data(), n, want);
        for (int threads : {1, 2, 0}) {
            rtb::RefBvh got;
            rtb::ref_bvh_partition(pts.data(), n, got, threads);
            bool same = got.nodes.size() == want.nodes.size() && got.leaf_of == want.leaf_of && got.degenerate == want.degenerate;
            for (size_t k = 0; same && k < want.nodes.size(); k++)
                same = std::memcmp(got.nodes[k].lo, want.nodes[k].lo, 12) == 0 && std::memcmp(got.nodes[k].hi, want.nodes[k].hi, 12) == 0 &&
                       got.nodes[k].parent == want.nodes[k].parent;
            if (!same) { mismatches++; printf("n %d threads %d differs\n", n, threads); }
        }
    }
    printf("mismatches %ld\n", mismatches);
    return mismatches ? 1 : 0;
}

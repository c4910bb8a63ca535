// rtb_refbvh.h -- host restatement of the reference's BVH partition (which mesh triangles share a
// box), needed only for the two degenerate cases in which the reference's box tests change the
// picture (see rtb_refbvh.cpp).
#pragma once
#include <vector>

namespace rtb {

struct RefBvhNode {
    float lo[3], hi[3];
    int parent;  // -1 at the root
};

struct RefBvh {
    std::vector<RefBvhNode> nodes;
    std::vector<int> leaf_of;      // per mesh triangle: node index of its leaf
    std::vector<char> degenerate;  // per mesh triangle: 1 if its leaf box has zero extent on an axis
};

// points: 9 floats per triangle (p0 p1 p2), n triangles of ONE mesh object.
// threads: host threads near the root of the recursion (1 = serial, 0 = all hardware threads); the
// result does not depend on it.
void ref_bvh_partition(const float *points, int n, RefBvh &out, int threads = 0);

}  // namespace rtb

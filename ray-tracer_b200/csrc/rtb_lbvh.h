// rtb_lbvh.h -- GPU LBVH builder interface (see rtb_lbvh.cu)
#pragma once
#include "rtb_device.cuh"

namespace rtb {

struct LbvhInput {
    int num_prims;
    const int *leaf_refs;    // device: encoded leaf refs (~((index << 3) | kind)), creation order
    const float4 *tris;      // device primitive records (layout in rtb_device.cuh)
    const float4 *spheres;
};

// Builds the tree on `st`.  *nodes_out: cudaMalloc'd (n-1)*4 float4 (caller frees), nullptr when n < 2.
// *root_ref: the root's child reference (an internal node index or, for a single primitive, its leaf ref).
// *height_out: height of the tree (a single node over two leaves = 1): the traversal needs one stack entry per level.
// builder: 0 = Karras hierarchy over the Morton codes (LBVH), 1 = PLOC (agglomerative, near-SAH
// quality) with search radius `ploc_radius` over the same Morton order.  PLOC returns cudaErrorNotSupported if a
// round merges nothing (cannot happen with finite boxes): the caller falls back to builder 0.
cudaError_t lbvh_build(const LbvhInput &in, float4 **nodes_out, int *root_ref, int *height_out, cudaStream_t st, int builder, int ploc_radius);

}  // namespace rtb

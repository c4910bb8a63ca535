// rtb_meshprep.h -- the flatten of a big mesh on the device (see rtb_meshprep.cu): the reference-BVH partition
// (which triangles share which box, rtb_refbvh.cpp on the host) and the per-triangle records.
#pragma once
#include "rtb_device.cuh"
#include <vector>

namespace rtb {

struct MeshPrep {
    int n = 0;                 // triangles of the mesh
    int kept = 0;              // triangles that are not culled (degenerate reference leaves are invisible in the reference)
    int *d_leaf_of = nullptr;  // [n] node index (inside this mesh's partition) of each triangle's leaf; nullptr = no partition
    int *d_slot = nullptr;     // [n] position among the kept triangles, -1 = culled; nullptr = all kept, slot = index
    std::vector<float4> nodes; // host copy, 2 per node: {lo.xyz, parent as int bits (-1 = root)} {hi.xyz, 0}
    void release();
};

// The reference's BVH partition of one mesh (src/objects.cu:602-770; the host restatement is rtb_refbvh.cpp) on the
// device.  d_pts: 9 floats per triangle.  *nonfinite = 1 when a coordinate is NaN / inf: the result is then
// unspecified and the caller takes the host path (whose merge sort handles NaN keys the reference's way).
cudaError_t mesh_partition_device(const float *d_pts, int n, cudaStream_t st, MeshPrep &out, int *nonfinite);

// Writes the records of the kept triangles of one mesh into the scene arrays (the arithmetic of make_tri / set_tri
// in rtb_api.cu, rounding by rounding): triangle j of the kept ones goes to tri index tri_base + j, leaf position
// leaf_base + j.  d_uv: 6 floats per triangle or nullptr (zeros).  bounds: 6 order-preserving ints, grown.
struct MeshFill {
    const float *d_pts;
    const float *d_uv;
    int n;
    const int *d_slot, *d_leaf_of;
    int obj, tri_base, leaf_base, node_base;
    float4 *tris;
    int2 *tri_ids;
    float2 *tri_uv;
    int *tri_refleaf;
    int *leaf_refs;
    int *bounds;
};
cudaError_t mesh_fill_device(const MeshFill &f, cudaStream_t st);
void mesh_bounds_init(int *d_bounds6, cudaStream_t st);
void mesh_bounds_decode(const int *h_bounds6, float *lo, float *hi);

// Leaves of a device-flattened mesh whose box area exceeds `threshold` (candidates for the large-primitive list):
// (position among the mesh's kept triangles, area) pairs, at most `cap`; *d_count counts all of them.
cudaError_t mesh_large_candidates(const float4 *tris, int tri_base, int count, float threshold, int *d_count, int2 *d_out, int cap, cudaStream_t st);

}  // namespace rtb

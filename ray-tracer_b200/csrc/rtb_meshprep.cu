// rtb_meshprep.cu -- the flatten of a big mesh on the device (dynamic million-triangle meshes).
//
// rtb_scene_commit has to know which boxes the REFERENCE puts around the triangles of a mesh (src/objects.cu:602-770:
// fixed depth 10; at each level the triangles are ordered by the distance of their first vertex to a reference point
// on the node's box with a merge sort that takes from the right run on ties, the first n/2+1 go left), because
// triangles in leaves with a zero-extent box are invisible in the reference and rays with a zero direction component
// see its box chain (rtb_refbvh.cpp explains both).  The host restatement (rtb_refbvh.cpp) was 68 ms of a 187 ms
// commit of the 1 012 200-triangle mesh, and the per-triangle records another 85 ms.  Here both run on the GPU:
//
//   partition   level by level, ALL nodes of a level at once.  The shape of the tree (how many triangles each node
//               holds, the preorder node numbers) follows from the triangle count alone, so the host builds the
//               segment table of every level up front.  Per level: segment boxes (warp-reduced atomic min / max on
//               order-preserving ints), reference points, distances, then ONE stable radix sort of the whole array by
//               (segment, distance bits) of the items written in REVERSED order inside their segment -- a merge sort
//               that takes from the right on ties orders by (distance ascending, input position descending), which is
//               what a stable sort of the reversed input gives (same argument as the host version's radix sort).
//               Every float operation is the host's, rounding by rounding (__fsub_rn / __fmul_rn / __fadd_rn /
//               __fsqrt_rn: no FMA contraction), so the order and the boxes are the host's bit for bit; the one
//               freedom is the sign of a zero in a box plane (min / max of -0 and +0 depend on the order on the host),
//               which no comparison downstream can see.
//   records     one thread per triangle: make_tri / set_tri of rtb_api.cu with the same roundings.
//
// tests/test_gpu_dynamic.py compares both paths (partition arrays, images, ids) on meshes with ties, duplicates and
// axis-aligned faces.
#include "rtb_meshprep.h"
#include <cuda_runtime.h>
#include <cstring>
#include <vector>

namespace rtb {
namespace {

#include "rtb_sort.cuh"

constexpr int PART_DEPTH = 10;  // src/objects.cu:786

// order-preserving float <-> int (signed compare); -0 < +0
__device__ __forceinline__ int enc_i(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__host__ __device__ __forceinline__ float dec_i(int i) {
    int b = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
    return __int_as_float(b);
#else
    float f; std::memcpy(&f, &b, 4); return f;
#endif
}
__device__ __forceinline__ bool finite_f(float x) { return fabsf(x) <= 3.4028234e38f; }  // false for NaN and inf

// last segment whose start is <= i (segments are contiguous and ascending; num >= 1)
__device__ __forceinline__ int seg_of(const int *seg_start, int num, int i) {
    int lo = 0, hi = num - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (__ldg(seg_start + mid) <= i) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void iota_kernel(int n, int *p) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}
__global__ void box_init_kernel(int num_seg, int *box) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < num_seg * 6) box[i] = (i % 6) < 3 ? 0x7fffffff : (int)0x80000000;
}

// min / max of six encoded values over the warp, lane 0 (or every lane when the warp spans segments) commits
__device__ __forceinline__ void commit_box(int *dst, bool valid, bool uniform, int v[6]) {
    if (uniform) {
#pragma unroll
        for (int k = 0; k < 6; k++) {
            int x = v[k];
            x = k < 3 ? __reduce_min_sync(0xffffffffu, x) : __reduce_max_sync(0xffffffffu, x);
            v[k] = x;
        }
        if ((threadIdx.x & 31) == 0) {
#pragma unroll
            for (int k = 0; k < 3; k++) { atomicMin(dst + k, v[k]); atomicMax(dst + 3 + k, v[3 + k]); }
        }
    } else if (valid) {
#pragma unroll
        for (int k = 0; k < 3; k++) { atomicMin(dst + k, v[k]); atomicMax(dst + 3 + k, v[3 + k]); }
    }
}

// boxes of the segments of one level: min / max over the three vertices of every triangle (build(), rtb_refbvh.cpp)
__global__ void part_box_kernel(int n, const float *pts, const int *items, const int *seg_start, int num_seg, int *box, int *nonfinite) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = i < n;
    int seg = 0;
    int v[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    if (valid) {
        seg = seg_of(seg_start, num_seg, i);
        const float *p = pts + 9 * (size_t)__ldg(items + i);
        bool bad = false;
#pragma unroll
        for (int k = 0; k < 9; k++) {
            float x = __ldg(p + k);
            if (!finite_f(x)) { bad = true; continue; }
            int e = enc_i(x);
            v[k % 3] = min(v[k % 3], e);
            v[3 + k % 3] = max(v[3 + k % 3], e);
        }
        if (bad) *nonfinite = 1;
    }
    unsigned m = __ballot_sync(0xffffffffu, valid);
    int seg0 = __shfl_sync(0xffffffffu, seg, m ? __ffs(m) - 1 : 0);
    bool uniform = __all_sync(0xffffffffu, !valid || seg == seg0);
    commit_box(box + 6 * (uniform ? seg0 : seg), valid, uniform && m != 0, v);
}

// per segment: the node record, and the reference point of the next split (build(), rtb_refbvh.cpp; src/objects.cu:708-719)
// or, at the leaf level, whether the box is flat on an axis
__global__ void part_ref_kernel(int num_seg, const int *box, const int *seg_node, const int *seg_parent, float4 *nodes, float *ref, unsigned char *flat, int leaf_level) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= num_seg) return;
    float lo[3], hi[3];
    for (int k = 0; k < 3; k++) { lo[k] = dec_i(box[6 * s + k]); hi[k] = dec_i(box[6 * s + 3 + k]); }
    int node = seg_node[s];
    nodes[2 * node] = make_float4(lo[0], lo[1], lo[2], __int_as_float(seg_parent[s]));
    nodes[2 * node + 1] = make_float4(hi[0], hi[1], hi[2], 0.f);
    if (leaf_level) {
        flat[s] = (lo[0] == hi[0] || lo[1] == hi[1] || lo[2] == hi[2]) ? 1 : 0;
        return;
    }
    float w = __fsub_rn(hi[0], lo[0]), h = __fsub_rn(hi[1], lo[1]), d = __fsub_rn(hi[2], lo[2]);
    float r0, r1, r2;
    if (w >= h && w >= d) { r0 = __fadd_rn(lo[0], __fdiv_rn(w, 2.f)); r1 = lo[1]; r2 = __fadd_rn(lo[2], __fdiv_rn(d, 2.f)); }
    else if (h >= w && h >= d) { r0 = lo[0]; r1 = __fadd_rn(lo[1], __fdiv_rn(h, 2.f)); r2 = __fadd_rn(lo[2], __fdiv_rn(d, 2.f)); }
    else { r0 = __fadd_rn(lo[0], __fdiv_rn(w, 2.f)); r1 = __fadd_rn(lo[1], __fdiv_rn(h, 2.f)); r2 = lo[2]; }
    ref[3 * s] = r0; ref[3 * s + 1] = r1; ref[3 * s + 2] = r2;
}

// sort keys of one level: (segment, bits of the distance of points[0] to the segment's reference point), written
// at the REVERSED position inside the segment
__global__ void part_dist_kernel(int n, const float *pts, const int *items, const int *seg_start, const int *seg_count, int num_seg, const float *ref,
                                 unsigned long long *keys, int *vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int seg = seg_of(seg_start, num_seg, i);
    int tri = __ldg(items + i);
    const float *p = pts + 9 * (size_t)tri;
    float x = __fsub_rn(__ldg(p), ref[3 * seg]), y = __fsub_rn(__ldg(p + 1), ref[3 * seg + 1]), z = __fsub_rn(__ldg(p + 2), ref[3 * seg + 2]);
    float mag_sq = __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z));
    float dist = __fsqrt_rn(mag_sq);
    int start = __ldg(seg_start + seg), count = __ldg(seg_count + seg);
    int rev = start + (count - 1 - (i - start));
    keys[rev] = ((unsigned long long)(unsigned int)seg << 32) | (unsigned long long)__float_as_uint(dist);
    vals[rev] = tri;
}

__global__ void part_leaf_kernel(int n, const int *items, const int *seg_start, const int *seg_node, int num_seg, const unsigned char *flat, int *leaf_of, unsigned char *degenerate) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int seg = seg_of(seg_start, num_seg, i);
    int tri = __ldg(items + i);
    leaf_of[tri] = __ldg(seg_node + seg);
    degenerate[tri] = flat[seg];
}

// slots of the kept triangles, in triangle order: per-block counts, one-block scan, per-thread ranks
constexpr int SLOT_T = 1024;
__global__ void slot_count_kernel(int n, const unsigned char *degenerate, unsigned int *block_count) {
    int i = blockIdx.x * SLOT_T + threadIdx.x;
    int c = __syncthreads_count(i < n && !degenerate[i]);
    if (threadIdx.x == 0) block_count[blockIdx.x] = (unsigned int)c;
}
__global__ void slot_write_kernel(int n, const unsigned char *degenerate, const unsigned int *block_offset, int *slot, int *total) {
    __shared__ unsigned int warp_count[SLOT_T / 32];
    int i = blockIdx.x * SLOT_T + threadIdx.x;
    bool keep = i < n && !degenerate[i];
    unsigned int m = __ballot_sync(0xffffffffu, keep);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warp_count[warp] = __popc(m);
    __syncthreads();
    unsigned int before = 0;
    for (int w = 0; w < warp; w++) before += warp_count[w];
    unsigned int pos = block_offset[blockIdx.x] + before + __popc(m & ((1u << lane) - 1u));
    if (i < n) slot[i] = keep ? (int)pos : -1;
    if (i == n - 1) *total = (int)(pos + (keep ? 1u : 0u));
}

// ---- the records of the kept triangles ------------------------------------------------------------------------------
__global__ void mesh_fill_kernel(MeshFill f) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = k < f.n;
    int j = valid ? (f.d_slot ? f.d_slot[k] : k) : -1;
    valid = j >= 0;
    int v[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    if (valid) {
        const float *p = f.d_pts + 9 * (size_t)k;
        float p0x = p[0], p0y = p[1], p0z = p[2];
        // make_tri (rtb_api.cu; Triangle::precompute, objects.cu:175-186)
        float s1x = __fsub_rn(p[3], p0x), s1y = __fsub_rn(p[4], p0y), s1z = __fsub_rn(p[5], p0z);
        float s2x = __fsub_rn(p[6], p0x), s2y = __fsub_rn(p[7], p0y), s2z = __fsub_rn(p[8], p0z);
        float cx = __fsub_rn(__fmul_rn(s1y, s2z), __fmul_rn(s1z, s2y));
        float cy = __fsub_rn(__fmul_rn(s1z, s2x), __fmul_rn(s1x, s2z));
        float cz = __fsub_rn(__fmul_rn(s1x, s2y), __fmul_rn(s1y, s2x));
        float mag_sq = __fadd_rn(__fadd_rn(__fmul_rn(cx, cx), __fmul_rn(cy, cy)), __fmul_rn(cz, cz));
        float inv = __fdiv_rn(1.f, __fsqrt_rn(mag_sq));
        float nx = __fmul_rn(cx, inv), ny = __fmul_rn(cy, inv), nz = __fmul_rn(cz, inv);
        size_t t = (size_t)f.tri_base + j;
        f.tris[3 * t] = make_float4(p0x, p0y, p0z, s1x);
        f.tris[3 * t + 1] = make_float4(s1y, s1z, s2x, s2y);
        f.tris[3 * t + 2] = make_float4(s2z, nx, ny, nz);
        f.tri_ids[t] = make_int2(f.obj, k);
        const float *uv = f.d_uv ? f.d_uv + 6 * (size_t)k : nullptr;
        for (int q = 0; q < 3; q++) f.tri_uv[3 * t + q] = uv ? make_float2(uv[2 * q], uv[2 * q + 1]) : make_float2(0.f, 0.f);
        f.tri_refleaf[t] = f.d_leaf_of ? f.node_base + f.d_leaf_of[k] : -1;
        f.leaf_refs[(size_t)f.leaf_base + j] = ~(((int)t << RTB_KIND_BITS) | KIND_TRI);
        // tri_box (rtb_api.cu): the vertices as the record gives them back, NaN-skipping min / max
        float q[9] = {p0x, p0y, p0z, __fadd_rn(p0x, s1x), __fadd_rn(p0y, s1y), __fadd_rn(p0z, s1z), __fadd_rn(p0x, s2x), __fadd_rn(p0y, s2y), __fadd_rn(p0z, s2z)};
#pragma unroll
        for (int c = 0; c < 9; c++) {
            if (q[c] != q[c]) continue;
            int e = enc_i(q[c]);
            v[c % 3] = min(v[c % 3], e);
            v[3 + c % 3] = max(v[3 + c % 3], e);
        }
    }
    commit_box(f.bounds, valid, true, v);
}

__global__ void mesh_large_kernel(const float4 *tris, int tri_base, int count, float threshold, int *out_count, int2 *out, int cap) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    size_t t = (size_t)tri_base + j;
    float4 a = tris[3 * t], b = tris[3 * t + 1], c = tris[3 * t + 2];
    float q[9] = {a.x, a.y, a.z, __fadd_rn(a.x, a.w), __fadd_rn(a.y, b.x), __fadd_rn(a.z, b.y), __fadd_rn(a.x, b.z), __fadd_rn(a.y, b.w), __fadd_rn(a.z, c.x)};
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
    for (int k = 0; k < 9; k++) {  // tri_box of rtb_api.cu
        float x = q[k];
        lo[k % 3] = (lo[k % 3] < x || x != x) ? lo[k % 3] : x;
        hi[k % 3] = (hi[k % 3] > x || x != x) ? hi[k % 3] : x;
    }
    float e0 = __fsub_rn(hi[0], lo[0]), e1 = __fsub_rn(hi[1], lo[1]), e2 = __fsub_rn(hi[2], lo[2]);
    float area = __fmul_rn(2.f, __fadd_rn(__fadd_rn(__fmul_rn(e0, e1), __fmul_rn(e1, e2)), __fmul_rn(e2, e0)));
    if (area > threshold) {
        int slot = atomicAdd(out_count, 1);
        if (slot < cap) out[slot] = make_int2(j, __float_as_int(area));
    }
}

int left_count(int n) { return n / 2 + 1 > n ? n : n / 2 + 1; }  // src/objects.cu:641-650

struct Seg {
    int start, count, node, parent;
};
// preorder numbering of the reference's recursion (a node, its left subtree, its right subtree; empty ranges make
// no node): the segments of every level, in position order
void plan(std::vector<Seg> *levels, int start, int count, int depth, int parent, int &next_node) {
    if (count == 0) return;
    int me = next_node++;
    levels[PART_DEPTH - depth].push_back(Seg{start, count, me, parent});
    if (depth <= 0) return;
    int nl = left_count(count);
    plan(levels, start, nl, depth - 1, me, next_node);
    plan(levels, start + nl, count - nl, depth - 1, me, next_node);
}

}  // namespace

#define MP_CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)

void MeshPrep::release() {
    cudaFree(d_leaf_of); cudaFree(d_slot);
    d_leaf_of = nullptr; d_slot = nullptr;
}

cudaError_t mesh_partition_device(const float *d_pts, int n, cudaStream_t st, MeshPrep &out, int *nonfinite) {
    out.n = n; out.kept = n;
    out.nodes.clear();
    *nonfinite = 0;
    if (n <= 0) return cudaSuccess;
    std::vector<Seg> levels[PART_DEPTH + 1];
    int num_nodes = 0;
    plan(levels, 0, n, PART_DEPTH, -1, num_nodes);
    // one table for all levels: start | count | node | parent, each num_segs_total long
    std::vector<int> table;
    int level_off[PART_DEPTH + 2] = {0};
    for (int L = 0; L <= PART_DEPTH; L++) level_off[L + 1] = level_off[L] + (int)levels[L].size();
    const int S = level_off[PART_DEPTH + 1];
    table.resize(4 * (size_t)S);
    for (int L = 0; L <= PART_DEPTH; L++)
        for (size_t s = 0; s < levels[L].size(); s++) {
            const Seg &g = levels[L][s];
            size_t at = level_off[L] + s;
            table[at] = g.start; table[S + at] = g.count; table[2 * (size_t)S + at] = g.node; table[3 * (size_t)S + at] = g.parent;
        }
    cudaError_t err = cudaSuccess;
    const int T = 256, B = (n + T - 1) / T;
    int *d_table = nullptr, *d_box = nullptr, *d_flag = nullptr, *vals = nullptr, *vals2 = nullptr, *d_total = nullptr;
    unsigned long long *keys = nullptr, *keys2 = nullptr;
    unsigned int *hist = nullptr, *d_block = nullptr;
    float *d_ref = nullptr;
    float4 *d_nodes = nullptr;
    unsigned char *d_flat = nullptr, *d_degenerate = nullptr;
    int max_seg = 0;
    for (int L = 0; L <= PART_DEPTH; L++) max_seg = std::max(max_seg, (int)levels[L].size());
    const int slot_blocks = (n + SLOT_T - 1) / SLOT_T;
    int h_out[2] = {0, 0};
    MP_CHECK(cudaMalloc(&d_table, sizeof(int) * table.size()));
    MP_CHECK(cudaMalloc(&d_box, sizeof(int) * 6 * (size_t)max_seg));
    MP_CHECK(cudaMalloc(&d_flag, sizeof(int) * 2));
    d_total = d_flag + 1;
    MP_CHECK(cudaMalloc(&d_ref, sizeof(float) * 3 * (size_t)max_seg));
    MP_CHECK(cudaMalloc(&d_flat, (size_t)max_seg));
    MP_CHECK(cudaMalloc(&d_nodes, sizeof(float4) * 2 * (size_t)num_nodes));
    MP_CHECK(cudaMalloc(&keys, sizeof(unsigned long long) * (size_t)n));
    MP_CHECK(cudaMalloc(&keys2, sizeof(unsigned long long) * (size_t)n));
    MP_CHECK(cudaMalloc(&vals, sizeof(int) * (size_t)n));
    MP_CHECK(cudaMalloc(&vals2, sizeof(int) * (size_t)n));
    MP_CHECK(cudaMalloc(&hist, sizeof(unsigned int) * 256 * (size_t)sort_blocks_for(n)));
    MP_CHECK(cudaMalloc(&d_block, sizeof(unsigned int) * (size_t)slot_blocks));
    MP_CHECK(cudaMalloc(&d_degenerate, (size_t)n));
    MP_CHECK(cudaMalloc(&out.d_leaf_of, sizeof(int) * (size_t)n));
    MP_CHECK(cudaMalloc(&out.d_slot, sizeof(int) * (size_t)n));
    MP_CHECK(cudaMemcpyAsync(d_table, table.data(), sizeof(int) * table.size(), cudaMemcpyHostToDevice, st));
    MP_CHECK(cudaMemsetAsync(d_flag, 0, sizeof(int) * 2, st));
    iota_kernel<<<B, T, 0, st>>>(n, vals);
    for (int L = 0; L <= PART_DEPTH; L++) {
        const int num_seg = (int)levels[L].size();
        const int *seg_start = d_table + level_off[L], *seg_count = d_table + S + level_off[L];
        const int *seg_node = d_table + 2 * (size_t)S + level_off[L], *seg_parent = d_table + 3 * (size_t)S + level_off[L];
        box_init_kernel<<<(6 * num_seg + T - 1) / T, T, 0, st>>>(num_seg, d_box);
        part_box_kernel<<<B, T, 0, st>>>(n, d_pts, vals, seg_start, num_seg, d_box, d_flag);
        part_ref_kernel<<<(num_seg + T - 1) / T, T, 0, st>>>(num_seg, d_box, seg_node, seg_parent, d_nodes, d_ref, d_flat, L == PART_DEPTH);
        if (L == PART_DEPTH) {
            part_leaf_kernel<<<B, T, 0, st>>>(n, vals, seg_start, seg_node, num_seg, d_flat, out.d_leaf_of, d_degenerate);
            break;
        }
        part_dist_kernel<<<B, T, 0, st>>>(n, d_pts, vals, seg_start, seg_count, num_seg, d_ref, keys2, vals2);
        {   // the keys were written into the second buffers: sort from there
            unsigned long long *tk = keys; keys = keys2; keys2 = tk;
            int *tv = vals; vals = vals2; vals2 = tv;
        }
        int seg_bits = 0;
        while ((1 << seg_bits) < num_seg) seg_bits++;
        radix_sort_pairs(n, keys, vals, keys2, vals2, hist, 4 + (seg_bits + 7) / 8, st);
    }
    slot_count_kernel<<<slot_blocks, SLOT_T, 0, st>>>(n, d_degenerate, d_block);
    sort_scan_kernel<<<1, 1024, 0, st>>>(d_block, slot_blocks);
    slot_write_kernel<<<slot_blocks, SLOT_T, 0, st>>>(n, d_degenerate, d_block, out.d_slot, d_total);
    MP_CHECK(cudaGetLastError());
    out.nodes.resize(2 * (size_t)num_nodes);
    MP_CHECK(cudaMemcpyAsync(out.nodes.data(), d_nodes, sizeof(float4) * out.nodes.size(), cudaMemcpyDeviceToHost, st));
    MP_CHECK(cudaMemcpyAsync(h_out, d_flag, sizeof(int) * 2, cudaMemcpyDeviceToHost, st));
    MP_CHECK(cudaStreamSynchronize(st));
    *nonfinite = h_out[0];
    out.kept = h_out[1];
done:
    cudaFree(d_table); cudaFree(d_box); cudaFree(d_flag); cudaFree(d_ref); cudaFree(d_flat); cudaFree(d_nodes);
    cudaFree(keys); cudaFree(keys2); cudaFree(vals); cudaFree(vals2); cudaFree(hist); cudaFree(d_block); cudaFree(d_degenerate);
    if (err != cudaSuccess) out.release();
    return err;
}

cudaError_t mesh_fill_device(const MeshFill &f, cudaStream_t st) {
    if (f.n <= 0) return cudaSuccess;
    mesh_fill_kernel<<<(f.n + 255) / 256, 256, 0, st>>>(f);
    return cudaGetLastError();
}

void mesh_bounds_init(int *d_bounds6, cudaStream_t st) { box_init_kernel<<<1, 32, 0, st>>>(1, d_bounds6); }

void mesh_bounds_decode(const int *h, float *lo, float *hi) {
    for (int k = 0; k < 3; k++) { lo[k] = dec_i(h[k]); hi[k] = dec_i(h[3 + k]); }
}

cudaError_t mesh_large_candidates(const float4 *tris, int tri_base, int count, float threshold, int *d_count, int2 *d_out, int cap, cudaStream_t st) {
    if (count <= 0) return cudaSuccess;
    mesh_large_kernel<<<(count + 255) / 256, 256, 0, st>>>(tris, tri_base, count, threshold, d_count, d_out, cap);
    return cudaGetLastError();
}

}  // namespace rtb

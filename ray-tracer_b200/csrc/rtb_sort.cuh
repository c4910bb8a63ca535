// rtb_sort.cuh -- stable LSD radix sort of (64-bit key, int value) pairs, 8 bits per pass: per-block digit
// histograms, one-block exclusive scan of the digit-major table, stable scatter.  Hand-written (no cub / thrust);
// shared by the tree build (rtb_lbvh.cu: Morton codes) and the device mesh flatten (rtb_meshprep.cu: the
// reference's merge-sort order).  Include inside `namespace rtb { namespace {`.
#pragma once

// ---- stable LSD radix sort, 8 bits per pass ------------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 8;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;

__global__ void sort_hist_kernel(int n, const unsigned long long *keys, int shift, unsigned int *block_hist, int num_blocks) {
    __shared__ unsigned int hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    int base = blockIdx.x * SORT_TILE;
    for (int k = 0; k < SORT_ITEMS; k++) {
        int i = base + k * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[(unsigned int)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    block_hist[threadIdx.x * num_blocks + blockIdx.x] = hist[threadIdx.x];  // digit-major
}

// exclusive scan of the digit-major histogram by ONE block (<= a few hundred thousand entries)
__global__ void sort_scan_kernel(unsigned int *data, int n) {
    __shared__ unsigned int warp_sums[32];
    __shared__ unsigned int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += blockDim.x) {
        int i = base + threadIdx.x;
        unsigned int v = i < n ? data[i] : 0, x = v;
        int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sums[warp] = x;
        __syncthreads();
        if (warp == 0) {
            unsigned int w = lane < (int)(blockDim.x >> 5) ? warp_sums[lane] : 0, ws = w;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned int y = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= o) ws += y;
            }
            warp_sums[lane] = ws - w;  // exclusive
        }
        __syncthreads();
        unsigned int excl = carry + warp_sums[warp] + (x - v);
        if (i < n) data[i] = excl;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
}

__global__ void sort_scatter_kernel(int n, const unsigned long long *keys_in, const int *vals_in, unsigned long long *keys_out, int *vals_out,
                                    int shift, const unsigned int *block_offsets, int num_blocks) {
    __shared__ unsigned int next_pos[256];  // next output slot of this block for each digit
    next_pos[threadIdx.x] = block_offsets[threadIdx.x * num_blocks + blockIdx.x];
    __syncthreads();
    int base = blockIdx.x * SORT_TILE;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = 0; k < SORT_ITEMS; k++) {
        int i = base + k * SORT_THREADS + threadIdx.x;
        bool valid = i < n;
        unsigned long long key = valid ? keys_in[i] : ~0ull;
        unsigned int digit = (unsigned int)(key >> shift) & 255u;
        // stable rank: warps take their turn in order, lanes in order inside a warp
        unsigned int peers = __match_any_sync(0xffffffffu, valid ? digit : 256u);
        unsigned int rank = __popc(peers & ((1u << lane) - 1u));
        int leader = __ffs(peers) - 1;
        unsigned int pos = 0;
        for (int w = 0; w < SORT_THREADS / 32; w++) {
            if (warp == w && valid) {
                unsigned int start = 0;
                if (lane == leader) {
                    start = next_pos[digit];
                    next_pos[digit] = start + __popc(peers);
                }
                start = __shfl_sync(peers, start, leader);
                pos = start + rank;
            }
            __syncthreads();
        }
        if (valid) {
            keys_out[pos] = key;
            vals_out[pos] = vals_in[i];
        }
    }
}

// `passes` passes of 8 bits starting at bit 0 over (keys, vals); every pass swaps the pointer pairs, so the result
// is in (keys, vals) as the caller sees them afterwards.  hist: 256 * sort_blocks_for(n) entries.
inline int sort_blocks_for(int n) { return (n + SORT_TILE - 1) / SORT_TILE; }
inline void radix_sort_pairs(int n, unsigned long long *&keys, int *&vals, unsigned long long *&keys2, int *&vals2, unsigned int *hist, int passes, cudaStream_t st) {
    const int sort_blocks = sort_blocks_for(n);
    for (int pass = 0; pass < passes; pass++) {
        const int shift = 8 * pass;
        sort_hist_kernel<<<sort_blocks, SORT_THREADS, 0, st>>>(n, keys, shift, hist, sort_blocks);
        sort_scan_kernel<<<1, 1024, 0, st>>>(hist, 256 * sort_blocks);
        sort_scatter_kernel<<<sort_blocks, SORT_THREADS, 0, st>>>(n, keys, vals, keys2, vals2, shift, hist, sort_blocks);
        unsigned long long *tk = keys; keys = keys2; keys2 = tk;
        int *tv = vals; vals = vals2; vals2 = tv;
    }
}


// rtb_launch.h -- kernel parameter blocks and the host-callable launchers each render variant
// (fast: -fmad=true, strict: -fmad=false) exports.  Shared between rtb_api.cu and rtb_render_*.cu.
#pragma once
#include "rtb_device.cuh"

namespace rtb {

struct RenderParams {
    DScene scene;
    DCamera cam;
    DRender rd;
    int width, height;
    int time_ms, frame_num;
    const float *prev;  // previous running mean (may alias out)
    float *out;
    PeerOut peers;      // fused tile exchange: every rank's image (n == 0: none)
    // screen-tile sharding: this launch renders rows y with (y / band_rows) % world == rank
    int shard_rank, shard_world, band_rows;
    int local_rows;                  // rows owned by this rank
    int refill;                      // SHADE phase starts when this many lanes of a warp wait (1..32)
    int walk_min;                   // NODE phase wants this many walking lanes (rtb_trace.cuh)
    unsigned int *work_counter;      // zeroed before launch
    unsigned long long *counters;    // 6 slots (RtbCounters order) or nullptr
};

struct IdsParams {
    DScene scene;
    DCamera cam;
    int width, height;
    int32_t *obj_id, *sub_id;
    float *t;
};

// ---- wavefront variant (rtb_wavefront.cuh) -----------------------------------------------------
#ifndef RTB_SHADE_THREADS
#define RTB_SHADE_THREADS 256   // block size of wf_shade (grid-stride over the live queue)
#endif
struct WfState {
    float4 *A, *B, *C, *O, *D, *H;
};

struct WfParams {
    DScene scene;
    DCamera cam;
    DRender rd;
    int width, height;
    int time_ms, frame_num;
    const float *prev;
    float *out;
    PeerOut peers;
    int shard_rank, shard_world, band_rows, local_rows;
    int slot_base, slot_count;  // the row group this launch works on: slots [slot_base, slot_base + slot_count)
    WfState st;
    unsigned char *alive;       // per slot: 1 while the pixel still has samples to trace
    const int *queue_in;        // live pixel slots of this iteration (nullptr = identity 0..n-1)
    int *queue_out;             // live slots of the next iteration
    const unsigned int *n_in;   // device counters
    unsigned int *n_out;        // wf_shade: appended to; wf_traverse: zeroed for the next wf_shade
    unsigned int *fetch;        // wf_traverse: next queue position to hand out (zeroed by wf_shade)
    int *tqueue;                // slots whose ray has to enter the tree (written by wf_shade, read by wf_traverse)
    unsigned int *tn, *tn_next; // its length (wf_shade appends, wf_traverse reads) / the counter wf_traverse zeroes for the next wf_shade
    int first;                  // wf_shade: 1 = no hit to consume yet (start of the frame)
    int refill, walk_min;
    unsigned long long *counters;
};

// ---- pooled persistent kernel (rtb_pool.cuh) -----------------------------------------------------
struct PoolParams {
    DScene scene;
    DCamera cam;
    DRender rd;
    int width, height;
    int time_ms, frame_num;
    const float *prev;
    float *out;
    PeerOut peers;
    int shard_rank, shard_world, band_rows, local_rows;
    int refill, walk_min;
    float4 *pool;                 // 6 planes of 64 float4 per warp of the grid
    unsigned int *work_counter;   // zeroed before launch
    unsigned long long *counters;
};

// ---- streaming kernel (rtb_stream.cuh) -------------------------------------------------------------
#ifndef RTB_STREAM_SHARDS
#define RTB_STREAM_SHARDS 8   // sub-queues per queue (a power of two)
#endif
struct StreamQueue {
    int *ring;            // RTB_STREAM_SHARDS rings of (mask + 1) entries: 0 = empty, else (slot + 1) | lap << 31
    unsigned int *head;   // per shard, 32 words apart
    unsigned int *tail;
    unsigned int mask;
};

struct StreamParams {
    DScene scene;
    DCamera cam;
    DRender rd;
    int width, height;
    int time_ms, frame_num;
    const float *prev;
    float *out;
    PeerOut peers;
    int shard_rank, shard_world, band_rows, local_rows;
    int refill, walk_min;
    WfState st;                  // six planes of num_slots float4
    unsigned int *pix;           // per slot: x | y << 16
    StreamQueue rq, hq;          // rays to traverse / hits to shade
    unsigned int *ctl;           // control words (rtb_stream.cuh), zeroed before launch
    int *sm_role;                // per %smid: 0 = not drawn yet, 1 = SHADE, 2 = TRAVERSE; zeroed before launch
    unsigned long long *stats;   // [0] shade busy lane-cycles, [1] shade lane-cycles, [2] traverse busy, [3] traverse total
    int num_slots;
    unsigned int total_pixels;
    int shade_sms, total_sms;
    unsigned long long *counters;
};

struct LaunchConfig {
    int blocks, threads;
};

#define RTB_DECLARE_VARIANT(ns)                                                                                     \
    namespace ns {                                                                                                  \
    cudaError_t launch_render(const RenderParams &p, bool brute, bool count, bool early_pop, LaunchConfig cfg, cudaStream_t st); \
    cudaError_t launch_primary_ids(const IdsParams &p, bool brute, cudaStream_t st);                                \
    cudaError_t render_occupancy(int threads, bool early_pop, int *blocks_per_sm);                                  \
    cudaError_t launch_wf_shade(const WfParams &p, bool brute, bool count, int blocks, cudaStream_t st);                     \
    cudaError_t launch_wf_traverse(const WfParams &p, bool brute, bool count, int blocks, cudaStream_t st);         \
    cudaError_t wf_occupancy(int *shade_blocks_per_sm, int *traverse_blocks_per_sm);                                \
    cudaError_t launch_wf_rebuild(int slot_base, int n_slots, const unsigned char *alive, int *block_count, int *queue, unsigned int *total, cudaStream_t st); \
    cudaError_t launch_pool(const PoolParams &p, bool brute, bool count, int blocks, cudaStream_t st);              \
    cudaError_t pool_occupancy(int *blocks_per_sm);                                                                 \
    int pool_block_threads();                                                                                       \
    cudaError_t launch_stream(const StreamParams &p, bool brute, bool count, int blocks, cudaStream_t st);          \
    cudaError_t stream_occupancy(int *blocks_per_sm);                                                               \
    int stream_block_threads();                                                                                     \
    }
RTB_DECLARE_VARIANT(fast)
RTB_DECLARE_VARIANT(strict)

}  // namespace rtb

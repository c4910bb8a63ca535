// rtb_ctx.h -- the context behind the opaque rtb_ctx of include/rtb.h, shared by the translation units of the
// launch layer (rtb_api.cu: scene, kernels, render entry points; rtb_comm.cu: multi-GPU exchange).
#pragma once
#include "../../include/rtb.h"
#include "rtb_launch.h"

#include <string>
#include <vector>

using namespace rtb;

std::string &rtb_create_error();  // message of a failed rtb_create (no context yet)

#define RTB_WF_MAX_LANES 8   // row groups of the wavefront kernels in flight at once, each on its own stream

struct rtb_ctx {
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;

    int width = 0, height = 0;
    bool have_cam = false, have_rd = false;
    RtbCamData cam{};
    RtbRenderData rd{};

    // host-side scene description (owned copies)
    std::vector<RtbObject> objects;
    std::vector<RtbMaterial> materials;
    std::vector<float> tri_points, tri_uv;
    std::vector<std::vector<float>> image_data;
    std::vector<RtbImage> images;
    bool committed = false;
    bool mesh_uv_given = false;      // some mesh came with texture points (else the device flatten writes zeros instead of uploading them)
    int opt_mega_early_pop = -1, tuned_mega_early_pop = 0;   // megakernel traversal variant (trav_run EARLY_POP): option (-1 = timed choice) / what the timed choice kept
    int opt_device_flatten = 2;      // meshes partitioned + flattened on the device: 0 never, 1 always, 2 = from 32768 triangles up

    // device scene
    DScene d{};
    std::vector<void *> scene_allocs;
    double build_ms = 0, commit_ms = 0;  // device time of the tree build / wall time of the whole commit
    double commit_partition_ms = 0, commit_flatten_ms = 0, commit_upload_ms = 0;  // wall-time parts of the last commit: reference-BVH partition (inside the flatten), whole flatten, uploads
    size_t scene_bytes = 0;
    int culled_tris = 0, tree_height = 0;

    // frame state
    float *d_image = nullptr;
    unsigned char *d_rgba = nullptr;  // display conversion buffer (rtb_display_rgba8)
    size_t rgba_pixels = 0;
    size_t image_floats = 0;
    int frame_num = 0;
    unsigned int *d_work = nullptr;
    unsigned long long *d_counters = nullptr;
    RtbCounters last_counters{};
    unsigned long long last_lane_stats[8] = {};  // traversal lane occupancy of the last counted render (wavefront)
    bool timing_pending = false;
    cudaEvent_t frame_done[4] = {nullptr, nullptr, nullptr, nullptr};  // end of the last frames (bounded run-ahead)
    long long frames_enqueued = 0;
    int opt_run_ahead = 0;
    cudaEvent_t pace_events[RTB_WF_MAX_LANES][64] = {};   // wavefront loop: per stream, one event every 8 iterations (bounded run-ahead)
    int opt_wf_pace = 0;

    // wavefront variant: per-pixel records, two queues, counters {n[0], n[1], fetch}
    float4 *d_wf_state = nullptr;
    int *d_wf_queue = nullptr;
    unsigned char *d_wf_alive = nullptr;
    int *d_wf_block_count = nullptr;
    unsigned int *d_wf_counts = nullptr;
    size_t wf_slots = 0;
    int wf_shade_per_sm = 0, wf_trav_per_sm = 0;
    // option "wf_timing": CUDA events around every launch of the next frame (diagnostics only)
    int opt_wf_timing = 0;
    int opt_wf_rebuild = 16;  // iterations between queue rebuilds (0 = never)
    int opt_wf_groups = 2, opt_wf_streams = 2;  // row groups taken through the frame one after the other / concurrent groups
    cudaStream_t lane_stream[RTB_WF_MAX_LANES] = {};   // [0] unused (lane 0 runs on `stream`); created at first use
    cudaEvent_t ev_fork = nullptr, ev_join[RTB_WF_MAX_LANES] = {};
    std::vector<cudaEvent_t> wf_events;
    size_t wf_events_used = 0;
    double wf_shade_ms = 0, wf_traverse_ms = 0, wf_total_ms = 0;

    // pooled persistent kernel: per-warp slot records (6 KB per warp of the grid)
    float4 *d_pool = nullptr;
    size_t pool_warps = 0;
    int pool_per_sm = 0;

    // streaming kernel: records of the slots in flight, two sharded ring queues, control words
    float4 *d_stream_state = nullptr;
    unsigned int *d_stream_pix = nullptr;
    int *d_stream_rings = nullptr;          // [2 queues][8 shards][cap]
    unsigned int *d_stream_ctl = nullptr;   // [ctl 8][head/tail: 2 queues x 2 x 8 shards x 32][sm_role 1024] + stats (4 x u64)
    size_t stream_slots = 0, stream_cap = 0;
    int stream_per_sm = 0;
    int opt_stream_slots = 2;               // slots in flight = SMs x 1024 x this (capped by the pixel count)
    int opt_stream_shade_sms = 0;           // 0 = balanced from the kernel's own utilisation counters
    int stream_shade_sms = 0;               // the split in use (0 = not chosen yet)
    double stream_util[2] = {0, 0};         // last frame: busy fraction of the SHADE / TRAVERSE lanes
    int stream_last_shade_sms = 0;          // split of the frame those counters belong to
    long long stream_frames_launched = 0;
    unsigned int *h_stream_error = nullptr; // pinned: the kernel's watchdog word of the last frame
    bool stream_error_check_pending = false;

    // options
    // 0 = persistent megakernel, 1 = wavefront, 3 = pooled persistent kernel, 2 = auto: the variants are timed on
    // the first frame after the scene / resolution / settings / shard change and the fastest is kept (bit-identical)
    int opt_kernel = 2;
    int auto_choice = -1;                  // -1 = not tuned yet
    double tune_ms[5] = {0, 0, 0, 0, 0};   // device ms of the tuning frame, by kernel id ([0] megakernel, [1] wavefront, [3] pooled, [4] streaming)
    int tuned_prims = -1;                  // primitive count the choice was tuned on
    // kernels the timed choice tries: bit k = kernel id k.  The pooled (3) and streaming (4) kernels are selectable but
    // not tried by default: measured slower than the wavefront pair on every workload
    int opt_tune_set = 2 | 1;
    float *d_tune_scratch = nullptr;       // scratch image of the tuning frames (kept: dynamic scenes re-tune)
    size_t tune_scratch_floats = 0;
    int opt_counters = 0, opt_traversal = 0, opt_strict = 0, opt_ref_quirks = 1, opt_large_split = 1, opt_max_large = 24, opt_refill = 24;
    // traversal policy knobs; 0 = automatic: megakernel walk_min 8; wavefront (walk_min, wf_refill) = (16, 8) or (12, 16),
    // whichever the tuning frame measures faster (closed rooms prefer the second, the 1 M-triangle grid the first)
    int opt_walk_min = 0, opt_wf_refill = 0;
    int tuned_walk_min = 16, tuned_refill = 8;
    int opt_builder = 1, opt_ploc_radius = 16;  // 0 = Karras LBVH, 1 = PLOC; take effect at the next commit
    int opt_threads = RTB_BLOCK_THREADS, opt_blocks_per_sm = 0;
    int shard_rank = 0, shard_world = 1, band_rows = 8;

    // multi-GPU exchange (rtb_comm.cu)
    void *comm = nullptr;            // ncclComm_t, or nullptr: the context is alone (or the caller runs the exchange)
    int comm_world = 1, comm_rank = 0;
    // 0 = the exchange is the caller's business (rtb_shard_pack / unpack, rtb_device_image), 1 = NCCL inside the render
    // calls (all-gather of the tile bands, or all-reduce of sample slices), 2 = tiles with the exchange fused into the
    // render kernels (peer stores) and an NCCL barrier at the end of the frame
    int opt_exchange = 0;
    int opt_slices = 0;              // 1 = sample slices: every rank renders whole frames with its own seeds
    float *d_dense = nullptr;        // [my bands | world x bands] for the NCCL gather
    size_t exch_floats = 0;
    int *d_barrier = nullptr;
    float *peer_image[RTB_MAX_PEERS] = {};  // every rank's accumulation image mapped into this context (fused exchange)
    int n_peers = 0;
    int peer_self = -1;              // index of this context's own image in peer_image (not an IPC mapping: never closed)
    bool peers_ipc = false;          // mappings came from CUDA IPC handles (one process per GPU): closed at destroy
    int peer_access[RTB_MAX_PEERS] = {};    // rtb_create_multi: this device can address device k directly
};

// internal entry points of rtb_comm.cu / rtb_api.cu
int rtb_exchange_frame(rtb_ctx *c);
int rtb_exchange_after_local_write(rtb_ctx *c);
void rtb_comm_release(rtb_ctx *c);
int rtb_exchange_frame_grouped(rtb_ctx **ctxs, int n);
int rtb_exchange_after_local_write_grouped(rtb_ctx **ctxs, int n, int i);
int rtb_set_accumulation_async(rtb_ctx *c, const float *rgb_host, int frame_num);
int rtb_render_frames_no_exchange(rtb_ctx *c, int n_frames, int time0_ms, int time_step_ms);

inline int fail(rtb_ctx *ctx, const char *where, const char *msg) {
    std::string s = std::string("Error from CUDA (") + where + "): " + msg;  // check_cuda_error shape, utils.cu:5-10
    if (ctx) ctx->err = s; else rtb_create_error() = s;
    return 1;
}

#define RTB_CUDA(ctx, call, where)                                           \
    do {                                                                     \
        cudaError_t e_ = (call);                                             \
        if (e_ != cudaSuccess) return fail(ctx, where, cudaGetErrorString(e_)); \
    } while (0)

struct Bind {  // make ctx's device current for the duration of a call, then give the caller's device back
    int prev = -1, dev = -1;
    explicit Bind(rtb_ctx *c) : dev(c->device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) cudaSetDevice(dev);
    }
    ~Bind() {
        if (prev >= 0 && prev != dev) cudaSetDevice(prev);
    }
    Bind(const Bind &) = delete;
    Bind &operator=(const Bind &) = delete;
};


// rtb_comm.cu -- the multi-GPU exchange behind the C ABI (include/rtb.h, "multi-GPU" section).
//
// The reference is single-device (src/dispatch.cu:127-153 launches on the current device and nothing else); the
// path shards with no data-path collective (pixels are independent, src/raytracer.cu:127) and ONE exchange per frame:
//   tiles (strong scaling)   rank r renders the row bands (y / band_rows) % world == r of the same frame; the image is
//                            completed either by NCCL (pack -> ncclAllGather -> unpack, option "exchange" = 1) or by the
//                            render kernels themselves (option "exchange" = 2): a finished pixel is stored into every
//                            GPU's accumulation image through peer mappings (NVLink / NVSwitch), so the transfer
//                            overlaps the rendering and the frame ends with a barrier instead of a 100 MB gather;
//   sample slices (weak)     every rank renders whole frames with its own seeds (frame k of rank r is seeded
//                            time0 + (k * world + r) * step), then ncclAllReduce(sum) of the accumulation buffer and a
//                            scale: the running mean over world x frames samples (option "slices" = 1).
// Two ways to get a communicator: one process per GPU (rtb_comm_unique_id on rank 0, the 128 bytes travel over the
// caller's own plumbing -- MPI, torch.distributed, a file -- then rtb_comm_init on every rank), or one process for
// all GPUs (rtb_create_multi: ncclCommInitAll).  NCCL is loaded at first use (dlopen), so librtb.so itself has no
// dependency on it and a process that already holds an NCCL (torch) shares that one.
#include "rtb_ctx.h"

#include <dlfcn.h>
#include <nccl.h>

#include <cstring>

namespace {

struct Nccl {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};

Nccl *nccl() {
    static Nccl n;
    if (n.handle || !n.error.empty()) return &n;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *name : names) {
        n.handle = dlopen(name, RTLD_NOW | RTLD_LOCAL);
        if (n.handle) break;
    }
    if (!n.handle) {
        const char *why = dlerror();  // a second call would return NULL: the first one clears the message
        n.error = std::string("cannot load NCCL (libnccl.so.2): ") + (why ? why : "not found");
        return &n;
    }
#define RTB_NCCL_SYM(field, symbol)                                              \
    n.field = reinterpret_cast<decltype(n.field)>(dlsym(n.handle, symbol));      \
    if (!n.field && n.error.empty()) n.error = std::string("NCCL symbol missing: ") + symbol;
    RTB_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    RTB_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    RTB_NCCL_SYM(CommInitAll, "ncclCommInitAll")
    RTB_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    RTB_NCCL_SYM(AllGather, "ncclAllGather")
    RTB_NCCL_SYM(AllReduce, "ncclAllReduce")
    RTB_NCCL_SYM(GroupStart, "ncclGroupStart")
    RTB_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    RTB_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef RTB_NCCL_SYM
    return &n;
}

#define RTB_NCCL(ctx, call, where)                                                 \
    do {                                                                           \
        ncclResult_t r_ = (call);                                                  \
        if (r_ != ncclSuccess) return fail(ctx, where, nccl()->GetErrorString(r_)); \
    } while (0)

int need_nccl(rtb_ctx *c, const char *where) {
    Nccl *n = nccl();
    if (!n->error.empty()) return fail(c, where, n->error.c_str());
    return 0;
}

__global__ void scale_kernel(float *v, size_t n, float s) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] *= s;
}

int ensure_exchange_buffers(rtb_ctx *c) {
    int padded = 0;
    rtb_shard_rows(c, nullptr, &padded);
    size_t per_rank = (size_t)padded * c->width * 3;
    size_t need = per_rank * (size_t)(c->shard_world + 1);
    if (c->exch_floats < need) {
        if (c->d_dense) cudaFree(c->d_dense);
        c->d_dense = nullptr; c->exch_floats = 0;
        RTB_CUDA(c, cudaMalloc((void **)&c->d_dense, need * sizeof(float)), "allocating exchange buffers");
        c->exch_floats = need;
    }
    return 0;
}

// a barrier of the job on the context's stream: nothing after it starts before every rank got here
int comm_barrier(rtb_ctx *c) {
    if (!c->comm) return 0;
    if (!c->d_barrier) {
        RTB_CUDA(c, cudaMalloc((void **)&c->d_barrier, sizeof(int)), "allocating barrier word");
        RTB_CUDA(c, cudaMemsetAsync(c->d_barrier, 0, sizeof(int), c->stream), "allocating barrier word");
    }
    RTB_NCCL(c, nccl()->AllReduce(c->d_barrier, c->d_barrier, 1, ncclInt, ncclSum, (ncclComm_t)c->comm, c->stream), "barrier");
    return 0;
}

void close_peers(rtb_ctx *c) {
    if (c->peers_ipc)
        for (int r = 0; r < c->n_peers; r++)
            if (r != c->peer_self && c->peer_image[r]) cudaIpcCloseMemHandle(c->peer_image[r]);
    for (int r = 0; r < RTB_MAX_PEERS; r++) c->peer_image[r] = nullptr;
    c->n_peers = 0;
    c->peers_ipc = false;
    c->peer_self = -1;
}

}  // namespace

// ---- used by rtb_api.cu ------------------------------------------------------------------------------------------
// the exchange step that closes a frame (on the context's stream); 0 launches when the context is alone
int rtb_exchange_frame(rtb_ctx *c) {
    if (!c->comm || c->comm_world <= 1 || c->opt_exchange == 0) return 0;
    Nccl *n = nccl();
    if (c->opt_slices) {  // sum of the ranks' running means -> mean over world x spp samples
        RTB_NCCL(c, n->AllReduce(c->d_image, c->d_image, c->image_floats, ncclFloat, ncclSum, (ncclComm_t)c->comm, c->stream), "all-reduce of the accumulation buffer");
        scale_kernel<<<(unsigned int)((c->image_floats + 255) / 256), 256, 0, c->stream>>>(c->d_image, c->image_floats, 1.0f / (float)c->comm_world);
        RTB_CUDA(c, cudaGetLastError(), "scaling the accumulation buffer");
        return 0;
    }
    if (c->shard_world != c->comm_world || c->shard_rank != c->comm_rank)
        return fail(c, "exchange", "the tile shard (rtb_set_shard) does not match the communicator's rank / size");
    if (c->opt_exchange == 2 && c->n_peers == c->comm_world) return comm_barrier(c);  // the kernels stored every pixel everywhere
    if (ensure_exchange_buffers(c) != 0) return 1;
    int padded = 0;
    rtb_shard_rows(c, nullptr, &padded);
    size_t per_rank = (size_t)padded * c->width * 3;
    float *dense = c->d_dense, *gathered = c->d_dense + per_rank;
    if (rtb_shard_pack(c, dense) != 0) return 1;
    RTB_NCCL(c, n->AllGather(dense, gathered, per_rank, ncclFloat, (ncclComm_t)c->comm, c->stream), "all-gather of the tile bands");
    return rtb_shard_unpack(c, gathered);
}

// after the accumulation image was changed locally (reset / upload) in fused mode: peers may only start storing into
// it once the change is done, and this rank may only start once the peers' images are ready as well
int rtb_exchange_after_local_write(rtb_ctx *c) {
    if (c->comm && c->comm_world > 1 && c->opt_exchange == 2 && c->n_peers == c->comm_world && !c->opt_slices) return comm_barrier(c);
    return 0;
}

void rtb_comm_release(rtb_ctx *c) {
    close_peers(c);
    if (c->comm && nccl()->CommDestroy) nccl()->CommDestroy((ncclComm_t)c->comm);
    c->comm = nullptr;
    c->comm_world = 1; c->comm_rank = 0;
    if (c->d_dense) cudaFree(c->d_dense);
    if (c->d_barrier) cudaFree(c->d_barrier);
    c->d_dense = nullptr; c->d_barrier = nullptr; c->exch_floats = 0;
}

extern "C" {

int rtb_comm_unique_id(unsigned char id_out[128]) {
    if (!id_out) return 1;
    if (need_nccl(nullptr, "rtb_comm_unique_id") != 0) return 1;
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    RTB_NCCL(nullptr, nccl()->GetUniqueId(&id), "rtb_comm_unique_id");
    std::memcpy(id_out, &id, 128);
    return 0;
}

int rtb_comm_init(rtb_ctx *c, const unsigned char id[128], int world, int rank) {
    if (!c || !id) return 1;
    if (world < 1 || world > RTB_MAX_PEERS || rank < 0 || rank >= world) return fail(c, "rtb_comm_init", "bad world size or rank (at most 8 GPUs)");
    if (need_nccl(c, "rtb_comm_init") != 0) return 1;
    Bind b(c);
    rtb_comm_release(c);
    ncclUniqueId nid;
    std::memcpy(&nid, id, 128);
    ncclComm_t comm = nullptr;
    RTB_NCCL(c, nccl()->CommInitRank(&comm, world, nid, rank), "rtb_comm_init");
    c->comm = comm;
    c->comm_world = world; c->comm_rank = rank;
    if (c->opt_exchange == 0) c->opt_exchange = 1;
    return 0;
}

int rtb_comm_info(rtb_ctx *c, int *world, int *rank) {
    if (!c) return 1;
    if (world) *world = c->comm ? c->comm_world : 1;
    if (rank) *rank = c->comm ? c->comm_rank : 0;
    return 0;
}

// ---- peer mappings of the accumulation images (fused exchange) ---------------------------------------------------
int rtb_peer_export(rtb_ctx *c, unsigned char handle_out[64]) {
    if (!c || !handle_out) return 1;
    Bind b(c);
    if (!rtb_device_image(c)) return 1;
    cudaIpcMemHandle_t h;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    RTB_CUDA(c, cudaIpcGetMemHandle(&h, c->d_image), "exporting the accumulation image");
    std::memcpy(handle_out, &h, 64);
    return 0;
}

int rtb_peer_open(rtb_ctx *c, const unsigned char *handles, int world, int rank) {
    if (!c || !handles) return 1;
    if (world < 1 || world > RTB_MAX_PEERS || rank < 0 || rank >= world) return fail(c, "rtb_peer_open", "bad world size or rank (at most 8 GPUs)");
    Bind b(c);
    if (!rtb_device_image(c)) return 1;
    close_peers(c);
    c->peers_ipc = true;
    c->n_peers = world;
    c->peer_self = rank;
    for (int r = 0; r < world; r++) {
        if (r == rank) { c->peer_image[r] = c->d_image; continue; }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + 64 * (size_t)r, 64);
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            c->n_peers = r;  // close what was opened
            close_peers(c);
            return fail(c, "opening a peer's accumulation image", cudaGetErrorString(e));
        }
        c->peer_image[r] = (float *)p;
    }
    return 0;
}

int rtb_peer_close(rtb_ctx *c) {
    if (!c) return 1;
    Bind b(c);
    cudaStreamSynchronize(c->stream);
    close_peers(c);
    return 0;
}

// ---- one process, several GPUs -----------------------------------------------------------------------------------
int rtb_create_multi(rtb_ctx **out, const int *devices, int n) {
    if (!out || !devices) return fail(nullptr, "rtb_create_multi", "null argument");
    if (n < 1 || n > RTB_MAX_PEERS) return fail(nullptr, "rtb_create_multi", "between 1 and 8 devices");
    for (int i = 0; i < n; i++) out[i] = nullptr;
    for (int i = 0; i < n; i++) {
        if (rtb_create(&out[i], devices[i]) != 0) {
            for (int k = 0; k < i; k++) { rtb_destroy(out[k]); out[k] = nullptr; }
            return 1;
        }
    }
    if (n == 1) return 0;
    auto undo = [&](const char *where, const char *msg) {
        for (int k = 0; k < n; k++) { rtb_destroy(out[k]); out[k] = nullptr; }
        return fail(nullptr, where, msg);
    };
    if (need_nccl(nullptr, "rtb_create_multi") != 0) {
        std::string m = rtb_last_error(nullptr);
        for (int k = 0; k < n; k++) { rtb_destroy(out[k]); out[k] = nullptr; }
        rtb_create_error() = m;
        return 1;
    }
    std::vector<ncclComm_t> comms(n);
    ncclResult_t r = nccl()->CommInitAll(comms.data(), n, devices);
    if (r != ncclSuccess) return undo("rtb_create_multi", nccl()->GetErrorString(r));
    int prev = -1;
    cudaGetDevice(&prev);
    for (int i = 0; i < n; i++) {
        out[i]->comm = comms[i];
        out[i]->comm_world = n; out[i]->comm_rank = i;
        out[i]->opt_exchange = 1;
        out[i]->shard_rank = i; out[i]->shard_world = n;
        cudaSetDevice(devices[i]);
        for (int k = 0; k < n; k++) {
            if (k == i) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[i], devices[k]);
            if (can) {
                cudaError_t e = cudaDeviceEnablePeerAccess(devices[k], 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) can = 0;
                cudaGetLastError();
            }
            out[i]->peer_access[k] = can;
        }
    }
    if (prev >= 0) cudaSetDevice(prev);
    return 0;
}

// the images of all contexts exist (camera set): point every context at every image (same process: plain pointers)
static int multi_link(rtb_ctx **ctxs, int n) {
    for (int i = 0; i < n; i++)
        if (!rtb_device_image(ctxs[i])) return 1;
    bool all = true;
    for (int i = 0; i < n; i++)
        for (int k = 0; k < n; k++)
            if (i != k && !ctxs[i]->peer_access[k]) all = false;
    for (int i = 0; i < n; i++) {
        rtb_ctx *c = ctxs[i];
        for (int r = 0; r < RTB_MAX_PEERS; r++) c->peer_image[r] = nullptr;
        c->n_peers = 0;
        c->peers_ipc = false;
        if (!all) continue;  // no NVLink / P2P between some pair: the fused mode falls back to the NCCL gather
        for (int k = 0; k < n; k++) c->peer_image[k] = ctxs[k]->d_image;
        c->n_peers = n;
    }
    return 0;
}

// == render(VariableRenderData*, int) on n GPUs of this process: the scene, camera and settings were given to every
// context; rgb_inout (host) is the previous image in and the new running mean out.  Tiles or sample slices per the
// contexts' options; every context ends up with the whole image.
int rtb_render_multi(rtb_ctx **ctxs, int n, float *rgb_inout, int *frame_num_inout, int current_time_ms) {
    if (!ctxs || n < 1 || !rgb_inout || !frame_num_inout) return 1;
    if (n == 1) return rtb_render(ctxs[0], rgb_inout, frame_num_inout, current_time_ms);
    rtb_ctx *c0 = ctxs[0];
    for (int i = 0; i < n; i++) {
        if (!ctxs[i] || !ctxs[i]->comm || ctxs[i]->comm_world != n) return fail(c0, "rtb_render_multi", "contexts were not created by rtb_create_multi");
        if (ctxs[i]->width != c0->width || ctxs[i]->height != c0->height) return fail(c0, "rtb_render_multi", "contexts differ in resolution");
    }
    if (multi_link(ctxs, n) != 0) return 1;
    const int frame = *frame_num_inout;
    // upload the previous image to every GPU, then render + exchange; NCCL calls of one process are grouped
    for (int i = 0; i < n; i++)
        if (rtb_set_accumulation_async(ctxs[i], rgb_inout, frame) != 0) { c0->err = ctxs[i]->err; return 1; }
    for (int i = 0; i < n; i++) {
        int rc = rtb_exchange_after_local_write_grouped(ctxs, n, i);
        if (rc != 0) { c0->err = ctxs[i]->err; return 1; }
    }
    for (int i = 0; i < n; i++)
        if (rtb_render_frames_no_exchange(ctxs[i], 1, current_time_ms, 1) != 0) { c0->err = ctxs[i]->err; return 1; }
    if (rtb_exchange_frame_grouped(ctxs, n) != 0) return 1;
    if (rtb_download(c0, rgb_inout, frame_num_inout) != 0) return 1;
    for (int i = 1; i < n; i++)
        if (rtb_sync(ctxs[i]) != 0) { c0->err = ctxs[i]->err; return 1; }
    return 0;
}

void rtb_destroy_multi(rtb_ctx **ctxs, int n) {
    if (!ctxs) return;
    for (int i = 0; i < n; i++) {
        if (ctxs[i]) rtb_destroy(ctxs[i]);
        ctxs[i] = nullptr;
    }
}

}  // extern "C"

// single-process variants: the NCCL calls of all ranks must sit inside one group
int rtb_exchange_frame_grouped(rtb_ctx **ctxs, int n) {
    Nccl *nc = nccl();
    rtb_ctx *c0 = ctxs[0];
    const bool slices = c0->opt_slices != 0;
    const bool fused = !slices && c0->opt_exchange == 2 && c0->n_peers == n;
    std::vector<size_t> per_rank(n, 0);
    if (!slices && !fused) {
        for (int i = 0; i < n; i++) {
            Bind b(ctxs[i]);
            if (ensure_exchange_buffers(ctxs[i]) != 0) { c0->err = ctxs[i]->err; return 1; }
            int padded = 0;
            rtb_shard_rows(ctxs[i], nullptr, &padded);
            per_rank[i] = (size_t)padded * ctxs[i]->width * 3;
            if (rtb_shard_pack(ctxs[i], ctxs[i]->d_dense) != 0) { c0->err = ctxs[i]->err; return 1; }
        }
    }
    if (fused) {
        for (int i = 0; i < n; i++) {
            rtb_ctx *c = ctxs[i];
            Bind b(c);
            if (!c->d_barrier) {
                RTB_CUDA(c0, cudaMalloc((void **)&c->d_barrier, sizeof(int)), "allocating barrier word");
                RTB_CUDA(c0, cudaMemsetAsync(c->d_barrier, 0, sizeof(int), c->stream), "allocating barrier word");
            }
        }
    }
    RTB_NCCL(c0, nc->GroupStart(), "exchange");
    for (int i = 0; i < n; i++) {
        rtb_ctx *c = ctxs[i];
        ncclResult_t r;
        if (slices) r = nc->AllReduce(c->d_image, c->d_image, c->image_floats, ncclFloat, ncclSum, (ncclComm_t)c->comm, c->stream);
        else if (fused) r = nc->AllReduce(c->d_barrier, c->d_barrier, 1, ncclInt, ncclSum, (ncclComm_t)c->comm, c->stream);
        else r = nc->AllGather(c->d_dense, c->d_dense + per_rank[i], per_rank[i], ncclFloat, (ncclComm_t)c->comm, c->stream);
        if (r != ncclSuccess) { nc->GroupEnd(); return fail(c0, "exchange", nc->GetErrorString(r)); }
    }
    RTB_NCCL(c0, nc->GroupEnd(), "exchange");
    for (int i = 0; i < n; i++) {
        rtb_ctx *c = ctxs[i];
        Bind b(c);
        if (slices) {
            scale_kernel<<<(unsigned int)((c->image_floats + 255) / 256), 256, 0, c->stream>>>(c->d_image, c->image_floats, 1.0f / (float)n);
            RTB_CUDA(c0, cudaGetLastError(), "scaling the accumulation buffer");
        } else if (!fused) {
            if (rtb_shard_unpack(c, c->d_dense + per_rank[i]) != 0) { c0->err = c->err; return 1; }
        }
    }
    return 0;
}

int rtb_exchange_after_local_write_grouped(rtb_ctx **ctxs, int n, int i) {
    // fused mode only: one grouped barrier once every context has uploaded (called for i = 0 .. n-1; acts at the last)
    rtb_ctx *c0 = ctxs[0];
    if (i != n - 1) return 0;
    if (c0->opt_slices || c0->opt_exchange != 2 || c0->n_peers != n) return 0;
    Nccl *nc = nccl();
    for (int k = 0; k < n; k++) {
        rtb_ctx *c = ctxs[k];
        Bind b(c);
        if (!c->d_barrier) {
            RTB_CUDA(c0, cudaMalloc((void **)&c->d_barrier, sizeof(int)), "allocating barrier word");
            RTB_CUDA(c0, cudaMemsetAsync(c->d_barrier, 0, sizeof(int), c->stream), "allocating barrier word");
        }
    }
    RTB_NCCL(c0, nc->GroupStart(), "barrier");
    for (int k = 0; k < n; k++) {
        rtb_ctx *c = ctxs[k];
        ncclResult_t r = nc->AllReduce(c->d_barrier, c->d_barrier, 1, ncclInt, ncclSum, (ncclComm_t)c->comm, c->stream);
        if (r != ncclSuccess) { nc->GroupEnd(); return fail(c0, "barrier", nc->GetErrorString(r)); }
    }
    RTB_NCCL(c0, nc->GroupEnd(), "barrier");
    return 0;
}

// rtb_noalloc.cu -- rtb_primary_ids_noalloc: the primary-visibility pass for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid allocation).  No counterpart in the
// reference.
//
// rtb_primary_ids (rtb_api.cu) stages the kernel's three output planes {obj_id, sub_id, t} in device buffers that it
// cudaMallocs and frees on every call.  This variant runs the SAME kernel (primary_ids_kernel, rtb_render.cuh) with
// the planes in a file-scope static device array of RTB_NOALLOC_MAX_PIXELS entries per plane (24 MiB in all, held
// by every GPU that loads the library until the process exits), then copies them to the caller's buffers.  It calls
// no cudaMalloc / cudaFree and no allocating library routine.
//
// The array's name is a host-side stand-in in host code, not a device address: the launch layer gets the address
// with cudaGetSymbolAddress, once per device, with that device current, and hands THAT pointer to the kernel.
//
// Every call on a device shares the array, and the array keeps the last call's contents.  Nothing in it has to
// start at zero: the kernel writes entry i of all three planes for every pixel i < width * height, and only those
// entries are copied out, so no stale entry can reach the caller.  Two calls that overlap on one GPU (two contexts
// of the same device -- their streams are independent -- or two host threads) race on the array: serialise them.
#include "rtb_ctx.h"

#include <cstring>
#include <mutex>

#define RTB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) int32_t g_noalloc_planes[3 * (size_t)RTB_NOALLOC_MAX_PIXELS];

namespace {
std::mutex g_noalloc_mutex;
int32_t *g_noalloc_addr[RTB_NOALLOC_MAX_DEVICES] = {};  // device address of g_noalloc_planes on device k (0 = not asked yet)

// the array's address on the CURRENT device (the caller holds a Bind on c->device)
cudaError_t noalloc_planes(int device, int32_t **out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_addr[device]) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_planes);
        if (e != cudaSuccess) return e;
        if (!ptr) return cudaErrorInvalidSymbol;
        g_noalloc_addr[device] = (int32_t *)ptr;
    }
    *out = g_noalloc_addr[device];
    return cudaSuccess;
}
}  // namespace

extern "C" int rtb_noalloc_max_pixels(void) { return RTB_NOALLOC_MAX_PIXELS; }

extern "C" int rtb_primary_ids_noalloc(rtb_ctx *c, int32_t *obj_id, int32_t *sub_id, float *t) {
    if (!c || !obj_id || !sub_id) return 1;
    Bind b(c);
    if (!c->have_cam) return fail(c, "rtb_primary_ids_noalloc", "camera not set");
    if (!c->committed) return fail(c, "rtb_primary_ids_noalloc", "scene not committed");
    if (c->device < 0 || c->device >= RTB_NOALLOC_MAX_DEVICES) return fail(c, "rtb_primary_ids_noalloc", "device ordinal beyond the table of static scratch addresses (64)");
    const size_t n = (size_t)c->width * c->height;
    if (n > (size_t)RTB_NOALLOC_MAX_PIXELS) {
        std::string msg = "image of " + std::to_string(c->width) + " x " + std::to_string(c->height) + " = " + std::to_string(n) +
                          " pixels exceeds the static scratch of " + std::to_string((long long)RTB_NOALLOC_MAX_PIXELS) +
                          " pixels (RTB_NOALLOC_MAX_PIXELS): use rtb_primary_ids";
        return fail(c, "rtb_primary_ids_noalloc", msg.c_str());
    }
    if (n == 0) return 0;
    int32_t *planes = nullptr;
    RTB_CUDA(c, noalloc_planes(c->device, &planes), "address of the static scratch");
    IdsParams p;
    p.scene = c->d;
    std::memcpy(&p.cam, &c->cam, sizeof(DCamera));
    p.width = c->width; p.height = c->height;
    p.obj_id = planes;
    p.sub_id = planes + (size_t)RTB_NOALLOC_MAX_PIXELS;
    p.t = (float *)(planes + 2 * (size_t)RTB_NOALLOC_MAX_PIXELS);
    cudaError_t e = c->opt_strict ? strict::launch_primary_ids(p, c->opt_traversal == 1, c->stream) : fast::launch_primary_ids(p, c->opt_traversal == 1, c->stream);
    // cudaMemcpyDefault: the caller's buffers may be host memory (as for rtb_primary_ids) or device memory
    if (e == cudaSuccess) e = cudaMemcpyAsync(obj_id, p.obj_id, n * 4, cudaMemcpyDefault, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sub_id, p.sub_id, n * 4, cudaMemcpyDefault, c->stream);
    if (e == cudaSuccess && t) e = cudaMemcpyAsync(t, p.t, n * 4, cudaMemcpyDefault, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return fail(c, "rtb_primary_ids_noalloc", cudaGetErrorString(e));
    return 0;
}

/* rtb.h -- C ABI of the B200-native path tracer (librtb.so).
 *
 * Drop-in boundary for the per-pixel path-tracing hot path of Ben-Edwards44/Ray-Tracer
 * (SURVEY.md 8(b)).  The reference has no plugin interface: its seam is src/dispatch.cu, i.e. the
 * host functions Camera::assign_constant_mem(), allocate_constant_mem(RenderData, AllObjects) and
 * render(VariableRenderData*, int).  Every entry point below cites the reference interface it
 * replaces.  Plain pointers and sizes only; no exceptions cross this boundary (every function
 * returns 0 on success, non-zero on error, message via rtb_last_error).  There is NO CPU fallback:
 * rtb_create fails when no CUDA device is usable.
 *
 * The C++ header ray-tracer_b200/host/rtb_compat.hpp re-exposes the reference's own names
 * (Vec3, Texture, Material, Object, ObjFileMesh, Camera, RenderData, VariableRenderData, render)
 * on top of this ABI, so scene code written against the reference compiles unchanged.
 */
#ifndef RTB_H
#define RTB_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB_ABI_VERSION 1

typedef struct RtbVec3 { float x, y, z; } RtbVec3;

/* == DeviceCamData, src/camera.cu:12-18 (same 48-byte layout) */
typedef struct RtbCamData {
    RtbVec3 cam_pos;
    RtbVec3 tl_pixel_pos;
    RtbVec3 delta_u;
    RtbVec3 delta_v;
} RtbCamData;

/* == RenderData, src/raytracer.cu:4-12 (bool antialias widened to int) */
typedef struct RtbRenderData {
    int rays_per_pixel;
    int reflection_limit;
    int antialias;
    RtbVec3 sky_colour;
} RtbRenderData;

/* Texture::type, src/material.cu:7-10 */
enum { RTB_TEX_COLOUR = 0, RTB_TEX_GRADIENT = 1, RTB_TEX_CHECKERBOARD = 2, RTB_TEX_IMAGE = 3 };

/* == Texture, src/material.cu:4-125 (image pixels referenced by handle instead of device pointer) */
typedef struct RtbTexture {
    int type;
    RtbVec3 colour;      /* COLOUR */
    RtbVec3 light;       /* CHECKERBOARD */
    RtbVec3 dark;
    int num_squares;
    int img_width;       /* IMAGE */
    int img_height;
    int image;           /* handle returned by rtb_scene_add_image */
} RtbTexture;

/* Material::type, src/material.cu:131-133 */
enum { RTB_MAT_STANDARD = 0, RTB_MAT_EMISSIVE = 1, RTB_MAT_REFRACTIVE = 2 };

/* == Material, src/material.cu:128-186 */
typedef struct RtbMaterial {
    int type;
    RtbTexture texture;
    float smoothness;
    int need_uv;
    RtbVec3 emitted_light;
    float refractive_index;
} RtbMaterial;

/* Object::type, src/objects.cu:804-809 */
enum { RTB_OBJ_SPHERE = 0, RTB_OBJ_TRIANGLE = 1, RTB_OBJ_QUAD = 2, RTB_OBJ_ONE_WAY_QUAD = 3, RTB_OBJ_CUBOID = 4, RTB_OBJ_MESH = 5 };

/* One scene object in creation order (order defines the `<=` tie rule of src/raytracer.cu:36).
 *   SPHERE        p[0] = center, f[0] = radius                          (Object::create_sphere,  objects.cu:845)
 *   TRIANGLE      p[0..2] = points, uv[0..5] = texture points            (Object::create_triangle, :854/:863)
 *   QUAD          p[0..3] = point1..point4                               (Object::create_quad,    :872)
 *   ONE_WAY_QUAD  p[0..3], invert_normal                                 (Object::create_one_way_quad, :881)
 *   CUBOID        p[0] = tl_near_pos, f[0..2] = width, height, depth     (Object::create_cuboid,  :890)
 *   MESH          tri_first/tri_count into RtbSceneDesc.mesh_tri_*       (Object::create_mesh,    :899)
 */
typedef struct RtbObject {
    int type;
    int material;        /* index into RtbSceneDesc.materials */
    RtbVec3 p[4];
    float f[3];
    float uv[6];
    int invert_normal;
    int tri_first;
    int tri_count;
} RtbObject;

typedef struct RtbImage {
    int width, height;
    const float *rgb;    /* width*height*3, row-major, as ImageTexture holds it (src/main.cu:40-91) */
} RtbImage;

/* Read-only view of a whole scene (what AllObjects + the per-mesh Triangle arrays hold in the
 * reference, src/objects.cu:910-913, src/main.cu:290-295). */
typedef struct RtbSceneDesc {
    int num_objects;
    const RtbObject *objects;
    int num_materials;
    const RtbMaterial *materials;
    int num_mesh_tris;
    const float *mesh_tri_points;   /* 9 floats per triangle: p0 p1 p2 */
    const float *mesh_tri_uv;       /* 6 floats per triangle (Triangle::texture_points) or NULL */
    int num_images;
    const RtbImage *images;
} RtbSceneDesc;

/* Work counters of the last counted render (the model is BVH::debug_stats, src/objects.cu:534-584) */
typedef struct RtbCounters {
    unsigned long long segments;     /* get_ray_collision-equivalent calls = "rays" of the metric */
    unsigned long long hit_segments;
    unsigned long long box_tests;    /* child-box slab tests */
    unsigned long long tri_tests;
    unsigned long long sphere_tests;
    unsigned long long samples;      /* trace_ray-equivalent calls */
} RtbCounters;

typedef struct rtb_ctx rtb_ctx;

/* ---- lifetime -------------------------------------------------------------------------- */
int rtb_abi_version(void);
/* 1 when the library was built with -DRTB_CHECKS (device-side asserts on every queue / slot / stack / node index:
 * the checked build the GPU tests run once, ray-tracer_b200/librtb_checks.so), else 0 */
int rtb_build_has_checks(void);
/* one context per CUDA device (one process per GPU); fails if the device cannot be used */
int rtb_create(rtb_ctx **out, int device);
void rtb_destroy(rtb_ctx *ctx);
/* message of the last failed call on ctx (ctx may be NULL for rtb_create failures);
 * shape "Error from CUDA (<where>): <msg>" as check_cuda_error makes it, src/utils.cu:5-10 */
const char *rtb_last_error(rtb_ctx *ctx);

/* ---- camera and settings ----------------------------------------------------------------- */
/* replaces Camera::assign_constant_mem's cudaMemcpyToSymbol(const_cam_data) (src/camera.cu:59)
 * and the compile-time SCREEN_WIDTH/SCREEN_HEIGHT (src/camera.cu:4-5) */
int rtb_set_camera(rtb_ctx *ctx, const RtbCamData *cam, int width, int height);
/* replaces cudaMemcpyToSymbol(const_render_data), src/dispatch.cu:106 */
int rtb_set_render_data(rtb_ctx *ctx, const RtbRenderData *rd);

/* ---- scene ------------------------------------------------------------------------------- */
int rtb_scene_begin(rtb_ctx *ctx);
/* these return the new index/handle (>= 0) or a negative value on error */
int rtb_scene_add_material(rtb_ctx *ctx, const RtbMaterial *mat);
int rtb_scene_add_image(rtb_ctx *ctx, const float *rgb, int width, int height);          /* Texture::create_image, material.cu:42 */
int rtb_scene_add_sphere(rtb_ctx *ctx, const float center[3], float radius, int material);
int rtb_scene_add_triangle(rtb_ctx *ctx, const float points[9], const float uv[6], int material);
int rtb_scene_add_quad(rtb_ctx *ctx, const float points[12], int material);
int rtb_scene_add_one_way_quad(rtb_ctx *ctx, const float points[12], int invert_normal, int material);
int rtb_scene_add_cuboid(rtb_ctx *ctx, const float tl_near[3], float width, float height, float depth, int material);
int rtb_scene_add_mesh(rtb_ctx *ctx, const float *tri_points, const float *tri_uv, int num_tris, int material);
/* flatten, upload and build the LBVH on the GPU.  Replaces cudaMemcpyToSymbol(const_objects)
 * (src/dispatch.cu:107), ReadOnlyDeviceArray<Object>/<Triangle> (src/main.cu:145,292) and the
 * host BVH constructor (src/objects.cu:452-464). */
int rtb_scene_commit(rtb_ctx *ctx);
/* Dynamic scenes (the reference builds its scene once, src/main.cu:290-295, and
 * rebuilds every host BVH from scratch if it changes).  The context keeps the host description after a
 * commit: replace the parameters of object `index` (non-mesh: any of the five simple types; mesh: only
 * the material) or the triangles of a mesh (same count), then call rtb_scene_commit again: flatten +
 * upload + GPU tree rebuild ("commit_ms" / "build_ms" of rtb_get_info report the cost).  Object order,
 * hence the tie rule, is unchanged. */
int rtb_scene_update_object(rtb_ctx *ctx, int index, const RtbObject *obj);
int rtb_scene_update_mesh(rtb_ctx *ctx, int index, const float *tri_points, const float *tri_uv);
/* rtb_scene_begin + one add per object of `desc` (same order) + rtb_scene_commit, in one call */
int rtb_scene_set(rtb_ctx *ctx, const RtbSceneDesc *desc);
/* view of the host-side description held by ctx (valid until the next rtb_scene_begin/destroy) */
int rtb_scene_desc(rtb_ctx *ctx, RtbSceneDesc *out);

/* ---- render ------------------------------------------------------------------------------ */
/* == render(VariableRenderData *data, int current_time_ms), src/dispatch.cu:156-163.
 * rgb_inout: width*height*3 floats on the HOST = VariableRenderData::previous_render (read as the
 * previous frame, overwritten with the new running mean); *frame_num_inout is incremented. */
int rtb_render(rtb_ctx *ctx, float *rgb_inout, int *frame_num_inout, int current_time_ms);

/* Device-resident progressive rendering: n_frames frames accumulated on the
 * device starting at the context's frame counter, frame k seeded with time0_ms + k*time_step_ms.
 * Asynchronous on the context's stream. */
int rtb_render_device(rtb_ctx *ctx, int n_frames, int time0_ms, int time_step_ms);
int rtb_sync(rtb_ctx *ctx);
int rtb_reset_accumulation(rtb_ctx *ctx);                                   /* frame counter = 0, image = 0 */
int rtb_set_accumulation(rtb_ctx *ctx, const float *rgb_host, int frame_num); /* resume (checkpoint) */
int rtb_download(rtb_ctx *ctx, float *rgb_host, int *frame_num_out);
/* Display conversion of the accumulation image: == parse_pixel_colours,
 * src/main.cu:343-371 (colour = (int)(value * 255) clamped to [0, 255], alpha 255), as one kernel on
 * the device; rgba_host receives width*height*4 bytes.  No window: the bytes are what the reference
 * hands to sf::Texture::update (src/main.cu:380). */
int rtb_display_rgba8(rtb_ctx *ctx, uint8_t *rgba_host);
/* the same bytes without alpha as a binary PPM (P6) file: the headless stand-in for the window */
int rtb_write_ppm(rtb_ctx *ctx, const char *path);
/* device address of the accumulation image (width*height*3 floats) and the context's cudaStream_t,
 * for callers that run collectives on it (torch.distributed / NCCL) */
void *rtb_device_image(rtb_ctx *ctx);
void *rtb_stream(rtb_ctx *ctx);
/* milliseconds of device time (CUDA events on the context's stream) of the last rtb_render /
 * rtb_render_device call; blocks until that work is complete */
int rtb_last_device_ms(rtb_ctx *ctx, float *ms_out);

/* Screen-tile sharding (SURVEY.md 8(e) partitioning A): this context renders only the pixel rows y
 * with (y / band_rows) % world == rank; other rows of the image are left untouched.  world = 1
 * restores full-frame rendering. */
int rtb_set_shard(rtb_ctx *ctx, int rank, int world, int band_rows);
/* pack this rank's rows into a dense buffer / scatter a gathered [world][rows_per_rank] buffer back
 * into the full image (device pointers; rows_per_rank = rtb_shard_rows(ctx) padded to the max over ranks) */
int rtb_shard_rows(rtb_ctx *ctx, int *rows_this_rank, int *rows_padded);
/* the index arithmetic of the band sharding, usable without a context or a device: the image row of
 * rank's local row (may be >= height in the padded tail), and how many rows of a height-row image
 * rank owns */
int rtb_shard_row_of(int local_row, int rank, int world, int band_rows);
int rtb_shard_row_count(int height, int rank, int world, int band_rows);
int rtb_shard_pack(rtb_ctx *ctx, void *dense_dev);
int rtb_shard_unpack(rtb_ctx *ctx, const void *gathered_dev);

/* ---- multi-GPU behind the C ABI -------------------------------------------------
 * The reference is single-device (src/dispatch.cu:127-153); its render(VariableRenderData*, int) has no
 * counterpart for several GPUs, so these entry points are the extension a maintainer binds next to
 * render().  NCCL (libnccl.so.2) is loaded at first use.  Two ways to a communicator:
 *   one process per GPU:  rank 0 calls rtb_comm_unique_id, the 128 bytes reach the other ranks through the
 *                         caller's own launcher (MPI, torchrun, a file), every rank calls rtb_comm_init;
 *   one process, n GPUs:  rtb_create_multi (ncclCommInitAll), then rtb_render_multi.
 * With a communicator, rtb_set_shard(rank, world, band_rows) and option "exchange" != 0, rtb_render /
 * rtb_render_device complete the image on every rank themselves, on the context's stream:
 *   "exchange" = 1  tiles: pack -> ncclAllGather -> unpack;  with option "slices" = 1 (sample slices: no shard,
 *                   frame k of rank r seeded time0 + (k * world + r) * step): ncclAllReduce(sum) + scale;
 *   "exchange" = 2  tiles with the exchange fused into the render kernels: every finished pixel is stored into
 *                   all ranks' accumulation images through peer mappings (rtb_peer_export / rtb_peer_open, or
 *                   direct peer access under rtb_create_multi), the frame ends with a barrier instead of a gather.
 * At most 8 ranks (one NVSwitch domain). */
int rtb_comm_unique_id(unsigned char id_out[128]);
int rtb_comm_init(rtb_ctx *ctx, const unsigned char id[128], int world, int rank);
int rtb_comm_info(rtb_ctx *ctx, int *world, int *rank);
/* CUDA IPC handle (64 bytes) of this context's accumulation image / map all ranks' images (world x 64 bytes,
 * rank order; the entry of `rank` itself is ignored).  The resolution must be set first (rtb_set_camera). */
int rtb_peer_export(rtb_ctx *ctx, unsigned char handle_out[64]);
int rtb_peer_open(rtb_ctx *ctx, const unsigned char *handles, int world, int rank);
int rtb_peer_close(rtb_ctx *ctx);
/* n contexts on n devices of this process, one communicator, shards preset to (i, n, 8 rows); peer access enabled
 * where the devices allow it.  Scene, camera and settings are given to every context. */
int rtb_create_multi(rtb_ctx **ctxs_out, const int *devices, int n);
/* == render(VariableRenderData*, int) on the n GPUs of rtb_create_multi: host image in / out like rtb_render */
int rtb_render_multi(rtb_ctx **ctxs, int n, float *rgb_inout, int *frame_num_inout, int current_time_ms);
void rtb_destroy_multi(rtb_ctx **ctxs, int n);

/* ---- introspection ----------------------------------------------------------------------- */
/* primary-hit ids with antialias off: obj_id = object index accepted by the
 * reference's `<=` rule or -1; sub_id = triangle inside it (mesh: triangle index; quad 0/1;
 * cuboid face*2+{0,1}; else 0); t = distance (INF = 2^30 on a miss).  HOST buffers, width*height. */
int rtb_primary_ids(rtb_ctx *ctx, int32_t *obj_id, int32_t *sub_id, float *t);
/* rtb_primary_ids for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid allocation).  No counterpart in the reference.
 * Same kernel, same arguments and results as rtb_primary_ids (buffers may be host or device memory; t may be
 * NULL), but the kernel's device-side staging planes live in a static device array of the library instead of
 * per-call cudaMalloc'ed buffers: the call allocates nothing.  Limit: width * height <= RTB_NOALLOC_MAX_PIXELS
 * (2 Mi pixels: 1920 x 1080 fits, 3840 x 2160 does not; 24 MiB held by every GPU that loads the library until the
 * process exits); a larger image is refused with an error that says so.  All calls on one GPU share the array:
 * calls that overlap on the GPU (two contexts of one device, two host threads) race -- serialise them. */
#define RTB_NOALLOC_MAX_PIXELS (2 * 1024 * 1024)
int rtb_primary_ids_noalloc(rtb_ctx *ctx, int32_t *obj_id, int32_t *sub_id, float *t);
int rtb_noalloc_max_pixels(void);
/* Test hook (no counterpart in the reference): the reference's BVH partition of one mesh (which triangles share which
 * box, src/objects.cu:602-770) computed by the host restatement (on_device = 0) or by the device version that
 * rtb_scene_commit uses for big meshes (on_device = 1; option "device_flatten").  tri_points: 9 floats per triangle.
 * Out: leaf_of[n] node index of each triangle's leaf, degenerate[n] 1 where that leaf's box is flat on an axis,
 * nodes[8 * nodes_cap] = {lo.xyz, parent, hi.xyz, 0} per node in the reference's preorder numbering. */
int rtb_debug_mesh_partition(int device, const float *tri_points, int n, int on_device, int32_t *leaf_of, uint8_t *degenerate, float *nodes,
                             int nodes_cap, int *num_nodes);
/* counters of the last render when option "counters" = 1 */
int rtb_counters(rtb_ctx *ctx, RtbCounters *out);
/* options: "counters" 0/1; "traversal" 0 = LBVH, 1 = brute force over the flattened primitives;
 * "kernel" 0 = persistent megakernel, 1 = wavefront pair, 3 = pooled persistent kernel (one launch per frame,
 * per-warp pools of 64 pixel slots), 4 = streaming kernel (one launch per frame, SMs specialised as shade /
 * traverse, ticket queues; "stream_slots", "stream_shade_sms" shape it), 2 = timed choice at the first frame
 * (default; bit-identical images; beyond the packing limits of kernels 1, 3, 4 -- reflection_limit > 255 or
 * rays_per_pixel >= 2^24 -- the choice is the megakernel, while pinning 1, 3 or 4 refuses such a frame);
 * "tune_set" (kernels the timed choice tries: 1 megakernel | 2 wavefront | 8 pooled | 16 streaming; default 3);
 * "mega_early_pop" -1 / 0 / 1 (the megakernel's traversal loop fetches the stack entries with the node: -1 = the timed
 * choice keeps the faster variant -- default; same image either way);
 * "strict_fp" 0/1 (kernels built with -fmad=false); "builder" 0 = Karras LBVH, 1 = PLOC (next commit; 2 = host
 * full-sweep SAH, a tree-quality yardstick for experiments only); "ploc_radius"; "device_flatten" 0 = meshes are
 * partitioned + flattened on the host, 1 = on the device, 2 = on the device from 32768 triangles up (default; next
 * commit; same arrays either way);
 * "large_split", "max_large", "reference_bvh_quirks" (next commit); "walk_min", "refill", "wf_refill",
 * "wf_rebuild", "wf_groups", "wf_streams" (traversal / wavefront policy, 0 = tuned; up to 8 row groups in flight); "block_threads";
 * "blocks_per_sm"; "wf_timing" (events around every wavefront launch of the next frame);
 * "run_ahead" (frames the host may enqueue ahead of the device, 0 = unbounded) and "wf_pace" (wavefront
 * iterations the host may run ahead, 0 = unbounded): host pacing knobs, measured without effect, off by default;
 * "exchange" 0/1/2 and "slices" 0/1 (multi-GPU section above).  A change of kernel-shaping options re-arms the
 * timed choice; a re-commit keeps it unless the primitive count changed by more than 2x.
 * info keys (rtb_get_info): "num_prims", "num_nodes", "tree_height", "build_ms", "commit_ms",
 * "commit_partition_ms", "commit_flatten_ms", "commit_upload_ms", "scene_bytes", "culled_tris", "kernel",
 * "tune_*_ms", "wf_*_ms", "exchange", "exchange_fused", "comm_world", "stream_*" (see rtb_api.cu). */
int rtb_set_option(rtb_ctx *ctx, const char *key, int value);
int rtb_get_info(rtb_ctx *ctx, const char *key, double *value);
/* sustained FP32 FMA throughput of the device (TFLOP/s, dependent-chain FFMA microbenchmark over
 * `ms_target` milliseconds): the measured denominator of the FP32 roofline */
int rtb_measure_fp32_peak(rtb_ctx *ctx, double ms_target, double *tflops_out);

#ifdef __cplusplus
}
#endif
#endif /* RTB_H */

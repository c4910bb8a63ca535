// BASELINE HARNESS (not product code): the reference's OWN CUDA source, unmodified except for the
// two resolution constants (src/camera.cu:4-5, patched by baseline/Makefile into a git-ignored
// copy under baseline/_ref/), built headless: everything below main.cu compiles without SFML
// (SURVEY.md F5).  Exposes extern "C" hooks so bench.py / tests can time the reference kernel
// get_pixel_colour (src/raytracer.cu:116) and its render() (src/dispatch.cu:156) on the same box.
// Scenes come from ray-tracer_b200/host/scenes.hpp compiled against the reference's classes.
#include <string>
#include <vector>
#include <cmath>
#include <cstdio>
#include <cstdint>
#include <fstream>
#include <stdexcept>
#include <random>
#include <chrono>

#define private public
#define protected public
#include "obj_read.cu"
#include "../ray-tracer_b200/host/scenes.hpp"
#undef private
#undef protected

namespace {
std::string g_err;
std::vector<Object> g_objects;
Object *g_dev_objects = nullptr;
float *g_prev = nullptr, *g_out = nullptr;
int *g_time = nullptr, *g_frame = nullptr;
cudaEvent_t g_e0, g_e1;
bool g_ready = false;
}

// primary-hit ids on the reference side: the 20 lines of get_ray_collision
// (src/raytracer.cu:24-46) recording the accepted object index; antialias off
__global__ void ref_primary_ids_kernel(int32_t *obj_id, float *t_out) {
    int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
    if (x >= SCREEN_WIDTH || y >= SCREEN_HEIGHT) return;
    uint rng_state = 0;
    Ray ray(x, y, &rng_state, false);
    int best = -1;
    float best_t = INF;
    for (int i = 0; i < const_objects.num_meshes; i++) {
        RayHitData hd = const_objects.meshes[i].hit(&ray);
        if (!hd.ray_hits) continue;
        if (hd.ray_travelled_dist <= best_t) { best = i; best_t = hd.ray_travelled_dist; }
    }
    obj_id[y * SCREEN_WIDTH + x] = best;
    t_out[y * SCREEN_WIDTH + x] = best_t;
}

// primary-hit sub ids: for the winning object, which of its triangles the reference's own
// Triangle::hit accepts -- quad 0/1 (t1 if it hits, else t2, src/objects.cu:229-233), cuboid face*2 + {0,1}
// (closest face by strict <, :315), mesh = brute force over its triangles in index order with strict <
__global__ void ref_primary_sub_ids_kernel(const int32_t *obj_id, int32_t *sub_id) {
    int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
    if (x >= SCREEN_WIDTH || y >= SCREEN_HEIGHT) return;
    int best = obj_id[y * SCREEN_WIDTH + x];
    int sub = -1;
    if (best >= 0) {
        uint rng_state = 0;
        Ray ray(x, y, &rng_state, false);
        Object &o = const_objects.meshes[best];
        sub = 0;
        if (o.type == Object::QUAD) {
            sub = o.quad.t1.hit(&ray).ray_hits ? 0 : 1;
        } else if (o.type == Object::ONE_WAY_QUAD) {
            sub = o.one_way_quad.t1.hit(&ray).ray_hits ? 0 : 1;
        } else if (o.type == Object::CUBOID) {
            float bt = INF;
            for (int f = 0; f < 6; f++) {
                RayHitData fh = o.cuboid.faces[f].hit(&ray);
                if (fh.ray_hits && fh.ray_travelled_dist < bt) {
                    bt = fh.ray_travelled_dist;
                    sub = f * 2 + (o.cuboid.faces[f].t1.hit(&ray).ray_hits ? 0 : 1);
                }
            }
        } else if (o.type == Object::MESH) {
            float bt = INF;
            sub = -1;
            for (int k = 0; k < o.mesh.num_triangles; k++) {
                RayHitData th = o.mesh.device_triangles[k].hit(&ray);
                if (th.ray_hits && th.ray_travelled_dist < bt) { bt = th.ray_travelled_dist; sub = k; }
            }
        }
    }
    sub_id[y * SCREEN_WIDTH + x] = sub;
}

// The Mrays/s denominator counted by the reference's own code on the GPU: the loops of get_ray_colour /
// trace_ray (src/raytracer.cu:67-105) with a segment and a hit counter, calling only reference functions.
// Never used for radiance.
__global__ void ref_count_segments_kernel(const int *current_time, unsigned long long *out) {
    int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
    if (x >= SCREEN_WIDTH || y >= SCREEN_HEIGHT) return;
    int array_index = (y * SCREEN_WIDTH + x) * 3;
    uint rng_state = array_index * 3145739 + *current_time * 6291469;
    Ray primary(x, y, &rng_state, const_render_data.antialias);
    unsigned long long segments = 0, hits = 0;
    for (int s = 0; s < const_render_data.rays_per_pixel; s++) {
        Ray ray = primary;
        Vec3 current_ray_colour(1, 1, 1);
        for (int b = 0; b < const_render_data.reflection_limit; b++) {
            ray.apply_antialias();
            RayCollision collision = get_ray_collision(&ray);
            segments++;
            if (!collision.hit_data.ray_hits) break;
            hits++;
            Material material = collision.hit_mesh_material;
            update_ray(&ray, &collision.hit_data, material);
        }
    }
    atomicAdd(out, segments);
    atomicAdd(out + 1, hits);
}

extern "C" {

const char *refc_last_error() { return g_err.c_str(); }
int refc_width() { return SCREEN_WIDTH; }
int refc_height() { return SCREEN_HEIGHT; }

int refc_init(int scene_id, int p0, int p1, const char *models_dir, int spp, int bounces, int antialias) {
    try {
        Camera camera;
        camera.assign_constant_mem();  // src/main.cu:389-392
        rtb_scenes::models_dir() = models_dir;
        rtb_scenes::Scene s = rtb_scenes::build(scene_id, p0, p1);
        g_objects = s.objects;
        ReadOnlyDeviceArray<Object> array(g_objects);  // src/main.cu:290-295
        g_dev_objects = array.device_pointer;
        AllObjects all{g_dev_objects, (int)g_objects.size()};
        RenderData rd = rtb_scenes::render_settings(s.use_sky, spp, bounces, antialias != 0);
        allocate_constant_mem(rd, all);  // src/dispatch.cu:104
        if (!g_ready) {
            check_cuda_error(cudaMalloc((void **)&g_prev, PIXEL_ARRAY_LEN * sizeof(float)), "harness buffers");
            check_cuda_error(cudaMalloc((void **)&g_out, PIXEL_ARRAY_LEN * sizeof(float)), "harness buffers");
            check_cuda_error(cudaMalloc((void **)&g_time, sizeof(int)), "harness buffers");
            check_cuda_error(cudaMalloc((void **)&g_frame, sizeof(int)), "harness buffers");
            cudaEventCreate(&g_e0);
            cudaEventCreate(&g_e1);
            g_ready = true;
        }
        check_cuda_error(cudaDeviceSynchronize(), "scene upload");
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

// the reference's own render() (per-frame malloc/upload/managed download included): wall-clock ms
int refc_render(float *rgb_inout, int *frame_num, int time_ms, double *wall_ms) {
    try {
        VariableRenderData data{*frame_num, std::vector<float>(rgb_inout, rgb_inout + PIXEL_ARRAY_LEN)};
        auto t0 = std::chrono::steady_clock::now();
        render(&data, time_ms);
        auto t1 = std::chrono::steady_clock::now();
        std::copy(data.previous_render.begin(), data.previous_render.end(), rgb_inout);
        *frame_num = data.frame_num;
        if (wall_ms) *wall_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

// the reference kernel alone, launch shape of src/dispatch.cu:136-139, device-timed with events
int refc_render_kernel(float *rgb_inout, int frame_num, int time_ms, float *kernel_ms) {
    try {
        check_cuda_error(cudaMemcpy(g_prev, rgb_inout, PIXEL_ARRAY_LEN * sizeof(float), cudaMemcpyHostToDevice), "prev upload");
        check_cuda_error(cudaMemcpy(g_time, &time_ms, sizeof(int), cudaMemcpyHostToDevice), "time upload");
        check_cuda_error(cudaMemcpy(g_frame, &frame_num, sizeof(int), cudaMemcpyHostToDevice), "frame upload");
        dim3 thread_dim(8, 8);
        dim3 block_dim = get_block_size(SCREEN_WIDTH, SCREEN_HEIGHT, thread_dim);
        cudaEventRecord(g_e0);
        get_pixel_colour<<<block_dim, thread_dim>>>(g_out, g_prev, g_time, g_frame);
        cudaEventRecord(g_e1);
        check_cuda_error(cudaDeviceSynchronize(), "reference kernel");
        check_cuda_error(cudaPeekAtLastError(), "reference kernel");
        if (kernel_ms) cudaEventElapsedTime(kernel_ms, g_e0, g_e1);
        check_cuda_error(cudaMemcpy(rgb_inout, g_out, PIXEL_ARRAY_LEN * sizeof(float), cudaMemcpyDeviceToHost), "image download");
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

int refc_primary_ids(int32_t *obj_id, float *t) {
    try {
        int32_t *d_obj;
        float *d_t;
        size_t n = (size_t)SCREEN_WIDTH * SCREEN_HEIGHT;
        check_cuda_error(cudaMalloc((void **)&d_obj, n * 4), "ids");
        check_cuda_error(cudaMalloc((void **)&d_t, n * 4), "ids");
        dim3 thread_dim(8, 8);
        dim3 block_dim = get_block_size(SCREEN_WIDTH, SCREEN_HEIGHT, thread_dim);
        ref_primary_ids_kernel<<<block_dim, thread_dim>>>(d_obj, d_t);
        check_cuda_error(cudaDeviceSynchronize(), "ids kernel");
        check_cuda_error(cudaMemcpy(obj_id, d_obj, n * 4, cudaMemcpyDeviceToHost), "ids");
        check_cuda_error(cudaMemcpy(t, d_t, n * 4, cudaMemcpyDeviceToHost), "ids");
        cudaFree(d_obj);
        cudaFree(d_t);
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

// like refc_primary_ids, plus the sub id of the winning primitive (brute force over a mesh's triangles: small scenes only)
int refc_primary_sub_ids(const int32_t *obj_id, int32_t *sub_id) {
    try {
        int32_t *d_obj, *d_sub;
        size_t n = (size_t)SCREEN_WIDTH * SCREEN_HEIGHT;
        check_cuda_error(cudaMalloc((void **)&d_obj, n * 4), "sub ids");
        check_cuda_error(cudaMalloc((void **)&d_sub, n * 4), "sub ids");
        check_cuda_error(cudaMemcpy(d_obj, obj_id, n * 4, cudaMemcpyHostToDevice), "sub ids");
        dim3 thread_dim(8, 8);
        dim3 block_dim = get_block_size(SCREEN_WIDTH, SCREEN_HEIGHT, thread_dim);
        ref_primary_sub_ids_kernel<<<block_dim, thread_dim>>>(d_obj, d_sub);
        check_cuda_error(cudaDeviceSynchronize(), "sub ids kernel");
        check_cuda_error(cudaMemcpy(sub_id, d_sub, n * 4, cudaMemcpyDeviceToHost), "sub ids");
        cudaFree(d_obj);
        cudaFree(d_sub);
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

// path segments ("rays") and hit segments of one frame at the current settings, counted by the reference's
// own functions on the GPU (the bench's Mrays/s denominator for the reference arm; untimed)
int refc_count_segments(int time_ms, unsigned long long *segments, unsigned long long *hits) {
    try {
        unsigned long long *d_out, h_out[2] = {0, 0};
        check_cuda_error(cudaMalloc((void **)&d_out, sizeof h_out), "segment counters");
        check_cuda_error(cudaMemcpy(d_out, h_out, sizeof h_out, cudaMemcpyHostToDevice), "segment counters");
        check_cuda_error(cudaMemcpy(g_time, &time_ms, sizeof(int), cudaMemcpyHostToDevice), "time upload");
        dim3 thread_dim(8, 8);
        dim3 block_dim = get_block_size(SCREEN_WIDTH, SCREEN_HEIGHT, thread_dim);
        ref_count_segments_kernel<<<block_dim, thread_dim>>>(g_time, d_out);
        check_cuda_error(cudaDeviceSynchronize(), "segment counting kernel");
        check_cuda_error(cudaMemcpy(h_out, d_out, sizeof h_out, cudaMemcpyDeviceToHost), "segment counters");
        cudaFree(d_out);
        *segments = h_out[0];
        *hits = h_out[1];
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

}  // extern "C"

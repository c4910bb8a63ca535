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

}  // extern "C"

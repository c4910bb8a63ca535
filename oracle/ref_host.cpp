// TEST INFRASTRUCTURE (oracle). Not part of the product path; nothing in the product links this.
//
// ref_host.cpp -- the reference's OWN device source (/root/reference/src/raytracer.cu and the
// files it #includes) compiled for the host through oracle/ref_shim/cuda_shim.h, plus extern "C"
// entry points so the tests can drive it with ctypes.  Built by oracle/Makefile into
// oracle/_ref/libref_host.so (git-ignored; travels to the GPU box).  This is SURVEY.md 8(c)'s
// "host scalar oracle = the reference source compiled for the CPU" and the thing that pins
// oracle/pt_oracle.c and the golden fixtures under tests/golden/.
//
// What is reference code here: everything reached through raytracer.cu (Ray, Object::hit, BVH,
// get_ray_collision, update_ray, trace_ray, get_ray_colour, get_pixel_colour, Camera, ...), the
// host-side pieces pulled in by LINE RANGE at build time into oracle/_ref/gen/ (never committed):
//   dispatch.cu:8-115   ReadOnlyDeviceArray & co, allocate_constant_mem, VariableRenderData
//   obj_read.cu:8-147   read_file, split_string, ObjFileMesh
//   main.cu:11-16,28-37,40-331   constants, host_rng, ImageTexture, SceneObjects, RenderSettings
// What is harness code: the pixel loop at a RUN-TIME resolution (the reference's resolution is
// a compile-time constant, src/camera.cu:4-5), which re-states only the ten lines of
// get_pixel_colour (src/raytracer.cu:116-136) and Camera::assign_constant_mem with W/H as
// variables; refh_render_kernel() runs the TRUE kernel function at the compiled 1000x800 so the
// two can be compared (tests/test_oracle_ref.py does).
#include "ref_shim/cuda_shim.h"
#include <thread>
#include <atomic>
#include <cstdint>
#include <unistd.h>

#define private public
#define protected public
#include "raytracer.cu"
#include "gen/dispatch_part.inc"
#include "gen/obj_read_part.inc"
#include "gen/main_part.inc"
namespace sf { typedef unsigned char Uint8; }  // the one SFML name parse_pixel_colours uses
#include "gen/display_part.inc"
#include "../ray-tracer_b200/host/scenes.hpp"
#undef private
#undef protected

namespace {

std::vector<Object> g_objects;   // host "device" copy the const_objects pointer refers to
RenderData g_render{1, 5, true, Vec3(0, 0, 0)};
bool g_use_sky = false;
std::string g_err;

void upload_scene() {
    AllObjects all{g_objects.data(), (int)g_objects.size()};
    allocate_constant_mem(g_render, all);  // dispatch.cu:104
}

// Camera::assign_constant_mem (src/camera.cu:46-108) with W/H as run-time values and the
// default pose (all rotations 0: Rx*Ry*Rz*p is the identity product evaluated in float).
DeviceCamData camera_for(int w, int h) {
    // volatile: the reference evaluates these from non-constexpr class members at run time
    // (glibc tanf), so the compiler must not fold tanf() with MPFR here either
    volatile float fov_v = 60 * (PI / 180);
    volatile float focal_v = 0.1;
    const float FOV = fov_v;
    const float FOCAL_LEN = focal_v;
    float aspect = static_cast<float>(w) / static_cast<float>(h);
    float viewport_width = 2 * FOCAL_LEN * tan(FOV / 2);
    float viewport_height = viewport_width / aspect;
    Vec3 u(1, 0, 0);
    u.set_mag(viewport_width / static_cast<float>(w));
    Vec3 v(0, -1, 0);
    v.set_mag(viewport_height / static_cast<float>(h));
    Vec3 n = v.cross(u).normalised();
    Vec3 u_step = u * -w / 2;
    Vec3 v_step = v * -h / 2;
    Vec3 cam_pos(0, 0, 0);
    Vec3 focal = n * FOCAL_LEN + cam_pos;
    Vec3 tl = u_step + v_step + focal;
    return DeviceCamData{cam_pos, tl, u, v};
}

struct SegCount {
    unsigned long long segments = 0, hits = 0;
};

// trace_ray (src/raytracer.cu:67-94) with the two counters the metric needs; calls only
// reference functions.  Used by refh_count_segments, never for radiance.
Vec3 trace_ray_counted(Ray ray, SegCount &c) {
    Vec3 final_colour(0, 0, 0);
    Vec3 current_ray_colour(1, 1, 1);
    for (int _ = 0; _ < const_render_data.reflection_limit; _++) {
        ray.apply_antialias();
        RayCollision collision = get_ray_collision(&ray);
        c.segments++;
        if (!collision.hit_data.ray_hits) {
            final_colour += const_render_data.sky_colour * current_ray_colour;
            break;
        }
        c.hits++;
        Material material = collision.hit_mesh_material;
        update_ray(&ray, &collision.hit_data, material);
        if (material.type == Material::EMISSIVE) {
            final_colour += material.emitted_light * current_ray_colour;
        } else {
            current_ray_colour *= material.texture.get_texture_colour(collision.hit_data.texture_uv);
        }
    }
    return final_colour;
}

template <typename F> void parallel_rows(int h, int nthreads, F f) {
    if (nthreads <= 1) {
        for (int y = 0; y < h; y++) f(y);
        return;
    }
    std::atomic<int> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; t++) {
        pool.emplace_back([&]() {
            for (;;) {
                int y = next.fetch_add(1);
                if (y >= h) break;
                f(y);
            }
        });
    }
    for (auto &t : pool) t.join();
}

}  // namespace

extern "C" {

const char *refh_last_error() { return g_err.c_str(); }

int refh_compiled_width() { return SCREEN_WIDTH; }
int refh_compiled_height() { return SCREEN_HEIGHT; }

// Build scene `scene_id` of ray-tracer_b200/host/scenes.hpp through the reference's classes.
// from_main != 0: use the reference's own SceneObjects(scene_id) (ids 0,1,3 only; src/main.cu:100-122)
// with the F11 patch (emissive smoothness/need_uv set explicitly) instead of scenes.hpp.
int refh_build_scene(int scene_id, int p0, int p1, const char *models_dir, int from_main) {
    try {
        if (from_main) {
            // main.cu opens "models/..." relative to cwd: models_dir must be ".../models"
            std::string dir(models_dir);
            std::string parent = dir.substr(0, dir.find_last_of('/'));
            char cwd[4096];
            if (!getcwd(cwd, sizeof cwd)) throw std::runtime_error("getcwd");
            if (chdir(parent.c_str()) != 0) throw std::runtime_error("chdir to models parent failed");
            try {
                SceneObjects so(scene_id);
                g_objects = so.objects;
                g_use_sky = so.use_sky;
            } catch (...) {
                if (chdir(cwd) != 0) {}
                throw;
            }
            if (chdir(cwd) != 0) {}
            for (Object &o : g_objects) {
                if (o.material.type == Material::EMISSIVE) {
                    o.material.smoothness = 0;
                    o.material.need_uv = false;
                }
            }
        } else {
            rtb_scenes::models_dir() = models_dir;
            rtb_scenes::Scene s = rtb_scenes::build(scene_id, p0, p1);
            g_objects = s.objects;
            g_use_sky = s.use_sky;
        }
        upload_scene();
        return 0;
    } catch (std::exception &e) {
        g_err = e.what();
        return 1;
    }
}

int refh_num_objects() { return (int)g_objects.size(); }
int refh_use_sky() { return g_use_sky ? 1 : 0; }

void refh_set_render(int spp, int bounces, int antialias, float sky_r, float sky_g, float sky_b) {
    g_render = RenderData{spp, bounces, antialias != 0, Vec3(sky_r, sky_g, sky_b)};
    upload_scene();
}

// reference camera at the compiled resolution (the real Camera::assign_constant_mem)
void refh_camera_reference(float *out12) {
    Camera cam;
    cam.assign_constant_mem();
    memcpy(out12, &const_cam_data, sizeof(DeviceCamData));
}

// restated camera at a run-time resolution; also installs it
void refh_set_camera(int w, int h, float *out12) {
    DeviceCamData d = camera_for(w, h);
    cudaMemcpyToSymbol(const_cam_data, &d, sizeof(d));
    if (out12) memcpy(out12, &d, sizeof(d));
}

void refh_set_camera_raw(const float *in12) { cudaMemcpyToSymbol(const_cam_data, in12, sizeof(DeviceCamData)); }

// One frame at run-time resolution: per pixel the body of get_pixel_colour
// (src/raytracer.cu:123-135) with SCREEN_WIDTH -> w.  Rows [y0, y1).
int refh_render(int w, int h, int y0, int y1, int time_ms, int frame_num, const float *prev, float *out, int nthreads) {
    parallel_rows(y1 - y0, nthreads, [&](int yy) {
        int y = y0 + yy;
        for (int x = 0; x < w; x++) {
            int array_index = (y * w + x) * 3;
            Vec3 previous_colour(prev ? prev[array_index] : 0.f, prev ? prev[array_index + 1] : 0.f, prev ? prev[array_index + 2] : 0.f);
            uint rng_state = array_index * 3145739 + time_ms * 6291469;
            Ray ray(x, y, &rng_state, const_render_data.antialias);
            Vec3 colour = get_ray_colour(previous_colour, ray, frame_num);
            out[array_index] = colour.x;
            out[array_index + 1] = colour.y;
            out[array_index + 2] = colour.z;
        }
    });
    return 0;
}

// The TRUE kernel function (src/raytracer.cu:116-136) executed thread by thread at the compiled
// resolution with the launch shape of src/dispatch.cu:136-139.  Camera must be the compiled one.
int refh_render_kernel(int time_ms, int frame_num, const float *prev, float *out, int nthreads) {
    int w = SCREEN_WIDTH, h = SCREEN_HEIGHT;
    std::vector<float> zero;
    if (!prev) {
        zero.assign((size_t)w * h * 3, 0.f);
        prev = zero.data();
    }
    int gx = w / 8 + 1, gy = h / 8 + 1;
    parallel_rows(gy, nthreads, [&](int by) {
        blockDim = dim3(8, 8);
        gridDim = dim3(gx, gy);
        for (int bx = 0; bx < gx; bx++)
            for (int ty = 0; ty < 8; ty++)
                for (int tx = 0; tx < 8; tx++) {
                    blockIdx = {(unsigned)bx, (unsigned)by, 0};
                    threadIdx = {(unsigned)tx, (unsigned)ty, 0};
                    get_pixel_colour(out, (float *)prev, &time_ms, &frame_num);
                }
    });
    return 0;
}

// Primary-hit ids, antialias off: obj_id = index accepted by the `<=` loop of
// get_ray_collision (src/raytracer.cu:31-43); sub_id = winning triangle inside the object
// (mesh: brute force in index order with strict <; quad 0/1; cuboid face*2+{0,1}); t = distance.
int refh_primary_ids(int w, int h, int32_t *obj_id, int32_t *sub_id, float *t_out, int nthreads) {
    parallel_rows(h, nthreads, [&](int y) {
        for (int x = 0; x < w; x++) {
            uint rng_state = 0;
            Ray ray(x, y, &rng_state, false);
            int best = -1;
            float best_t = INF;
            for (int i = 0; i < const_objects.num_meshes; i++) {
                RayHitData hd = const_objects.meshes[i].hit(&ray);
                if (!hd.ray_hits) continue;
                if (hd.ray_travelled_dist <= best_t) {
                    best = i;
                    best_t = hd.ray_travelled_dist;
                }
            }
            int sub = 0;
            if (best >= 0) {
                Object &o = const_objects.meshes[best];
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
                    int n = o.mesh.num_triangles;
                    Triangle *tris = o.mesh.device_triangles;
                    sub = -1;
                    for (int k = 0; k < n; k++) {
                        RayHitData th = tris[k].hit(&ray);
                        if (th.ray_hits && th.ray_travelled_dist < bt) {
                            bt = th.ray_travelled_dist;
                            sub = k;
                        }
                    }
                }
            }
            size_t p = (size_t)y * w + x;
            obj_id[p] = best;
            sub_id[p] = best >= 0 ? sub : -1;
            if (t_out) t_out[p] = best >= 0 ? best_t : (float)INF;
        }
    });
    return 0;
}

// segments ("rays") and hit segments of one frame, for the Mrays/s denominator
int refh_count_segments(int w, int h, int y0, int y1, int time_ms, unsigned long long *segments, unsigned long long *hits, int nthreads) {
    std::vector<SegCount> rows(h);
    parallel_rows(y1 - y0, nthreads, [&](int yy) {
        int y = y0 + yy;
        for (int x = 0; x < w; x++) {
            int array_index = (y * w + x) * 3;
            uint rng_state = array_index * 3145739 + time_ms * 6291469;
            Ray ray(x, y, &rng_state, const_render_data.antialias);
            for (int s = 0; s < const_render_data.rays_per_pixel; s++) trace_ray_counted(ray, rows[y]);
        }
    });
    unsigned long long a = 0, b = 0;
    for (auto &r : rows) {
        a += r.segments;
        b += r.hits;
    }
    *segments = a;
    *hits = b;
    return 0;
}

// ---- known-answer hooks ----
unsigned int refh_seed(int array_index, int time_ms) {
    uint s = array_index * 3145739 + time_ms * 6291469;  // src/raytracer.cu:127
    return s;
}
void refh_rng(unsigned int state, int n, float *vals, unsigned int *states) {
    for (int i = 0; i < n; i++) {
        vals[i] = pseudorandom_num(&state);
        states[i] = state;
    }
}
float refh_rng_value(unsigned int result_bits) { return result_bits / 4294967295.0; }  // src/utils.cu:228
float refh_normal(unsigned int state, unsigned int *state_out) {
    float v = normally_dist_num(&state);
    if (state_out) *state_out = state;
    return v;
}
void refh_primary_dir(int x, int y, float *dir3) {
    uint st = 0;
    Ray r(x, y, &st, false);
    dir3[0] = r.direction.x;
    dir3[1] = r.direction.y;
    dir3[2] = r.direction.z;
}
int refh_inf() { return INF; }
// parse_pixel_colours (src/main.cu:343-371) at the compiled resolution: in = W*H*3 floats, out = W*H*4 bytes
void refh_display(const float *in, unsigned char *out) {
    std::vector<float> v(in, in + (size_t)SCREEN_WIDTH * SCREEN_HEIGHT * 3);
    std::vector<sf::Uint8> r = parse_pixel_colours(v);
    std::memcpy(out, r.data(), r.size());
}
// struct sizes, for the record in DESIGN.md
void refh_sizes(int *out, int n) {
    int v[] = {(int)sizeof(Vec3), (int)sizeof(Texture), (int)sizeof(Material), (int)sizeof(Sphere), (int)sizeof(Triangle),
               (int)sizeof(Quad), (int)sizeof(Cuboid), (int)sizeof(BoundingBoxTris), (int)sizeof(Object), (int)sizeof(RayHitData), (int)sizeof(Ray)};
    for (int i = 0; i < n && i < (int)(sizeof v / sizeof v[0]); i++) out[i] = v[i];
}

// Flat dump of the scene as the reference's classes hold it, so that the flattening done by the
// product's compat layer can be compared field by field (tests/test_scene_parity.py).
// Per object 64 floats: [type, mat.type, tex.type, colour(3), light(3), dark(3), num_squares, img_w, img_h,
//   smoothness, need_uv, emitted(3), refr_index, n_tris, then type-specific geometry]
int refh_dump_object(int i, float *out64) {
    if (i < 0 || i >= (int)g_objects.size()) return 1;
    Object &o = g_objects[i];
    for (int k = 0; k < 64; k++) out64[k] = 0;
    Material &m = o.material;
    float *p = out64;
    *p++ = o.type; *p++ = m.type; *p++ = m.texture.type;
    *p++ = m.texture.colour.x; *p++ = m.texture.colour.y; *p++ = m.texture.colour.z;
    *p++ = m.texture.light.x; *p++ = m.texture.light.y; *p++ = m.texture.light.z;
    *p++ = m.texture.dark.x; *p++ = m.texture.dark.y; *p++ = m.texture.dark.z;
    *p++ = m.texture.num_squares; *p++ = m.texture.img_tex_width; *p++ = m.texture.img_tex_height;
    *p++ = m.smoothness; *p++ = m.need_uv; *p++ = m.emitted_light.x; *p++ = m.emitted_light.y; *p++ = m.emitted_light.z;
    *p++ = m.refractive_index;
    auto put3 = [&](Vec3 v) { *p++ = v.x; *p++ = v.y; *p++ = v.z; };
    switch (o.type) {
        case Object::SPHERE: *p++ = 0; put3(o.sphere.center); *p++ = o.sphere.radius; break;
        case Object::TRIANGLE: *p++ = 1; put3(o.triangle.points[0]); put3(o.triangle.points[1]); put3(o.triangle.points[2]);
            put3(o.triangle.normal_vec); for (int k = 0; k < 3; k++) { *p++ = o.triangle.texture_points[k].x; *p++ = o.triangle.texture_points[k].y; } break;
        case Object::QUAD: *p++ = 2; put3(o.quad.point1); put3(o.quad.point2); put3(o.quad.point3); put3(o.quad.point4);
            put3(o.quad.t1.normal_vec); put3(o.quad.t2.normal_vec); break;
        case Object::ONE_WAY_QUAD: *p++ = 2; put3(o.one_way_quad.point1); put3(o.one_way_quad.point2); put3(o.one_way_quad.point3); put3(o.one_way_quad.point4);
            put3(o.one_way_quad.t1.normal_vec); put3(o.one_way_quad.t2.normal_vec); put3(o.one_way_quad.normal_vec); break;
        case Object::CUBOID: *p++ = 12; for (int f = 0; f < 6; f++) put3(o.cuboid.faces[f].point1); put3(o.cuboid.faces[1].point3); break;
        case Object::MESH: *p++ = o.mesh.num_triangles; break;
    }
    return 0;
}
// mesh triangles of object i: per triangle 12 floats p0,p1,p2,normal
int refh_dump_mesh(int i, float *out, int max_tris) {
    if (i < 0 || i >= (int)g_objects.size() || g_objects[i].type != Object::MESH) return -1;
    Object &o = g_objects[i];
    int n = o.mesh.num_triangles;
    for (int k = 0; k < n && k < max_tris; k++) {
        Triangle &t = o.mesh.device_triangles[k];
        float *p = out + (size_t)k * 12;
        for (int v = 0; v < 3; v++) { p[v * 3] = t.points[v].x; p[v * 3 + 1] = t.points[v].y; p[v * 3 + 2] = t.points[v].z; }
        p[9] = t.normal_vec.x; p[10] = t.normal_vec.y; p[11] = t.normal_vec.z;
    }
    return n;
}

}  // extern "C"

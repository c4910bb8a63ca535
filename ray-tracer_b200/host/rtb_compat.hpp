// rtb_compat.hpp -- host-side mirror of the reference's scene / camera / render interface on top
// of the C ABI in include/rtb.h.  Code written against Ben-Edwards44/Ray-Tracer's public names
// (the scene builders in its src/main.cu, or ray-tracer_b200/host/scenes.hpp) compiles against this
// header unchanged and renders through librtb.so on a B200.  Header-only; C++17; no CUDA headers.
//
// Mirrored interface (reference file:line):
//   Vec3, Vec2                       src/utils.cu:13-185
//   Matrix, EnlargementMatrix, RotationMatrix   src/matrix.cu:6-149
//   Texture, Material                src/material.cu:4-186
//   DeviceCamData, Camera            src/camera.cu:12-108   (resolution is a run-time value here)
//   Vertex, Triangle, Quad, OneWayQuad, Object, AllObjects   src/objects.cu:19-290,801-913
//   RenderData                       src/raytracer.cu:4-12
//   ReadOnlyDeviceArray, allocate_constant_mem, VariableRenderData, render   src/dispatch.cu:40-72,104-115,156-163
//   read_file, split_string, ObjFileMesh    src/obj_read.cu:8-147
//   ImageTexture                     src/main.cu:40-91
// Same names, argument meaning and error behaviour (std::runtime_error with the
// "Error from CUDA (<where>): <msg>" shape of check_cuda_error, src/utils.cu:5-10).
//
// All host arithmetic that shapes geometry (vertex transforms, triangle normals, the camera) is
// evaluated in float exactly as the reference's host code does; compile with -ffp-contract=off.
#pragma once
#include <cmath>
#include <math.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "rtb.h"

#ifndef __CUDACC__
#ifndef RTB_COMPAT_HAVE_FLOAT3
#define RTB_COMPAT_HAVE_FLOAT3
struct float3 {
    float x, y, z;
};
#endif
#endif

// ---------------------------------------------------------------------------------------------
// vectors (src/utils.cu:13-185): host operators only; the device math lives in csrc/
// ---------------------------------------------------------------------------------------------
class Vec3 {
  public:
    float x, y, z;
    Vec3() : x(0), y(0), z(0) {}
    Vec3(float vx, float vy, float vz) : x(vx), y(vy), z(vz) {}
    Vec3(float3 v) : x(v.x), y(v.y), z(v.z) {}
    Vec3 operator+(Vec3 o) const { return Vec3(x + o.x, y + o.y, z + o.z); }
    Vec3 operator-(Vec3 o) const { return Vec3(x - o.x, y - o.y, z - o.z); }
    Vec3 operator*(float s) const { return Vec3(x * s, y * s, z * s); }
    Vec3 operator/(float s) const { return Vec3(x / s, y / s, z / s); }
    float magnitude() const {
        float mag_sq = x * x + y * y + z * z;
        return sqrtf(mag_sq);
    }
    Vec3 normalised() const {
        float inv_mag = 1 / magnitude();
        return Vec3(x * inv_mag, y * inv_mag, z * inv_mag);
    }
    Vec3 cross(Vec3 o) const { return Vec3(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x); }
    void set_mag(float desired_mag) {
        float scale = desired_mag / magnitude();
        x *= scale;
        y *= scale;
        z *= scale;
    }
};

class Vec2 {
  public:
    float x, y;
    Vec2() : x(0), y(0) {}
    Vec2(float vx, float vy) : x(vx), y(vy) {}
};

inline void check_cuda_error(int error, std::string msg, const char *detail = "") {
    if (error != 0) throw std::runtime_error("Error from CUDA (" + msg + "): " + detail);
}

// ---------------------------------------------------------------------------------------------
// host matrices (src/matrix.cu): row-major float, products accumulated as `sum += a*b` from 0
// ---------------------------------------------------------------------------------------------
class Matrix {
  public:
    int rows = 0, cols = 0;
    std::vector<std::vector<float>> items;

    Matrix() {}
    Matrix(int r, int c) : rows(r), cols(c) { reset_items(); }
    Matrix(std::vector<std::vector<float>> m) : items(std::move(m)) {
        rows = (int)items.size();
        cols = rows == 0 ? 0 : (int)items[0].size();
    }
    Matrix operator*(const Matrix &o) const {
        if (cols != o.rows) throw std::runtime_error("Matrix dimensions do not match for multiplication.");
        Matrix out(rows, o.cols);
        for (int r = 0; r < rows; r++) {
            const std::vector<float> &row = items[r];
            std::vector<float> &dst = out.items[r];
            for (int c = 0; c < o.cols; c++) {
                float sum = 0;
                for (int i = 0; i < cols; i++) sum += row[i] * o.items[i][c];
                dst[c] = sum;
            }
        }
        return out;
    }
    Matrix transpose() const {
        Matrix t(cols, rows);
        for (int r = 0; r < rows; r++)
            for (int c = 0; c < cols; c++) t.items[c][r] = items[r][c];
        return t;
    }
    void reset_items() { items.assign(rows, std::vector<float>(cols, 0)); }
};

class EnlargementMatrix : public Matrix {
  public:
    float scale;
    EnlargementMatrix(float scale_fact, int dimensions) : Matrix(dimensions, dimensions), scale(scale_fact) {
        for (int i = 0; i < rows; i++) items[i][i] = scale;
    }
};

class RotationMatrix : public Matrix {
  public:
    static const int X_AXIS = 0;
    static const int Y_AXIS = 1;
    static const int Z_AXIS = 2;
    int axis;
    float angle;
    // sign convention of src/matrix.cu:119-135 (float sin/cos of the float angle)
    RotationMatrix(int rot_axis, float rot_angle) : axis(rot_axis), angle(rot_angle) {
        rows = cols = 3;
        float s = sinf(angle), c = cosf(angle);
        if (axis == X_AXIS) items = {{1, 0, 0}, {0, c, s}, {0, -s, c}};
        else if (axis == Y_AXIS) items = {{c, 0, -s}, {0, 1, 0}, {s, 0, c}};
        else items = {{c, -s, 0}, {s, c, 0}, {0, 0, 1}};
    }
};

// ---------------------------------------------------------------------------------------------
// process-wide state behind the reference's global-function interface (the reference keeps its
// equivalents in three __constant__ symbols: one scene / camera per process)
// ---------------------------------------------------------------------------------------------
namespace rtb_compat {

struct ImageStore {
    int width, height;
    std::vector<float> rgb;
};

struct State {
    rtb_ctx *ctx = nullptr;
    int device = 0;
    int width = 1000, height = 800;  // SCREEN_WIDTH / SCREEN_HEIGHT defaults, src/camera.cu:4-5
    bool have_cam = false;
    RtbCamData cam{};
    std::vector<ImageStore> images;
    ~State() {
        if (ctx) rtb_destroy(ctx);
    }
};

inline State &state() {
    static State s;
    return s;
}

inline void set_resolution(int w, int h) {
    state().width = w;
    state().height = h;
}
inline void set_device(int device) { state().device = device; }

inline RtbVec3 to_rtb(Vec3 v) { return RtbVec3{v.x, v.y, v.z}; }

}  // namespace rtb_compat

#define SCREEN_WIDTH (rtb_compat::state().width)
#define SCREEN_HEIGHT (rtb_compat::state().height)
#define PIXEL_ARRAY_LEN (SCREEN_WIDTH * SCREEN_HEIGHT * 3)

const float PI = 3.141592653589793;

// ---------------------------------------------------------------------------------------------
// Texture / Material (src/material.cu)
// ---------------------------------------------------------------------------------------------
class Texture {
  public:
    static const int COLOUR = 0;
    static const int GRADIENT = 1;
    static const int CHECKERBOARD = 2;
    static const int IMAGE = 3;

    int type = COLOUR;
    Vec3 colour;
    Vec3 light, dark;
    int num_squares = 0;
    int img_tex_width = 0, img_tex_height = 0;
    int image = -1;  // handle into the compat image store (the reference holds a device pointer)

    Texture() {}
    Texture(int texture_type) : type(texture_type) {}

    static Texture create_const_colour(Vec3 texture_colour) {
        Texture t(COLOUR);
        t.colour = texture_colour;
        return t;
    }
    static Texture create_gradient() { return Texture(GRADIENT); }
    static Texture create_checkerboard(Vec3 light_colour, Vec3 dark_colour, int num_sq) {
        Texture t(CHECKERBOARD);
        t.light = light_colour;
        t.dark = dark_colour;
        t.num_squares = num_sq;
        return t;
    }
    static Texture create_image(int width, int height, std::vector<Vec3> rgb_values) {
        if ((size_t)width * height != rgb_values.size())
            throw std::runtime_error("Error from CUDA (allocating texture rgb values): size mismatch");
        Texture t(IMAGE);
        t.img_tex_width = width;
        t.img_tex_height = height;
        rtb_compat::ImageStore img{width, height, {}};
        img.rgb.reserve(rgb_values.size() * 3);
        for (const Vec3 &p : rgb_values) {
            img.rgb.push_back(p.x);
            img.rgb.push_back(p.y);
            img.rgb.push_back(p.z);
        }
        rtb_compat::state().images.push_back(std::move(img));
        t.image = (int)rtb_compat::state().images.size() - 1;
        return t;
    }
};

class Material {
  public:
    static const int STANDARD = 0;
    static const int EMISSIVE = 1;
    static const int REFRACTIVE = 2;

    int type = STANDARD;
    Texture texture;
    float smoothness = 0;
    bool need_uv = false;
    Vec3 emitted_light;
    float refractive_index = 1;

    Material() {}
    Material(int mat_type) : type(mat_type) {}

    static Material create_standard(Texture mat_tex, float smoothness_val) {
        Material m(STANDARD);
        m.texture = mat_tex;
        m.smoothness = smoothness_val;
        m.need_uv = mat_tex.type != Texture::COLOUR;
        return m;
    }
    // NB: the reference leaves smoothness / need_uv / texture uninitialised here;
    // this mirror defines them (0 / false / black) -- callers that care set them explicitly.
    static Material create_emissive(Vec3 emit_colour, float emit_strength) {
        Material m(EMISSIVE);
        m.emitted_light = emit_colour * emit_strength;
        return m;
    }
    static Material create_refractive(Texture mat_tex, float n) {
        Material m(REFRACTIVE);
        m.texture = mat_tex;
        m.refractive_index = n;
        m.need_uv = mat_tex.type != Texture::COLOUR;
        m.smoothness = 1;
        return m;
    }
};

// ---------------------------------------------------------------------------------------------
// camera (src/camera.cu): DeviceCamData + Camera::assign_constant_mem with run-time W/H
// ---------------------------------------------------------------------------------------------
struct DeviceCamData {
    Vec3 cam_pos;
    Vec3 tl_pixel_pos;
    Vec3 delta_u;
    Vec3 delta_v;
};

class Camera {
  public:
    // "these values can be changed" in the reference (src/camera.cu:33-41): public here
    Vec3 CAM_POS = Vec3(0, 0, 0);
    float FOV = 60 * (PI / 180);
    float FOCAL_LEN = 0.1;
    float X_ROT = 0 * (PI / 180);
    float Y_ROT = 0 * (PI / 180);
    float Z_ROT = 0 * (PI / 180);

    Camera() {}

    DeviceCamData get_cam_data() const {
        int W = SCREEN_WIDTH, H = SCREEN_HEIGHT;
        float aspect = static_cast<float>(W) / static_cast<float>(H);
        volatile float half_fov = FOV / 2;  // keep tanf a run-time libm call, as in the reference
        float viewport_width = 2 * FOCAL_LEN * tanf(half_fov);
        float viewport_height = viewport_width / aspect;
        Vec3 delta_u = rotate_point(Vec3(1, 0, 0)) - Vec3(0, 0, 0);
        delta_u.set_mag(viewport_width / static_cast<float>(W));
        Vec3 delta_v = rotate_point(Vec3(0, -1, 0)) - Vec3(0, 0, 0);
        delta_v.set_mag(viewport_height / static_cast<float>(H));
        Vec3 plane_normal = delta_v.cross(delta_u).normalised();
        Vec3 u_step = delta_u * (float)(-W) / 2;
        Vec3 v_step = delta_v * (float)(-H) / 2;
        Vec3 focal_len_offset = plane_normal * FOCAL_LEN + CAM_POS;
        Vec3 tl_pos = u_step + v_step + focal_len_offset;
        return DeviceCamData{CAM_POS, tl_pos, delta_u, delta_v};
    }

    // == Camera::assign_constant_mem, src/camera.cu:46-60
    void assign_constant_mem() {
        DeviceCamData d = get_cam_data();
        rtb_compat::State &st = rtb_compat::state();
        st.cam = RtbCamData{rtb_compat::to_rtb(d.cam_pos), rtb_compat::to_rtb(d.tl_pixel_pos), rtb_compat::to_rtb(d.delta_u),
                            rtb_compat::to_rtb(d.delta_v)};
        st.have_cam = true;
        if (st.ctx && rtb_set_camera(st.ctx, &st.cam, st.width, st.height) != 0)
            throw std::runtime_error(rtb_last_error(st.ctx));
    }

  private:
    Vec3 rotate_point(Vec3 point) const {
        std::vector<std::vector<float>> p = {{point.x}, {point.y}, {point.z}};
        Matrix r = RotationMatrix(RotationMatrix::X_AXIS, X_ROT) * RotationMatrix(RotationMatrix::Y_AXIS, Y_ROT) *
                   RotationMatrix(RotationMatrix::Z_AXIS, Z_ROT) * Matrix(p);
        return Vec3(r.items[0][0], r.items[1][0], r.items[2][0]);
    }
};

// ---------------------------------------------------------------------------------------------
// geometry (src/objects.cu): lean records; the device layout is built by librtb.so
// ---------------------------------------------------------------------------------------------
struct Vertex {
    Vec3 world_point;
    Vec2 texture_point;
};

class Triangle {
  public:
    Vec3 points[3];
    Material material;
    Vec3 normal_vec;
    Vec2 texture_points[3];

    Triangle() {}
    Triangle(Vec3 p1, Vec3 p2, Vec3 p3, Material mat) : material(mat) {
        points[0] = p1;
        points[1] = p2;
        points[2] = p3;
        precompute();
    }
    Triangle(Vertex p1, Vertex p2, Vertex p3, Material mat) : material(mat) {
        points[0] = p1.world_point;
        points[1] = p2.world_point;
        points[2] = p3.world_point;
        texture_points[0] = p1.texture_point;
        texture_points[1] = p2.texture_point;
        texture_points[2] = p3.texture_point;
        precompute();
    }

  private:
    void precompute() {  // src/objects.cu:175-186 (side1/side2/normal are recomputed by librtb.so)
        Vec3 side1 = points[1] - points[0];
        Vec3 side2 = points[2] - points[0];
        normal_vec = side1.cross(side2).normalised();
    }
};

class Quad {
  public:
    Triangle t1, t2;
    Material material;
    Vec3 point1, point2, point3, point4;

    Quad() {}
    Quad(Vec3 p1, Vec3 p2, Vec3 p3, Vec3 p4, Material mat) : material(mat), point1(p1), point2(p2), point3(p3), point4(p4) {
        // src/objects.cu:244-253
        Vertex v1{p1, Vec2(0, 0)}, v2{p2, Vec2(1, 0)}, v3{p3, Vec2(1, 1)}, v4{p4, Vec2(0, 1)};
        t1 = Triangle(v1, v2, v3, mat);
        t2 = Triangle(v1, v4, v3, mat);
    }
};

class Object {
  public:
    static const int SPHERE = 0;
    static const int TRIANGLE = 1;
    static const int QUAD = 2;
    static const int ONE_WAY_QUAD = 3;
    static const int CUBOID = 4;
    static const int MESH = 5;

    int type;
    Material material;

    // flat parameters (meaning per type as documented at RtbObject in rtb.h)
    Vec3 p[4];
    float f[3] = {0, 0, 0};
    Vec2 uv[3];
    bool invert_normal = false;
    std::shared_ptr<std::vector<Triangle>> mesh_triangles;

    Object(int t, Material mat) : type(t), material(mat) {}

    static Object create_sphere(Vec3 center, float radius, Material mat) {
        Object o(SPHERE, mat);
        o.p[0] = center;
        o.f[0] = radius;
        return o;
    }
    static Object create_triangle(Vec3 p1, Vec3 p2, Vec3 p3, Material mat) {
        Object o(TRIANGLE, mat);
        o.p[0] = p1; o.p[1] = p2; o.p[2] = p3;
        return o;
    }
    static Object create_triangle(Vertex p1, Vertex p2, Vertex p3, Material mat) {
        Object o(TRIANGLE, mat);
        o.p[0] = p1.world_point; o.p[1] = p2.world_point; o.p[2] = p3.world_point;
        o.uv[0] = p1.texture_point; o.uv[1] = p2.texture_point; o.uv[2] = p3.texture_point;
        return o;
    }
    static Object create_quad(Vec3 p1, Vec3 p2, Vec3 p3, Vec3 p4, Material mat) {
        Object o(QUAD, mat);
        o.p[0] = p1; o.p[1] = p2; o.p[2] = p3; o.p[3] = p4;
        return o;
    }
    static Object create_one_way_quad(Vec3 p1, Vec3 p2, Vec3 p3, Vec3 p4, bool invert, Material mat) {
        Object o(ONE_WAY_QUAD, mat);
        o.p[0] = p1; o.p[1] = p2; o.p[2] = p3; o.p[3] = p4;
        o.invert_normal = invert;
        return o;
    }
    static Object create_cuboid(Vec3 tl_near_pos, float width, float height, float depth, Material mat) {
        Object o(CUBOID, mat);
        o.p[0] = tl_near_pos;
        o.f[0] = width; o.f[1] = height; o.f[2] = depth;
        return o;
    }
    // device_triangle_array is what ReadOnlyDeviceArray<Triangle>::device_pointer returned; the
    // triangles are taken from host_triangle_array (the reference keeps both, src/objects.cu:780-787)
    static Object create_mesh(std::vector<Triangle> host_triangle_array, Triangle *device_triangle_array, Material mat) {
        (void)device_triangle_array;
        Object o(MESH, mat);
        o.mesh_triangles = std::make_shared<std::vector<Triangle>>(std::move(host_triangle_array));
        return o;
    }
};

struct AllObjects {
    Object *meshes;
    int num_meshes;
};

// == RenderData, src/raytracer.cu:4-12
struct RenderData {
    int rays_per_pixel;
    int reflection_limit;
    bool antialias;
    Vec3 sky_colour;
};

// ---------------------------------------------------------------------------------------------
// dispatch (src/dispatch.cu)
// ---------------------------------------------------------------------------------------------
// The reference copies the vector to the device and exposes the device pointer.  Here the "device
// pointer" is a stable host copy: librtb.so performs the real upload in allocate_constant_mem.
template <typename T> class ReadOnlyDeviceArray {
  public:
    T *device_pointer;
    ReadOnlyDeviceArray(std::vector<T> values) : host_values(std::make_shared<std::vector<T>>(std::move(values))) {
        device_pointer = host_values->data();
        keep_alive().push_back(host_values);
    }
    void free_memory() { host_values.reset(); }

  private:
    std::shared_ptr<std::vector<T>> host_values;
    static std::vector<std::shared_ptr<std::vector<T>>> &keep_alive() {
        static std::vector<std::shared_ptr<std::vector<T>>> v;  // the reference never frees scene arrays either
        return v;
    }
};

namespace rtb_compat {

// Host-side flat description of a scene: owns the arrays an RtbSceneDesc points into.
struct SceneDescStorage {
    std::vector<RtbObject> objects;
    std::vector<RtbMaterial> materials;
    std::vector<float> tri_points, tri_uv;
    std::vector<RtbImage> images;
    RtbSceneDesc desc{};
};

inline RtbMaterial to_rtb(const Material &m) {
    RtbMaterial r;
    std::memset(&r, 0, sizeof r);
    r.type = m.type;
    r.texture.type = m.texture.type;
    r.texture.colour = to_rtb(m.texture.colour);
    r.texture.light = to_rtb(m.texture.light);
    r.texture.dark = to_rtb(m.texture.dark);
    r.texture.num_squares = m.texture.num_squares;
    r.texture.img_width = m.texture.img_tex_width;
    r.texture.img_height = m.texture.img_tex_height;
    r.texture.image = m.texture.image;
    r.smoothness = m.smoothness;
    r.need_uv = m.need_uv ? 1 : 0;
    r.emitted_light = to_rtb(m.emitted_light);
    r.refractive_index = m.refractive_index;
    return r;
}

// Object[] (creation order preserved: it defines the tie rule) -> RtbSceneDesc; materials de-duplicated
inline void build_desc(const Object *objs, int n, SceneDescStorage &out) {
    out = SceneDescStorage();
    for (int i = 0; i < n; i++) {
        const Object &o = objs[i];
        RtbMaterial m = to_rtb(o.material);
        int mi = -1;
        for (size_t k = 0; k < out.materials.size(); k++)
            if (std::memcmp(&out.materials[k], &m, sizeof m) == 0) { mi = (int)k; break; }
        if (mi < 0) { out.materials.push_back(m); mi = (int)out.materials.size() - 1; }
        RtbObject r;
        std::memset(&r, 0, sizeof r);
        r.type = o.type;
        r.material = mi;
        for (int k = 0; k < 4; k++) r.p[k] = to_rtb(o.p[k]);
        for (int k = 0; k < 3; k++) { r.f[k] = o.f[k]; r.uv[2 * k] = o.uv[k].x; r.uv[2 * k + 1] = o.uv[k].y; }
        r.invert_normal = o.invert_normal ? 1 : 0;
        if (o.type == Object::MESH) {
            const std::vector<Triangle> &tris = *o.mesh_triangles;
            r.tri_first = (int)(out.tri_points.size() / 9);
            r.tri_count = (int)tris.size();
            for (const Triangle &t : tris) {
                for (int v = 0; v < 3; v++) {
                    out.tri_points.push_back(t.points[v].x);
                    out.tri_points.push_back(t.points[v].y);
                    out.tri_points.push_back(t.points[v].z);
                    out.tri_uv.push_back(t.texture_points[v].x);
                    out.tri_uv.push_back(t.texture_points[v].y);
                }
            }
        }
        out.objects.push_back(r);
    }
    for (const ImageStore &img : state().images) out.images.push_back(RtbImage{img.width, img.height, img.rgb.data()});
    out.desc.num_objects = (int)out.objects.size();
    out.desc.objects = out.objects.data();
    out.desc.num_materials = (int)out.materials.size();
    out.desc.materials = out.materials.data();
    out.desc.num_mesh_tris = (int)(out.tri_points.size() / 9);
    out.desc.mesh_tri_points = out.tri_points.data();
    out.desc.mesh_tri_uv = out.tri_uv.data();
    out.desc.num_images = (int)out.images.size();
    out.desc.images = out.images.data();
}

inline rtb_ctx *ensure_ctx() {
    State &st = state();
    if (!st.ctx) {
        if (rtb_create(&st.ctx, st.device) != 0) {
            std::string msg = rtb_last_error(nullptr);
            st.ctx = nullptr;
            throw std::runtime_error(msg);
        }
    }
    return st.ctx;
}

}  // namespace rtb_compat

// == allocate_constant_mem(RenderData, AllObjects), src/dispatch.cu:104-108
inline void allocate_constant_mem(RenderData render_data, AllObjects mesh_data) {
    rtb_compat::State &st = rtb_compat::state();
    rtb_ctx *ctx = rtb_compat::ensure_ctx();
    if (!st.have_cam) Camera().assign_constant_mem();
    if (rtb_set_camera(ctx, &st.cam, st.width, st.height) != 0) throw std::runtime_error(rtb_last_error(ctx));
    RtbRenderData rd{render_data.rays_per_pixel, render_data.reflection_limit, render_data.antialias ? 1 : 0,
                     rtb_compat::to_rtb(render_data.sky_colour)};
    if (rtb_set_render_data(ctx, &rd) != 0) throw std::runtime_error(rtb_last_error(ctx));
    rtb_compat::SceneDescStorage storage;
    rtb_compat::build_desc(mesh_data.meshes, mesh_data.num_meshes, storage);
    if (rtb_scene_set(ctx, &storage.desc) != 0) throw std::runtime_error(rtb_last_error(ctx));
}

// == VariableRenderData, src/dispatch.cu:111-115
struct VariableRenderData {
    int frame_num;
    std::vector<float> previous_render;
};

// == render(VariableRenderData*, int), src/dispatch.cu:156-163
inline void render(VariableRenderData *data, int current_time_ms) {
    rtb_ctx *ctx = rtb_compat::ensure_ctx();
    if ((int)data->previous_render.size() != PIXEL_ARRAY_LEN)
        throw std::runtime_error("Error from CUDA (copying real only device array): previous_render has the wrong length");
    if (rtb_render(ctx, data->previous_render.data(), &data->frame_num, current_time_ms) != 0)
        throw std::runtime_error(rtb_last_error(ctx));
}

// ---------------------------------------------------------------------------------------------
// OBJ loading contract (src/obj_read.cu): `v x y z` and `f a[/..] b[/..] c[/..] [d[/..]]` lines only;
// position index only; vt / vn / usemtl / mtllib / o / s are ignored, exactly like the reference.
// ---------------------------------------------------------------------------------------------
inline std::vector<std::string> read_file(std::string filename) {
    std::ifstream file(filename);
    if (!file) throw std::runtime_error("Could not find file to open.");
    std::vector<std::string> lines;
    // the reference's eof() loop also pushes the empty string after the final newline
    while (!file.eof()) {
        std::string line;
        std::getline(file, line);
        lines.push_back(line);
    }
    return lines;
}

// python-like split; every token after the first keeps a leading space (src/obj_read.cu:35)
inline std::vector<std::string> split_string(std::string str, char split_char) {
    std::vector<std::string> parts;
    std::string cur = "";
    for (char ch : str) {
        if (ch == split_char) {
            parts.push_back(cur);
            cur = " ";
        } else {
            cur += ch;
        }
    }
    parts.push_back(cur);
    return parts;
}

class ObjFileMesh {
  public:
    Matrix vertex_mat;                          // 3 x num_vertices, as the reference keeps it
    std::vector<std::vector<float3>> faces;

    ObjFileMesh(std::string filename) {
        std::vector<std::vector<float>> vertices;
        for (const std::string &line : read_file(filename)) {
            std::vector<std::string> tok = split_string(line, ' ');
            if (tok[0] == "v") {
                vertices.push_back({std::stof(tok.at(1)), std::stof(tok.at(2)), std::stof(tok.at(3))});
            } else if (tok[0] == "f") {
                std::vector<int> idx;
                for (size_t i = 1; i < tok.size(); i++) idx.push_back(std::stoi(split_string(tok[i], '/')[0]) - 1);
                face_indices.push_back(idx);
            }
        }
        vertex_mat = Matrix(vertices).transpose();
        extract_faces();
    }

    void enlarge(float scale_fact) {
        vertex_mat = EnlargementMatrix(scale_fact, 3) * vertex_mat;
        extract_faces();
    }

    void rotate(float x_angle, float y_angle, float z_angle) {
        RotationMatrix x_rot(RotationMatrix::X_AXIS, x_angle);
        RotationMatrix y_rot(RotationMatrix::Y_AXIS, y_angle);
        RotationMatrix z_rot(RotationMatrix::Z_AXIS, z_angle);
        vertex_mat = x_rot * y_rot * z_rot * vertex_mat;  // ((Rx*Ry)*Rz)*V, float accumulation
        extract_faces();
    }

    void translate(float offset_x, float offset_y, float offset_z) {
        for (int i = 0; i < vertex_mat.cols; i++) {
            vertex_mat.items[0][i] += offset_x;
            vertex_mat.items[1][i] += offset_y;
            vertex_mat.items[2][i] += offset_z;
        }
        extract_faces();
    }

  private:
    std::vector<std::vector<int>> face_indices;  // parsed once (the reference re-parses the text each time)

    void extract_faces() {
        faces.clear();
        faces.reserve(face_indices.size());
        for (const std::vector<int> &idx : face_indices) {
            std::vector<float3> face;
            for (int v : idx) {
                float3 p;
                p.x = vertex_mat.items[0].at(v);
                p.y = vertex_mat.items[1].at(v);
                p.z = vertex_mat.items[2].at(v);
                face.push_back(p);
            }
            faces.push_back(face);
        }
    }
};

// == ImageTexture, src/main.cu:40-91: reads textures/parsed_textures.txt as written by
// textures/parse_textures.py (count line; per image: name, width, height, one line of floats)
class ImageTexture {
  public:
    std::string PARSED_TEXTURE_FILENAME = "textures/parsed_textures.txt";

    ImageTexture(std::string filename) { parse_file(filename, PARSED_TEXTURE_FILENAME); }
    ImageTexture(std::string filename, std::string parsed_file) : PARSED_TEXTURE_FILENAME(parsed_file) {
        parse_file(filename, parsed_file);
    }
    Texture get_device_texture() { return Texture::create_image(width, height, rgb_values); }

    int width = 0, height = 0;
    std::vector<Vec3> rgb_values;

  private:
    void parse_file(const std::string &filename, const std::string &parsed_file) {
        std::vector<std::string> lines = read_file(parsed_file);
        for (size_t i = 0; i + 3 < lines.size(); i++) {
            if (lines[i] == filename) {
                width = std::stoi(lines[i + 1]);
                height = std::stoi(lines[i + 2]);
                // the line ends with a space, so the last token is empty (src/main.cu:81)
                std::vector<std::string> tok = split_string(lines[i + 3], ' ');
                for (size_t k = 0; k + 1 < tok.size() && k + 2 < tok.size(); k += 3)
                    rgb_values.push_back(Vec3(std::stof(tok[k]), std::stof(tok[k + 1]), std::stof(tok[k + 2])));
                return;
            }
        }
        throw std::runtime_error("Image file not found.\n");
    }
};

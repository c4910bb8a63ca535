// scenes.hpp -- the scene fixtures (C1..C5 + the reference's test scenes)
// written ONCE against the reference's public scene API names
//   Vec3, Texture::create_*, Material::create_*, Object::create_*, Triangle, Quad,
//   ObjFileMesh{faces, enlarge, rotate, translate}, ReadOnlyDeviceArray<T>
// (reference: src/objects.cu:845-906, src/material.cu:21-51,157-185, src/obj_read.cu:47-85,
//  src/dispatch.cu:40-72).  The same file is compiled
//   (a) against rtb_compat.hpp  -> the B200 product path (librtb.so underneath), and
//   (b) against the reference's own sources (oracle/ref_host.cpp, baseline/ref_cuda_main.cu),
// which is what "drop-in for the scene API" means in practice: no #ifdef separates the two.
//
// Scene ids 0,1,3 re-state src/main.cu:150-213 (+ create_cornell_box :252-288) so that the
// reference side can be cross-checked against its OWN SceneObjects(n) (oracle/ref_host.cpp does).
// Emissive materials get smoothness/need_uv/texture set explicitly (the
// reference leaves them uninitialised and update_ray lerps with smoothness on EMISSIVE hits).
#pragma once
#include <string>
#include <vector>
#include <stdexcept>
#include <cstdio>
#include <cstdint>

namespace rtb_scenes {

// where the OBJ fixtures live ("models/" relative to cwd in the reference, src/main.cu:156)
inline std::string &models_dir() {
    static std::string dir = "models";
    return dir;
}

struct Scene {
    std::vector<Object> objects;
    bool use_sky = true;
};

inline Material make_emissive(Vec3 colour, float strength) {
    Material m = Material::create_emissive(colour, strength);
    m.texture = Texture::create_const_colour(Vec3(0, 0, 0));
    m.smoothness = 0;
    m.need_uv = false;
    m.refractive_index = 1;
    return m;
}

inline Material make_standard(Texture tex, float smoothness) {
    Material m = Material::create_standard(tex, smoothness);
    m.emitted_light = Vec3(0, 0, 0);
    m.refractive_index = 1;
    return m;
}

inline Material make_refractive(Texture tex, float n) {
    Material m = Material::create_refractive(tex, n);
    m.emitted_light = Vec3(0, 0, 0);
    return m;
}

}  // namespace rtb_scenes

#include "obj_mtl.hpp"  // OBJ + MTL + vt loader, parsed-texture files (uses the helpers above)

namespace rtb_scenes {

// faces -> Triangle list, as SceneObjects::create_mesh does (src/main.cu:127-148):
// 3-vertex faces become one Triangle, 4-vertex faces the two triangles of a Quad.
inline void append_faces(std::vector<Triangle> &triangles, ObjFileMesh &obj, Material mat) {
    for (std::vector<float3> &face : obj.faces) {
        if (face.size() == 3) {
            triangles.push_back(Triangle(Vec3(face[0]), Vec3(face[1]), Vec3(face[2]), mat));
        } else if (face.size() == 4) {
            Quad quad(Vec3(face[0]), Vec3(face[1]), Vec3(face[2]), Vec3(face[3]), mat);
            triangles.push_back(quad.t1);
            triangles.push_back(quad.t2);
        } else {
            throw std::logic_error("Only triangle or quad meshes are supported.\n");
        }
    }
}

inline Object mesh_from_triangles(std::vector<Triangle> &triangles, Material mat) {
    ReadOnlyDeviceArray<Triangle> device_array(triangles);
    return Object::create_mesh(triangles, device_array.device_pointer, mat);
}

inline Object create_mesh(ObjFileMesh &obj, Material mat) {
    std::vector<Triangle> triangles;
    append_faces(triangles, obj, mat);
    return mesh_from_triangles(triangles, mat);
}

// src/main.cu:252-288
inline void create_cornell_box(Scene &s, Vec3 tl_near_pos, float width, float height, float depth, float light_width) {
    s.use_sky = false;

    Material floor = make_standard(Texture::create_checkerboard(Vec3(0.1, 0.8, 0.1), Vec3(0.1, 0.5, 0.1), 8), 0);
    Material l_wall = make_standard(Texture::create_const_colour(Vec3(1, 0.2, 0.2)), 0);
    Material r_wall = make_standard(Texture::create_const_colour(Vec3(0.3, 0.3, 1)), 0);
    Material back = make_standard(Texture::create_const_colour(Vec3(0.2, 0.2, 0.2)), 0);
    Material roof = make_standard(Texture::create_const_colour(Vec3(0.9, 0.9, 0.9)), 0);
    Material front = make_standard(Texture::create_const_colour(Vec3(1, 1, 1)), 0);

    Vec3 w(width, 0, 0);
    Vec3 h(0, height, 0);
    Vec3 d(0, 0, depth);
    Vec3 p = tl_near_pos;

    s.objects.push_back(Object::create_quad(p - h, p - h + w, p - h + w + d, p - h + d, floor));
    s.objects.push_back(Object::create_quad(p, p - h, p - h + d, p + d, l_wall));
    s.objects.push_back(Object::create_quad(p + w, p + w - h, p + w - h + d, p + w + d, r_wall));
    s.objects.push_back(Object::create_quad(p + d, p + w + d, p + w - h + d, p - h + d, back));
    s.objects.push_back(Object::create_quad(p, p + d, p + w + d, p + w, roof));
    s.objects.push_back(Object::create_one_way_quad(p, p + w, p + w - h, p - h, false, front));

    Material light_mat = make_emissive(Vec3(1, 1, 1), 6);
    Vec3 light_pos(p.x + width / 2 - light_width / 2, p.y, p.z + depth / 2 - light_width / 2);
    s.objects.push_back(Object::create_cuboid(light_pos, light_width, 0.04, light_width, light_mat));
}

// id 0 -- src/main.cu:150-170 (C1 / C3 / C5 geometry)
inline void monkey_scene(Scene &s) {
    create_cornell_box(s, Vec3(-0.5, 0.5, 1.2), 1, 1, 1, 0.5);

    Material monkey_mat = make_standard(Texture::create_const_colour(Vec3(1, 1, 1)), 0);
    ObjFileMesh m(models_dir() + "/low_poly_monkey.obj");
    m.enlarge(0.3);
    m.rotate(0, 2.3, 0);
    m.translate(0.1, -0.1, 1.6);
    s.objects.push_back(create_mesh(m, monkey_mat));

    Material sphere_mat = make_standard(Texture::create_const_colour(Vec3(0.8, 0.8, 0.8)), 1);
    s.objects.push_back(Object::create_sphere(Vec3(-0.25, -0.25, 1.95), 0.25, sphere_mat));
}

// id 1 -- src/main.cu:172-187
inline void reflection_scene(Scene &s) {
    create_cornell_box(s, Vec3(-0.5, 0.5, 1.2), 1, 1, 1, 0.5);
    Texture tex = Texture::create_const_colour(Vec3(1, 1, 1));
    s.objects.push_back(Object::create_sphere(Vec3(-0.2, 0.2, 1.7), 0.15, make_standard(tex, 0)));
    s.objects.push_back(Object::create_sphere(Vec3(0.2, 0.2, 1.7), 0.15, make_standard(tex, 0.33)));
    s.objects.push_back(Object::create_sphere(Vec3(-0.2, -0.2, 1.7), 0.15, make_standard(tex, 0.66)));
    s.objects.push_back(Object::create_sphere(Vec3(0.2, -0.2, 1.7), 0.15, make_standard(tex, 1)));
}

// A deterministic stand-in for earth.png (not shipped): values are k/256 like
// textures/parse_textures.py:32 writes them.
inline std::vector<Vec3> synthetic_image(int w, int h) {
    std::vector<Vec3> rgb;
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            int r = (x * 7 + y * 3) % 256;
            int g = ((x / 4 + y / 4) % 2) ? 200 : 40;
            int b = (x * y) % 256;
            rgb.push_back(Vec3(r / 256.0f, g / 256.0f, b / 256.0f));
        }
    }
    return rgb;
}

// id 2 -- src/main.cu:189-204 with the synthetic image in place of earth.png
inline void texture_scene(Scene &s) {
    create_cornell_box(s, Vec3(-0.5, 0.5, 1.2), 1, 1, 1, 0.5);

    Texture img = Texture::create_image(64, 32, synthetic_image(64, 32));
    s.objects.push_back(Object::create_sphere(Vec3(0, 0, 1.7), 0.25, make_standard(img, 0)));

    Material tri_mat = make_standard(Texture::create_checkerboard(Vec3(1, 1, 1), Vec3(0, 0, 0), 4), 0);
    s.objects.push_back(Object::create_triangle(Vertex{Vec3(0.1, 0, 1.7), Vec2(0, 0)}, Vertex{Vec3(0.6, 0.5, 1.9), Vec2(0, 1)},
                                                Vertex{Vec3(0.8, 0.4, 2), Vec2(1, 1)}, tri_mat));
    // a gradient-textured sphere exercises Texture::GRADIENT and sphere UVs (src/objects.cu:82-97)
    s.objects.push_back(Object::create_sphere(Vec3(-0.3, -0.3, 1.6), 0.12, make_standard(Texture::create_gradient(), 0.2)));
}

// id 3 -- src/main.cu:206-213
inline void refract_scene(Scene &s) {
    create_cornell_box(s, Vec3(-0.5, 0.5, 1.2), 1, 1, 1, 0.5);
    Material glass = make_refractive(Texture::create_const_colour(Vec3(1, 1, 1)), 1.5);
    s.objects.push_back(Object::create_sphere(Vec3(0, -0.1, 1.7), 0.3, glass));
}

// small seeded generator replacing std::random_device (src/main.cu:28-37), SURVEY.md 8(f) row 4
struct SceneRng {
    uint32_t s;
    explicit SceneRng(uint32_t seed) : s(seed) {}
    float next(float lo, float hi) {
        s = s * 1664525u + 1013904223u;
        float u = (float)(s >> 8) / 16777216.0f;
        return lo + u * (hi - lo);
    }
};

// id 4 -- src/main.cu:215-250, seeded, every material initialised
inline void rand_sphere_scene(Scene &s, int num_spheres, uint32_t seed) {
    s.use_sky = true;
    SceneRng rng(seed);
    float floor_y = -1, floor_width = 10, floor_depth = 10;
    for (int i = 0; i < num_spheres; i++) {
        float cr = rng.next(0, 1), cg = rng.next(0, 1), cb = rng.next(0, 1);
        Texture tex = Texture::create_const_colour(Vec3(cr, cg, cb));
        float mat_num = rng.next(0, 1);
        Material material;
        if (mat_num < 0.3f) {
            material = make_standard(tex, rng.next(0, 1));
        } else if (mat_num < 0.6f) {
            material = make_refractive(tex, rng.next(0.5, 2));
        } else if (mat_num < 0.7f) {
            material = make_emissive(Vec3(cr, cg, cb), 3);
        } else {
            material = make_standard(tex, 0);
        }
        float radius = rng.next(0.1, 0.5);
        float cx = rng.next(-floor_width / 2, floor_width / 2);
        float cz = rng.next(0, floor_depth);
        s.objects.push_back(Object::create_sphere(Vec3(cx, floor_y + radius, cz), radius, material));
    }
    Material floor_mat = make_standard(Texture::create_checkerboard(Vec3(0.7, 0.7, 0.7), Vec3(0.4, 0.4, 0.4), 10), 0);
    s.objects.push_back(Object::create_quad(Vec3(-floor_width / 2, floor_y, 0), Vec3(floor_width / 2, floor_y, 0),
                                            Vec3(floor_width / 2, floor_y, floor_depth), Vec3(-floor_width / 2, floor_y, floor_depth), floor_mat));
}

// id 5 -- C2: cube.obj + three spheres over a checker floor, sky on
inline void cube_spheres_scene(Scene &s) {
    s.use_sky = true;
    Material floor_mat = make_standard(Texture::create_checkerboard(Vec3(0.7, 0.7, 0.7), Vec3(0.4, 0.4, 0.4), 10), 0);
    s.objects.push_back(Object::create_quad(Vec3(-5, -1, 0), Vec3(5, -1, 0), Vec3(5, -1, 10), Vec3(-5, -1, 10), floor_mat));

    ObjFileMesh cube(models_dir() + "/cube.obj");
    cube.enlarge(0.5);
    cube.rotate(0, 0.6, 0);
    cube.translate(0, -0.5, 4);
    s.objects.push_back(create_mesh(cube, make_standard(Texture::create_const_colour(Vec3(1, 1, 1)), 0)));

    s.objects.push_back(Object::create_sphere(Vec3(-1.5, -0.5, 3.5), 0.5, make_standard(Texture::create_const_colour(Vec3(0.8, 0.3, 0.3)), 0)));
    s.objects.push_back(Object::create_sphere(Vec3(1.5, -0.5, 3.5), 0.5, make_standard(Texture::create_const_colour(Vec3(0.8, 0.8, 0.8)), 1)));
    s.objects.push_back(Object::create_sphere(Vec3(0, 1.5, 4), 0.5, make_emissive(Vec3(1, 1, 1), 6)));
}

// id 6 -- C4: nx * ny monkey instances merged into ONE mesh object
// (40 x 35 = 1400 instances = 1 012 200 triangles at full size), back wall, sky on.
inline void monkey_grid_scene(Scene &s, int nx, int ny) {
    s.use_sky = true;
    Material white = make_standard(Texture::create_const_colour(Vec3(1, 1, 1)), 0);
    std::vector<Triangle> triangles;
    for (int j = 0; j < ny; j++) {
        for (int i = 0; i < nx; i++) {
            ObjFileMesh m(models_dir() + "/low_poly_monkey.obj");
            m.enlarge(0.02);
            m.rotate(0, 2.3f + 0.1f * (float)(i + j), 0);
            m.translate(-1.1f + 2.2f * ((float)i + 0.5f) / (float)nx, -0.6f + 1.2f * ((float)j + 0.5f) / (float)ny, 2.0f);
            append_faces(triangles, m, white);
        }
    }
    s.objects.push_back(mesh_from_triangles(triangles, white));
    Material wall = make_standard(Texture::create_const_colour(Vec3(0.5, 0.5, 0.5)), 0);
    s.objects.push_back(Object::create_quad(Vec3(-3, 2, 2.5), Vec3(3, 2, 2.5), Vec3(3, -2, 2.5), Vec3(-3, -2, 2.5), wall));
}

// id 7 -- SURVEY.md 8(f) row 1: an OBJ with an MTL library (four materials, two of them on quads with
// `vt`, an image texture through the parsed-texture file) inside the Cornell box
inline void studio_scene(Scene &s) {
    create_cornell_box(s, Vec3(-0.5, 0.5, 1.2), 1, 1, 1, 0.5);
    ObjMtlMesh m(assets_dir() + "/models/studio.obj");
    m.enlarge(0.3);
    m.rotate(0.05, 0.4, 0.03);
    m.translate(0.0, -0.19, 1.7);
    Material fallback = make_standard(Texture::create_const_colour(Vec3(1, 1, 1)), 0);
    for (Object &o : m.create_objects(fallback, assets_dir() + "/textures")) s.objects.push_back(o);
}

// p0/p1: scene parameters (id 4: num_spheres, seed; id 6: nx, ny); 0 = default
inline Scene build(int scene_id, int p0 = 0, int p1 = 0) {
    Scene s;
    switch (scene_id) {
        case 0: monkey_scene(s); break;
        case 1: reflection_scene(s); break;
        case 2: texture_scene(s); break;
        case 3: refract_scene(s); break;
        case 4: rand_sphere_scene(s, p0 > 0 ? p0 : 100, p1 > 0 ? (uint32_t)p1 : 2024u); break;
        case 5: cube_spheres_scene(s); break;
        case 6: monkey_grid_scene(s, p0 > 0 ? p0 : 40, p1 > 0 ? p1 : 35); break;
        case 7: studio_scene(s); break;
        default: throw std::domain_error("Test scene must be a number between 0 and 7 (inclusive).\n");
    }
    return s;
}

// src/main.cu:13 and :318-330
inline RenderData render_settings(bool use_sky, int rays_per_pixel = 100, int reflect_limit = 5, bool antialias = true) {
    Vec3 sky = use_sky ? Vec3(0.8, 1, 1) : Vec3(0, 0, 0);
    return RenderData{rays_per_pixel, reflect_limit, antialias, sky};
}

}  // namespace rtb_scenes

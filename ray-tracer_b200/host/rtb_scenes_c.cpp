// rtb_scenes_c.cpp -- C entry points over the scene fixtures (scenes.hpp compiled against
// rtb_compat.hpp) so that Python (tests, bench.py) can obtain an RtbSceneDesc / camera for a scene
// id and hand it to librtb.so.  Host-only; built into librtb_scenes.so.
#include "rtb_compat.hpp"
#include "scenes.hpp"

namespace {
struct Handle {
    rtb_compat::SceneDescStorage storage;
    bool use_sky = false;
};
std::string g_err;
}  // namespace

extern "C" {

const char *rtbs_last_error() { return g_err.c_str(); }

void *rtbs_build(int scene_id, int p0, int p1, const char *models_dir) {
    try {
        rtb_compat::state().images.clear();
        rtb_scenes::models_dir() = models_dir;
        rtb_scenes::Scene s = rtb_scenes::build(scene_id, p0, p1);
        Handle *h = new Handle();
        rtb_compat::build_desc(s.objects.data(), (int)s.objects.size(), h->storage);
        // the image store is process-wide: take private copies so the handle stays valid
        static std::vector<std::vector<float>> keep;
        for (RtbImage &img : h->storage.images) {
            keep.emplace_back(img.rgb, img.rgb + (size_t)img.width * img.height * 3);
            img.rgb = keep.back().data();
        }
        h->use_sky = s.use_sky;
        return h;
    } catch (std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}

// directory of this repo's own fixtures (default: $RTB_ASSETS_DIR, else ray-tracer_b200/assets under cwd)
void rtbs_set_assets_dir(const char *dir) { rtb_scenes::assets_dir() = dir; }

// any OBJ (+ the MTL library it names, + map_Kd images from
// <texture_dir>/parsed_textures.{bin,txt}) as a scene: the mesh is scaled / rotated / translated like
// ObjFileMesh does (src/obj_read.cu:58-85) and split into one MESH object per material.
void *rtbs_load_obj_mtl(const char *obj_path, const char *texture_dir, float scale, const float rotate_xyz[3], const float translate_xyz[3]) {
    try {
        rtb_compat::state().images.clear();
        rtb_scenes::ObjMtlMesh m(obj_path);
        m.enlarge(scale);
        m.rotate(rotate_xyz[0], rotate_xyz[1], rotate_xyz[2]);
        m.translate(translate_xyz[0], translate_xyz[1], translate_xyz[2]);
        Material fallback = rtb_scenes::make_standard(Texture::create_const_colour(Vec3(1, 1, 1)), 0);
        std::vector<Object> objects = m.create_objects(fallback, texture_dir);
        Handle *h = new Handle();
        rtb_compat::build_desc(objects.data(), (int)objects.size(), h->storage);
        static std::vector<std::vector<float>> keep;
        for (RtbImage &img : h->storage.images) {
            keep.emplace_back(img.rgb, img.rgb + (size_t)img.width * img.height * 3);
            img.rgb = keep.back().data();
        }
        h->use_sky = true;
        return h;
    } catch (std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}

int rtbs_desc(void *handle, RtbSceneDesc *out) {
    if (!handle) return 1;
    *out = static_cast<Handle *>(handle)->storage.desc;
    return 0;
}

int rtbs_use_sky(void *handle) { return handle && static_cast<Handle *>(handle)->use_sky ? 1 : 0; }

void rtbs_free(void *handle) { delete static_cast<Handle *>(handle); }

// Camera().get_cam_data() at the default pose for a run-time resolution (src/camera.cu:46-108)
void rtbs_default_camera(int width, int height, RtbCamData *out) {
    rtb_compat::set_resolution(width, height);
    DeviceCamData d = Camera().get_cam_data();
    *out = RtbCamData{rtb_compat::to_rtb(d.cam_pos), rtb_compat::to_rtb(d.tl_pixel_pos), rtb_compat::to_rtb(d.delta_u),
                      rtb_compat::to_rtb(d.delta_v)};
}

// Camera at any pose: the members the reference marks "these values can be changed" (src/camera.cu:33-41)
// -- position, rotations about x / y / z in radians, field of view in radians, focal length
void rtbs_camera(int width, int height, const float pos[3], const float rot_xyz[3], float fov, float focal_len, RtbCamData *out) {
    rtb_compat::set_resolution(width, height);
    Camera cam;
    cam.CAM_POS = Vec3(pos[0], pos[1], pos[2]);
    cam.X_ROT = rot_xyz[0]; cam.Y_ROT = rot_xyz[1]; cam.Z_ROT = rot_xyz[2];
    if (fov > 0) cam.FOV = fov;
    if (focal_len > 0) cam.FOCAL_LEN = focal_len;
    DeviceCamData d = cam.get_cam_data();
    *out = RtbCamData{rtb_compat::to_rtb(d.cam_pos), rtb_compat::to_rtb(d.tl_pixel_pos), rtb_compat::to_rtb(d.delta_u),
                      rtb_compat::to_rtb(d.delta_v)};
}

// src/main.cu:13,318-330
void rtbs_render_settings(int use_sky, int spp, int bounces, int antialias, RtbRenderData *out) {
    RenderData r = rtb_scenes::render_settings(use_sky != 0, spp, bounces, antialias != 0);
    *out = RtbRenderData{r.rays_per_pixel, r.reflection_limit, r.antialias ? 1 : 0, rtb_compat::to_rtb(r.sky_colour)};
}

}  // extern "C"

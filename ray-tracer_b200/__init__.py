"""ray-tracer_b200 -- Python binding (ctypes) of the B200-native path tracer's C ABI.

Thin plumbing over ``librtb.so`` (include/rtb.h) and ``librtb_scenes.so`` (the scene fixtures built
through the C++ compat header that mirrors the reference's scene API).  Used by tests/, bench.py and
__graft_entry__.py.  There is NO fallback: importing works without a GPU (so the CPU suite can check
that the library loads and exports every symbol), but creating a :class:`Tracer` raises unless a
CUDA device is usable, and a missing ``librtb.so`` raises at import.

The directory name contains a hyphen (the repository layout asks for it); import it through the
``raytracer_b200`` shim module at the repository root.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)


from .abi import (ASSETS_DIR, MAT_EMISSIVE, MAT_REFRACTIVE, MAT_STANDARD, OBJ_CUBOID, OBJ_MESH, OBJ_ONE_WAY_QUAD, OBJ_QUAD,  # noqa: F401
                  OBJ_SPHERE, OBJ_TRIANGLE, REF_MODELS_DIR, TEX_CHECKERBOARD, TEX_COLOUR, TEX_GRADIENT, TEX_IMAGE, CamData, Counters,
                  Image, Material, Object, RenderData, RtbError, SceneDesc, Texture, Vec3, declared_symbols, models_dir)


def _load(name):
    # RTB_LIBRTB names an alternative build of librtb.so (A/B experiments of build-time knobs)
    path = os.environ.get("RTB_LIBRTB") if name == "librtb.so" and os.environ.get("RTB_LIBRTB") else os.path.join(_HERE, name)
    if not os.path.exists(path):
        raise RtbError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       f"(there is no CPU fallback for the render path)")
    return ctypes.CDLL(path)


lib = _load("librtb.so")
scenes_lib = _load("librtb_scenes.so")


_vp, _i, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
_fp, _ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)
lib.rtb_last_error.restype = ctypes.c_char_p
lib.rtb_last_error.argtypes = [_vp]
lib.rtb_create.argtypes = [ctypes.POINTER(_vp), _i]
lib.rtb_destroy.argtypes = [_vp]
lib.rtb_destroy.restype = None
lib.rtb_set_camera.argtypes = [_vp, ctypes.POINTER(CamData), _i, _i]
lib.rtb_set_render_data.argtypes = [_vp, ctypes.POINTER(RenderData)]
lib.rtb_scene_set.argtypes = [_vp, ctypes.POINTER(SceneDesc)]
lib.rtb_scene_desc.argtypes = [_vp, ctypes.POINTER(SceneDesc)]
lib.rtb_scene_commit.argtypes = [_vp]
lib.rtb_scene_update_object.argtypes = [_vp, _i, ctypes.POINTER(Object)]
lib.rtb_scene_update_mesh.argtypes = [_vp, _i, _vp, _vp]
lib.rtb_scene_begin.argtypes = [_vp]
lib.rtb_scene_commit.argtypes = [_vp]
lib.rtb_scene_desc.argtypes = [_vp, ctypes.POINTER(SceneDesc)]
lib.rtb_scene_add_material.argtypes = [_vp, ctypes.POINTER(Material)]
lib.rtb_scene_add_image.argtypes = [_vp, _vp, _i, _i]
lib.rtb_scene_add_sphere.argtypes = [_vp, _vp, _f, _i]
lib.rtb_scene_add_triangle.argtypes = [_vp, _vp, _vp, _i]
lib.rtb_scene_add_quad.argtypes = [_vp, _vp, _i]
lib.rtb_scene_add_one_way_quad.argtypes = [_vp, _vp, _i, _i]
lib.rtb_scene_add_cuboid.argtypes = [_vp, _vp, _f, _f, _f, _i]
lib.rtb_scene_add_mesh.argtypes = [_vp, _vp, _vp, _i, _i]
lib.rtb_render.argtypes = [_vp, _vp, _ip, _i]
lib.rtb_render_device.argtypes = [_vp, _i, _i, _i]
lib.rtb_sync.argtypes = [_vp]
lib.rtb_reset_accumulation.argtypes = [_vp]
lib.rtb_set_accumulation.argtypes = [_vp, _vp, _i]
lib.rtb_download.argtypes = [_vp, _vp, _ip]
lib.rtb_display_rgba8.argtypes = [_vp, _vp]
lib.rtb_write_ppm.argtypes = [_vp, ctypes.c_char_p]
lib.rtb_device_image.argtypes = [_vp]
lib.rtb_device_image.restype = _vp
lib.rtb_stream.argtypes = [_vp]
lib.rtb_stream.restype = _vp
lib.rtb_last_device_ms.argtypes = [_vp, _fp]
lib.rtb_set_shard.argtypes = [_vp, _i, _i, _i]
lib.rtb_shard_rows.argtypes = [_vp, _ip, _ip]
lib.rtb_shard_row_of.argtypes = [_i, _i, _i, _i]
lib.rtb_shard_row_count.argtypes = [_i, _i, _i, _i]
lib.rtb_shard_pack.argtypes = [_vp, _vp]
lib.rtb_shard_unpack.argtypes = [_vp, _vp]
lib.rtb_primary_ids.argtypes = [_vp, _vp, _vp, _vp]
lib.rtb_primary_ids_noalloc.argtypes = [_vp, _vp, _vp, _vp]
lib.rtb_noalloc_max_pixels.argtypes = []
lib.rtb_debug_mesh_partition.argtypes = [_i, _vp, _i, _i, _vp, _vp, _vp, _i, _ip]
lib.rtb_counters.argtypes = [_vp, ctypes.POINTER(Counters)]
lib.rtb_set_option.argtypes = [_vp, ctypes.c_char_p, _i]
lib.rtb_get_info.argtypes = [_vp, ctypes.c_char_p, ctypes.POINTER(ctypes.c_double)]
lib.rtb_measure_fp32_peak.argtypes = [_vp, ctypes.c_double, ctypes.POINTER(ctypes.c_double)]
lib.rtb_comm_unique_id.argtypes = [_vp]
lib.rtb_comm_init.argtypes = [_vp, _vp, _i, _i]
lib.rtb_comm_info.argtypes = [_vp, _ip, _ip]
lib.rtb_peer_export.argtypes = [_vp, _vp]
lib.rtb_peer_open.argtypes = [_vp, _vp, _i, _i]
lib.rtb_peer_close.argtypes = [_vp]
lib.rtb_create_multi.argtypes = [ctypes.POINTER(_vp), _ip, _i]
lib.rtb_render_multi.argtypes = [ctypes.POINTER(_vp), _i, _vp, _ip, _i]
lib.rtb_destroy_multi.argtypes = [ctypes.POINTER(_vp), _i]
lib.rtb_destroy_multi.restype = None

scenes_lib.rtbs_build.restype = _vp
scenes_lib.rtbs_build.argtypes = [_i, _i, _i, ctypes.c_char_p]
scenes_lib.rtbs_desc.argtypes = [_vp, ctypes.POINTER(SceneDesc)]
scenes_lib.rtbs_set_assets_dir.argtypes = [ctypes.c_char_p]
scenes_lib.rtbs_set_assets_dir.restype = None
scenes_lib.rtbs_load_obj_mtl.restype = _vp
scenes_lib.rtbs_load_obj_mtl.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_float, ctypes.POINTER(ctypes.c_float * 3),
                                         ctypes.POINTER(ctypes.c_float * 3)]
scenes_lib.rtbs_use_sky.argtypes = [_vp]
scenes_lib.rtbs_free.argtypes = [_vp]
scenes_lib.rtbs_free.restype = None
scenes_lib.rtbs_last_error.restype = ctypes.c_char_p
scenes_lib.rtbs_default_camera.argtypes = [_i, _i, ctypes.POINTER(CamData)]
scenes_lib.rtbs_camera.argtypes = [_i, _i, ctypes.POINTER(ctypes.c_float * 3), ctypes.POINTER(ctypes.c_float * 3), ctypes.c_float,
                                   ctypes.c_float, ctypes.POINTER(CamData)]
scenes_lib.rtbs_camera.restype = None
scenes_lib.rtbs_render_settings.argtypes = [_i, _i, _i, _i, ctypes.POINTER(RenderData)]


class Scene:
    """One of the scene fixtures of host/scenes.hpp, built through the compat mirror of the reference API.

    ids: 0 Cornell+monkey+sphere (C1/C3/C5), 1 reflection, 2 textures, 3 glass, 4 seeded random spheres
    (p0 = count, p1 = seed), 5 cube+spheres (C2), 6 monkey grid (C4; p0 x p1 instances), 7 studio.obj + studio.mtl
    (per-face materials, vt, image texture from the parsed-texture file).
    """

    def __init__(self, scene_id, p0=0, p1=0, models=None, assets=None):
        scenes_lib.rtbs_set_assets_dir((assets or os.environ.get("RTB_ASSETS_DIR") or ASSETS_DIR).encode())
        needs_models = scene_id in (0, 5, 6)
        mdir = (models or models_dir()) if needs_models else (models or "models")
        self.handle = scenes_lib.rtbs_build(scene_id, p0, p1, mdir.encode())
        if not self.handle:
            raise RtbError(scenes_lib.rtbs_last_error().decode())
        self.desc = SceneDesc()
        scenes_lib.rtbs_desc(self.handle, ctypes.byref(self.desc))
        self.use_sky = bool(scenes_lib.rtbs_use_sky(self.handle))
        self.scene_id = scene_id

    @classmethod
    def from_obj(cls, obj_path, texture_dir=None, scale=1.0, rotate=(0.0, 0.0, 0.0), translate=(0.0, 0.0, 0.0)):
        """An OBJ file with its MTL library (per-face materials, vt, map_Kd images from the parsed-texture
        file in texture_dir) as a scene of MESH objects, one per material (host/obj_mtl.hpp)."""
        self = cls.__new__(cls)
        rot, tra = (ctypes.c_float * 3)(*rotate), (ctypes.c_float * 3)(*translate)
        tdir = texture_dir or os.path.join(ASSETS_DIR, "textures")
        self.handle = scenes_lib.rtbs_load_obj_mtl(str(obj_path).encode(), str(tdir).encode(), scale, ctypes.byref(rot), ctypes.byref(tra))
        if not self.handle:
            raise RtbError(scenes_lib.rtbs_last_error().decode())
        self.desc = SceneDesc()
        scenes_lib.rtbs_desc(self.handle, ctypes.byref(self.desc))
        self.use_sky = True
        self.scene_id = -1
        return self

    def render_settings(self, spp, bounces, antialias=True):
        rd = RenderData()
        scenes_lib.rtbs_render_settings(int(self.use_sky), spp, bounces, int(antialias), ctypes.byref(rd))
        return rd

    def __del__(self):
        if getattr(self, "handle", None) and scenes_lib is not None:   # module globals are gone at interpreter shutdown
            scenes_lib.rtbs_free(self.handle)
            self.handle = None


def default_camera(width, height):
    cam = CamData()
    scenes_lib.rtbs_default_camera(width, height, ctypes.byref(cam))
    return cam


def camera(width, height, pos=(0.0, 0.0, 0.0), rot=(0.0, 0.0, 0.0), fov=0.0, focal_len=0.0):
    """Camera::assign_constant_mem's data (src/camera.cu:46-108) for any pose; rot in radians about x, y, z;
    fov / focal_len 0 = the reference's defaults (60 degrees, 0.1)"""
    cam = CamData()
    p, r = (ctypes.c_float * 3)(*pos), (ctypes.c_float * 3)(*rot)
    scenes_lib.rtbs_camera(width, height, ctypes.byref(p), ctypes.byref(r), fov, focal_len, ctypes.byref(cam))
    return cam


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class Tracer:
    """One rendering context on one CUDA device (rtb_ctx)."""

    def __init__(self, device=0, _ctx=None):
        self.ctx = ctypes.c_void_p()
        self._owned = _ctx is None
        if _ctx is not None:
            self.ctx = _ctx
        elif lib.rtb_create(ctypes.byref(self.ctx), device) != 0:
            msg = lib.rtb_last_error(None).decode()
            self.ctx = None
            raise RtbError(msg)
        self.width = self.height = 0

    def _check(self, rc):
        if rc != 0:
            raise RtbError(lib.rtb_last_error(self.ctx).decode())

    def close(self):
        if self.ctx and self._owned:
            lib.rtb_destroy(self.ctx)
        self.ctx = None

    # ---- multi-GPU: one process per GPU (the 128-byte id travels over the caller's own launcher) ----
    @staticmethod
    def comm_unique_id():
        buf = (ctypes.c_ubyte * 128)()
        if lib.rtb_comm_unique_id(buf) != 0:
            raise RtbError(lib.rtb_last_error(None).decode())
        return bytes(buf)

    def comm_init(self, unique_id, world, rank):
        buf = (ctypes.c_ubyte * 128).from_buffer_copy(unique_id)
        self._check(lib.rtb_comm_init(self.ctx, buf, world, rank))

    def peer_export(self):
        buf = (ctypes.c_ubyte * 64)()
        self._check(lib.rtb_peer_export(self.ctx, buf))
        return bytes(buf)

    def peer_open(self, handles, world, rank):
        """handles: world x 64 bytes in rank order (rtb_peer_export of every rank)"""
        buf = (ctypes.c_ubyte * (64 * world)).from_buffer_copy(handles)
        self._check(lib.rtb_peer_open(self.ctx, buf, world, rank))

    def peer_close(self):
        self._check(lib.rtb_peer_close(self.ctx))

    def __del__(self):
        self.close()

    def set_camera(self, cam, width, height):
        self._check(lib.rtb_set_camera(self.ctx, ctypes.byref(cam), width, height))
        self.width, self.height = width, height

    def set_render_data(self, rd):
        self._check(lib.rtb_set_render_data(self.ctx, ctypes.byref(rd)))
        self.rd = rd

    def set_scene(self, scene):
        desc = scene.desc if isinstance(scene, Scene) else scene
        self._check(lib.rtb_scene_set(self.ctx, ctypes.byref(desc)))

    def setup(self, scene, width, height, spp, bounces, antialias=True):
        self.set_camera(default_camera(width, height), width, height)
        self.set_render_data(scene.render_settings(spp, bounces, antialias))
        self.set_scene(scene)
        return self

    # dynamic scenes: edit the committed description in place, commit again
    def scene_desc(self):
        """view of the description the context holds (valid until the next scene begin / set)"""
        d = SceneDesc()
        self._check(lib.rtb_scene_desc(self.ctx, ctypes.byref(d)))
        return d

    def update_object(self, index, obj):
        self._check(lib.rtb_scene_update_object(self.ctx, index, ctypes.byref(obj)))

    def update_mesh(self, index, tri_points, tri_uv=None):
        tri_points = np.ascontiguousarray(tri_points, np.float32)
        if tri_uv is not None:
            tri_uv = np.ascontiguousarray(tri_uv, np.float32)
        self._check(lib.rtb_scene_update_mesh(self.ctx, index, _ptr(tri_points), _ptr(tri_uv) if tri_uv is not None else None))

    def commit(self):
        self._check(lib.rtb_scene_commit(self.ctx))

    def set_option(self, key, value):
        self._check(lib.rtb_set_option(self.ctx, key.encode(), int(value)))

    def info(self, key):
        v = ctypes.c_double()
        self._check(lib.rtb_get_info(self.ctx, key.encode(), ctypes.byref(v)))
        return v.value

    def render(self, image, frame_num, time_ms):
        """== render(VariableRenderData*, int) (src/dispatch.cu:156): image (H*W*3 float32, host) in/out."""
        assert image.dtype == np.float32 and image.size == self.width * self.height * 3 and image.flags.c_contiguous
        f = ctypes.c_int(frame_num)
        self._check(lib.rtb_render(self.ctx, _ptr(image), ctypes.byref(f), time_ms))
        return f.value

    def render_device(self, n_frames, time0_ms, time_step_ms=1):
        self._check(lib.rtb_render_device(self.ctx, n_frames, time0_ms, time_step_ms))

    def sync(self):
        self._check(lib.rtb_sync(self.ctx))

    def reset(self):
        self._check(lib.rtb_reset_accumulation(self.ctx))

    def set_accumulation(self, image, frame_num):
        self._check(lib.rtb_set_accumulation(self.ctx, _ptr(np.ascontiguousarray(image, np.float32)), frame_num))

    def download(self):
        img = np.empty((self.height, self.width, 3), np.float32)
        f = ctypes.c_int()
        self._check(lib.rtb_download(self.ctx, _ptr(img), ctypes.byref(f)))
        return img, f.value

    def display_rgba8(self):
        """== parse_pixel_colours (src/main.cu:343-371) of the accumulation image, on the device."""
        out = np.empty((self.height, self.width, 4), np.uint8)
        self._check(lib.rtb_display_rgba8(self.ctx, _ptr(out)))
        return out

    def write_ppm(self, path):
        self._check(lib.rtb_write_ppm(self.ctx, str(path).encode()))

    def device_ms(self):
        ms = ctypes.c_float()
        self._check(lib.rtb_last_device_ms(self.ctx, ctypes.byref(ms)))
        return ms.value

    def device_image_ptr(self):
        p = lib.rtb_device_image(self.ctx)
        if not p:
            raise RtbError(lib.rtb_last_error(self.ctx).decode())
        return p

    def stream(self):
        return lib.rtb_stream(self.ctx)

    def set_shard(self, rank, world, band_rows=8):
        self._check(lib.rtb_set_shard(self.ctx, rank, world, band_rows))

    def shard_rows(self):
        a, b = ctypes.c_int(), ctypes.c_int()
        self._check(lib.rtb_shard_rows(self.ctx, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def shard_pack(self, dense_dev_ptr):
        self._check(lib.rtb_shard_pack(self.ctx, dense_dev_ptr))

    def shard_unpack(self, gathered_dev_ptr):
        self._check(lib.rtb_shard_unpack(self.ctx, gathered_dev_ptr))

    def primary_ids(self):
        n = self.width * self.height
        obj, sub, t = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.float32)
        self._check(lib.rtb_primary_ids(self.ctx, _ptr(obj), _ptr(sub), _ptr(t)))
        shape = (self.height, self.width)
        return obj.reshape(shape), sub.reshape(shape), t.reshape(shape)

    def primary_ids_noalloc(self):
        """rtb_primary_ids_noalloc: the same pass with the staging planes in the library's static device array (no
        cudaMalloc in the call); refused beyond rtb_noalloc_max_pixels()"""
        n = self.width * self.height
        obj, sub, t = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.float32)
        self._check(lib.rtb_primary_ids_noalloc(self.ctx, _ptr(obj), _ptr(sub), _ptr(t)))
        shape = (self.height, self.width)
        return obj.reshape(shape), sub.reshape(shape), t.reshape(shape)

    def fp32_peak_tflops(self, ms_target=50.0):
        v = ctypes.c_double()
        self._check(lib.rtb_measure_fp32_peak(self.ctx, ms_target, ctypes.byref(v)))
        return v.value

    def counters(self):
        c = Counters()
        self._check(lib.rtb_counters(self.ctx, ctypes.byref(c)))
        return c.as_dict()


class MultiTracer:
    """n rendering contexts on n CUDA devices of THIS process with one NCCL communicator (rtb_create_multi);
    render() is the n-GPU form of render(VariableRenderData*, int)."""

    def __init__(self, devices):
        self.n = len(devices)
        self._ctxs = (ctypes.c_void_p * self.n)()
        devs = (ctypes.c_int * self.n)(*devices)
        if lib.rtb_create_multi(self._ctxs, devs, self.n) != 0:
            raise RtbError(lib.rtb_last_error(None).decode())
        self.tracers = [Tracer(_ctx=ctypes.c_void_p(self._ctxs[i])) for i in range(self.n)]

    def setup(self, scene, width, height, spp, bounces, antialias=True):
        for t in self.tracers:
            t.setup(scene, width, height, spp, bounces, antialias)
        self.width, self.height = width, height
        return self

    def set_option(self, key, value):
        for t in self.tracers:
            t.set_option(key, value)

    def render(self, image, frame_num, time_ms):
        assert image.dtype == np.float32 and image.size == self.width * self.height * 3 and image.flags.c_contiguous
        f = ctypes.c_int(frame_num)
        if lib.rtb_render_multi(self._ctxs, self.n, _ptr(image), ctypes.byref(f), time_ms) != 0:
            raise RtbError(lib.rtb_last_error(self._ctxs[0]).decode())
        return f.value

    def close(self):
        if self._ctxs is not None:
            for t in self.tracers:
                t.ctx = None
            lib.rtb_destroy_multi(self._ctxs, self.n)
            self._ctxs = None

    def __del__(self):
        self.close()


def shard_rows_host(height, rank, world, band_rows):
    """Rows owned by `rank` under the row-cyclic band sharding (host mirror of rtb_set_shard)."""
    return [y for y in range(height) if (y // band_rows) % world == rank]

"""TEST INFRASTRUCTURE: ctypes bindings of the two checkers.

  Oracle  -> oracle/libpt_oracle.so   (plain-C restatement, always buildable)
  RefHost -> oracle/_ref/libref_host.so (the reference's own source compiled for the host; built only
             where /root/reference exists, travels to the GPU box as a binary)
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this module.
"""
import ctypes
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load_abi():
    """The ctypes mirror of include/rtb.h (ray-tracer_b200/abi.py), loaded WITHOUT the package around it: importing
    the package loads librtb.so, and the checkers must also work in processes that may not contain the product
    (bench.py --impl reference).  Registered under the package's own submodule name, so a later
    `import raytracer_b200` shares these classes."""
    name = "raytracer_b200.abi"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "ray-tracer_b200", "abi.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


rtb = _load_abi()  # structure definitions and fixture paths only
_ORACLE = os.path.join(ROOT, "oracle", "libpt_oracle.so")
_REF = os.path.join(ROOT, "oracle", "_ref", "libref_host.so")
NTHREADS = max(1, min(os.cpu_count() or 1, 32))


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


class Oracle:
    lib = None

    def __init__(self, scene):
        if Oracle.lib is None:
            L = ctypes.CDLL(_ORACLE)
            L.pto_scene_create.restype = ctypes.c_void_p
            L.pto_scene_create.argtypes = [ctypes.POINTER(rtb.SceneDesc)]
            L.pto_scene_destroy.argtypes = [ctypes.c_void_p]
            L.pto_render.argtypes = [ctypes.c_void_p, ctypes.POINTER(rtb.CamData), ctypes.c_int, ctypes.c_int,
                                     ctypes.POINTER(rtb.RenderData), ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                     ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(rtb.Counters)]
            L.pto_primary_ids.argtypes = [ctypes.c_void_p, ctypes.POINTER(rtb.CamData), ctypes.c_int, ctypes.c_int,
                                          ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
            L.pto_seed.restype = ctypes.c_uint
            L.pto_rng.restype = ctypes.c_float
            L.pto_rng.argtypes = [ctypes.POINTER(ctypes.c_uint)]
            L.pto_normal.restype = ctypes.c_float
            L.pto_normal.argtypes = [ctypes.POINTER(ctypes.c_uint)]
            L.pto_default_camera.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(rtb.CamData)]
            L.pto_display_rgba8.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
            L.pto_primary_dir.argtypes = [ctypes.POINTER(rtb.CamData), ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
            Oracle.lib = L
        desc = scene.desc if hasattr(scene, "desc") else scene
        self._keep = scene
        self.scene = Oracle.lib.pto_scene_create(ctypes.byref(desc))

    def __del__(self):
        if getattr(self, "scene", None):
            Oracle.lib.pto_scene_destroy(self.scene)
            self.scene = None

    def render(self, cam, w, h, rd, time_ms, frame_num, prev, y0=0, y1=None, nthreads=NTHREADS, counters=None):
        out = np.zeros((h, w, 3), np.float32)
        if prev is not None:
            prev = np.ascontiguousarray(prev, np.float32)
        Oracle.lib.pto_render(self.scene, ctypes.byref(cam), w, h, ctypes.byref(rd), time_ms, frame_num, _p(prev), _p(out),
                              y0, h if y1 is None else y1, nthreads, ctypes.byref(counters) if counters is not None else None)
        return out

    def primary_ids(self, cam, w, h, nthreads=NTHREADS):
        obj, sub, t = np.empty((h, w), np.int32), np.empty((h, w), np.int32), np.empty((h, w), np.float32)
        Oracle.lib.pto_primary_ids(self.scene, ctypes.byref(cam), w, h, _p(obj), _p(sub), _p(t), nthreads)
        return obj, sub, t


def oracle_display_rgba8(image):
    """parse_pixel_colours (src/main.cu:343-371) through the C oracle; image H x W x 3 float32"""
    if Oracle.lib is None:
        Oracle(rtb.SceneDesc())
    image = np.ascontiguousarray(image, np.float32)
    h, w = image.shape[:2]
    out = np.empty((h, w, 4), np.uint8)
    Oracle.lib.pto_display_rgba8(_p(image), w, h, _p(out))
    return out


def have_ref_host():
    return os.path.exists(_REF)


class RefHost:
    """The reference's own source on the host.  Process-global state (like the reference's
    __constant__ symbols): one scene at a time."""
    lib = None

    def __init__(self):
        if RefHost.lib is None:
            L = ctypes.CDLL(_REF)
            L.refh_last_error.restype = ctypes.c_char_p
            L.refh_seed.restype = ctypes.c_uint
            L.refh_normal.restype = ctypes.c_float
            L.refh_rng_value.restype = ctypes.c_float
            L.refh_set_render.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_float]
            RefHost.lib = L
        self.L = RefHost.lib

    def build_scene(self, scene_id, p0=0, p1=0, from_main=False):
        rc = self.L.refh_build_scene(scene_id, p0, p1, rtb.models_dir().encode(), int(from_main))
        if rc != 0:
            raise RuntimeError(self.L.refh_last_error().decode())

    def set_render(self, rd):
        self.L.refh_set_render(rd.rays_per_pixel, rd.reflection_limit, rd.antialias, rd.sky_colour.x, rd.sky_colour.y, rd.sky_colour.z)

    def set_camera(self, w, h):
        cam = (ctypes.c_float * 12)()
        self.L.refh_set_camera(w, h, cam)
        return np.array(cam, np.float32)

    def render(self, w, h, time_ms, frame_num, prev, nthreads=NTHREADS):
        out = np.zeros((h, w, 3), np.float32)
        if prev is not None:
            prev = np.ascontiguousarray(prev, np.float32)
        self.L.refh_render(w, h, 0, h, time_ms, frame_num, _p(prev), _p(out), nthreads)
        return out

    def primary_ids(self, w, h, nthreads=NTHREADS):
        obj, sub, t = np.empty((h, w), np.int32), np.empty((h, w), np.int32), np.empty((h, w), np.float32)
        self.L.refh_primary_ids(w, h, _p(obj), _p(sub), _p(t), nthreads)
        return obj, sub, t


def ref_cuda_path(w, h, nofma=False):
    """the reference's own CUDA build at w x h; nofma = the same sources compiled with -fmad=false (SURVEY.md H2 i)"""
    return os.path.join(ROOT, "baseline", "_ref", f"libref_cuda_{'nofma_' if nofma else ''}{w}x{h}.so")


class RefCuda:
    """The reference's own CUDA build (baseline/ref_cuda.cu), one library per resolution.
    Process-global state (its three __constant__ symbols)."""
    _libs = {}

    def __init__(self, w, h, nofma=False):
        path = ref_cuda_path(w, h, nofma)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        if (w, h, nofma) not in RefCuda._libs:
            L = ctypes.CDLL(path)
            L.refc_last_error.restype = ctypes.c_char_p
            L.refc_init.argtypes = [ctypes.c_int] * 3 + [ctypes.c_char_p] + [ctypes.c_int] * 3
            RefCuda._libs[(w, h, nofma)] = L
        self.L = RefCuda._libs[(w, h, nofma)]
        self.w, self.h = w, h
        assert self.L.refc_width() == w and self.L.refc_height() == h

    def init(self, scene_id, p0, p1, spp, bounces, antialias=True):
        if self.L.refc_init(scene_id, p0, p1, rtb.models_dir().encode(), spp, bounces, int(antialias)) != 0:
            raise RuntimeError(self.L.refc_last_error().decode())

    def render_kernel(self, image, frame_num, time_ms):
        """kernel-only device time (ms); image in/out like render()"""
        ms = ctypes.c_float()
        if self.L.refc_render_kernel(_p(image), frame_num, time_ms, ctypes.byref(ms)) != 0:
            raise RuntimeError(self.L.refc_last_error().decode())
        return ms.value

    def render(self, image, frame_num, time_ms):
        """the reference's own render(): returns (new frame_num, wall ms)"""
        f, ms = ctypes.c_int(frame_num), ctypes.c_double()
        if self.L.refc_render(_p(image), ctypes.byref(f), time_ms, ctypes.byref(ms)) != 0:
            raise RuntimeError(self.L.refc_last_error().decode())
        return f.value, ms.value

    def primary_ids(self):
        obj, t = np.empty((self.h, self.w), np.int32), np.empty((self.h, self.w), np.float32)
        if self.L.refc_primary_ids(_p(obj), _p(t)) != 0:
            raise RuntimeError(self.L.refc_last_error().decode())
        return obj, t

    def primary_sub_ids(self, obj):
        """sub id of every pixel's winning primitive (brute force over a mesh's triangles: small scenes only)"""
        sub = np.empty((self.h, self.w), np.int32)
        if self.L.refc_primary_sub_ids(_p(np.ascontiguousarray(obj, np.int32)), _p(sub)) != 0:
            raise RuntimeError(self.L.refc_last_error().decode())
        return sub

    def count_segments(self, time_ms):
        """(segments, hit segments) of one frame at the current settings, counted by the reference's own code on the GPU"""
        seg, hits = ctypes.c_ulonglong(), ctypes.c_ulonglong()
        if self.L.refc_count_segments(time_ms, ctypes.byref(seg), ctypes.byref(hits)) != 0:
            raise RuntimeError(self.L.refc_last_error().decode())
        return seg.value, hits.value

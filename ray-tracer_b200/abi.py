"""Plain-Python mirror of include/rtb.h: the ctypes structures, enums and fixture paths of the C ABI.

Loads no library, so the test infrastructure (tests/oracle_binding.py) and bench.py's reference arm can use the
structure definitions without pulling librtb.so into a process that must not contain the product.  The package
__init__ re-exports everything defined here."""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)


class RtbError(RuntimeError):
    """Mirrors the std::runtime_error the reference's check_cuda_error throws (src/utils.cu:5-10)."""


class Vec3(ctypes.Structure):
    _fields_ = [("x", ctypes.c_float), ("y", ctypes.c_float), ("z", ctypes.c_float)]


class CamData(ctypes.Structure):  # RtbCamData == DeviceCamData (src/camera.cu:12-18)
    _fields_ = [("cam_pos", Vec3), ("tl_pixel_pos", Vec3), ("delta_u", Vec3), ("delta_v", Vec3)]


class RenderData(ctypes.Structure):  # RtbRenderData == RenderData (src/raytracer.cu:4-12)
    _fields_ = [("rays_per_pixel", ctypes.c_int), ("reflection_limit", ctypes.c_int), ("antialias", ctypes.c_int),
                ("sky_colour", Vec3)]


class Texture(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int), ("colour", Vec3), ("light", Vec3), ("dark", Vec3), ("num_squares", ctypes.c_int),
                ("img_width", ctypes.c_int), ("img_height", ctypes.c_int), ("image", ctypes.c_int)]


class Material(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int), ("texture", Texture), ("smoothness", ctypes.c_float), ("need_uv", ctypes.c_int),
                ("emitted_light", Vec3), ("refractive_index", ctypes.c_float)]


class Object(ctypes.Structure):
    _fields_ = [("type", ctypes.c_int), ("material", ctypes.c_int), ("p", Vec3 * 4), ("f", ctypes.c_float * 3),
                ("uv", ctypes.c_float * 6), ("invert_normal", ctypes.c_int), ("tri_first", ctypes.c_int),
                ("tri_count", ctypes.c_int)]


class Image(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int), ("height", ctypes.c_int), ("rgb", ctypes.POINTER(ctypes.c_float))]


class SceneDesc(ctypes.Structure):
    _fields_ = [("num_objects", ctypes.c_int), ("objects", ctypes.POINTER(Object)),
                ("num_materials", ctypes.c_int), ("materials", ctypes.POINTER(Material)),
                ("num_mesh_tris", ctypes.c_int), ("mesh_tri_points", ctypes.POINTER(ctypes.c_float)),
                ("mesh_tri_uv", ctypes.POINTER(ctypes.c_float)),
                ("num_images", ctypes.c_int), ("images", ctypes.POINTER(Image))]


class Counters(ctypes.Structure):
    _fields_ = [(n, ctypes.c_ulonglong) for n in ("segments", "hit_segments", "box_tests", "tri_tests", "sphere_tests", "samples")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def declared_symbols(header=os.path.join(REPO_ROOT, "include", "rtb.h")):
    """Every function name include/rtb.h declares (used by the symbol-export test)."""
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtb_[a-z_0-9]+)\s*\(", text)))


# enums of include/rtb.h
TEX_COLOUR, TEX_GRADIENT, TEX_CHECKERBOARD, TEX_IMAGE = 0, 1, 2, 3
MAT_STANDARD, MAT_EMISSIVE, MAT_REFRACTIVE = 0, 1, 2
OBJ_SPHERE, OBJ_TRIANGLE, OBJ_QUAD, OBJ_ONE_WAY_QUAD, OBJ_CUBOID, OBJ_MESH = 0, 1, 2, 3, 4, 5

# this repo's own fixtures (OBJ + MTL, images, parsed-texture files); host/obj_mtl.hpp reads the variable
ASSETS_DIR = os.path.join(_HERE, "assets")
os.environ.setdefault("RTB_ASSETS_DIR", ASSETS_DIR)
# cube.obj / low_poly_monkey.obj are reference DATA: never committed; __graft_entry__.build() copies them here
# (git-ignored, travels to the GPU box with the snapshot)
REF_MODELS_DIR = os.path.join(ASSETS_DIR, "_ref_models")


def models_dir():
    """Directory holding cube.obj / low_poly_monkey.obj (reference DATA, never committed here)."""
    cands = [os.environ.get("RTB_MODELS_DIR"), REF_MODELS_DIR, "/root/reference/models"]
    for c in cands:
        if c and os.path.exists(os.path.join(c, "low_poly_monkey.obj")):
            return c
    raise RtbError("OBJ fixtures not found (set RTB_MODELS_DIR, or run __graft_entry__.build() where /root/reference exists)")

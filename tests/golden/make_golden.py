"""TEST INFRASTRUCTURE: generates the golden fixtures under tests/golden/ from the REFERENCE ITSELF.

The reference has no tests and no golden vectors of its own, so the fixtures are
outputs of the reference's own source compiled for the host (oracle/_ref/libref_host.so, built by
oracle/Makefile from /root/reference/src in place; nothing is copied).  Run in the build container
(where /root/reference exists):

    python tests/golden/make_golden.py            # everything
    python tests/golden/make_golden.py 7          # only the fixtures of the listed scene ids

Per scene fixture (0..7): primary-hit ids / sub ids / distances with antialias off, and two
progressive frames of radiance at matched RNG streams.  kat.json holds the scalar known answers
(seed, RNG stream, Box-Muller, camera, primary directions, struct sizes, INF).
"""
import ctypes
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_binding as ob  # noqa: E402
import raytracer_b200 as rtb  # noqa: E402

# (scene id, p0, p1): the same list the GPU parity tests use
SCENES = [(0, 0, 0), (1, 0, 0), (2, 0, 0), (3, 0, 0), (4, 60, 7), (5, 0, 0), (6, 4, 3), (7, 0, 0)]
W, H, SPP, BOUNCES = 96, 80, 4, 5
TIMES = (12345, 777)


def main():
    ref = ob.RefHost()
    L = ref.L
    only = [int(a) for a in sys.argv[1:]]
    for sid, p0, p1 in SCENES:
        if only and sid not in only:
            continue
        ref.build_scene(sid, p0, p1)
        scene = rtb.Scene(sid, p0, p1)
        rd = scene.render_settings(SPP, BOUNCES)
        ref.set_render(rd)
        cam = ref.set_camera(W, H)
        obj, sub, t = ref.primary_ids(W, H)
        f0 = ref.render(W, H, TIMES[0], 0, None)
        f1 = ref.render(W, H, TIMES[1], 1, f0)
        np.savez_compressed(os.path.join(HERE, f"scene{sid}.npz"), scene=np.array([sid, p0, p1], np.int32),
                            size=np.array([W, H, SPP, BOUNCES], np.int32), times=np.array(TIMES, np.int32), cam=cam,
                            obj=obj.astype(np.int16), sub=sub, t=t, frame0=f0, frame1=f1)
        print(f"scene {sid}: {int((obj >= 0).sum())} hit pixels, mean radiance {f1.mean():.6f}")

    if only:
        return
    kat = {}
    kat["inf"] = L.refh_inf()
    sizes = (ctypes.c_int * 32)()
    L.refh_sizes(sizes, 32)
    kat["sizes"] = list(sizes)
    kat["seed"] = {"0,12345": L.refh_seed(0, 12345), "2998500,12345": L.refh_seed(2998500, 12345),
                   "2399997,-5": L.refh_seed(2399997, -5)}
    vals, states = (ctypes.c_float * 16)(), (ctypes.c_uint * 16)()
    L.refh_rng.argtypes = [ctypes.c_uint, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    L.refh_rng(358773477, 16, vals, states)
    kat["rng_from_358773477"] = {"values_hex": [np.float32(v).tobytes().hex() for v in vals], "states": list(states)}
    L.refh_rng_value.argtypes = [ctypes.c_uint]
    edge = [0, 1, 127, 128, 129, 255, 4294967166, 4294967167, 4294967168, 4294967294, 4294967295, 2147483648, 16777216, 16777217]
    kat["rng_value"] = {str(r): np.float32(L.refh_rng_value(r)).tobytes().hex() for r in edge}
    L.refh_normal.argtypes = [ctypes.c_uint, ctypes.POINTER(ctypes.c_uint)]
    normals = {}
    for s0 in (1, 12345, 358773477, 4000000000):
        so = ctypes.c_uint()
        v = L.refh_normal(s0, ctypes.byref(so))
        normals[str(s0)] = {"value": float(v), "state": so.value}
    kat["normal"] = normals  # depends on the host libm (logf, cosf): compared with a tolerance
    cams = {}
    for (w, h) in ((1000, 800), (1920, 1080), (96, 80)):
        cam = ref.set_camera(w, h)
        d = (ctypes.c_float * 3)()
        dirs = {}
        for (x, y) in ((0, 0), (w - 1, h - 1), (w // 2, h // 2), (7, 3)):
            L.refh_primary_dir(x, y, d)
            dirs[f"{x},{y}"] = [np.float32(v).tobytes().hex() for v in d]
        cams[f"{w}x{h}"] = {"cam_hex": [np.float32(v).tobytes().hex() for v in cam], "dirs": dirs}
    kat["camera"] = cams
    json.dump(kat, open(os.path.join(HERE, "kat.json"), "w"), indent=1)
    print("kat.json written")


if __name__ == "__main__":
    main()

"""commit cost of dynamic scenes (diagnostic; run on the GPU box): python tools/dyn_probe.py"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracer_b200 as rtb

for name, args in (("scene 0 (733 prims)", (0, 0, 0)), ("scene 4, 2000 spheres", (4, 2000, 3)), ("scene 4, 20000 spheres", (4, 20000, 3)),
                   ("scene 6, 10x10 monkeys (72k tris)", (6, 10, 10)), ("scene 6, 40x35 monkeys (1.01M tris)", (6, 40, 35))):
    sc = rtb.Scene(*args)
    tr = rtb.Tracer(0).setup(sc, 640, 360, 1, 4)
    c, b = [], []
    for k in range(4):
        tr.commit()
        c.append(tr.info("commit_ms")); b.append(tr.info("build_ms"))
    print(f"{name}: prims {int(tr.info('num_prims'))}, commit {min(c):.2f} ms wall (flatten + upload + build), of which tree build {min(b):.2f} ms on the device", flush=True)
    tr.close()

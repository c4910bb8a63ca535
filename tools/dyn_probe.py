"""commit cost of dynamic scenes (diagnostic; run on the GPU box): python tools/dyn_probe.py
Per scene: the commit (wall) and its parts, the steady frame, and commit + the first frame after it -- the first
frame after a commit would carry the timed kernel choice if the commit re-armed it (it does not, unless the
primitive count changed by more than 2x: ADVICE round 1)."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracer_b200 as rtb

for name, args in (("scene 0 (733 prims)", (0, 0, 0)), ("scene 4, 2000 spheres", (4, 2000, 3)), ("scene 4, 20000 spheres", (4, 20000, 3)),
                   ("scene 6, 10x10 monkeys (72k tris)", (6, 10, 10)), ("scene 6, 40x35 monkeys (1.01M tris)", (6, 40, 35))):
    sc = rtb.Scene(*args)
    tr = rtb.Tracer(0).setup(sc, 640, 360, 1, 4)
    tr.reset(); tr.render_device(1, 1); tr.sync()          # first frame ever: includes the timed kernel choice
    tr.reset(); tr.render_device(1, 2); tr.sync()
    steady = tr.device_ms()
    rows = []
    for k in range(4):
        t0 = time.perf_counter()
        tr.commit()
        t1 = time.perf_counter()
        tr.reset(); tr.render_device(1, 10 + k); tr.sync()
        t2 = time.perf_counter()
        rows.append(dict(commit=tr.info("commit_ms"), build=tr.info("build_ms"), partition=tr.info("commit_partition_ms"), flatten=tr.info("commit_flatten_ms"),
                         upload=tr.info("commit_upload_ms"), commit_wall=(t1 - t0) * 1e3, first_frame_wall=(t2 - t1) * 1e3, first_frame_device=tr.device_ms()))
    best = min(rows, key=lambda r: r["commit"])
    print(json.dumps(dict(scene=name, prims=int(tr.info("num_prims")), steady_frame_ms=round(steady, 3), **{k: round(v, 3) for k, v in best.items()})), flush=True)
    tr.close()

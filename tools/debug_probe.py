"""Ad-hoc GPU investigation script (not a test)."""
import sys, json
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import raytracer_b200 as rtb, oracle_binding as ob

W, H = 320, 256
for sid, p0, p1 in [(5, 0, 0), (4, 60, 7), (3, 0, 0)]:
    sc = rtb.Scene(sid, p0, p1)
    orc = ob.Oracle(sc)
    cam = rtb.default_camera(W, H)
    ref = orc.render(cam, W, H, sc.render_settings(8, 5), 12345, 0, None)
    o_obj, _, _ = orc.primary_ids(cam, W, H)
    for strict in (0, 1):
        tr = rtb.Tracer(0).setup(sc, W, H, 8, 5)
        tr.set_option("strict_fp", strict)
        img = np.zeros((H, W, 3), np.float32); tr.render(img, 0, 12345)
        d = np.abs(img - ref).max(axis=2)
        hist = [int((d > th).sum()) for th in (1e-6, 1e-4, 1e-3, 1e-2, 2e-2, 1e-1)]
        ys, xs = np.nonzero(d > 0.02)
        objs = collections = {}
        import collections as C
        cnt = C.Counter(int(o_obj[y, x]) for y, x in zip(ys, xs))
        print(f"scene {sid} strict={strict}: pixels with diff > [1e-6,1e-4,1e-3,1e-2,2e-2,1e-1] = {hist}; primary obj of bad pixels {dict(cnt)}; sample {list(zip(xs[:6].tolist(), ys[:6].tolist()))}", flush=True)
        tr.close()

# C4 regression: large split on/off
for split in (0, 1):
    sc = rtb.Scene(6, 40, 35)
    tr = rtb.Tracer(0)
    tr.set_option("large_split", split)
    tr.setup(sc, 1920, 1080, 1, 8)
    tr.set_option("counters", 1); tr.render_device(1, 7); tr.sync(); c = tr.counters(); tr.set_option("counters", 0)
    tr.reset(); tr.render_device(1, 8); tr.sync(); ms = tr.device_ms()
    print("C4 split", split, "num_large", tr.info("num_large"), "ms", ms, {k: v / c["segments"] for k, v in c.items()}, flush=True)
    tr.close()

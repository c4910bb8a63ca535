"""Ad-hoc GPU probe: C4 full frame at 16 spp, 40 frames in a row: with a sync after every frame, then enqueued back to back."""
import sys, json, time
sys.path.insert(0, ".")
import raytracer_b200 as rtb
sc = rtb.Scene(6, 40, 35)
tr = rtb.Tracer(0)
tr.set_option("kernel", 1); tr.set_option("walk_min", 12); tr.set_option("wf_refill", 16)
tr.setup(sc, 3840, 2160, 16, 8)
tr.reset(); tr.render_device(1, 999); tr.sync()
ms = []
for k in range(40):
    tr.reset(); tr.render_device(1, 1000 + k); tr.sync(); ms.append(round(tr.device_ms(), 1))
print(json.dumps(dict(mode="sync after every frame", ms=ms[::4])), flush=True)
for n in (5, 40):
    t0 = time.perf_counter()
    for k in range(n):
        tr.reset(); tr.render_device(1, 1000 + k)
    tr.sync()
    print(json.dumps(dict(mode=f"{n} frames back to back", wall_ms_per_frame=round((time.perf_counter() - t0) * 1e3 / n, 1), last_frame_ms=round(tr.device_ms(), 1))), flush=True)
for ra in (1, 3):
  tr.set_option("run_ahead", ra)
  for n in (40,):
    t0 = time.perf_counter()
    for k in range(n):
        tr.reset(); tr.render_device(1, 1000 + k)
    tr.sync()
    print(json.dumps(dict(mode=f"{n} frames back to back, run-ahead {ra}", wall_ms_per_frame=round((time.perf_counter() - t0) * 1e3 / n, 1), last_frame_ms=round(tr.device_ms(), 1))), flush=True)
tr.close()

"""Ad-hoc GPU probe: C4 full frame at 16 spp, 40 frames enqueued back to back, per-frame device intervals (events on the
library's stream).  Found the NaN-ray stall: one frame in forty took 1.2 s before ray_has_nan (rtb_trace.cuh)."""
import sys, json, time
sys.path.insert(0, ".")
import torch
import raytracer_b200 as rtb
sc = rtb.Scene(6, 40, 35)
tr = rtb.Tracer(0)
tr.set_option("kernel", 1); tr.set_option("walk_min", 12); tr.set_option("wf_refill", 16)
tr.setup(sc, 3840, 2160, 16, 8)
ext = torch.cuda.ExternalStream(tr.stream(), device=torch.device("cuda", 0))
tr.reset(); tr.render_device(1, 999); tr.sync()
n = 40
evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
t0 = time.perf_counter()
with torch.cuda.stream(ext):
    evs[0].record(ext)
    for k in range(n):
        tr.reset(); tr.render_device(1, 1000 + k)
        evs[k + 1].record(ext)
    tr.sync()
wall = (time.perf_counter() - t0) * 1e3 / n
iv = [round(evs[k].elapsed_time(evs[k + 1]), 1) for k in range(n)]
print(json.dumps(dict(mode="back to back", wall_ms_per_frame=round(wall, 1), max_interval=max(iv), min_interval=min(iv))), flush=True)
tr.close()
del sc

"""Single-GPU probe: the megakernel (both traversal variants) on the C4 eighth and on C2; for build A/Bs via RTB_LIBRTB."""
import json, sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb
from bench import WORKLOADS
for wl, spp, world in (("C4", 8, 8), ("C2", 64, 1)):
    sid, p0, p1, W, H, _, bounces = WORKLOADS[wl]
    tr = rtb.Tracer(0).setup(rtb.Scene(sid, p0, p1), W, H, spp, bounces)
    tr.set_shard(0, world, 8)
    tr.set_option("kernel", 0)
    row = dict(workload=wl, share=f"1/{world}")
    for v in (0, 1):
        tr.set_option("mega_early_pop", v)
        best = 1e9
        for rep in range(5):
            tr.reset(); tr.render_device(1, 1000 + rep); tr.sync()
            if rep:
                best = min(best, tr.device_ms())
        row["early" if v else "late"] = round(best, 3)
    print(json.dumps(row), flush=True)
    tr.close()

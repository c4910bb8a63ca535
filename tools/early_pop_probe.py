"""Single-GPU probe: the megakernel's two traversal-loop variants (option mega_early_pop) on full frames and tile shares."""
import json, sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb
from bench import WORKLOADS
for wl, spp, world in (("C4", 8, 8), ("C4", 8, 4), ("C4", 8, 1), ("C5", 8, 8), ("C2", 64, 1), ("C3", 16, 1)):
    sid, p0, p1, W, H, _, bounces = WORKLOADS[wl]
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0).setup(sc, W, H, spp, bounces)
    tr.set_shard(0, world, 8)
    tr.set_option("kernel", 0)
    row = dict(workload=wl, spp=spp, share=f"1/{world}")
    for v in (0, 1):
        tr.set_option("mega_early_pop", v)
        best = 1e9
        for rep in range(4):
            tr.reset(); tr.render_device(1, 1000 + rep); tr.sync()
            if rep:
                best = min(best, tr.device_ms())
        row["early_pop_ms" if v else "late_pop_ms"] = round(best, 3)
    tr.set_option("mega_early_pop", -1); tr.set_option("kernel", 2)
    tr.reset(); tr.render_device(1, 999); tr.sync()
    tr.reset(); tr.render_device(1, 1000); tr.sync()
    row.update(auto_kernel=int(tr.info("kernel")), auto_early_pop=int(tr.info("mega_early_pop")), auto_ms=round(tr.device_ms(), 3))
    print(json.dumps(row), flush=True)
    tr.close()

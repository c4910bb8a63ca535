"""Single-GPU probe: megakernel policy knobs on the C4 tile share (rank 0 of 8)."""
import json, sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb
from bench import WORKLOADS
wl, spp, world = "C4", 8, int(sys.argv[1]) if len(sys.argv) > 1 else 8
sid, p0, p1, W, H, _, bounces = WORKLOADS[wl]
sc = rtb.Scene(sid, p0, p1)
tr = rtb.Tracer(0).setup(sc, W, H, spp, bounces)
tr.set_shard(0, world, 8)
tr.set_option("kernel", 0)
tr.set_option("mega_early_pop", 1)
def t():
    best = 1e9
    for rep in range(3):
        tr.reset(); tr.render_device(1, 1000 + rep); tr.sync()
        if rep:
            best = min(best, tr.device_ms())
    return round(best, 3)
for walk_min in (4, 8, 12, 16, 24):
    for refill in (4, 8, 16, 24, 32):
        tr.set_option("walk_min", walk_min); tr.set_option("refill", refill)
        print(json.dumps(dict(share=f"1/{world}", walk_min=walk_min, refill=refill, ms=t())), flush=True)
tr.set_option("walk_min", 0); tr.set_option("refill", 0)
for bps in (3, 4, 5, 6):
    tr.set_option("blocks_per_sm", bps)
    print(json.dumps(dict(share=f"1/{world}", blocks_per_sm=bps, ms=t())), flush=True)
tr.close()

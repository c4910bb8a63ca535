"""Single-GPU probe: wavefront kernels with several row groups in flight (options wf_groups / wf_streams), on the
full frame and on rank 0's share of a tile split.  usage: lanes_probe.py <workload> <spp> <world> [blocks_per_sm]"""
import json, sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb
from bench import WORKLOADS

wl = sys.argv[1] if len(sys.argv) > 1 else "C4"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
world = int(sys.argv[3]) if len(sys.argv) > 3 else 8
bps = int(sys.argv[4]) if len(sys.argv) > 4 else 0
sid, p0, p1, W, H, _, bounces = WORKLOADS[wl]
sc = rtb.Scene(sid, p0, p1)
tr = rtb.Tracer(0).setup(sc, W, H, spp, bounces)
tr.set_shard(0, world, 8)
ref = None
for kernel, groups, streams in ((0, 1, 1), (1, 1, 1), (1, 2, 2), (1, 4, 2), (1, 4, 4), (1, 8, 4), (1, 8, 8), (1, 16, 8)):
    tr.set_option("kernel", kernel)
    tr.set_option("wf_groups", groups)
    tr.set_option("wf_streams", streams)
    tr.set_option("blocks_per_sm", bps)
    best = 1e9
    for rep in range(4):
        tr.reset(); tr.render_device(1, 1000); tr.sync()
        if rep:
            best = min(best, tr.device_ms())
    img, _ = tr.download()
    same = True if ref is None else bool((img == ref).all())
    if ref is None:
        ref = img
    print(json.dumps(dict(workload=wl, world=world, spp=spp, kernel=kernel, groups=groups, streams=streams, blocks_per_sm=bps, ms=round(best, 3), identical=same)), flush=True)
tr.close()

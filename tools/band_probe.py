"""Single-GPU probe: C4 (4K, 1 012 200 triangles) cut into 8 tile shares, for several band heights.
One GPU renders each rank's share in turn; the N = 8 frame time is the slowest share.  Small bands (8 rows)
balance the load but every GPU walks the whole 158 MB scene; tall bands (contiguous strips) keep a GPU's rays in a
part of the scene that fits its L2 but may be unbalanced.  usage: band_probe.py [spp] [world]"""
import json, sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
W, H = 3840, 2160
sc = rtb.Scene(6, 40, 35)
tr = rtb.Tracer(0).setup(sc, W, H, spp, 8)
for kernel in (0, 1):
    tr.set_option("kernel", kernel)
    for band in (8, 16, 32, 68, 135, 270):
        times = []
        for rank in range(world):
            tr.set_shard(rank, world, band)
            best = 1e9
            for rep in range(3):
                tr.reset(); tr.render_device(1, 1000 + rep); tr.sync()
                if rep:
                    best = min(best, tr.device_ms())
            times.append(round(best, 2))
        print(json.dumps(dict(kernel=kernel, band_rows=band, spp=spp, world=world, max_ms=max(times), mean_ms=round(sum(times) / len(times), 2), per_rank=times)), flush=True)
tr.close()

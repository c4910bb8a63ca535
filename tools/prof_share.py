"""Ad-hoc GPU probe (not a test): rank 0's share of an 8-way tile split of C4 through the wavefront kernels, one row
group on one stream, two 4-spp frames -- for an ncu capture of one small wf_traverse launch
(ncu -k regex:wf_traverse --launch-skip 42 --launch-count 1)."""
import sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb
from bench import WORKLOADS

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sid, p0, p1, W, H, _, bounces = WORKLOADS["C4"]
sc = rtb.Scene(sid, p0, p1)
tr = rtb.Tracer(0)
for k, v in (("kernel", 1), ("wf_groups", 1), ("wf_streams", 1)):
    tr.set_option(k, v)
tr.setup(sc, W, H, 4, bounces)
if world > 1:
    tr.set_shard(0, world, 8)
for k in range(2):
    tr.reset()
    tr.render_device(1, 1000 + k)
    tr.sync()
    print(f"C4 share 1/{world}, 4 spp, frame {k}: {tr.device_ms():.3f} ms")
tr.close()

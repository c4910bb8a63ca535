"""Ad-hoc GPU probe (not a test): two launches of the render kernel for ncu (profile the second).
usage: prof_one.py [workload=C3] [spp=8] [key=value ...options]"""
import sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb
sys.path.insert(0, ".")
from bench import WORKLOADS

wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
sid, p0, p1, W, H, _, bounces = WORKLOADS[wl]
sc = rtb.Scene(sid, p0, p1)
tr = rtb.Tracer(0)
for kv in sys.argv[3:]:
    k, v = kv.split("=")
    tr.set_option(k, int(v))
tr.setup(sc, W, H, spp, bounces)
for k in range(2):
    tr.reset()
    tr.render_device(1, 1000 + k)
    tr.sync()
    print(f"{wl} {spp} spp launch {k}: {tr.device_ms():.3f} ms")
tr.close()

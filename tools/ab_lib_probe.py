"""A/B of two builds of the library on the same box (RTB_LIBRTB selects the build): frame times of the wavefront pair
and the megakernel on C3 / C4 / C1 / C2.  usage: RTB_LIBRTB=<path> python tools/ab_lib_probe.py <label>"""
import json, os, sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb
from bench import WORKLOADS

label = sys.argv[1] if len(sys.argv) > 1 else os.environ.get("RTB_LIBRTB", "default")
for wl, spp in (("C3", 32), ("C4", 8), ("C1", 32), ("C2", 64)):
    sid, p0, p1, W, H, _, bounces = WORKLOADS[wl]
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0).setup(sc, W, H, spp, bounces)
    for kernel in (1, 0):
        tr.set_option("kernel", kernel)
        best = 1e9
        for rep in range(5):
            tr.reset(); tr.render_device(1, 1000 + rep); tr.sync()
            if rep:
                best = min(best, tr.device_ms())
        print(json.dumps(dict(build=label, workload=wl, spp=spp, kernel=kernel, ms=round(best, 3))), flush=True)
    tr.close()

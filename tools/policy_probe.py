"""Ad-hoc GPU probe (not a test): frame time of the wavefront kernels over the traversal policy (walk_min, wf_refill)."""
import sys, json
sys.path.insert(0, ".")
import raytracer_b200 as rtb


def sweep(name, sid, p0, p1, W, H, spp, bounces, pairs):
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    tr.set_option("kernel", 1)
    tr.setup(sc, W, H, spp, bounces)
    tr.render_device(1, 1); tr.sync()
    res = {}
    for wm, rf in pairs:
        tr.set_option("walk_min", wm); tr.set_option("wf_refill", rf)
        best = 1e30
        for _ in range(3):
            tr.reset(); tr.render_device(1, 100); tr.sync(); best = min(best, tr.device_ms())
        res[f"{wm},{rf}"] = round(best, 2)
    print(name, json.dumps(res), flush=True)
    tr.close()


if __name__ == "__main__":
    pairs = [(16, 8), (12, 16), (12, 12), (12, 20), (12, 24), (10, 16), (14, 16), (8, 16), (8, 24), (16, 16), (10, 20), (6, 20)]
    sweep("C3-32", 0, 0, 0, 1920, 1080, 32, 8, pairs)
    sweep("C4-8", 6, 40, 35, 3840, 2160, 8, 8, pairs)
    sweep("C1-25", 0, 0, 0, 1000, 800, 25, 5, pairs)

"""Ad-hoc GPU probe (not a test): frame times of the tuned default on the bench workloads, for A/B runs of two
builds of the library (RTB_LIBRTB=<path> selects the build)."""
import sys, json, os
sys.path.insert(0, ".")
import raytracer_b200 as rtb


def run(name, sid, p0, p1, W, H, spp, bounces, opts):
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    for k, v in opts.items():
        tr.set_option(k, v)
    tr.setup(sc, W, H, spp, bounces)
    tr.render_device(1, 1); tr.sync()
    best = 1e30
    for _ in range(3):
        tr.reset(); tr.render_device(1, 100); tr.sync(); best = min(best, tr.device_ms())
    out = dict(lib=os.environ.get("RTB_LIBRTB", "default"), name=name, opts=opts, ms=best, kernel=int(tr.info("kernel")), walk_min=int(tr.info("walk_min")), wf_refill=int(tr.info("wf_refill")))
    if int(tr.info("kernel")) == 1:
        tr.set_option("kernel", 1); tr.set_option("wf_groups", 1); tr.set_option("wf_timing", 1); tr.reset(); tr.render_device(1, 100); tr.sync()
        out["wf_ms"] = dict(shade=tr.info("wf_shade_ms"), traverse=tr.info("wf_traverse_ms"))
    print(json.dumps(out), flush=True)
    tr.close()


if __name__ == "__main__":
    cfgs = (("C3-32", (0, 0, 0, 1920, 1080, 32, 8)), ("C4-8", (6, 40, 35, 3840, 2160, 8, 8)), ("C2-16", (5, 0, 0, 1920, 1080, 16, 8)), ("C1", (0, 0, 0, 1000, 800, 100, 5)))
    for name, cfg in cfgs:
        run(name, *cfg, opts={})
        run(name, *cfg, opts={"kernel": 0})

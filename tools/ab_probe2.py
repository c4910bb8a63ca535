"""Ad-hoc GPU probe (not a test): A/B of library builds (RTB_LIBRTB) on C3 / C4 full frames and on a 1/8 tile share of C4."""
import sys, json, os
sys.path.insert(0, ".")
import raytracer_b200 as rtb


def run(name, world, sid, p0, p1, W, H, spp, bounces, opts):
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    for k, v in opts.items():
        tr.set_option(k, v)
    tr.setup(sc, W, H, spp, bounces)
    if world > 1:
        tr.set_shard(0, world, 8)
    tr.render_device(1, 1); tr.sync()
    best = 1e30
    for _ in range(3):
        tr.reset(); tr.render_device(1, 100); tr.sync(); best = min(best, tr.device_ms())
    print(json.dumps(dict(lib=os.path.basename(os.environ.get("RTB_LIBRTB", "default")), name=name, world=world, opts=opts, ms=round(best, 2), kernel=int(tr.info("kernel")))), flush=True)
    tr.close()


if __name__ == "__main__":
    C3 = (0, 0, 0, 1920, 1080, 32, 8)
    C4 = (6, 40, 35, 3840, 2160, 8, 8)
    run("C3-32", 1, *C3, opts={"kernel": 1})
    run("C3-32", 1, *C3, opts={"kernel": 0})
    run("C4-8", 1, *C4, opts={"kernel": 1})
    run("C4-8", 1, *C4, opts={"kernel": 0})
    run("C4-8", 8, *C4, opts={"kernel": 1})
    run("C4-8", 8, *C4, opts={"kernel": 0})

"""Diagnostic (run on the GPU box): per-GPU work of C4 at N = 8 emulated on one GPU with rtb_set_shard(0, 8)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_b200 as rtb

def run(world, opts, sid=6, p0=40, p1=35, W=3840, H=2160, spp=16, bounces=8, reps=3):
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    for k, v in opts.items():
        tr.set_option(k, v)
    tr.setup(sc, W, H, spp, bounces)
    tr.set_shard(0, world, 8)
    tr.render_device(1, 1); tr.sync()
    best = 1e9
    for r in range(reps):
        tr.reset(); tr.render_device(1, 100 + r); tr.sync(); best = min(best, tr.device_ms())
    print(json.dumps(dict(world=world, opts=opts, ms=best, launches=tr.info("launches_per_frame"))), flush=True)
    tr.close()

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "c4"
    if which == "c4":
        run(1, {})
        for o in ({}, {"wf_groups": 1}, {"wf_groups": 2, "wf_streams": 1}, {"wf_groups": 4}, {"kernel": 0}, {"wf_rebuild": 8}, {"wf_rebuild": 32},
                  {"wf_groups": 1, "wf_rebuild": 8}):
            run(8, o)
    if which == "c3":
        kw = dict(sid=0, p0=0, p1=0, W=1920, H=1080, spp=64, bounces=8)
        run(1, {}, **kw)
        for o in ({}, {"wf_groups": 1}, {"kernel": 0}):
            run(8, o, **kw)
    if which == "auto":
        def show(name, world, **kw):
            sc = rtb.Scene(kw.get("sid", 0), kw.get("p0", 0), kw.get("p1", 0))
            out = {}
            for kern in (0, 1, 2):
                tr = rtb.Tracer(0)
                tr.set_option("kernel", kern)
                tr.setup(sc, kw["W"], kw["H"], kw["spp"], kw["bounces"])
                tr.set_shard(0, world, 8)
                tr.render_device(1, 1); tr.sync()
                best = 1e9
                for r in range(2):
                    tr.reset(); tr.render_device(1, 100 + r); tr.sync(); best = min(best, tr.device_ms())
                out[kern] = round(best, 2)
                if kern == 2:
                    out["choice"] = int(tr.info("kernel")); out["tune"] = (round(tr.info("tune_megakernel_ms"), 2), round(tr.info("tune_wavefront_ms"), 2))
                tr.close()
            print(name, "world", world, json.dumps(out), flush=True)
        show("C1", 1, sid=0, W=1000, H=800, spp=100, bounces=5)
        show("C2", 1, sid=5, W=1920, H=1080, spp=64, bounces=8)
        show("C3/64spp", 1, sid=0, W=1920, H=1080, spp=64, bounces=8)
        show("C3/64spp", 8, sid=0, W=1920, H=1080, spp=64, bounces=8)
        show("C4", 1, sid=6, p0=40, p1=35, W=3840, H=2160, spp=16, bounces=8)
        show("C4", 2, sid=6, p0=40, p1=35, W=3840, H=2160, spp=16, bounces=8)
        show("C4", 8, sid=6, p0=40, p1=35, W=3840, H=2160, spp=16, bounces=8)
        show("C5", 1, sid=0, W=3840, H=2160, spp=32, bounces=8)
    if which == "mega":
        for o in ({"kernel": 0}, {"kernel": 0, "refill": 8}, {"kernel": 0, "refill": 16}, {"kernel": 0, "refill": 32}, {"kernel": 0, "walk_min": 8},
                  {"kernel": 0, "walk_min": 24}, {"kernel": 0, "refill": 16, "walk_min": 8}, {"kernel": 0, "blocks_per_sm": 2}, {"kernel": 0, "blocks_per_sm": 3},
                  {"kernel": 0, "builder": 0}, {"kernel": 0, "large_split": 0}, {"kernel": 0, "ploc_radius": 32}, {"kernel": 0, "ploc_radius": 8}):
            run(8, o)
    if which == "wm":
        for o in ({"kernel": 0, "walk_min": 4}, {"kernel": 0, "walk_min": 8, "refill": 20}, {"kernel": 0, "walk_min": 12}, {"kernel": 0, "walk_min": 1},
                  {"kernel": 1, "walk_min": 8}, {"kernel": 1, "walk_min": 4}, {"kernel": 1, "walk_min": 12}):
            run(8, o)
        for o in ({"kernel": 1}, {"kernel": 1, "walk_min": 8}, {"kernel": 1, "walk_min": 4}, {"kernel": 1, "walk_min": 12}, {"kernel": 0, "walk_min": 8}):
            run(1, o)
        kw = dict(sid=0, p0=0, p1=0, W=1920, H=1080, spp=64, bounces=8)
        for o in ({"kernel": 1}, {"kernel": 1, "walk_min": 12}, {"kernel": 1, "walk_min": 8}, {"kernel": 1, "walk_min": 20}):
            run(1, o, **kw)
        for o in ({"kernel": 0}, {"kernel": 0, "walk_min": 8}, {"kernel": 0, "walk_min": 12}):
            run(8, o, **kw)
    if which == "c3wm":
        kw = dict(sid=0, p0=0, p1=0, W=1920, H=1080, spp=256, bounces=8, reps=2)
        for o in ({"kernel": 1}, {"kernel": 1, "walk_min": 12}, {"kernel": 1, "walk_min": 8}, {"kernel": 1, "walk_min": 8, "wf_refill": 12}, {"kernel": 1, "walk_min": 12, "wf_refill": 12}):
            run(1, o, **kw)
    if which == "ab":
        kw = dict(sid=0, p0=0, p1=0, W=1920, H=1080, spp=64, bounces=8, reps=3)
        run(1, {"kernel": 1}, **kw)
        run(1, {"kernel": 1, "walk_min": 8, "wf_refill": 12}, **kw)
        run(1, {"kernel": 0}, **kw)
        run(1, {"kernel": 1})
        run(8, {"kernel": 0})
    if which == "mega2":
        run(8, {"kernel": 0})
        run(1, {"kernel": 0})
        run(8, {"kernel": 0}, sid=0, p0=0, p1=0, W=1920, H=1080, spp=64, bounces=8)
        run(1, {"kernel": 0}, sid=5, p0=0, p1=0, W=1920, H=1080, spp=64, bounces=8)
    if which == "grid":
        kw = dict(sid=0, p0=0, p1=0, W=1920, H=1080, spp=128, bounces=8, reps=2)
        for wm in (6, 8, 12, 16):
            for rf in (8, 12, 16):
                run(1, {"kernel": 1, "walk_min": wm, "wf_refill": rf}, **kw)
        for rb in (8, 32):
            run(1, {"kernel": 1, "walk_min": 8, "wf_refill": 12, "wf_rebuild": rb}, **kw)
        run(1, {"kernel": 1, "walk_min": 8, "wf_refill": 12, "wf_groups": 3}, **kw)

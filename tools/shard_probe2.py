"""Diagnostic (run on the GPU box): where the wavefront kernels lose time on a 1/8 tile share of C4 (emulated on one
GPU with rtb_set_shard(0, 8)): per-kernel totals of a frame (one row group, events around every launch)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_b200 as rtb


def run(world, opts, sid=6, p0=40, p1=35, W=3840, H=2160, spp=16, bounces=8):
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    for k, v in opts.items():
        tr.set_option(k, v)
    tr.setup(sc, W, H, spp, bounces)
    tr.set_shard(0, world, 8)
    tr.render_device(1, 1); tr.sync()
    best = 1e9
    for r in range(3):
        tr.reset(); tr.render_device(1, 100 + r); tr.sync(); best = min(best, tr.device_ms())
    out = dict(world=world, opts=opts, ms=best, kernel=int(tr.info("kernel")))
    if int(tr.info("kernel")) == 1 and opts.get("wf_groups") == 1:
        tr.set_option("wf_timing", 1); tr.reset(); tr.render_device(1, 100); tr.sync()
        out["wf_ms"] = dict(shade=round(tr.info("wf_shade_ms"), 2), traverse=round(tr.info("wf_traverse_ms"), 2), total=round(tr.info("wf_total_ms"), 2))
    print(json.dumps(out), flush=True)
    tr.close()


if __name__ == "__main__":
    for world in (1, 8):
        run(world, {"kernel": 1, "wf_groups": 1, "wf_streams": 1})
        run(world, {"kernel": 1})
    for o in ({"kernel": 1, "wf_groups": 1, "wf_streams": 1, "blocks_per_sm": 4}, {"kernel": 1, "wf_groups": 1, "wf_streams": 1, "blocks_per_sm": 2},
              {"kernel": 1, "blocks_per_sm": 4}, {"kernel": 1, "wf_groups": 4, "wf_streams": 2, "blocks_per_sm": 4},
              {"kernel": 1, "wf_groups": 1, "wf_streams": 1, "wf_rebuild": 0}, {"kernel": 0}):
        run(8, o)

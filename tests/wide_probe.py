"""Ad-hoc GPU probe (not a test): wf_traverse over the binary nodes against their 4-wide view."""
import sys, json
sys.path.insert(0, ".")
import raytracer_b200 as rtb


def run(name, sid, p0, p1, W, H, spp, bounces, opts):
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    for k, v in opts.items():
        tr.set_option(k, v)
    tr.setup(sc, W, H, spp, bounces)
    tr.set_render_data(sc.render_settings(1, bounces)); tr.set_option("counters", 1)
    tr.render_device(1, 7); tr.sync(); c = tr.counters(); tr.set_option("counters", 0)
    ls = [tr.info(f"lane_stat{k}") for k in range(7)]
    tr.set_render_data(sc.render_settings(spp, bounces)); tr.reset()
    tr.render_device(1, 1); tr.sync()
    best = 1e30
    for _ in range(2):
        tr.reset(); tr.render_device(1, 100); tr.sync(); best = min(best, tr.device_ms())
    out = dict(name=name, opts=opts, ms=best, width=int(tr.info("bvh_width")), kernel=int(tr.info("kernel")),
               box_per_ray=c["box_tests"] / c["segments"], tri_per_ray=c["tri_tests"] / c["segments"],
               node_iters_per_ray=ls[0] * 32 / max(1, c["segments"]), node_walk=ls[1] / max(1, ls[0]), leaf_iters_per_ray=ls[4] * 32 / max(1, c["segments"]))
    if int(tr.info("kernel")) == 1:
        tr.set_option("wf_groups", 1); tr.set_option("wf_timing", 1); tr.reset(); tr.render_device(1, 100); tr.sync()
        out["wf_ms"] = dict(shade=tr.info("wf_shade_ms"), traverse=tr.info("wf_traverse_ms"))
    out["tune"] = dict(bvh2=tr.info("tune_bvh2_ms"), bvh4=tr.info("tune_bvh4_ms"), mega=tr.info("tune_megakernel_ms"))
    print(json.dumps(out), flush=True)
    tr.close()


if __name__ == "__main__":
    for name, cfg in (("C3-32", (0, 0, 0, 1920, 1080, 32, 8)), ("C4-8", (6, 40, 35, 3840, 2160, 8, 8)), ("C2-16", (5, 0, 0, 1920, 1080, 16, 8))):
        for wm, rf in ((16, 8), (12, 16)):
            for width in (2, 4):
                run(name, *cfg, opts={"kernel": 1, "bvh_width": width, "walk_min": wm, "wf_refill": rf})
        run(name, *cfg, opts={})  # the tuned default

"""Ad-hoc GPU probe (not a test): times the megakernel on the BASELINE configs."""
import sys, time, json
import numpy as np
sys.path.insert(0, ".")
import raytracer_b200 as rtb

def run(name, sid, p0, p1, W, H, spp, bounces, frames=1, opts=None):
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    for k, v in (opts or {}).items():
        tr.set_option(k, v)
    tr.setup(sc, W, H, spp, bounces)
    info = {k: tr.info(k) for k in ("num_prims", "build_ms", "scene_bytes", "blocks_per_sm", "block_threads", "culled_tris")}
    # counted pass at 1 spp for the op mix
    tr.set_render_data(sc.render_settings(1, bounces)); tr.set_option("counters", 1)
    tr.render_device(1, 7); tr.sync(); c = tr.counters(); tr.set_option("counters", 0)
    ls = [tr.info(f"lane_stat{k}") for k in range(7)]
    if ls[0] > 0:
        info["lanes"] = dict(node_iters_per_ray=ls[0] * 32 / c["segments"], node_walk=ls[1] / ls[0], node_blocked=ls[2] / ls[0], node_done=ls[3] / ls[0],
                             node_mask=ls[6] / ls[0], leaf_iters_per_ray=ls[4] * 32 / c["segments"], leaf_lanes=ls[5] / max(1, ls[4]))
    tr.set_render_data(sc.render_settings(spp, bounces)); tr.reset()
    tr.render_device(1, 1); tr.sync()  # warm-up
    tr.reset(); tr.render_device(frames, 100); tr.sync(); ms = tr.device_ms()
    if int(tr.info("kernel")) == 1:
        tr.set_option("wf_timing", 1); tr.reset(); tr.render_device(1, 100); tr.sync(); tr.set_option("wf_timing", 0)
        info["wf_ms"] = dict(shade=tr.info("wf_shade_ms"), traverse=tr.info("wf_traverse_ms"), total=tr.info("wf_total_ms"))
    seg_per_sample = c["segments"] / c["samples"]
    rays = seg_per_sample * W * H * spp * frames
    out = dict(name=name, opts=opts, ms=ms, mrays_s=rays / ms / 1e3, spp_s=spp * frames / (ms / 1e3), seg_per_sample=seg_per_sample,
               box_per_ray=c["box_tests"] / c["segments"], tri_per_ray=c["tri_tests"] / c["segments"], sph_per_ray=c["sphere_tests"] / c["segments"], **info)
    print(json.dumps(out), flush=True)
    tr.close()

if __name__ == "__main__":
    C3 = ("C3", 0, 0, 0, 1920, 1080, 32, 8)
    which = sys.argv[1] if len(sys.argv) > 1 else ""
    if which == "groups":
        for g, st in ((1, 1), (2, 1), (2, 2), (4, 1), (4, 2), (8, 2), (16, 2)):
            run("C3-256", 0, 0, 0, 1920, 1080, 256, 8, opts={"wf_groups": g, "wf_streams": st})
        sys.exit(0)
    if which == "rebuild":
        for rb in (0, 4, 16, 64):
            run("C3-256", 0, 0, 0, 1920, 1080, 256, 8, opts={"wf_rebuild": rb})
        run(*C3, opts={"wf_rebuild": 0})
        run(*C3, opts={"wf_rebuild": 16})
        sys.exit(0)
    if which == "one":
        run(*C3, opts={})
        sys.exit(0)
    if which == "split":
        run(*C3, opts={})
        run("C3-256", 0, 0, 0, 1920, 1080, 256, 8, opts={})
        run("C1", 0, 0, 0, 1000, 800, 100, 5, opts={})
        run("C2", 5, 0, 0, 1920, 1080, 64, 8, opts={})
        sys.exit(0)
    if which == "quick":
        run(*C3, opts={"kernel": 1, "builder": 1, "wf_refill": 8, "walk_min": 16})
        run(*C3, opts={"kernel": 1, "builder": 1, "wf_refill": 8, "walk_min": 20})
        run(*C3, opts={"kernel": 1, "builder": 1, "wf_refill": 4, "walk_min": 16})
        run(*C3, opts={"kernel": 0, "builder": 1, "walk_min": 12})
        run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"kernel": 1, "builder": 0, "walk_min": 16})
        sys.exit(0)
    if which == "lanes2":
        for wm in (8, 12, 16, 20, 24, 28):
            run(*C3, opts={"kernel": 1, "builder": 1, "wf_refill": 8, "walk_min": wm})
        for r in (2, 4, 16):
            run(*C3, opts={"kernel": 1, "builder": 1, "wf_refill": r, "walk_min": 20})
        for wm in (12, 20, 28):
            run(*C3, opts={"kernel": 0, "builder": 1, "walk_min": wm})
        run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"kernel": 1, "builder": 0, "walk_min": 20})
        sys.exit(0)

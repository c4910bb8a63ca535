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
    tr.set_render_data(sc.render_settings(spp, bounces)); tr.reset()
    tr.render_device(1, 1); tr.sync()  # warm-up
    tr.reset(); tr.render_device(frames, 100); tr.sync(); ms = tr.device_ms()
    seg_per_sample = c["segments"] / c["samples"]
    rays = seg_per_sample * W * H * spp * frames
    out = dict(name=name, opts=opts, ms=ms, mrays_s=rays / ms / 1e3, spp_s=spp * frames / (ms / 1e3), seg_per_sample=seg_per_sample,
               box_per_ray=c["box_tests"] / c["segments"], tri_per_ray=c["tri_tests"] / c["segments"], sph_per_ray=c["sphere_tests"] / c["segments"], **info)
    print(json.dumps(out), flush=True)
    tr.close()

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "sweep":
        for r in (1, 4, 8, 12, 16, 20, 24, 28, 32):
            run("C3", 0, 0, 0, 1920, 1080, 32, 8, opts={"refill": r})
        run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"refill": 8})
        run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"refill": 16})
        run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"refill": 32})
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "sweep4":
        C3 = ("C3", 0, 0, 0, 1920, 1080, 32, 8)
        run(*C3, opts={"kernel": 0})
        run(*C3, opts={"kernel": 1})
        for b, ls, rad in ((1, 1, 16), (1, 0, 16), (1, 0, 8), (1, 0, 32), (0, 0, 16)):
            for kern in (0, 1):
                run(*C3, opts={"kernel": kern, "builder": b, "large_split": ls, "ploc_radius": rad})
        for r in (2, 4, 16):
            run(*C3, opts={"kernel": 1, "builder": 1, "large_split": 0, "wf_refill": r})
        for ne in (0, 2, 8):
            run(*C3, opts={"kernel": 1, "builder": 1, "large_split": 0, "node_exit": ne})
        for kern in (0, 1):
            for b, ls in ((0, 1), (1, 0)):
                run("C1", 0, 0, 0, 1000, 800, 32, 5, opts={"kernel": kern, "builder": b, "large_split": ls})
                run("C2", 5, 0, 0, 1920, 1080, 16, 8, opts={"kernel": kern, "builder": b, "large_split": ls})
                run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"kernel": kern, "builder": b, "large_split": ls})
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "sweep3":
        run("C3", 0, 0, 0, 1920, 1080, 32, 8, opts={"kernel": 0})
        for r in (2, 4, 8, 16, 32):
            for ne in (0, 4):
                run("C3", 0, 0, 0, 1920, 1080, 32, 8, opts={"kernel": 1, "wf_refill": r, "node_exit": ne})
        run("C3", 0, 0, 0, 1920, 1080, 32, 8, opts={"kernel": 1, "wf_refill": 8, "node_exit": 4, "blocks_per_sm": 4})
        run("C3", 0, 0, 0, 1920, 1080, 32, 8, opts={"kernel": 1, "wf_refill": 8, "node_exit": 4, "blocks_per_sm": 6})
        run("C1", 0, 0, 0, 1000, 800, 32, 5, opts={"kernel": 0})
        run("C1", 0, 0, 0, 1000, 800, 32, 5, opts={"kernel": 1})
        run("C2", 5, 0, 0, 1920, 1080, 16, 8, opts={"kernel": 0})
        run("C2", 5, 0, 0, 1920, 1080, 16, 8, opts={"kernel": 1})
        run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"kernel": 0})
        run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"kernel": 1})
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "sweep2":
        for r in (4, 8, 12, 16, 24):
            for ne in (0, 2, 4, 8):
                run("C3", 0, 0, 0, 1920, 1080, 32, 8, opts={"refill": r, "node_exit": ne})
        for r in (8, 16):
            for ne in (0, 4):
                run("C4-small", 6, 40, 35, 1920, 1080, 4, 8, opts={"refill": r, "node_exit": ne})
        sys.exit(0)
    run("C1", 0, 0, 0, 1000, 800, 100, 5)
    run("C2", 5, 0, 0, 1920, 1080, 64, 8)
    run("C3", 0, 0, 0, 1920, 1080, 64, 8)
    run("C4-small", 6, 40, 35, 1920, 1080, 4, 8)

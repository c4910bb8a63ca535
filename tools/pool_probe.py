"""Ad-hoc GPU probe (not a test): frame times of the three render kernels (0 megakernel, 1 wavefront, 3 pooled) on
the bench workloads, full frames and 1/8 tile shares, with the pooled kernel's policy knobs swept.
RTB_LIBRTB=<path> selects an A/B build of the library."""
import json
import os
import sys
sys.path.insert(0, ".")
import raytracer_b200 as rtb

CFG = {"C3-32": (0, 0, 0, 1920, 1080, 32, 8), "C3-256": (0, 0, 0, 1920, 1080, 256, 8), "C4-8": (6, 40, 35, 3840, 2160, 8, 8),
       "C2-16": (5, 0, 0, 1920, 1080, 16, 8), "C1": (0, 0, 0, 1000, 800, 100, 5), "C5-16": (0, 0, 0, 3840, 2160, 16, 8)}


def run(name, opts, share=1, reps=3):
    sid, p0, p1, W, H, spp, bounces = CFG[name]
    sc = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    for k, v in opts.items():
        tr.set_option(k, v)
    tr.setup(sc, W, H, spp, bounces)
    if share > 1:
        tr.set_shard(0, share, 8)
    tr.render_device(1, 1); tr.sync()
    best = 1e30
    for _ in range(reps):
        tr.reset(); tr.render_device(1, 100); tr.sync(); best = min(best, tr.device_ms())
    out = dict(lib=os.path.basename(os.environ.get("RTB_LIBRTB", "default")), name=name, share=share, opts=opts, ms=round(best, 3),
               kernel=int(tr.info("kernel")), walk_min=int(tr.info("walk_min")), refill=int(tr.info("wf_refill")))
    if "counters" in sys.argv:
        tr.set_option("counters", 1); tr.reset(); tr.render_device(1, 100); tr.sync(); c = tr.counters()
        out["mrays"] = round(c["segments"] / best / 1e3, 1)
        out["box"] = round(c["box_tests"] / c["segments"], 2); out["tri"] = round(c["tri_tests"] / c["segments"], 2)
        ls = [tr.info(f"lane_stat{k}") for k in range(7)]
        if ls[0] > 0:
            out["lanes"] = dict(node_walk=round(ls[1] / ls[0], 1), blocked=round(ls[2] / ls[0], 1), done=round(ls[3] / ls[0], 1),
                                leaf_lanes=round(ls[5] / max(1, ls[4]), 1), node_iters_per_ray=round(ls[0] * 32 / c["segments"], 2),
                                leaf_iters_per_ray=round(ls[4] * 32 / c["segments"], 2))
    print(json.dumps(out), flush=True)
    tr.close()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "kernels"
    if which == "kernels":
        for name, share in (("C3-32", 1), ("C3-32", 8), ("C4-8", 1), ("C4-8", 8), ("C2-16", 1), ("C1", 1), ("C5-16", 1)):
            for k in (0, 1, 3):
                run(name, {"kernel": k}, share)
        run("C3-256", {"kernel": 1}); run("C3-256", {"kernel": 3})
    elif which == "policy":
        for name, share in (("C3-32", 1), ("C4-8", 8)):
            for wm, rf in ((16, 8), (12, 16), (16, 4), (20, 8), (12, 8), (8, 8), (16, 12), (24, 4)):
                run(name, {"kernel": 3, "walk_min": wm, "wf_refill": rf}, share)
            for bps in (4, 5, 6):
                run(name, {"kernel": 3, "blocks_per_sm": bps}, share)
    elif which == "one":
        for name, share in (("C3-32", 1), ("C4-8", 1), ("C4-8", 8), ("C2-16", 1)):
            run(name, {"kernel": 3}, share)
    elif which == "c3":
        for k in (1, 3, 0):
            run("C3-32", {"kernel": k})
        run("C3-32", {"kernel": 3, "blocks_per_sm": 4}); run("C3-32", {"kernel": 3, "blocks_per_sm": 5})
        run("C4-8", {"kernel": 3}, 8); run("C4-8", {"kernel": 1}, 8); run("C2-16", {"kernel": 3})
    elif which == "sah":
        for name, share in (("C3-32", 1), ("C1", 1), ("C2-16", 1)):
            for b, ls in ((1, 1), (2, 1), (2, 0), (1, 0), (0, 1)):
                run(name, {"kernel": 1, "builder": b, "large_split": ls}, share)
    elif which == "wf":
        run("C3-32", {"kernel": 1}); run("C1", {"kernel": 1}); run("C4-8", {"kernel": 1}); run("C2-16", {"kernel": 0}); run("C5-16", {"kernel": 1})
    elif which == "stream":
        run("C3-32", {"kernel": 1})
        for ns in (0, 48, 56, 64, 72):
            sid, p0, p1, W, H, spp, bounces = CFG["C3-32"]
            sc = rtb.Scene(sid, p0, p1)
            tr = rtb.Tracer(0)
            tr.set_option("kernel", 4)
            if ns:
                tr.set_option("stream_shade_sms", ns)
            tr.setup(sc, W, H, spp, bounces)
            ms = []
            for k in range(6):
                tr.reset(); tr.render_device(1, 100 + k); tr.sync(); ms.append(round(tr.device_ms(), 2))
            print(json.dumps(dict(name="C3-32", kernel=4, shade_sms_opt=ns, ms=ms, shade_sms=tr.info("stream_shade_sms"),
                                  util_shade=round(tr.info("stream_util_shade"), 3), util_trav=round(tr.info("stream_util_traverse"), 3),
                                  slots=tr.info("stream_slots"))), flush=True)
            tr.close()
        for name, share in (("C3-32", 8), ("C4-8", 1), ("C4-8", 8), ("C2-16", 1), ("C1", 1), ("C5-16", 1)):
            run(name, {"kernel": 4}, share, reps=6)
            run(name, {}, share, reps=3)

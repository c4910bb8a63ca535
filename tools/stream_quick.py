import sys, json, os
sys.path.insert(0, ".")
import raytracer_b200 as rtb
sc = rtb.Scene(0)
for opts in ({}, {"wf_refill": 16, "walk_min": 12}):
    tr = rtb.Tracer(0)
    tr.set_option("kernel", 4)
    for k, v in opts.items(): tr.set_option(k, v)
    tr.setup(sc, 1920, 1080, 32, 8)
    ms = []
    for k in range(8):
        tr.reset(); tr.render_device(1, 100 + k); tr.sync(); ms.append(round(tr.device_ms(), 2))
    print(json.dumps(dict(lib=os.path.basename(os.environ.get("RTB_LIBRTB", "default")), opts=opts, ms=ms, shade_sms=tr.info("stream_shade_sms"), util_shade=round(tr.info("stream_util_shade"), 3), util_trav=round(tr.info("stream_util_traverse"), 3))), flush=True)
    tr.close()

"""Ad-hoc GPU probe: does a C4 frame run slower after a C3 job in the same process?"""
import sys, json, time
sys.path.insert(0, ".")
import raytracer_b200 as rtb

def c4(tag, n=8):
    sc = rtb.Scene(6, 40, 35)
    tr = rtb.Tracer(0)
    tr.setup(sc, 3840, 2160, 16, 8)
    tr.reset(); tr.render_device(1, 999); tr.sync()
    ms = []
    for k in range(n):
        tr.reset(); tr.render_device(1, 1000 + k); tr.sync(); ms.append(round(tr.device_ms(), 1))
    print(json.dumps(dict(tag=tag, ms=ms, kernel=int(tr.info("kernel")), walk_min=int(tr.info("walk_min")))), flush=True)
    tr.close()

c4("fresh process")
sc = rtb.Scene(0)
tr = rtb.Tracer(0).setup(sc, 1920, 1080, 256, 8)
t0 = time.perf_counter()
for k in range(8):
    tr.reset(); tr.render_device(1, 1000 + k)
tr.sync()
print(json.dumps(dict(tag="C3 burn", s=round(time.perf_counter() - t0, 2))), flush=True)
if "keep" in sys.argv:
    c4("after C3, C3 tracer still open")
tr.close()
c4("after C3, C3 tracer closed")
time.sleep(5)
c4("5 s later")

#!/usr/bin/env python
"""bench.py -- the headline benchmark: Mrays/s (and 1080p spp/s) of the per-pixel path-tracing hot
path on N B200s, measured on the workload BASELINE.json quotes its target on.

    python bench.py --gpus N --steps K --warmup W [--workload C3] [--impl reference]

A "step" is one frame (one pass of the hot path over every pixel: spp x bounces segments).
Workloads:  C1 1000x800/100spp/5b scene 0 | C2 1080p/64/8 cube+spheres |
C3 1080p/256/8 Cornell+monkey (DEFAULT: the config the >=20x target is quoted on) |
C4 4K/1024/8 1M-triangle grid (spp bounded per step) | C5 = C3 geometry at 4K, sample slices.

value  = Mrays/s, device-timed (CUDA events on the render stream), scene and accumulation buffer
         resident in HBM; whole job (all ranks), time = max over ranks.
e2e    = the same metric through the reference-facing entry point rtb_render() == render(
         VariableRenderData*, int): host image in, host image out, H2D and D2H inside the timed region.
N > 1  : one process per GPU (torchrun).  Two partitionings, --sharding:
         tiles   row-cyclic bands of the SAME frame (strong scaling, image bit-identical to 1 GPU),
                 the exchange of the bands inside the timed region;
         slices  sample slices of a progressive render: every rank renders the whole frame for its
                 own frames (current_time_ms = base + rank), NCCL all-reduce of the accumulation
                 buffer inside the timed region; a step yields N x spp samples (weak scaling).
         auto    (default) tiles for C4 (BASELINE.json's tile-sharded config), slices otherwise (C5 is
                 the survey's sample-slice config; a 1080p frame cut in 8 leaves 0.26 M pixels per GPU,
                 too few rays per launch to fill 148 SMs: measured at N = 2 on C3, tiles 83 % efficient).
         Both exchanges run INSIDE the library (rtb_comm.cu); tiles default to the fused peer-store exchange.
--impl reference : the reference's OWN CUDA build (baseline/_ref, unmodified sources, resolution
         constants patched) on the same workload at a bounded spp per step (a rate, spp independent);
         if that library is absent, the reference source on the host cores (oracle/_ref), else the C port.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    #        scene p0 p1   W     H    spp  bounces
    "C1": (0, 0, 0, 1000, 800, 100, 5),
    "C2": (5, 0, 0, 1920, 1080, 64, 8),
    "C3": (0, 0, 0, 1920, 1080, 256, 8),
    "C4": (6, 40, 35, 3840, 2160, 16, 8),   # 1024 spp in the config; 16 spp per step keeps a step ~1 s
    "C5": (0, 0, 0, 3840, 2160, 128, 8),
}
WORKLOAD_DESC = {
    "C1": "reference default scene 0 (Cornell box + Suzanne 723 tris + mirror sphere) 1000x800 100spp 5 bounces",
    "C2": "cube.obj (12 tris) + spheres 1920x1080 64spp 8 bounces",
    "C3": "low_poly_monkey.obj scene (Cornell box + Suzanne + sphere) 1920x1080 256spp 8 bounces",
    "C4": "1,012,200-triangle monkey grid 3840x2160 8 bounces, 16 spp per step (of 1024)",
    "C5": "scene 0 geometry 3840x2160 8 bounces, 128 spp per step (progressive to 16384)",
}


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, interval_ms=200):
        self.proc = None
        self.lines = []
        self.gpu = gpu_index
        self.interval_ms = interval_ms

    def start(self):
        if os.environ.get("RTB_BENCH_NO_SAMPLER"):   # diagnostics only: does the nvidia-smi loop perturb the run?
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", str(self.interval_ms),
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(power), "samples": len(sm),
                "reasons": sorted(reasons)}


def flops_per_ray(c):
    """canonical cost of the reference's formulas, weighted by the measured op mix."""
    seg = max(1, c["segments"])
    return (24.0 * c["box_tests"] + 48.0 * c["tri_tests"] + 30.0 * c["sphere_tests"] + 250.0 * c["hit_segments"]) / seg + 40.0


def bytes_per_ray(c):
    seg = max(1, c["segments"])
    return (32.0 * c["box_tests"] + 48.0 * c["tri_tests"] + 16.0 * c["sphere_tests"]) / seg


SKY = (0.8, 1.0, 1.0)  # SKY_COLOUR, src/main.cu:13


def cpu_baseline_leg(workload, threads):
    """The reference's own source on the host cores (oracle/_ref), else the C port: a bounded band of
    the same workload at a few spp.  Returns the cpu_baseline object.  Loads the product only for the port."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ctypes
    import numpy as np
    import oracle_binding as ob
    sid, p0, p1, W, H, spp, bounces = WORKLOADS[workload]
    if sid == 6:
        p0, p1 = 6, 5  # the host BVH of the full grid takes minutes to build; a 6x5 grid is the sample
    # 8 bands of H/128 rows spread evenly over the frame (1/16 of the pixels)
    nb, band = 8, max(1, H // 128)
    bands = [((2 * b + 1) * H // (2 * nb) - band // 2, (2 * b + 1) * H // (2 * nb) - band // 2 + band) for b in range(nb)]
    sample_spp = 4  # about 10 s of single-threaded CPU work on the C3 sample
    rays, dt = 0, 0.0
    if ob.have_ref_host():
        ref = ob.RefHost()
        ref.build_scene(sid, p0, p1)
        sky = SKY if ref.L.refh_use_sky() else (0.0, 0.0, 0.0)
        ref.L.refh_set_render(sample_spp, bounces, 1, *sky)
        ref.set_camera(W, H)
        out = np.zeros((H, W, 3), np.float32)
        for y0, y1 in bands:
            t0 = time.perf_counter()
            ref.L.refh_render(W, H, y0, y1, 12345, 0, None, out.ctypes.data_as(ctypes.c_void_p), threads)
            dt += time.perf_counter() - t0
            seg, hits = ctypes.c_ulonglong(), ctypes.c_ulonglong()
            ref.L.refh_count_segments(W, H, y0, y1, 12345, ctypes.byref(seg), ctypes.byref(hits), os.cpu_count() or 1)
            rays += seg.value
        kind = "reference"
    else:
        import raytracer_b200 as rtb
        scene = rtb.Scene(sid, p0, p1)
        rd = scene.render_settings(sample_spp, bounces)
        orc = ob.Oracle(scene)
        for y0, y1 in bands:
            cnt = rtb.Counters()
            t0 = time.perf_counter()
            orc.render(rtb.default_camera(W, H), W, H, rd, 12345, 0, None, y0=y0, y1=y1, nthreads=threads, counters=cnt)
            dt += time.perf_counter() - t0
            rays += cnt.segments
        kind = "port"
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
            "sample": f"{workload} scene at {W}x{H}, {nb} bands of {band} rows spread over the frame, {sample_spp} spp, {bounces} bounces: {rays} rays in {dt:.2f} s"
                      + (" (6x5 instance grid)" if sid == 6 else "")}


KERNEL_NAMES = {0: ["render_megakernel"], 1: ["wf_traverse", "wf_shade"], 3: ["pool_render"], 4: ["stream_render"]}


def ncu_traffic(workload, kernel_kind):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the render kernels.  NOT measured by this run (a
    number taken under a profiler is never a bench value, and ncu cannot run inside the timed process): read from the
    committed `ncu --set full` summaries of the same kernels on the same workload, newest round first, and labelled
    "source": "static ..."; None when no capture of this workload / kernel is committed."""
    out = {}
    for kern in KERNEL_NAMES.get(kernel_kind, []):
        path = None
        for rnd in ("r2", "r1e"):
            cand = os.path.join(ROOT, "profiles", f"{rnd}_ncu_{kern}_{workload}.txt")
            if os.path.exists(cand):
                path = cand
                break
        if path is None:
            return None
        total = 0.0
        for ln in open(path):
            if ln.startswith(("dram__bytes_read.sum,", "dram__bytes_write.sum,")):
                val, unit = ln.split(",", 1)[1].split()[:2]
                total += float(val) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
        out[kern + "_kernel"] = total
        out.setdefault("files", []).append(os.path.relpath(path, ROOT))
    if not out:
        return None
    out["unit"] = "bytes per launch"
    out["source"] = "static: committed ncu --set full summaries of the same kernels on this workload (see files), not measured in this run"
    return out


def workload_config(workload):
    """config of the JSON line: what one step computes.  Identical in the product arm and the reference arm."""
    sid, p0, p1, W, H, spp, bounces = WORKLOADS[workload]
    return {"workload": f"{workload}: {WORKLOAD_DESC[workload]}", "scene_id": sid, "width": W, "height": H, "spp_per_step": spp,
            "bounces": bounces,
            "l2": "scene is L2/L1-resident by design (not an HBM-bound path); the accumulation image is reset each step"}


def reference_step_spp(workload, requested):
    """spp per step of the reference arm: the workload's own (same config as the product arm) whenever a step of the
    reference then stays under ~10 s; the 4K configs, where its fixed-depth BVH degenerates, are bounded
    and reported as a rate."""
    if requested:
        return requested
    spp = WORKLOADS[workload][5]
    return {"C4": 1, "C5": min(spp, 32)}.get(workload, spp)


def reference_arm(args):
    """--impl reference: the reference's own CUDA build (baseline/_ref: unmodified sources, only the two resolution
    constants patched) on the product arm's config.  Rank 0 only.  Nothing of the product is loaded: the rays of a step
    are counted by the reference's own functions (refc_count_segments)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_binding as ob
    sid, p0, p1, W, H, spp, bounces = WORKLOADS[args.workload]
    base = {"impl": "reference", "metric": "Mrays/s", "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic"}
    config = workload_config(args.workload)
    threads = os.cpu_count() or 1
    if os.path.exists(ob.ref_cuda_path(W, H)):
        ref_spp = reference_step_spp(args.workload, args.ref_spp)
        config["spp_per_step"] = ref_spp
        ref = ob.RefCuda(W, H)
        ref.init(sid, p0, p1, ref_spp, bounces, True)
        rays, _hits = ref.count_segments(1000)   # same inputs as the first timed step, untimed
        img = np.zeros((H, W, 3), np.float32)
        sampler = ClockSampler(0)
        for k in range(args.warmup):
            ref.render_kernel(img, 0, 900 + k)
        sampler.start()
        kernel_ms = [ref.render_kernel(img, 0, 1000 + k) for k in range(args.steps)]
        clocks = sampler.stop()
        ms_step = sum(kernel_ms) / len(kernel_ms)
        # end to end through the reference's own render() (src/dispatch.cu:156), host image in and out: wall clock,
        # same spp per step; as many of the K steps as fit in about a minute (a C3 step of the reference is 6.3 s: with
        # the driver's K = 20 the kernel-timed part alone is 2.4 minutes)
        e2e_steps = max(1, min(args.steps, int(60.0 / max(ms_step / 1e3, 1e-3))))
        img2 = np.zeros((H, W, 3), np.float32)
        ref.render(img2, 0, 899)
        wall = []
        for k in range(e2e_steps):
            img2[:] = 0
            _, ms = ref.render(img2, 0, 1000 + k)
            wall.append(ms)
        value = rays / ms_step / 1e3
        e2e = rays / (sum(wall) / len(wall)) / 1e3
        nbytes = W * H * 3 * 4
        line = dict(base, value=value, ms_per_step=ms_step, clocks=clocks, gpu_launches=args.steps, rays_per_step=rays,
                    spp_per_s_1080p=(ref_spp / (ms_step / 1e3)) if (W, H) == (1920, 1080) else None,
                    e2e={"value": e2e, "unit": "Mrays/s", "h2d_bytes_per_step": nbytes + 8, "d2h_bytes_per_step": nbytes,
                         "ms_per_step": sum(wall) / len(wall), "steps": e2e_steps,
                         "note": "reference render(): per-frame cudaMalloc + H2D of previous_render, managed-memory read-back"},
                    config=config,
                    reference_build="the reference's own CUDA kernel get_pixel_colour, nvcc -arch=sm_100, default flags; stock launch shape and render()",
                    cpu_baseline=None if args.no_cpu_baseline else cpu_baseline_leg(args.workload, threads))
    else:
        cb = cpu_baseline_leg(args.workload, threads)
        line = dict(base, value=cb["value"], ms_per_step=None, cpu_baseline=cb, gpu_launches=0,
                    e2e={"value": cb["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    config=config, reference_build="reference CUDA library absent; host build of the reference source")
    emit(line)


_REAL_STDOUT = None


def emit(line):
    """the run's one JSON line, on the process's real stdout (see main)"""
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


KERNEL_DESC = {0: "persistent megakernel", 1: "wavefront: wf_shade + persistent wf_traverse per path segment",
               3: "pooled persistent kernel: one launch per frame, per-warp slot pools",
               4: "streaming kernel: one launch per frame, SMs specialised into shading and traversal roles"}


class Job:
    """One workload on this rank's GPU: scene, tracer, sharding and the exchange step of a frame."""

    def __init__(self, args, workload, sharding, rank, world, local_rank, dist):
        import torch
        import raytracer_b200 as rtb
        self.torch, self.dist, self.rank, self.world, self.local_rank = torch, dist, rank, world, local_rank
        self.workload = workload
        sid, p0, p1, W, H, spp, bounces = WORKLOADS[workload]
        self.W, self.H, self.spp, self.bounces = W, H, spp, bounces
        if sharding == "auto":
            # C4 is the tile-sharded config (partitioning A), C5 the sample-slice one (B: GPU g renders
            # the frames f with f mod N == g, its own time seeds, one all-reduce); the 1080p-and-below configs leave too
            # few pixels per GPU for tiles
            sharding = "tiles" if workload == "C4" else "slices"
        if world == 1:
            sharding = "none"
        self.sharding = sharding
        self.tiles, self.slices = sharding == "tiles", sharding == "slices"
        self.band_rows = args.band_rows
        self.scene = rtb.Scene(sid, p0, p1)
        self.tr = tr = rtb.Tracer(local_rank)
        if args.kernel is not None:
            tr.set_option("kernel", args.kernel)
        if args.builder is not None:
            tr.set_option("builder", args.builder)
        tr.setup(self.scene, W, H, spp, bounces)
        self.exchange_desc = "none"
        self.exchange_launches = 0
        if world > 1:
            # the exchange runs INSIDE the library (rtb_comm.cu) on its own NCCL communicator; torch.distributed is
            # only the plumbing that carries the 128-byte NCCL id and the CUDA IPC handles between the ranks
            ids = [rtb.Tracer.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            tr.comm_init(ids[0], world, rank)
            if self.tiles:
                tr.set_shard(rank, world, args.band_rows)
                fused = False
                if args.exchange != 1:
                    handles = [None] * world
                    dist.all_gather_object(handles, tr.peer_export())
                    try:
                        tr.peer_open(b"".join(handles), world, rank)
                        fused = True
                    except rtb.RtbError:
                        fused = False
                    flags = [None] * world
                    dist.all_gather_object(flags, fused)
                    fused = all(flags)
                tr.set_option("exchange", 2 if fused else 1)
                self.exchange_desc = ("fused into the render kernels: finished pixels are stored into every GPU's image over NVLink peer mappings, "
                                      "NCCL barrier per frame" if fused else "pack + ncclAllGather + unpack per frame")
                self.exchange_launches = 0 if fused else 2
            else:
                tr.set_option("slices", 1)
                tr.set_option("exchange", 1)
                self.exchange_desc = "ncclAllReduce(sum) of the accumulation buffer + scale per step"
                self.exchange_launches = 1
        self.ext = torch.cuda.ExternalStream(tr.stream(), device=torch.device("cuda", local_rank))

    def sharding_desc(self):
        if self.world == 1:
            return "none"
        if self.tiles:
            return f"tiles: row-cyclic bands of {self.band_rows} rows of one frame; exchange inside the library ({self.exchange_desc})"
        return (f"sample slices: {self.world} frames of {self.spp} spp per step (one per rank, own time seed); exchange inside the library "
                f"({self.exchange_desc})")

    def exchange(self):
        """the exchange of a frame is part of rtb_render_device / rtb_render; returns OUR kernels it launched"""
        return self.exchange_launches

    def frame_time(self, k):
        """time0 of step k; in sample-slice mode the library seeds this rank's frame with time0 + rank"""
        return 1000 + k * (self.world if self.slices else 1)

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, values, op="max"):
        t = self.torch.tensor([float(v) for v in values], device="cuda", dtype=self.torch.float64)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return t.tolist()

    def count(self):
        """op mix / ray count of one step: counter build, same inputs, untimed; summed over ranks"""
        tr = self.tr
        names = ("segments", "hit_segments", "box_tests", "tri_tests", "sphere_tests", "samples")
        with self.torch.cuda.stream(self.ext):
            tr.set_option("counters", 1)
            tr.reset()
            tr.render_device(1, self.frame_time(0))
            tr.sync()
            cnt = tr.counters()
            tr.set_option("counters", 0)
            vals = self.reduce([cnt[k] for k in names], "sum")
        return dict(zip(names, [int(x) for x in vals]))

    def device_step(self, k):
        """one step, everything resident: render my share, then the exchange; returns launched kernels"""
        self.tr.reset()
        self.tr.render_device(1, self.frame_time(k))
        return self.exchange()

    def timed(self, steps, sample_ms=200, first_k=0, sync_each_step=False):
        """`steps` steps bracketed by barrier + synchronize, CUDA events on the library's stream, max over ranks.
        Returns (ms per step, ms per step of the render kernels alone, launches, clocks of rank 0).
        sync_each_step: the host waits for every step before it enqueues the next (inside the timed region)."""
        torch = self.torch
        with torch.cuda.stream(self.ext):
            per_frame = int(self.tr.info("launches_per_frame"))
            self.barrier()
            sampler = ClockSampler(self.local_rank, sample_ms)
            if self.rank == 0:
                sampler.start()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            launches = 0
            ev0.record(self.ext)
            for k in range(steps):
                kev[k][0].record(self.ext)
                self.tr.reset()
                self.tr.render_device(1, self.frame_time(first_k + k))
                kev[k][1].record(self.ext)
                launches += per_frame + self.exchange()
                if sync_each_step:
                    self.tr.sync()
            ev1.record(self.ext)
            self.barrier()
            clocks = sampler.stop() if self.rank == 0 else None
            ms_total = ev0.elapsed_time(ev1)
            kernel_ms = sum(a.elapsed_time(b) for a, b in kev) / steps
            ms_total, kernel_ms = self.reduce([ms_total, kernel_ms], "max")
        return ms_total / steps, kernel_ms, launches, clocks

    def kernel_desc(self):
        tr = self.tr
        kind = int(tr.info("kernel"))
        tuned = (f"auto: {KERNEL_DESC[kind].split(':')[0]} kept after timing the variants on a {min(self.spp, 8)}-spp frame "
                 f"(wavefront {tr.info('tune_wavefront_ms'):.2f} ms, megakernel {tr.info('tune_megakernel_ms'):.2f} ms)"
                 if int(tr.info("kernel_tuned")) else "fixed by --kernel")
        policy = (f", traversal policy walk_min {int(tr.info('walk_min'))} / refill {int(tr.info('wf_refill'))}" if kind != 0 else
                  f", walk_min {int(tr.info('walk_min'))}" + (", stack entries fetched with the node (early-pop variant)" if int(tr.info("mega_early_pop")) else ""))
        return kind, KERNEL_DESC[kind] + (", PLOC tree" if int(tr.info("builder")) == 1 else ", Karras LBVH") + policy, tuned

    def close(self):
        if self.world > 1 and self.tiles:
            self.tr.peer_close()
        self.barrier()
        self.tr.close()


def c4_tiles_record(args, rank, world, local_rank, dist):
    """The strong-scaling configuration of BASELINE.json (config 4: 1 012 200 triangles at 3840x2160, screen tiles
    with the image gathered over NCCL), measured in THIS run at THIS world size next to the headline: the 1-GPU line
    is the denominator of the N = 2/4/8 lines.  16 spp per step (of the config's 1024: a rate)."""
    import math
    note("c4: building the job")
    job = Job(args, "C4", "tiles", rank, world, local_rank, dist)
    note("c4: job ready:", job.sharding_desc())
    cnt = job.count()
    note("c4: counted", cnt["segments"])
    with job.torch.cuda.stream(job.ext):
        for k in range(3):
            job.device_step(k)
    ms_warm, _, _, _ = job.timed(2)
    # >= 5 s timed: >= 20 clock samples at 200 ms
    steps = max(5, min(600, int(math.ceil(5200.0 / max(ms_warm, 1e-3)))))
    ms_step, kernel_ms, launches, clocks = job.timed(steps, sample_ms=200)
    kind, kdesc, tuned = job.kernel_desc()
    rec = None
    if rank == 0:
        rec = {"workload": f"C4: {WORKLOAD_DESC['C4']}", "metric": "Mrays/s", "value": cnt["segments"] / ms_step / 1e3, "unit": "Mrays/s",
               "n_gpus": world, "scaling": "strong", "steps": steps, "warmup": 5, "ms_per_step": ms_step, "render_ms_per_step": kernel_ms,
               "rays_per_step": cnt["segments"], "spp_per_step": job.spp, "sharding": job.sharding_desc(), "kernel": kdesc, "kernel_choice": tuned,
               "prims": int(job.tr.info("num_prims")), "lbvh_build_ms": job.tr.info("build_ms"), "commit_ms": job.tr.info("commit_ms"),
               "gather_bytes_per_step": (job.W * job.H * 3 * 4) if world > 1 else 0, "gpu_launches": launches, "clocks": clocks,
               "op_mix_per_ray": {"box": cnt["box_tests"] / cnt["segments"], "tri": cnt["tri_tests"] / cnt["segments"]}}
    job.close()
    return rec


def note(*a):
    """progress notes on stderr (RTB_BENCH_VERBOSE=1)"""
    if os.environ.get("RTB_BENCH_VERBOSE"):
        print(f"[bench rank {os.environ.get('RANK', '0')} t={time.perf_counter():.1f}]", *a, file=sys.stderr, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="rtb")
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--band-rows", type=int, default=8)
    ap.add_argument("--sharding", default="auto", choices=["auto", "tiles", "slices"])
    ap.add_argument("--ref-spp", type=int, default=0, help="spp per step of the reference arm (0 = the workload's own where a step stays under ~10 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c4-tiles", action="store_true", help="skip the C4 screen-tile record that rides along with the C3 headline")
    ap.add_argument("--exchange", type=int, default=2, help="tiles at N > 1: 2 = fused peer-store exchange (falls back to 1 without peer access), 1 = NCCL all-gather")
    ap.add_argument("--kernel", type=int, default=None, help="0 = persistent megakernel, 1 = wavefront, 3 = pooled, 4 = streaming, 2 = timed choice (library default)")
    ap.add_argument("--builder", type=int, default=None, help="0 = Karras LBVH, 1 = PLOC (library default)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    # The contract is ONE JSON line on stdout.  Libraries print on fd 1 as well (NCCL's version banner at the first
    # communicator, for one): from here on fd 1 is stderr, and the line goes to the real stdout through emit().
    sys.stdout.flush()
    global _REAL_STDOUT
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return reference_arm(args)

    import numpy as np
    import torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the render path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    note("building the job")
    job = Job(args, args.workload, args.sharding, rank, world, local_rank, dist)
    note("job ready:", job.sharding_desc())
    tr, W, H, spp = job.tr, job.W, job.H, job.spp
    cnt = job.count()
    rays_per_step = cnt["segments"]
    with torch.cuda.stream(job.ext):
        for k in range(args.warmup):
            job.device_step(k)
    # with option "kernel" = 2 (the default) the first warm-up frame timed the variants on this rank's share of the
    # workload and kept the fastest; from here on the choice is fixed
    note("warm-up done")
    ms_step, kernel_ms, launches, clocks = job.timed(args.steps)
    note("timed:", ms_step)
    kernel_kind, kernel_desc, tuned = job.kernel_desc()

    # ---- end to end through the reference-facing entry point, host buffers ----
    with torch.cuda.stream(job.ext):
        host = torch.zeros(H * W * 3, dtype=torch.float32).pin_memory()
        himg = host.numpy()
        nbytes = H * W * 3 * 4

        def e2e_step(k):
            if world == 1:
                himg[:] = 0  # the caller's previous_render (frame 0)
                tr.render(himg, 0, 1000 + k)  # H2D + kernel + D2H + sync
            else:
                himg[:] = 0
                tr.render(himg, 0, job.frame_time(k))  # H2D + kernel + exchange + D2H + sync, every rank gets the image

        e2e_step(0)
        job.barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            e2e_step(k)
        job.barrier()
        e2e_ms = job.reduce([(time.perf_counter() - t0) * 1e3], "max")[0]
        checksum = float(himg.astype(np.float64).sum())
        fp32_peak = tr.fp32_peak_tflops(40.0) if rank == 0 else None
    info = {k: tr.info(k) for k in ("builder", "num_prims", "build_ms")}
    sharding_desc = job.sharding_desc()
    slices = job.slices
    note("e2e done")
    job.close()
    note("job closed")

    c4 = None
    if args.workload == "C3" and not args.no_c4_tiles:
        c4 = c4_tiles_record(args, rank, world, local_rank, dist)
    note("c4 record done")

    if rank == 0:
        value = rays_per_step / ms_step / 1e3                       # Mrays/s, whole job
        e2e_value = rays_per_step / (e2e_ms / args.steps) / 1e3
        fpr = flops_per_ray(cnt)
        peaks = load_peaks()
        nominal_peak = 148 * 128 * 2 * (peaks["sm_max_mhz"] if peaks else 1965.0) * 1e6 / 1e12
        kernel_rate = rays_per_step / world / kernel_ms / 1e3      # Mrays/s of one GPU's render kernels
        achieved_tf = kernel_rate * 1e6 * fpr / 1e12
        hbm_peak = peaks["hbm_gbs"] if peaks else 6650.0
        achieved_gbs = (kernel_rate * 1e6 * bytes_per_ray(cnt) + (W * H // world) * 24 / (kernel_ms / 1e3)) / 1e9
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak" if slices else "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "spp_per_s_1080p": (spp * (world if slices else 1) / (ms_step / 1e3)) if (W, H) == (1920, 1080) else None,
            "config": workload_config(args.workload),
            "rays_per_step": rays_per_step, "segments_per_sample": cnt["segments"] / max(1, cnt["samples"]),
            "sharding": sharding_desc, "kernel": kernel_desc, "kernel_choice": tuned, "prims": int(info["num_prims"]),
            "lbvh_build_ms": info["build_ms"],
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes,
                    "ms_per_step": e2e_ms / args.steps, "checksum": checksum},
            "roofline": {"bound": "fp32", "kernel": " + ".join(n + "_kernel" for n in KERNEL_NAMES[kernel_kind]) + " (all launches of a step)",
                         "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved_tf / fp32_peak if fp32_peak else None,
                         "peak_kind": "measured FFMA microbenchmark on this GPU (rtb_measure_fp32_peak)", "peak_nominal": nominal_peak,
                         "frac_of_nominal": achieved_tf / nominal_peak, "traffic": ncu_traffic(args.workload, kernel_kind),
                         "flops_per_ray": fpr, "kernel_ms": kernel_ms,
                         "op_mix_per_ray": {"box": cnt["box_tests"] / cnt["segments"], "tri": cnt["tri_tests"] / cnt["segments"],
                                            "sphere": cnt["sphere_tests"] / cnt["segments"], "hit": cnt["hit_segments"] / cnt["segments"]},
                         "why": "no dense contraction and an L2-resident working set: the FP32 pipe bounds this path (SURVEY.md 8(d))",
                         "algorithmic_bytes": {"achieved": achieved_gbs, "unit": "GB/s", "peak_hbm": hbm_peak,
                                               "note": "node/primitive bytes the traversal asks for (32 B per box, 48 B per triangle): served by L1/L2, NOT HBM traffic; "
                                                       "measured dram bytes are in profiles/ (ncu)",
                                               "peak_kind": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"}},
        }
        if c4 is not None:
            line["c4_tiles"] = c4
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_leg(args.workload, 1)
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

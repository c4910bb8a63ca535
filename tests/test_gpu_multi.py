"""Multi-GPU behind the C ABI: tiles (bit-identical to one GPU) and sample slices (the same
running mean), through NCCL and through the fused peer-store exchange.

The single-GPU tests check the arithmetic of the sample-slice partitioning against the 1-GPU progressive render
(src/raytracer.cu:107-112 is the running mean it must reproduce).  The tests that need two devices are skipped on a
one-GPU box; `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu` runs them
(profiles/r2_multi_gpu_tests.log is that run)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import raytracer_b200 as rtb

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


two_gpus = pytest.mark.skipif(_ngpus() < 2, reason="needs two CUDA devices (gpurun --gpus 2)")
W, H = 200, 152


def test_sample_slices_equal_the_progressive_render():
    """C5's partitioning: frame f of a progressive render goes to GPU f mod world with
    current_time_ms = 1000 + f; the mean of the ranks' running means is the 1-GPU running mean over all frames
    (src/raytracer.cu:107-112), up to the rounding order of the float sums.  Ranks emulated one after the other on
    one device; the noise of a 4-spp frame is ~0.3 per pixel, so a wrong seed or weight would show as ~1e-1."""
    frames, world, spp, bounces = 16, 8, 4, 5
    scene = rtb.Scene(0)
    tr = rtb.Tracer(0).setup(scene, W, H, spp, bounces)
    tr.reset()
    tr.render_device(frames, 1000, 1)
    tr.sync()
    whole, f = tr.download()
    assert f == frames
    acc = np.zeros((H, W, 3), np.float64)
    for rank in range(world):
        tr.reset()
        tr.render_device(frames // world, 1000 + rank, world)   # frames rank, rank + world, ...
        tr.sync()
        part, f = tr.download()
        assert f == frames // world
        acc += part
    tr.close()
    mean = (acc / world).astype(np.float32)
    err = float(np.abs(mean - whole).max())
    rmse = float(np.sqrt(((mean - whole) ** 2).mean()))
    print(f"sample slices vs progressive render: max abs diff {err:.3e}, RMSE {rmse:.3e}")
    assert err < 2e-5 * max(1.0, float(whole.max())) and rmse < 1e-6


def test_single_context_ignores_the_exchange_options():
    """a context without a communicator renders the same bits whatever the exchange options say"""
    scene = rtb.Scene(3)
    tr = rtb.Tracer(0).setup(scene, 96, 64, 2, 4)
    a = np.zeros((64, 96, 3), np.float32)
    tr.render(a, 0, 5)
    tr.set_option("exchange", 2)
    tr.set_option("slices", 1)
    b = np.zeros((64, 96, 3), np.float32)
    tr.render(b, 0, 5)
    tr.close()
    assert (a == b).all()


@two_gpus
@pytest.mark.parametrize("exchange", [1, 2])
def test_one_process_two_gpus_tiles_bit_identical(exchange):
    """rtb_create_multi + rtb_render_multi: the two-GPU frame (tiles; NCCL all-gather, then the fused peer-store
    exchange) equals the one-GPU frame bit for bit, over two progressive frames"""
    scene = rtb.Scene(0)
    one = rtb.Tracer(0).setup(scene, W, H, 3, 5)
    ref = np.zeros((H, W, 3), np.float32)
    f = one.render(ref, 0, 77)
    one.render(ref, f, 78)
    one.close()
    mt = rtb.MultiTracer([0, 1]).setup(scene, W, H, 3, 5)
    mt.set_option("exchange", exchange)
    img = np.zeros((H, W, 3), np.float32)
    f = mt.render(img, 0, 77)
    f = mt.render(img, f, 78)
    fused = int(mt.tracers[0].info("exchange_fused"))
    mt.close()
    assert f == 2
    assert (img == ref).all()
    if exchange == 2:
        print(f"fused exchange active: {fused}")


@two_gpus
def test_one_process_two_gpus_sample_slices():
    scene = rtb.Scene(1)
    one = rtb.Tracer(0).setup(scene, W, H, 4, 5)
    one.reset()
    one.render_device(2, 500, 1)   # frames seeded 500, 501
    one.sync()
    ref, _ = one.download()
    one.close()
    mt = rtb.MultiTracer([0, 1]).setup(scene, W, H, 4, 5)
    mt.set_option("slices", 1)
    for t in mt.tracers:
        t.set_shard(0, 1, 8)   # sample slices render whole frames
    img = np.zeros((H, W, 3), np.float32)
    assert mt.render(img, 0, 500) == 1   # rank r renders the frame seeded 500 + r; all-reduce; / 2
    mt.close()
    assert float(np.abs(img - ref).max()) < 2e-5 * max(1.0, float(ref.max()))


_WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, os.environ["RTB_ROOT"])
import raytracer_b200 as rtb
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
W, H = 200, 152
scene = rtb.Scene(0)
tr = rtb.Tracer(rank).setup(scene, W, H, 3, 5)
ids = [rtb.Tracer.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
tr.comm_init(ids[0], world, rank)
tr.set_shard(rank, world, 8)
handles = [None] * world
dist.all_gather_object(handles, tr.peer_export())
tr.peer_open(b"".join(handles), world, rank)
out = {}
for exchange in (1, 2):
    tr.set_option("exchange", exchange)
    tr.reset()
    tr.render_device(2, 77, 1)
    tr.sync()
    img, f = tr.download()
    out[exchange] = img
    assert f == 2
    if exchange == 2:
        assert int(tr.info("exchange_fused")) == 1
# host-buffer entry point too
tr.set_option("exchange", 2)
host = np.zeros((H, W, 3), np.float32)
f = tr.render(host, 0, 77)
tr.render(host, f, 78)
tr.peer_close()
tr.close()
if rank == 0:
    one = rtb.Tracer(0).setup(scene, W, H, 3, 5)
    ref = np.zeros((H, W, 3), np.float32)
    f = one.render(ref, 0, 77)
    one.render(ref, f, 78)
    one.close()
    assert (out[1] == ref).all(), "NCCL all-gather tiles differ from one GPU"
    assert (out[2] == ref).all(), "fused peer-store tiles differ from one GPU"
    assert (host == ref).all(), "rtb_render with the fused exchange differs from one GPU"
    print("MULTI_PROCESS_OK")
dist.barrier()
dist.destroy_process_group()
'''


@two_gpus
def test_one_process_per_gpu_tiles_bit_identical(tmp_path):
    """the torchrun shape: one process per GPU, the NCCL id and the CUDA IPC handles of the accumulation images
    travel over the launcher's own plumbing (gloo here), the exchange runs inside the library"""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, RTB_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29577", str(script)], capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert r.returncode == 0 and "MULTI_PROCESS_OK" in r.stdout, (r.stdout + r.stderr)[-3000:]

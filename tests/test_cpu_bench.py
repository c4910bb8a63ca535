"""Host logic of bench.py that needs no GPU: the reference arm's rank rule, the loud failure without a CUDA
device, the roofline inputs."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _run(args, env_extra):
    env = dict(os.environ, **env_extra)
    env.pop("CUDA_VISIBLE_DEVICES", None)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)


def test_reference_arm_runs_on_rank_zero_only():
    """Under torchrun every rank is started with --impl reference; ranks other than 0 exit 0 without work or output."""
    r = _run(["--impl", "reference", "--gpus", "2"], {"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0, r.stderr
    assert r.stdout.strip() == ""


def test_product_arm_fails_loudly_without_a_gpu():
    """No CPU fallback: on a machine without a CUDA device the product arm stops with a message, it does not
    route through the oracle."""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("this machine has a CUDA device")
    r = _run(["--steps", "1"], {})
    assert r.returncode != 0
    assert "no CPU fallback" in (r.stderr + r.stdout)
    assert not any(ln.startswith("{") for ln in r.stdout.splitlines())


def test_roofline_inputs():
    import bench
    cnt = {"segments": 1000, "hit_segments": 900, "box_tests": 27000, "tri_tests": 4000, "sphere_tests": 250, "samples": 160}
    # flops/ray = 24 n_box + 48 n_tri + 30 n_sph + 250 [hit] + 40
    assert abs(bench.flops_per_ray(cnt) - (24 * 27 + 48 * 4 + 30 * 0.25 + 250 * 0.9 + 40)) < 1e-9
    assert bench.bytes_per_ray(cnt) > 0
    t = bench.ncu_traffic("C3", 1)
    assert t is not None and t["wf_traverse_kernel"] > 1e6 and t["wf_shade_kernel"] > 1e6
    assert bench.ncu_traffic("C3", 0) is None and bench.ncu_traffic("C4", 1) is None
    for name, (sid, p0, p1, W, H, spp, bounces) in bench.WORKLOADS.items():
        assert name in bench.WORKLOAD_DESC and W > 0 and H > 0 and spp > 0 and bounces > 0
    json.dumps(t)

"""The checked build (librtb_checks.so: -DRTB_CHECKS, device-side asserts on every queue / slot / stack / node
index of the three render kernels, see RTB_ASSERT in csrc/rtb_device.cuh) runs the kernel-variant parity tests on
all eight scenes plus the sharded, edge-case and deep-tree cases.  A failed assert traps the kernel, the C ABI call
returns an error and the inner pytest run fails.  This is the stand-in for compute-sanitizer's memcheck, which is
closed on this GPU pool (gpurun_out/memcheck_smoke.log, round 1)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECKED = os.path.join(ROOT, "ray-tracer_b200", "librtb_checks.so")


@pytest.mark.skipif(not os.path.exists(CHECKED), reason="checked build not present (make -C ray-tracer_b200 checks)")
def test_parity_under_device_side_index_asserts():
    env = dict(os.environ, RTB_LIBRTB=CHECKED)
    sel = ("wavefront_equals_megakernel or lbvh_equals_brute_force or ploc_tree_equals_brute_force or pool_kernel_edge_cases "
           "or wavefront_sharded_and_odd_sizes or wavefront_row_groups_and_streams or edge_cases or deep_tree")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_parity.py"), "-x", "-q", "-m", "gpu", "-k", sel,
                        "-p", "no:cacheprovider"], capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    tail = (r.stdout + r.stderr)[-2000:]
    assert r.returncode == 0, tail
    assert " passed" in r.stdout and "librtb_checks.so" in env["RTB_LIBRTB"], tail


@pytest.mark.skipif(not os.path.exists(CHECKED), reason="checked build not present")
def test_checked_build_is_the_one_loaded():
    """the inner run really uses the checked library: it exports the marker symbol the plain build does not"""
    code = ("import os, ctypes; import raytracer_b200 as rtb; "
            "print(int(rtb.lib.rtb_build_has_checks()))")
    out = {}
    for name, path in (("checked", CHECKED), ("plain", os.path.join(ROOT, "ray-tracer_b200", "librtb.so"))):
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=dict(os.environ, RTB_LIBRTB=path), cwd=ROOT, timeout=300)
        assert r.returncode == 0, r.stderr[-1000:]
        out[name] = r.stdout.strip().splitlines()[-1]
    assert out == {"checked": "1", "plain": "0"}

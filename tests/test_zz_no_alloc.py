"""rtb_primary_ids_noalloc (include/rtb.h): the primary-visibility pass with its device staging planes in the
library's static device array -- no cudaMalloc in the call.  No counterpart in the reference;
the checker is the existing entry point, rtb_primary_ids, on the same context: ids and t are
integers / the output of the same kernel, so the bar is bit-exact, as in the other id tests.

Collected last (file name): a failure here cannot hide another test under `pytest -x`."""
import numpy as np
import pytest

import raytracer_b200 as rtb

pytestmark = pytest.mark.gpu


def _same(a, b):
    for x, y in zip(a, b):
        assert x.dtype == y.dtype and x.shape == y.shape
        assert (x.view(np.uint32) == y.view(np.uint32)).all()


def test_noalloc_equals_primary_ids_two_calls_in_a_row():
    """two calls in a row on two different scenes and sizes: the second call must not see the first call's
    contents of the shared array (a smaller image after a larger one, other ids)"""
    assert rtb.lib.rtb_noalloc_max_pixels() == 2 * 1024 * 1024
    big = rtb.Tracer(0).setup(rtb.Scene(0), 320, 256, 1, 2)
    small = rtb.Tracer(0).setup(rtb.Scene(3), 200, 152, 1, 2)
    want_big, want_small = big.primary_ids(), small.primary_ids()
    assert (want_big[0] >= 0).any() and (want_small[0] >= 0).any()
    got_big = big.primary_ids_noalloc()        # call 1
    got_small = small.primary_ids_noalloc()    # call 2, same array, other context of the same device
    _same(got_big, want_big)
    _same(got_small, want_small)
    # and the same context twice in a row, strict-fp kernels too
    _same(big.primary_ids_noalloc(), want_big)
    big.set_option("strict_fp", 1)
    want_strict = big.primary_ids()
    _same(big.primary_ids_noalloc(), want_strict)
    _same(big.primary_ids_noalloc(), want_strict)
    big.close()
    small.close()


def test_noalloc_accepts_device_buffers_and_null_t():
    """the caller's buffers may be device memory (what a caller that cannot allocate inside the call passes); t may
    be NULL as for rtb_primary_ids"""
    import torch
    tr = rtb.Tracer(0).setup(rtb.Scene(1), 200, 152, 1, 2)
    want = tr.primary_ids()
    n = 200 * 152
    obj = torch.full((n,), -7, dtype=torch.int32, device="cuda:0")
    sub = torch.full((n,), -7, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    rc = rtb.lib.rtb_primary_ids_noalloc(tr.ctx, obj.data_ptr(), sub.data_ptr(), None)
    assert rc == 0, rtb.lib.rtb_last_error(tr.ctx).decode()
    assert (obj.cpu().numpy().reshape(152, 200) == want[0]).all()
    assert (sub.cpu().numpy().reshape(152, 200) == want[1]).all()
    tr.close()


def test_noalloc_refuses_an_image_beyond_the_static_array():
    """3840 x 2160 = 8.3 M pixels > 2 Mi: a clear error, nothing written"""
    tr = rtb.Tracer(0).setup(rtb.Scene(0), 3840, 2160, 1, 2)
    obj = np.full(4, 123, np.int32)
    sub = np.full(4, 123, np.int32)
    rc = rtb.lib.rtb_primary_ids_noalloc(tr.ctx, obj.ctypes.data, sub.ctypes.data, None)
    msg = rtb.lib.rtb_last_error(tr.ctx).decode()
    assert rc != 0 and "exceeds the static scratch" in msg and "2097152" in msg, msg
    assert (obj == 123).all() and (sub == 123).all()
    tr.close()
    # 1920 x 1080 fits (the C3 frame)
    tr = rtb.Tracer(0).setup(rtb.Scene(0), 1920, 1080, 1, 2)
    _same(tr.primary_ids_noalloc(), tr.primary_ids())
    tr.close()

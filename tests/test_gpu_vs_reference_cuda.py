"""GPU parity against the reference's OWN CUDA build (baseline/_ref/libref_cuda_320x256.so, built
here from the unmodified reference sources with only the resolution constants patched): the strict
comparison the north_star names.  Skipped when the prebuilt library is absent."""
import os

import numpy as np
import pytest

import oracle_binding as ob
import raytracer_b200 as rtb

pytestmark = pytest.mark.gpu
W, H = 320, 256
needs_ref = pytest.mark.skipif(not os.path.exists(ob.ref_cuda_path(W, H)), reason="reference CUDA build not present")

SCENES = [(0, 0, 0), (1, 0, 0), (2, 0, 0), (3, 0, 0), (4, 60, 7), (5, 0, 0), (6, 4, 3), (7, 0, 0)]


@needs_ref
@pytest.mark.parametrize("sid,p0,p1", SCENES)
def test_ids_and_radiance_vs_reference_cuda(sid, p0, p1):
    spp, bounces = 8, 5
    ref = ob.RefCuda(W, H)
    ref.init(sid, p0, p1, spp, bounces, True)
    r_obj, r_t = ref.primary_ids()
    r_img = np.zeros((H, W, 3), np.float32)
    ref.render_kernel(r_img, 0, 12345)
    r_img2 = r_img.copy()
    ref.render_kernel(r_img2, 1, 777)

    scene = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0).setup(scene, W, H, spp, bounces)
    obj, sub, t = tr.primary_ids()
    img = np.zeros((H, W, 3), np.float32)
    tr.render(img, 0, 12345)
    img2 = img.copy()
    tr.render(img2, 1, 777)
    tr.close()

    mism = int((obj != r_obj).sum())
    rmse = float(np.sqrt(((img - r_img) ** 2).mean()))
    rmse2 = float(np.sqrt(((img2 - r_img2) ** 2).mean()))
    ident = float((img == r_img).mean())
    print(f"scene {sid}: id mismatches {mism}/{W*H}, t bit-equal {float((t == r_t).mean()):.6f}, "
          f"RMSE {rmse:.3e} / frame1 {rmse2:.3e}, bit-identical floats {ident:.4f}")
    # bit-exact ids for identical camera rays (north_star); the default-flag builds of both sides
    # contract FMAs, so a few silhouette pixels may still differ -- measured and bounded
    assert mism <= W * H // 2000, f"{mism} id mismatches vs the reference CUDA build"
    # 8 spp is far from converged: ONE path that takes a different discrete turn under a light of strength 6
    # moves its pixel by 0.75, an RMSE of 2.6e-3 over the frame.  So: RMSE < 1e-3, or at most 3 differing pixels
    # (scene 7, glass gem under two lights: 1 pixel measured); test_converged_radiance_vs_reference_cuda holds
    # the 1e-3 bar itself at 128 spp.
    bad = int((np.abs(img - r_img).max(axis=2) > 1e-5).sum())
    bad2 = int((np.abs(img2 - r_img2).max(axis=2) > 1e-5).sum())
    assert (rmse < 1e-3 or bad <= 3) and (rmse2 < 1e-3 or bad2 <= 3), f"RMSE {rmse} / {rmse2}, differing pixels {bad} / {bad2}"


@needs_ref
@pytest.mark.parametrize("kernel", [0, 1])
def test_converged_radiance_vs_reference_cuda(kernel):
    """north_star acceptance: converged radiance within RMSE < 1e-3 (PSNR > 45 dB) of the reference
    render at matched spp -- default build of both sides, glass scene and monkey scene, 128 spp,
    megakernel and wavefront."""
    for sid in (3, 0, 7):   # 7: OBJ + MTL scene (image-textured, glass, emissive and glossy meshes)
        spp, bounces = 128, 5
        ref = ob.RefCuda(W, H)
        ref.init(sid, 0, 0, spp, bounces, True)
        r_img = np.zeros((H, W, 3), np.float32)
        ref.render_kernel(r_img, 0, 2024)
        scene = rtb.Scene(sid, 0, 0)
        tr = rtb.Tracer(0)
        tr.set_option("kernel", kernel)
        tr.setup(scene, W, H, spp, bounces)
        img = np.zeros((H, W, 3), np.float32)
        tr.render(img, 0, 2024)
        tr.close()
        rmse = float(np.sqrt(((img - r_img) ** 2).mean()))
        psnr = 20 * np.log10(max(1.0, float(r_img.max())) / max(rmse, 1e-12))
        print(f"scene {sid} kernel {kernel} at {spp} spp: RMSE {rmse:.3e}, PSNR {psnr:.1f} dB, bit-identical {float((img == r_img).mean()):.4f}")
        assert rmse < 1e-3 and psnr > 45


@needs_ref
def test_reference_render_entry_point_matches_its_kernel():
    """render() (src/dispatch.cu:156) and the harness launch of the same kernel agree bit for bit."""
    ref = ob.RefCuda(W, H)
    ref.init(1, 0, 0, 2, 5, True)
    a = np.zeros((H, W, 3), np.float32)
    b = np.zeros((H, W, 3), np.float32)
    ref.render_kernel(a, 0, 42)
    f, _ = ref.render(b, 0, 42)
    assert f == 1 and (a == b).all()

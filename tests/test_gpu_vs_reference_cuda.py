"""GPU parity against the reference's OWN CUDA build (baseline/_ref/libref_cuda_320x256.so, built
here from the unmodified reference sources with only the resolution constants patched): the strict
comparison the north_star names.  Skipped when the prebuilt library is absent."""
import json
import os

import numpy as np
import pytest

import oracle_binding as ob
import raytracer_b200 as rtb

pytestmark = pytest.mark.gpu
W, H = 320, 256
needs_ref = pytest.mark.skipif(not os.path.exists(ob.ref_cuda_path(W, H)), reason="reference CUDA build not present")

SCENES = [(0, 0, 0), (1, 0, 0), (2, 0, 0), (3, 0, 0), (4, 60, 7), (5, 0, 0), (6, 4, 3), (7, 0, 0)]

# Measured parity numbers of this run go to gpurun_out/r2_parity.json (the copy under profiles/ is the one
# measured on the B200 box of round 2): what the bounds asserted below were tightened to.
_RECORD = {}


def _record(key, **values):
    _RECORD[key] = values
    out = os.path.join(os.environ.get("GRAFT_REPO_ROOT", ob.ROOT), "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        path = os.path.join(out, "r2_parity.json")
        old = json.load(open(path)) if os.path.exists(path) else {}
        old.update(_RECORD)
        json.dump(old, open(path, "w"), indent=1, sort_keys=True)
    except OSError:
        pass


# id mismatches of the default (-fmad=true) kernels against the reference's default CUDA build, 81 920 pixels.
# Measured on B200 (profiles/r2_parity.json): 0 on every scene, distances bit-equal on every pixel -- nvcc contracts
# the primitive tests of both programs the same way -- so the bound is the north_star's own: bit-exact.
MAX_ID_MISMATCH = {0: 0, 1: 0, 2: 0, 3: 0, 4: 0, 5: 0, 6: 0, 7: 0}


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
    _record(f"fast_vs_reference_cuda_320x256_scene{sid}", id_mismatches=mism, pixels=W * H, t_bit_equal=float((t == r_t).mean()),
            rmse_8spp=rmse, rmse_8spp_frame1=rmse2, bit_identical_floats=ident)
    # bit-exact ids for identical camera rays (north_star); the default-flag builds of both sides
    # contract FMAs, so a few silhouette pixels may still differ -- measured and bounded per scene
    assert mism <= MAX_ID_MISMATCH[sid], f"{mism} id mismatches vs the reference CUDA build"
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
    render at matched spp -- default build of both sides, 128 spp, megakernel and wavefront."""
    # wavefront pair: all eight scenes; megakernel: glass, monkey room and the OBJ + MTL scene (image-textured, glass,
    # emissive and glossy meshes)
    for sid, p0, p1 in (SCENES if kernel == 1 else [(3, 0, 0), (0, 0, 0), (7, 0, 0)]):
        spp, bounces = 128, 5
        ref = ob.RefCuda(W, H)
        ref.init(sid, p0, p1, spp, bounces, True)
        r_img = np.zeros((H, W, 3), np.float32)
        ref.render_kernel(r_img, 0, 2024)
        scene = rtb.Scene(sid, p0, p1)
        tr = rtb.Tracer(0)
        tr.set_option("kernel", kernel)
        tr.setup(scene, W, H, spp, bounces)
        img = np.zeros((H, W, 3), np.float32)
        tr.render(img, 0, 2024)
        tr.close()
        rmse = float(np.sqrt(((img - r_img) ** 2).mean()))
        _record(f"converged_128spp_scene{sid}_kernel{kernel}", rmse=rmse, bit_identical_floats=float((img == r_img).mean()))
        psnr = 20 * np.log10(max(1.0, float(r_img.max())) / max(rmse, 1e-12))
        print(f"scene {sid} kernel {kernel} at {spp} spp: RMSE {rmse:.3e}, PSNR {psnr:.1f} dB, bit-identical {float((img == r_img).mean()):.4f}")
        assert rmse < 1e-3 and psnr > 45


@needs_ref
def test_c5_sample_slices_vs_the_reference_progressive_render():
    """C5 as SURVEY.md 8(d) specifies it (at 320x256, the size the reference library here is built for): a progressive
    render of 128 frames x 128 spp = 16 384 spp; GPU g of 8 renders the frames f with f mod 8 == g, current_time_ms =
    1000 + f; the per-GPU running means are summed and divided by 8 (the all-reduce of partitioning B, emulated rank
    by rank on one device: the arithmetic of rtb_comm.cu's slice exchange).  The checker is the reference's OWN CUDA
    build rendering the same 128 frames one after the other through its running mean (src/raytracer.cu:107-112).
    Bars: RMSE < 1e-3 and PSNR > 45 dB (the north_star's), with the noise of the estimate itself far below (16 384
    spp); a wrong seed, weight or frame count shows as ~1e-2."""
    frames, spp, bounces, world = 128, 128, 8, 8
    ref = ob.RefCuda(W, H)
    ref.init(0, 0, 0, spp, bounces, True)
    r_img = np.zeros((H, W, 3), np.float32)
    for f in range(frames):
        ref.render_kernel(r_img, f, 1000 + f)
    scene = rtb.Scene(0)
    tr = rtb.Tracer(0).setup(scene, W, H, spp, bounces)
    acc = np.zeros((H, W, 3), np.float64)
    for rank in range(world):
        tr.reset()
        tr.render_device(frames // world, 1000 + rank, world)   # frames rank, rank + 8, ...
        tr.sync()
        part, n = tr.download()
        assert n == frames // world
        acc += part
    sliced = (acc / world).astype(np.float32)
    # and the same frames on ONE context, in order: what the slices must reproduce up to the order of the float sums
    tr.reset()
    tr.render_device(frames, 1000, 1)
    tr.sync()
    whole, n = tr.download()
    tr.close()
    assert n == frames
    rmse = float(np.sqrt(((sliced - r_img) ** 2).mean()))
    psnr = 20 * np.log10(max(1.0, float(r_img.max())) / max(rmse, 1e-12))
    own = float(np.abs(sliced - whole).max())
    print(f"C5 at {W}x{H}: sliced vs reference progressive render RMSE {rmse:.3e}, PSNR {psnr:.1f} dB; sliced vs own 1-GPU render max abs diff {own:.3e}")
    _record("c5_slices_320x256_vs_reference_cuda", frames=frames, spp_per_frame=spp, ranks=world, rmse=rmse, psnr_db=float(psnr), max_abs_diff_vs_own_progressive=own)
    assert rmse < 1e-3 and psnr > 45
    assert own < 5e-5 * max(1.0, float(whole.max()))


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


needs_nofma = pytest.mark.skipif(not os.path.exists(ob.ref_cuda_path(W, H, nofma=True)), reason="-fmad=false reference CUDA build not present")


@needs_nofma
@pytest.mark.parametrize("sid,p0,p1", SCENES)
def test_strict_ids_vs_reference_cuda_nofma(sid, p0, p1):
    """the reference's CUDA sources built with -fmad=false against the strict:: kernels (also
    -fmad=false): primary-hit object ids, sub ids and distances must agree on EVERY pixel, bit for bit -- the GPU
    form of the statement the host oracle already proves for the g++ build of the same source
    (src/raytracer.cu:24-46, src/objects.cu:135-163, 40-79)."""
    ref = ob.RefCuda(W, H, nofma=True)
    ref.init(sid, p0, p1, 4, 5, True)
    r_obj, r_t = ref.primary_ids()
    r_sub = ref.primary_sub_ids(r_obj)
    r_img = np.zeros((H, W, 3), np.float32)
    ref.render_kernel(r_img, 0, 4321)
    scene = rtb.Scene(sid, p0, p1)
    tr = rtb.Tracer(0)
    tr.set_option("strict_fp", 1)
    tr.setup(scene, W, H, 4, 5)
    obj, sub, t = tr.primary_ids()
    img = np.zeros((H, W, 3), np.float32)
    tr.render(img, 0, 4321)
    tr.close()
    mism_obj, mism_sub, mism_t = int((obj != r_obj).sum()), int((sub != r_sub).sum()), int((t != r_t).sum())
    differ = int((np.abs(img - r_img).max(axis=2) > 0).sum())
    rmse = float(np.sqrt(((img - r_img) ** 2).mean()))
    print(f"scene {sid} strict vs -fmad=false reference: obj {mism_obj} sub {mism_sub} t {mism_t} mismatches; radiance: {differ} pixels differ, RMSE {rmse:.3e}")
    _record(f"strict_vs_reference_cuda_nofma_320x256_scene{sid}", obj_mismatches=mism_obj, sub_mismatches=mism_sub, t_mismatches=mism_t,
            pixels=W * H, radiance_pixels_differing=differ, rmse_4spp=rmse)
    assert mism_obj == 0 and mism_sub == 0 and mism_t == 0
    # same libm, same rounding of + - * / sqrt on both sides: the radiance agrees too, up to the few places where
    # nvcc evaluates a libm call or an FP64 expression differently inside the two programs
    assert rmse < 1e-3 or differ <= 3


# ---- parity at the sizes the bench numbers are quoted on (BASELINE.json configs 3 and 4) -----------------
C3W, C3H = 1920, 1080
C4W, C4H = 3840, 2160


@pytest.mark.skipif(not os.path.exists(ob.ref_cuda_path(C3W, C3H)), reason="1080p reference CUDA build not present")
def test_c3_config_size_vs_reference_cuda():
    """C3 (scene 0 at 1920x1080, 8 bounces): primary-hit ids and an 8-spp frame against the reference's own CUDA
    build compiled for 1920x1080 (its resolution is a compile-time constant, src/camera.cu:4-5), every render
    kernel, default row groups / streams / policies -- the launch shapes the bench runs."""
    spp, bounces = 8, 8
    ref = ob.RefCuda(C3W, C3H)
    ref.init(0, 0, 0, spp, bounces, True)
    r_obj, r_t = ref.primary_ids()
    r_img = np.zeros((C3H, C3W, 3), np.float32)
    ref.render_kernel(r_img, 0, 1000)
    scene = rtb.Scene(0, 0, 0)
    tr = rtb.Tracer(0).setup(scene, C3W, C3H, spp, bounces)
    obj, sub, t = tr.primary_ids()
    mism = int((obj != r_obj).sum())
    imgs = {}
    for kernel in (4, 3, 1, 0):
        tr.set_option("kernel", kernel)
        img = np.zeros((C3H, C3W, 3), np.float32)
        tr.render(img, 0, 1000)
        imgs[kernel] = img
    tr.set_option("kernel", 2)
    auto = np.zeros((C3H, C3W, 3), np.float32)
    tr.render(auto, 0, 1000)
    rec = dict(id_mismatches=mism, pixels=C3W * C3H, t_bit_equal=float((t == r_t).mean()))
    if os.path.exists(ob.ref_cuda_path(C3W, C3H, nofma=True)):
        refn = ob.RefCuda(C3W, C3H, nofma=True)
        refn.init(0, 0, 0, spp, bounces, True)
        n_obj, n_t = refn.primary_ids()
        tr.set_option("strict_fp", 1)
        s_obj, s_sub, s_t = tr.primary_ids()
        tr.set_option("strict_fp", 0)
        rec.update(strict_obj_mismatches=int((s_obj != n_obj).sum()), strict_t_mismatches=int((s_t != n_t).sum()))
        assert rec["strict_obj_mismatches"] == 0 and rec["strict_t_mismatches"] == 0
    tr.close()
    assert (imgs[0] == imgs[1]).all() and (imgs[0] == imgs[3]).all() and (imgs[0] == imgs[4]).all() and (imgs[0] == auto).all()
    rmse = float(np.sqrt(((imgs[3] - r_img) ** 2).mean()))
    bad = int((np.abs(imgs[3] - r_img).max(axis=2) > 1e-5).sum())
    rec.update(rmse_8spp=rmse, pixels_differing=bad, bit_identical_floats=float((imgs[3] == r_img).mean()))
    print(f"C3 1920x1080: {rec}")
    _record("c3_1920x1080_vs_reference_cuda", **rec)
    assert mism == 0, f"{mism} id mismatches"   # measured: 0 of 2 073 600, distances bit-equal
    assert rmse < 1e-3


@pytest.mark.skipif(not os.path.exists(ob.ref_cuda_path(C4W, C4H)), reason="4K reference CUDA build not present")
def test_c4_config_size_vs_reference_cuda():
    """C4 (scene 6 at 40 x 35 instances = 1 012 200 triangles, 3840x2160): primary-hit ids and a 1-spp frame
    against the reference's own CUDA build at 4K (its fixed-depth BVH holds ~990 triangles per leaf there, so a
    sample takes it minutes at 8 bounces: 3 here), the pooled and wavefront kernels full frame and the 1/8 tile share
    of rank 3."""
    spp, bounces = 1, 3
    ref = ob.RefCuda(C4W, C4H)
    ref.init(6, 40, 35, spp, bounces, True)
    r_obj, r_t = ref.primary_ids()
    r_img = np.zeros((C4H, C4W, 3), np.float32)
    ref.render_kernel(r_img, 0, 1000)
    scene = rtb.Scene(6, 40, 35)
    tr = rtb.Tracer(0).setup(scene, C4W, C4H, spp, bounces)
    obj, sub, t = tr.primary_ids()
    mism = int((obj != r_obj).sum())
    imgs = {}
    for kernel in (4, 3, 1):
        tr.set_option("kernel", kernel)
        img = np.zeros((C4H, C4W, 3), np.float32)
        tr.render(img, 0, 1000)
        imgs[kernel] = img
    tr.set_option("kernel", 2)
    tr.set_shard(3, 8, 8)
    part = np.zeros((C4H, C4W, 3), np.float32)
    tr.render(part, 0, 1000)
    rows = rtb.shard_rows_host(C4H, 3, 8, 8)
    tr.close()
    assert (imgs[3] == imgs[1]).all() and (imgs[4] == imgs[1]).all()
    assert (part[rows] == imgs[3][rows]).all()
    rmse = float(np.sqrt(((imgs[3] - r_img) ** 2).mean()))
    bad = int((np.abs(imgs[3] - r_img).max(axis=2) > 1e-5).sum())
    rec = dict(id_mismatches=mism, pixels=C4W * C4H, t_bit_equal=float((t == r_t).mean()), rmse_1spp=rmse, pixels_differing=bad,
               bit_identical_floats=float((imgs[3] == r_img).mean()), triangles=1012200)
    print(f"C4 3840x2160: {rec}")
    _record("c4_3840x2160_vs_reference_cuda", **rec)
    assert mism == 0, f"{mism} id mismatches"   # measured: 0 of 8 294 400, distances bit-equal
    # 1 spp: a path that takes a different discrete turn moves its pixel by O(1); bound the number of such pixels
    assert rmse < 1e-3 or bad <= C4W * C4H // 20000, f"RMSE {rmse}, {bad} pixels differ"

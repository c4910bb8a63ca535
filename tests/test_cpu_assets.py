"""SURVEY.md 8(f) rows 1 and 3 on the CPU: the texture pre-parser against the reference's own script,
the binary fast path, and the OBJ + MTL loader (per-face materials, vt, map_Kd)."""
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

import oracle_binding as ob
import raytracer_b200 as rtb
from raytracer_b200 import textures

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TEX = os.path.join(rtb.ASSETS_DIR, "textures")
GOLDEN = os.path.join(ROOT, "tests", "golden", "parsed_textures_reference.txt")
REF_SCRIPT = "/root/reference/textures/parse_textures.py"


def _copy_images(dst):
    os.makedirs(dst, exist_ok=True)
    for n in textures.list_images(TEX):
        shutil.copy(os.path.join(TEX, n), dst)


def test_preparser_text_equals_reference_script_output(tmp_path):
    """byte for byte the file textures/parse_textures.py wrote for the same two images (golden)"""
    d = str(tmp_path / "tex")
    _copy_images(d)
    # listdir order decides the order of the blocks: compare block-wise so the test does not depend on the file system
    textures.parse_directory(d)
    mine = open(os.path.join(d, textures.OUTPUT_TEXT)).read().split("\n")
    gold = open(GOLDEN).read().split("\n")
    assert mine[0] == gold[0] == "2"
    blocks = lambda L: {L[i]: L[i:i + 4] for i in range(1, len(L) - 1, 4)}
    assert blocks(mine) == blocks(gold)
    assert open(os.path.join(TEX, textures.OUTPUT_TEXT)).read().split("\n")[1:5] in (gold[1:5], gold[5:9])


@pytest.mark.skipif(not os.path.exists(REF_SCRIPT), reason="reference script only exists in the build container")
def test_preparser_against_reference_script_run_here(tmp_path):
    a, b = str(tmp_path / "a"), str(tmp_path / "b")
    _copy_images(a)
    _copy_images(b)
    subprocess.run([sys.executable, REF_SCRIPT], cwd=a, check=True)
    textures.parse_directory(b, binary=False)
    assert open(os.path.join(a, "parsed_textures.txt"), "rb").read() == open(os.path.join(b, "parsed_textures.txt"), "rb").read()


def test_binary_fast_path_holds_the_same_values(tmp_path):
    t = textures.read_text(os.path.join(TEX, textures.OUTPUT_TEXT))
    b = textures.read_binary(os.path.join(TEX, textures.OUTPUT_BINARY))
    assert set(t) == set(b) == {"earthlike.png", "checker.png"}
    for k in t:
        assert t[k].shape == b[k].shape and (t[k] == b[k]).all()
    assert t["earthlike.png"].shape == (32, 64, 3) and t["checker.png"].shape == (8, 8, 3)
    # all values are k/256
    v = t["earthlike.png"] * 256
    assert (v == np.rint(v)).all() and v.max() <= 255


def test_ppm_input(tmp_path):
    img = (np.arange(5 * 3 * 3) * 5 % 256).astype(np.uint8).reshape(3, 5, 3)
    p = tmp_path / "x.ppm"
    p.write_bytes(b"P6\n# comment\n5 3\n255\n" + img.tobytes())
    assert (textures.read_image(str(p)) == img).all()


def _desc_arrays(scene):
    d = scene.desc
    tris = np.ctypeslib.as_array(d.mesh_tri_points, (d.num_mesh_tris * 9,)).copy()
    uv = np.ctypeslib.as_array(d.mesh_tri_uv, (d.num_mesh_tris * 6,)).copy()
    imgs = [np.ctypeslib.as_array(d.images[i].rgb, (d.images[i].width * d.images[i].height * 3,)).copy() for i in range(d.num_images)]
    return tris, uv, imgs


def test_loader_text_and_binary_paths_give_the_same_scene(tmp_path):
    full = rtb.Scene(7)
    alt = tmp_path / "assets"
    shutil.copytree(rtb.ASSETS_DIR, alt)
    os.remove(alt / "textures" / textures.OUTPUT_BINARY)       # forces the text path of load_parsed_texture
    text = rtb.Scene(7, assets=str(alt))
    a, b = _desc_arrays(full), _desc_arrays(text)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all() and len(a[2]) == len(b[2]) == 1 and (a[2][0] == b[2][0]).all()
    ref = textures.read_binary(os.path.join(TEX, textures.OUTPUT_BINARY))["earthlike.png"]
    assert (a[2][0] == ref.reshape(-1)).all()
    os.remove(alt / "textures" / textures.OUTPUT_TEXT)
    with pytest.raises(rtb.RtbError, match="Image file not found"):
        rtb.Scene(7, assets=str(alt))


def test_mtl_mapping_and_groups():
    s = rtb.Scene.from_obj(os.path.join(rtb.ASSETS_DIR, "models", "studio.obj"))
    d = s.desc
    assert d.num_objects == 4 and d.num_mesh_tris == 28 and d.num_images == 1
    objs = [d.objects[i] for i in range(4)]
    mats = [d.materials[o.material] for o in objs]
    assert [o.type for o in objs] == [rtb.OBJ_MESH] * 4
    # crate: 6 quads + the lid declared later in the file under the same usemtl -> 14 triangles, IMAGE texture, rough
    assert objs[0].tri_count == 14 and mats[0].type == rtb.MAT_STANDARD and mats[0].texture.type == rtb.TEX_IMAGE
    assert (mats[0].texture.img_width, mats[0].texture.img_height) == (64, 32) and mats[0].smoothness == 0 and mats[0].need_uv == 1
    # glass: d < 1 -> REFRACTIVE with Ni
    assert objs[1].tri_count == 8 and mats[1].type == rtb.MAT_REFRACTIVE and mats[1].refractive_index == pytest.approx(1.5)
    # panel: Ke -> EMISSIVE, colour * strength == Ke
    assert objs[2].tri_count == 2 and mats[2].type == rtb.MAT_EMISSIVE
    e = mats[2].emitted_light
    assert (e.x, e.y, e.z) == pytest.approx((5.0, 4.5, 3.5), rel=1e-6)
    # lacquer: Ns 810 -> smoothness sqrt(0.81)
    assert objs[3].tri_count == 4 and mats[3].smoothness == pytest.approx(0.9, rel=1e-6)
    c = mats[3].texture.colour
    assert (c.x, c.y, c.z) == pytest.approx((0.8, 0.15, 0.1))
    tris, uv, _ = _desc_arrays(s)
    # quads split (1,2,3),(1,4,3) with their vt: first crate face
    assert uv[:12].tolist() == [0, 0, 1, 0, 1, 1, 0, 0, 0, 1, 1, 1]
    # the pyramid's `-1` is the last vertex of the file (its apex)
    pyr = tris[(14 + 8 + 2) * 9:].reshape(4, 3, 3)
    assert np.allclose(pyr[:, 2], [0.8, 0.1, 0.8])
    # faces without vt get (0, 0)
    assert (uv[14 * 6:] == 0).all()


def test_reference_cube_obj_with_its_mtl():
    """models/cube.obj + cube.mtl (the only MTL content the reference ships, models/cube.mtl:5-12):
    Kd 0.8, Ns 250 -> smoothness 0.5; its vt reach Triangle::texture_points"""
    s = rtb.Scene.from_obj(os.path.join(rtb.models_dir(), "cube.obj"))
    d = s.desc
    assert d.num_objects == 1 and d.objects[0].tri_count == 12
    m = d.materials[0]
    assert m.type == rtb.MAT_STANDARD and m.smoothness == pytest.approx(0.5) and m.texture.colour.x == pytest.approx(0.8)
    _, uv, _ = _desc_arrays(s)
    assert uv[:6].tolist() == pytest.approx([0.625, 0.5, 0.875, 0.5, 0.875, 0.75])


def test_loader_errors(tmp_path):
    p = tmp_path / "bad.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv 0 0 1\nf 1 2 3 4 5\n")
    with pytest.raises(rtb.RtbError, match="Only triangle or quad meshes are supported"):
        rtb.Scene.from_obj(str(p))
    p.write_text("v 0 0 0\nv 1 0 0\nf 1 2 9\n")
    with pytest.raises(rtb.RtbError, match="does not exist"):
        rtb.Scene.from_obj(str(p))
    with pytest.raises(rtb.RtbError, match="Could not find file"):
        rtb.Scene.from_obj(str(tmp_path / "nope.obj"))


@pytest.mark.skipif(not ob.have_ref_host(), reason="oracle/_ref not built")
def test_studio_scene_in_the_reference_build_matches():
    """scene 7 compiled against the reference's own sources == compiled against the mirror: ids and radiance
    (the broader per-scene tests in test_cpu_oracle.py also cover scene 7)"""
    W, H = 64, 48
    sc = rtb.Scene(7)
    ref = ob.RefHost()
    ref.build_scene(7, 0, 0)
    rd = sc.render_settings(2, 6)
    ref.set_render(rd)
    ref.set_camera(W, H)
    orc = ob.Oracle(sc)
    cam = rtb.default_camera(W, H)
    for a, b in zip(ref.primary_ids(W, H), orc.primary_ids(cam, W, H)):
        assert (a == b).all()
    assert (ref.render(W, H, 4242, 0, None) == orc.render(cam, W, H, rd, 4242, 0, None)).all()

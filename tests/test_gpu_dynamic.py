"""SURVEY.md 8(f) row 4 on the GPU: dynamic scenes (objects and meshes that move between frames, tree
rebuilt on the device at every commit) and a moving camera, each frame checked against the oracle."""
import copy
import ctypes

import numpy as np
import pytest

import oracle_binding as ob
import raytracer_b200 as rtb

pytestmark = pytest.mark.gpu
W, H = 256, 192


def _object_copy(o):
    c = rtb.Object()
    ctypes.memmove(ctypes.byref(c), ctypes.byref(o), ctypes.sizeof(rtb.Object))
    return c


def _ids_equal(tr, desc, cam):
    """strict-fp primary ids / distances of the committed scene == the oracle's on `desc`"""
    obj, sub, t = tr.primary_ids()
    o_obj, o_sub, o_t = ob.Oracle(desc).primary_ids(cam, W, H)
    return (obj == o_obj).all() and (sub == o_sub).all() and (t == o_t).all()


def test_moving_spheres_rebuild_each_frame():
    """scene 4 (60 seeded spheres over a floor): every frame a third of the spheres move; the context's
    description is edited in place (rtb_scene_update_object) and committed again.  Each frame: ids
    bit-exact against the oracle on the same description, radiance bit-identical to a context that was
    handed the whole modified scene from scratch (rtb_scene_set)."""
    scene = rtb.Scene(4, 60, 7)
    tr = rtb.Tracer(0).setup(scene, W, H, 4, 5)
    tr.set_option("strict_fp", 1)
    fresh = rtb.Tracer(0)
    fresh.set_option("strict_fp", 1)
    cam = rtb.default_camera(W, H)
    rng = np.random.default_rng(11)
    try:
        for frame in range(4):
            desc = tr.scene_desc()
            for i in range(frame % 3, desc.num_objects - 1, 3):
                o = _object_copy(desc.objects[i])
                assert o.type == rtb.OBJ_SPHERE
                o.p[0].x += np.float32(rng.uniform(-0.3, 0.3))
                o.p[0].z += np.float32(rng.uniform(-0.3, 0.3))
                o.f[0] = np.float32(o.f[0] * rng.uniform(0.8, 1.2))
                o.p[0].y = np.float32(-1.0) + np.float32(o.f[0])
                tr.update_object(i, o)
            tr.commit()
            assert tr.info("commit_ms") > 0 and tr.info("build_ms") > 0
            desc = tr.scene_desc()
            assert _ids_equal(tr, desc, cam)
            img = np.zeros((H, W, 3), np.float32)
            tr.render(img, 0, 500 + frame)
            fresh.set_camera(cam, W, H)
            fresh.set_render_data(scene.render_settings(4, 5))
            fresh.set_scene(desc)
            ref = np.zeros((H, W, 3), np.float32)
            fresh.render(ref, 0, 500 + frame)
            assert (img == ref).all()
    finally:
        tr.close()
        fresh.close()


def test_rotating_mesh_and_material_change():
    """scene 0: the monkey mesh turns about the vertical axis through its centre (rtb_scene_update_mesh), the
    mirror sphere moves and changes material; ids bit-exact against the oracle every frame."""
    scene = rtb.Scene(0)
    tr = rtb.Tracer(0).setup(scene, W, H, 2, 4)
    tr.set_option("strict_fp", 1)
    cam = rtb.default_camera(W, H)
    try:
        desc = tr.scene_desc()
        mesh_index = [i for i in range(desc.num_objects) if desc.objects[i].type == rtb.OBJ_MESH][0]
        sphere_index = [i for i in range(desc.num_objects) if desc.objects[i].type == rtb.OBJ_SPHERE][0]
        m = desc.objects[mesh_index]
        base = np.ctypeslib.as_array(desc.mesh_tri_points, (desc.num_mesh_tris * 9,)).copy().reshape(-1, 3)[m.tri_first * 3:(m.tri_first + m.tri_count) * 3]
        centre = base.mean(axis=0, dtype=np.float64)
        previous = None
        for frame in range(3):
            a = 0.4 * (frame + 1)
            rot = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
            pts = ((base - centre) @ rot.T + centre).astype(np.float32)
            tr.update_mesh(mesh_index, pts)
            s = _object_copy(tr.scene_desc().objects[sphere_index])
            s.p[0].x = np.float32(-0.25 + 0.05 * frame)
            s.material = (s.material + frame) % desc.num_materials
            tr.update_object(sphere_index, s)
            tr.commit()
            d = tr.scene_desc()
            got = np.ctypeslib.as_array(d.mesh_tri_points, (d.num_mesh_tris * 9,)).reshape(-1, 3)[m.tri_first * 3:(m.tri_first + m.tri_count) * 3]
            assert (got == pts).all() and d.objects[sphere_index].material == s.material
            assert _ids_equal(tr, d, cam)
            obj, _, _ = tr.primary_ids()
            assert previous is None or (obj != previous).any()   # the picture really changes
            previous = obj
            img = np.zeros((H, W, 3), np.float32)
            tr.render(img, 0, 42)
            ref = ob.Oracle(d).render(cam, W, H, scene.render_settings(2, 4), 42, 0, None)
            assert float((np.abs(img - ref).max(axis=2) > 1e-5).mean()) <= 5e-4
    finally:
        tr.close()


def test_moving_camera():
    """the pose members the reference marks changeable (src/camera.cu:33-41): position, three rotations,
    field of view; ids bit-exact against the oracle for each pose"""
    scene = rtb.Scene(0)
    tr = rtb.Tracer(0).setup(scene, W, H, 2, 4)
    tr.set_option("strict_fp", 1)
    try:
        poses = [((0.0, 0.0, 0.0), (0.0, 0.0, 0.0), 0.0), ((0.15, 0.05, 0.2), (0.0, 0.25, 0.0), 0.0),
                 ((-0.2, -0.1, 0.4), (0.1, -0.3, 0.05), 1.2), ((0.0, 0.3, 1.3), (-0.4, 0.0, 0.0), 1.6)]
        seen = []
        for pos, rot, fov in poses:
            cam = rtb.camera(W, H, pos, rot, fov)
            tr.set_camera(cam, W, H)
            obj, sub, t = tr.primary_ids()
            o_obj, o_sub, o_t = ob.Oracle(scene).primary_ids(cam, W, H)
            assert (obj == o_obj).all() and (sub == o_sub).all() and (t == o_t).all()
            seen.append(obj)
        assert (seen[0] != seen[1]).any() and (seen[1] != seen[2]).any()
    finally:
        tr.close()


def test_update_errors():
    scene = rtb.Scene(1)
    tr = rtb.Tracer(0).setup(scene, 64, 48, 1, 2)
    try:
        o = _object_copy(tr.scene_desc().objects[0])
        with pytest.raises(rtb.RtbError, match="out of range"):
            tr.update_object(99, o)
        o.material = 1234
        with pytest.raises(rtb.RtbError, match="unknown material"):
            tr.update_object(0, o)
        with pytest.raises(rtb.RtbError, match="not a mesh"):
            tr.update_mesh(0, np.zeros(9, np.float32))
    finally:
        tr.close()

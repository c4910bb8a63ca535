"""The device version of the mesh flatten (csrc/rtb_meshprep.cu): the
reference's BVH partition of a mesh (src/objects.cu:602-770) and the per-triangle records are made on the GPU for big
meshes.  The checker is the host restatement (csrc/rtb_refbvh.cpp, itself compared with the literal recursion by
tests/refbvh_check.cpp): leaf assignment, degenerate flags and boxes array by array, then whole scenes through both
paths (ids, distances, images, counts) -- bit-exact everywhere."""
import ctypes

import numpy as np
import pytest

import raytracer_b200 as rtb

pytestmark = pytest.mark.gpu


def _partition(points, on_device):
    pts = np.ascontiguousarray(points, np.float32).reshape(-1, 9)
    n = pts.shape[0]
    leaf_of = np.empty(n, np.int32)
    degenerate = np.empty(n, np.uint8)
    nodes = np.zeros((2048, 8), np.float32)
    num = ctypes.c_int(0)
    rc = rtb.lib.rtb_debug_mesh_partition(0, pts.ctypes.data, n, int(on_device), leaf_of.ctypes.data, degenerate.ctypes.data, nodes.ctypes.data, 2048,
                                          ctypes.byref(num))
    assert rc == 0
    return leaf_of, degenerate, nodes[:num.value]


def _meshes():
    rng = np.random.default_rng(5)
    yield "one triangle", rng.normal(size=(1, 9))
    yield "two", rng.normal(size=(2, 9))
    yield "three", rng.normal(size=(3, 9))
    yield "seven", rng.normal(size=(7, 9))
    yield "1025 random", rng.normal(size=(1025, 9))
    yield "5000 random, wide range", rng.normal(size=(5000, 9)) * np.exp(rng.normal(size=(5000, 1)) * 3)
    # ties: every triangle four times (equal distances: the merge sort's take-from-the-right rule decides), shuffled
    base = rng.normal(size=(1500, 9))
    dup = np.repeat(base, 4, axis=0)
    yield "duplicates", dup[rng.permutation(len(dup))]
    # a grid of axis-aligned squares at a few heights: equal distances by symmetry, flat leaf boxes (degenerate leaves)
    g = []
    for z in (0.0, 0.5, -0.0):
        for ix in range(40):
            for iy in range(40):
                x, y = ix * 0.25 - 5, iy * 0.25 - 5
                g.append([x, y, z, x + 0.25, y, z, x + 0.25, y + 0.25, z])
                g.append([x, y, z, x + 0.25, y + 0.25, z, x, y + 0.25, z])
    yield "flat grid", np.array(g)
    yield "all the same point", np.zeros((3000, 9))
    yield "40000 random", rng.uniform(-50, 50, size=(40000, 9))


@pytest.mark.parametrize("name,points", list(_meshes()), ids=[m[0] for m in _meshes()])
def test_device_partition_equals_host_partition(name, points):
    h_leaf, h_deg, h_nodes = _partition(points, False)
    d_leaf, d_deg, d_nodes = _partition(points, True)
    assert h_nodes.shape == d_nodes.shape
    assert (h_leaf == d_leaf).all(), f"{name}: {(h_leaf != d_leaf).sum()} triangles in another leaf"
    assert (h_deg == d_deg).all()
    # boxes and parents equal as numbers (the sign of a zero plane is free: min / max of -0 and +0 are order dependent)
    assert (h_nodes == d_nodes).all()
    if name == "flat grid":
        assert h_deg.any()


def _render_both(scene, w, h, spp=2, bounces=4):
    out = []
    for mode in (0, 1):
        tr = rtb.Tracer(0)
        tr.set_option("device_flatten", mode)
        tr.set_option("strict_fp", 1)
        tr.setup(scene, w, h, spp, bounces)
        ids = tr.primary_ids()
        img = np.zeros((h, w, 3), np.float32)
        tr.render(img, 0, 4242)
        out.append((ids, img, int(tr.info("num_prims")), int(tr.info("culled_tris")), int(tr.info("num_large")), int(tr.info("num_nodes"))))
        tr.close()
    return out


@pytest.mark.parametrize("scene_args", [(0, 0, 0), (5, 0, 0), (6, 4, 3), (7, 0, 0)], ids=["monkey room", "cube.obj (degenerate leaves)", "monkey grid", "obj + mtl + uv"])
def test_scene_through_the_device_flatten_is_bit_identical(scene_args):
    """every mesh of the scene partitioned and flattened on the device (option device_flatten = 1) against the host
    path: ids, distances, radiance, primitive / culled / large counts"""
    host, dev = _render_both(rtb.Scene(*scene_args), 200, 152)
    for a, b in zip(host[0], dev[0]):
        assert (a.view(np.uint32) == b.view(np.uint32)).all()
    assert (host[1].view(np.uint32) == dev[1].view(np.uint32)).all()
    assert host[2:] == dev[2:]
    if scene_args[0] == 5:
        assert host[3] > 0   # cube.obj's axis-aligned faces sit in flat leaves: culled on both paths


def test_mesh_with_a_large_triangle_on_the_device_path():
    """a mesh that holds two huge floor triangles among small ones: the large-primitive list takes them out of the
    device-built primitive list (swap with the last entry), same picture as the host path"""
    rng = np.random.default_rng(9)
    small = rng.uniform(-1, 1, size=(3000, 9)).astype(np.float32) * 0.05 + np.repeat(rng.uniform(-2, 2, size=(3000, 3)), 3, axis=1).reshape(3000, 9).astype(np.float32)
    floor = np.array([[-8, -1.5, -2, 8, -1.5, -2, 8, -1.3, 14], [-8, -1.5, -2, 8, -1.3, 14, -8, -1.3, 14]], np.float32)
    pts = np.concatenate([small[:1000], floor[:1], small[1000:], floor[1:]]).astype(np.float32)
    pts[:, 2::3] += 6.0
    base = rtb.Scene(1)   # spheres, materials and a light; the mesh joins the committed description
    results = []
    for mode in (0, 1):
        tr = rtb.Tracer(0)
        tr.set_option("device_flatten", mode)
        tr.setup(base, 200, 152, 2, 4)
        assert rtb.lib.rtb_scene_add_mesh(tr.ctx, pts.ctypes.data, None, len(pts), 0) >= 0
        tr.commit()
        ids = tr.primary_ids()
        img = np.zeros((152, 200, 3), np.float32)
        tr.render(img, 0, 77)
        results.append((ids, img, int(tr.info("num_large")), int(tr.info("num_prims")), int(tr.info("num_nodes"))))
        tr.close()
    (h_ids, h_img, h_large, h_prims, h_nodes), (d_ids, d_img, d_large, d_prims, d_nodes) = results
    assert h_large == d_large and h_large >= 2 and h_prims == d_prims and h_nodes == d_nodes
    for a, b in zip(h_ids, d_ids):
        assert (a.view(np.uint32) == b.view(np.uint32)).all()
    assert (h_img.view(np.uint32) == d_img.view(np.uint32)).all()


def test_big_mesh_commit_takes_the_device_path_by_default():
    """72 300 triangles (10 x 10 monkeys): above the threshold of the default, so the commit runs on the device; same
    ids as the host path, and the commit reports its parts"""
    scene = rtb.Scene(6, 10, 10)
    dev = rtb.Tracer(0).setup(scene, 320, 256, 1, 2)
    assert int(dev.info("device_flatten")) == 2
    host = rtb.Tracer(0)
    host.set_option("device_flatten", 0)
    host.setup(scene, 320, 256, 1, 2)
    for a, b in zip(host.primary_ids(), dev.primary_ids()):
        assert (a.view(np.uint32) == b.view(np.uint32)).all()
    print(f"commit of 72 300 triangles: device path {dev.info('commit_ms'):.1f} ms, host path {host.info('commit_ms'):.1f} ms")
    assert int(dev.info("num_prims")) == int(host.info("num_prims"))
    host.close()
    dev.close()

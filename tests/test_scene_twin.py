"""CPU tests of the Python twin's scene-side helpers added for BASELINE.json config C5 (classroom): OBJ loader,
AliasTable2D construction, environment-map weights, matrix helpers; and of the oracle's vatan2."""
import ctypes as C
import math

import numpy as np

from vision_b200 import scene_py

OBJ = """# quad as one polygon + a triangle sharing an edge, with and without vt; a degenerate face; negative indices
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 2 0 0
vt 0 0
vt 1 0
vt 1 1
vt 0 0.25
f 1/1 2/2 3/3 4/4
f 2/2 5 3/3
f 1/1 1/1 2/2
f -5/1 -4/2 -3/3
"""


def test_obj_loader_face_order_dedupe_and_generated_normals(tmp_path):
    p = tmp_path / "m.obj"
    p.write_text(OBJ)
    m = scene_py.obj_mesh(p)
    # fan triangulation in file order; the degenerate face is dropped; the last face re-uses vertices 0,1,2
    assert m["tri"].tolist() == [[0, 1, 2], [0, 2, 3], [1, 4, 2], [0, 1, 2]]
    assert m["pos"].shape == (5, 3) and m["pos"][4].tolist() == [2.0, 0.0, 0.0]
    np.testing.assert_array_equal(m["nrm"], np.tile([0, 0, 1], (5, 1)).astype(np.float32))       # face normals
    assert m["uv"][3].tolist() == [0.0, 0.75] and m["uv"][4].tolist() == [0.0, 1.0]              # v -> 1 - v; missing vt = (0, 0)
    m2 = scene_py.obj_mesh(p, flip_uv=False)
    assert m2["uv"][3].tolist() == [0.0, 0.25]
    m3 = scene_py.obj_mesh(p, smooth=True)
    np.testing.assert_allclose(m3["nrm"], np.tile([0, 0, 1], (5, 1)), atol=1e-7)


def test_obj_loader_keeps_file_normals(tmp_path):
    p = tmp_path / "n.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 -1\nvn 0 1 0\nf 1//1 2//1 3//2\nf 1//2 2//1 3//2\n")
    m = scene_py.obj_mesh(p)
    assert m["tri"].tolist() == [[0, 1, 2], [3, 1, 2]]              # (v, vn) pairs: vertex 1 with normal 2 is a new vertex
    assert m["nrm"].tolist() == [[0, 0, -1], [0, 0, -1], [0, 1, 0], [0, 1, 0]]


def test_alias2d_tables_reproduce_the_weights():
    rng = np.random.default_rng(3)
    rx, ry = 8, 5
    w = rng.random(rx * ry).astype(np.float32)
    w[rx:2 * rx] = 0.0                                              # one all-zero row
    mt, mf, mi, ct, cf = scene_py.alias2d_build([float(x) for x in w], rx, ry)
    assert len(mt) == ry and len(ct) == rx * ry and cf == [float(x) for x in w]
    np.testing.assert_allclose(mf, w.reshape(ry, rx).astype(np.float64).mean(axis=1), rtol=1e-6)
    np.testing.assert_allclose(mi, np.mean(mf), rtol=1e-6)
    # probability of landing in cell (v, u) through the two alias tables == w / sum(w)
    def pmf(table):
        n = len(table)
        out = np.zeros(n)
        for i, (prob, alias) in enumerate(table):
            out[i] += prob / n
            out[alias] += (1.0 - prob) / n
        return out
    pv = pmf(mt)
    assert pv[1] == 0.0
    tot = np.zeros((ry, rx))
    for v in range(ry):
        tot[v] = pv[v] * pmf(ct[v * rx:(v + 1) * rx])
    np.testing.assert_allclose(tot[[0, 2, 3, 4]], (w.reshape(ry, rx) / w.sum())[[0, 2, 3, 4]], atol=1e-6)


def test_env_weights_follow_the_reference_index_quirk():
    tex = np.ones((4, 8, 4), np.float32)                            # h = 4, w = 8
    w = scene_py.env_weights(tex)
    lum = np.float32(np.float32(0.212671) + np.float32(0.715160)) + np.float32(0.072169)
    # spherical.cpp:175-177: v = idx / res.y + 0.5; theta = v / res.x  (rows of 4 texels share a theta, not rows of 8)
    for idx in (0, 3, 4, 31):
        theta = (idx // 4 + 0.5) / 8
        np.testing.assert_allclose(w[idx], lum * math.sin(math.pi * theta), rtol=1e-6)
    assert w[0] == w[3] and w[3] != w[4]


def test_matrix_helpers():
    m = scene_py.euler_matrix({"yaw": -30, "pitch": 10, "roll": 5, "position": [1, 2, 3]}).astype(np.float64)
    np.testing.assert_allclose(m[:3, :3] @ m[:3, :3].T, np.eye(3), atol=1e-6)
    assert m[:3, 3].tolist() == [1, 2, 3]
    np.testing.assert_allclose(scene_py.invert4(m) @ m, np.eye(4), atol=1e-12)
    o2w, w2o = scene_py.env_o2w({"flip_u": True, "o2w": {"type": "Euler", "param": {"yaw": -30}}})
    np.testing.assert_allclose(o2w.astype(np.float64) @ w2o.astype(np.float64), np.eye(4), atol=1e-6)
    # rotation_x(-90) maps the map's +z (theta = 0, top row) to world +y
    np.testing.assert_allclose(o2w[:3, :3] @ np.array([0, 0, 1.0]), [0, 1, 0], atol=1e-6)


def test_vatan2_matches_libm(oracle):
    f = oracle.lib.vo_vatan2
    f.restype = C.c_float
    f.argtypes = [C.c_float, C.c_float]
    rng = np.random.default_rng(5)
    ys, xs = rng.normal(size=4000).astype(np.float32), rng.normal(size=4000).astype(np.float32)
    got = np.array([f(float(y), float(x)) for y, x in zip(ys, xs)], np.float32)
    np.testing.assert_allclose(got, np.arctan2(ys.astype(np.float64), xs.astype(np.float64)), atol=4e-7)
    assert f(0.0, 1.0) == 0.0 and abs(f(0.0, -1.0) - math.pi) < 1e-6 and abs(f(1.0, 0.0) - math.pi / 2) < 1e-6
    assert f(0.0, 0.0) == 0.0

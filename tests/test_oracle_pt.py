"""CPU-only: known answers for the oracle's sampler / math restatements, and golden image stats."""
import ctypes as C
from pathlib import Path

import numpy as np

from vision_b200 import scene_py

ROOT = Path(__file__).resolve().parent.parent


def _bind(o):
    lib = o.lib
    lib.vo_tea4.argtypes = [C.c_uint32, C.c_uint32]; lib.vo_tea4.restype = C.c_uint32
    lib.vo_lcg.argtypes = [C.POINTER(C.c_uint32)]; lib.vo_lcg.restype = C.c_float
    lib.vo_sincos_2pi.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]; lib.vo_sincos_2pi.restype = None
    lib.vo_offset_ray_origin.argtypes = [C.c_void_p] * 3; lib.vo_offset_ray_origin.restype = None
    return lib


def _tea4_py(v0, v1):
    """independent python restatement of reference src/math/util.h:13-24 with its published constants"""
    M = 0xFFFFFFFF
    s0 = 0
    for _ in range(4):
        s0 = (s0 + 0x9E3779B9) & M
        v0 = (v0 + ((((v1 << 4) & M) + 0xA341316C) & M ^ ((v1 + s0) & M) ^ (((v1 >> 5) + 0xC8013EA4) & M))) & M
        v1 = (v1 + ((((v0 << 4) & M) + 0xAD90777D) & M ^ ((v0 + s0) & M) ^ (((v0 >> 5) + 0x7E95761E) & M))) & M
    return v0


def test_tea_and_lcg_known_answers(oracle):
    lib = _bind(oracle)
    for a, b in [(0, 0), (1, 2), (255, 255), (123456789, 0xFFFFFFFF), (0xDEADBEEF, 17)]:
        assert lib.vo_tea4(a, b) == _tea4_py(a, b)
    st = C.c_uint32(12345)
    s = 12345
    for _ in range(10):
        s = (1664525 * s + 1013904223) & 0xFFFFFFFF
        assert lib.vo_lcg(C.byref(st)) == np.float32(s & 0x00FFFFFF) * np.float32(1.0 / 16777216.0)
        assert st.value == s


def test_sincos_accuracy(oracle):
    lib = _bind(oracle)
    s, c = C.c_float(), C.c_float()
    worst = 0.0
    for u in np.linspace(0, 1, 4001, dtype=np.float32):
        lib.vo_sincos_2pi(float(u), C.byref(s), C.byref(c))
        worst = max(worst, abs(s.value - np.sin(2 * np.pi * float(u))), abs(c.value - np.cos(2 * np.pi * float(u))))
    assert worst < 3e-7


def test_offset_ray_origin(oracle):
    lib = _bind(oracle)
    out = np.zeros(3, np.float32)
    p = np.array([1.0, -2.0, 0.001], np.float32); n = np.array([0.0, 1.0, 0.0], np.float32)
    lib.vo_offset_ray_origin(p.ctypes.data, n.ctypes.data, out.ctypes.data)
    assert out[0] == p[0] and out[2] == p[2]
    assert out[1] > p[1] and out[1] - p[1] < 1e-4          # moved towards +n by 256 ulps
    assert np.float32(out[1]).view(np.uint32) == np.float32(-2.0).view(np.uint32) - 256
    p2 = np.array([0.001, 0.0, 0.0], np.float32); n2 = np.array([1.0, 0.0, 0.0], np.float32)
    lib.vo_offset_ray_origin(p2.ctypes.data, n2.ctypes.data, out.ctypes.data)
    assert out[0] == np.float32(p2[0] + np.float32(1.0 / 65536.0))


def test_alias_table_matches_reference_algorithm():
    table, func, integral = scene_py.alias_build([1.0, 3.0, 2.0, 2.0])
    assert abs(integral - 2.0) < 1e-7
    # every column's kept + donated mass reproduces the weights
    n = len(table)
    mass = [0.0] * n
    for i, (p, a) in enumerate(table):
        mass[i] += p / n
        mass[a] += (1 - p) / n
    np.testing.assert_allclose(mass, np.array(func) / sum(func), atol=1e-6)


def test_cbox_scene_records():
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json")
    assert sc.resolution == (256, 256) and sc.spp == 16 and sc.params.max_depth == 5
    assert len(sc.meshes) == 8 and sum(len(m["tri"]) for m in sc.meshes) == 36
    assert len(sc.lights) == 1 and sc.lights[0].inst_id == 7
    # light colour [17,12,4] is normalised by its max and folded into scale (light.cpp:19-24)
    assert abs(sc.lights[0].scale - 17.0) < 1e-6 and abs(sc.lights[0].color.value[0] - 1.0) < 1e-7
    assert sc.camera.filter_type == 1 and len(sc.textures) == 1


def test_oracle_cbox_golden_stats(oracle):
    """frozen statistics of the oracle's own C1 render (tests/golden/cbox_matte_oracle.json): guards the
    oracle against accidental edits; GPU parity is asserted against the live oracle in test_pt_gpu.py"""
    import json
    import oracle_lib
    gold = json.loads((ROOT / "tests" / "golden" / "cbox_matte_oracle.json").read_text())
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", dict(scene_py.CBOX_MATTE_OVERRIDES, resolution=[64, 64], spp=4))
    osc = oracle_lib.OracleScene(oracle, sc, threads=4)
    img, _, st = osc.render(0, 4, threads=4, stats=True)
    assert st["closest"] == gold["closest"] and st["any"] == gold["any"] and st["vertices"] == gold["vertices"]
    np.testing.assert_allclose(img[..., :3].mean(axis=(0, 1)), gold["mean_rgb"], rtol=1e-5)
    ph = osc.primary_hits(0)
    assert int(ph["prim"].astype(np.uint64).sum()) == gold["primary_prim_sum"]
    assert int(ph["inst"].astype(np.uint64).sum()) == gold["primary_inst_sum"]

"""CPU-only: the oracle against itself (BVH vs brute force), synth twins, C-ABI symbol export."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import synth

ROOT = Path(__file__).resolve().parent.parent


def test_synth_mesh_twin_bitexact(oracle):
    a = oracle.synth_mesh(20000, seed=1)
    b = synth.synth_mesh(20000, seed=1)
    assert a.tobytes() == b.tobytes()


def test_synth_rays_twin(oracle):
    a = oracle.synth_rays(50000)
    b = synth.synth_rays(50000)
    for f in ("ox", "oy", "oz", "dz", "tmin", "tmax", "tmax_any"):
        assert a[f].tobytes() == b[f].tobytes(), f
    for f in ("dx", "dy"):      # trig in double then rounded: allow 1 ulp between libm and numpy
        np.testing.assert_allclose(a[f], b[f], rtol=0, atol=1.2e-7)
    n = np.sqrt(b["dx"] ** 2 + b["dy"] ** 2 + b["dz"] ** 2)
    np.testing.assert_allclose(n, 1.0, atol=1e-6)


@pytest.mark.parametrize("ntri", [1, 2, 3, 7, 5000])
def test_oracle_bvh_equals_bruteforce(oracle, ntri):
    mesh = oracle.synth_mesh(ntri, seed=11)
    rays = oracle.synth_rays(4000, seed=12, seed_any=13)
    acc = oracle.accel()
    acc.add_soup(mesh)
    acc.build(threads=4)
    h_bvh = acc.trace_closest(rays, threads=8)
    h_bf = acc.trace_closest(rays, threads=8, brute=True)
    for f in ("inst", "prim", "u", "v", "t"):
        assert h_bvh[f].tobytes() == h_bf[f].tobytes(), f
    any_rays = dict(rays, tmax=rays["tmax_any"])
    o1, _ = acc.trace_any(any_rays)
    o2, _ = acc.trace_any(any_rays, brute=True)
    assert o1.tobytes() == o2.tobytes()
    if ntri >= 5000:
        assert (h_bvh["inst"] != 0xFFFFFFFF).sum() > 100


def test_oracle_known_answer_triangle(oracle):
    """hand-computed: ray down -z onto the unit right triangle at z=0"""
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    tri = np.array([[0, 1, 2]], np.uint32)
    acc = oracle.accel()
    acc.add_instance(pos, tri, None, inst_id=5)
    acc.build(1)
    rays = {"ox": np.array([0.25, 2.0], np.float32), "oy": np.array([0.5, 2.0], np.float32),
            "oz": np.array([3.0, 3.0], np.float32), "tmin": np.zeros(2, np.float32),
            "dx": np.zeros(2, np.float32), "dy": np.zeros(2, np.float32), "dz": np.array([-1, -1], np.float32),
            "tmax": np.full(2, 1e30, np.float32)}
    h = acc.trace_closest(rays, threads=1)
    assert h["inst"][0] == 5 and h["prim"][0] == 0
    assert h["t"][0] == 3.0 and h["u"][0] == 0.25 and h["v"][0] == 0.5     # u <-> vertex 1, v <-> vertex 2
    assert h["inst"][1] == 0xFFFFFFFF and h["prim"][1] == 0xFFFFFFFF


def test_oracle_tie_break_smallest_ids(oracle):
    """two coincident triangles: equal t -> smallest (inst, prim) wins regardless of insertion order"""
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    tri = np.array([[0, 1, 2], [0, 1, 2]], np.uint32)
    acc = oracle.accel()
    acc.add_instance(pos, tri, None, inst_id=3)
    acc.add_instance(pos, tri, None, inst_id=1)
    acc.build(1)
    rays = {"ox": np.array([0.25], np.float32), "oy": np.array([0.25], np.float32), "oz": np.array([1.0], np.float32),
            "tmin": np.zeros(1, np.float32), "dx": np.zeros(1, np.float32), "dy": np.zeros(1, np.float32),
            "dz": np.array([-1], np.float32), "tmax": np.full(1, 1e30, np.float32)}
    for brute in (False, True):
        h = acc.trace_closest(rays, threads=1, brute=brute)
        assert (h["inst"][0], h["prim"][0]) == (1, 0)


def test_cabi_exports_every_declared_symbol():
    """every VB_API function declared in include/*.h is exported by libvision_b200.so (no compute calls)"""
    from vision_b200 import LIB_PATH
    assert LIB_PATH.exists(), "libvision_b200.so not built: run __graft_entry__.build()"
    lib = C.CDLL(str(LIB_PATH))
    names = set()
    for h in (ROOT / "include").glob("*.h"):
        names |= set(re.findall(r"VB_API\s+[\w\s\*]+?\b(vb_\w+)\s*\(", h.read_text()))
    assert len(names) >= 15
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, missing


def test_no_gpu_fails_loudly():
    """without a CUDA device the product refuses to run (no CPU fallback)"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from vision_b200 import Context, VbError
    with pytest.raises(VbError):
        Context(0)

"""vb_dump_path_trace (SURVEY.md 8b debug / parity surface): the per-vertex record stream of one pixel-sample, GPU vs oracle byte for byte,
and its consistency with the film (the end record's radiance is the sample the film accumulates)."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import scene_py, capi

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _records(raw):
    n = len(raw) // C.sizeof(capi.vb_path_vertex)
    return (capi.vb_path_vertex * n).from_buffer_copy(raw)


@pytest.mark.parametrize("name,ov", [("cbox_matte", {}), ("cbox_materials", {}), ("cbox_fog", {}), ("cbox_lights", {}),
                                     ("cbox_principled", {}), ("cbox_dispersion", {}), ("cbox_matte", {"max_depth": 1}),
                                     ("cbox_lobesets", {})])
def test_path_dump_matches_oracle(vbctx, oracle, name, ov):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / f"{name}.json", dict(ov, resolution=[64, 64], spp=1))
    vbctx.upload_scene(sc)
    osc = oracle_lib.OracleScene(oracle, sc, threads=oracle.hw_threads())
    vbctx.film_clear()
    vbctx.render_frames(2, 1, 1, capi.VB_ACCUM_SUM)
    film = vbctx.film_download(capi.VB_FILM_RT)
    lengths = []
    for px, py in [(5, 7), (32, 32), (20, 50), (47, 12), (60, 60), (33, 5), (10, 31), (50, 40)]:
        g = vbctx.dump_path_trace(px, py, 2)
        o = oracle_lib.oracle_dump_path_trace(osc, px, py, 2)
        assert g == o, (name, px, py, len(g) // 96, len(o) // 96)
        r = _records(g)
        assert r[-1].inst == capi.VB_PATH_END and all(v.inst != capi.VB_PATH_END for v in r[:-1])
        assert tuple(r[-1].radiance) == tuple(film[py, px, :3])              # the film got exactly this sample
        assert r[0].bounces == 0 and tuple(r[0].throughput) == (1.0, 1.0, 1.0) and r[0].scatter_pdf == np.float32(1e16)
        for a, b in zip(r[:-2], r[1:-1]):                                     # bounces never decrease; surface vertices advance them by one
            assert b.bounces in (a.bounces, a.bounces + 1)
        lengths.append(len(r))
    assert max(lengths) > 2                                                   # some path scattered at least once


def test_path_dump_bounds(vbctx):
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", {"resolution": [32, 32], "spp": 1})
    vbctx.upload_scene(sc)
    n = C.c_uint32()
    assert vbctx.lib.vb_dump_path_trace(vbctx.h, 32, 0, 0, None, 0, C.byref(n)) != 0          # pixel outside the film
    assert vbctx.lib.vb_dump_path_trace(vbctx.h, 16, 16, 0, None, 0, C.byref(n)) == 0 and n.value >= 2   # counting call: nothing written
    buf = (capi.vb_path_vertex * 1)()
    assert vbctx.lib.vb_dump_path_trace(vbctx.h, 16, 16, 0, buf, 1, C.byref(n)) == 0 and n.value >= 2 and buf[0].bounces == 0

"""G-buffer planes of the camera stage (reference rt_geom, src/base/sensor/frame_buffer.cpp:156-219: shading normal, material albedo,
emission, linear depth / film position of the primary hit - the denoiser inputs): GPU vs oracle bit for bit, plus known answers."""
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import scene_py, capi

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.parametrize("name,ov", [("cbox_matte", {}), ("cbox_materials", {}), ("cbox_principled", {}), ("cbox_lights", {}),
                                     ("cbox_dispersion", {}), ("cbox_principled", {"spectrum": "hero"}), ("cbox_lobesets", {})])
def test_gbuffer_matches_oracle(vbctx, oracle, name, ov):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / f"{name}.json", dict(ov, resolution=[160, 160], spp=1))
    vbctx.upload_scene(sc)
    osc = oracle_lib.OracleScene(oracle, sc, threads=oracle.hw_threads())
    for frame in (0, 5):
        g = vbctx.gbuffer(frame)
        o = oracle_lib.oracle_gbuffer(osc, frame, threads=oracle.hw_threads())
        for k in ("normal", "albedo", "emission", "geom"):
            assert np.isfinite(g[k]).all(), k
            assert np.array_equal(g[k], o[k]), (name, k, float(np.abs(g[k] - o[k]).max()))
    hit = g["geom"][..., 3] == 1
    assert hit.mean() > 0.5
    assert np.allclose(np.linalg.norm(g["normal"][hit][:, :3], axis=-1), 1.0, atol=1e-5)
    assert (g["normal"][~hit] == 0).all() and (g["albedo"][..., 3] == 1).all() and (g["emission"][..., 3] == 1).all()
    assert (g["geom"][hit][:, 0] > 0).all()                               # in front of the camera


def test_gbuffer_known_answers(vbctx):
    """cbox matte under srgb: the floor's normal is +y and its albedo the material colour; the ceiling quad emits color * scale;
    depth of the back wall = its distance along the view axis; p_film stays inside its pixel (box filter radius 0.5)"""
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", {"resolution": [128, 128], "spp": 1})
    vbctx.upload_scene(sc)
    g = vbctx.gbuffer(0)
    n, a, e, geom = g["normal"], g["albedo"], g["emission"], g["geom"]
    def facing(v):
        return np.abs(n[..., :3] - np.asarray(v, np.float32)).max(-1) < 1e-5
    floor, red, green, back = facing([0, 1, 0]), facing([1, 0, 0]), facing([-1, 0, 0]), facing([0, 0, 1])
    assert floor.mean() > 0.05 and red.mean() > 0.02 and green.mean() > 0.02 and back.mean() > 0.02
    assert np.allclose(a[red][:, :3], [0.63, 0.065, 0.05], atol=1e-6)              # LeftWall / RightWall colours of the scene file
    assert np.allclose(a[green][:, :3], [0.14, 0.45, 0.091], atol=1e-6)
    wall_y = n[..., 1] > 0.999                                                       # floor + box tops: the white diffuse material
    assert np.allclose(a[wall_y][:, :3], [0.725, 0.71, 0.68], atol=1e-6)
    lit = e[..., :3].sum(-1) > 0
    assert 0 < lit.mean() < 0.1 and n[lit][:, 1].max() < -0.99                      # the emitter faces down
    le = sc.lights[0]
    assert np.allclose(e[lit][:, :3], np.array(le.color.value[:3]) * le.scale, rtol=1e-6)
    xs, ys = np.meshgrid(np.arange(128), np.arange(128))
    assert (np.abs(geom[..., 1] - (xs + 0.5)) <= 0.5).all() and (np.abs(geom[..., 2] - (ys + 0.5)) <= 0.5).all()
    # the back wall is a plane perpendicular to the view axis: one linear depth for all its pixels, and the largest in the image
    d = geom[back][:, 0]
    assert d.std() < 1e-3 and abs(d.mean() - geom[..., 0].max()) < 1e-2 and d.mean() > 6.0


def test_gbuffer_needs_render_first(vbctx):
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", {"resolution": [33, 17], "spp": 1})
    vbctx.upload_scene(sc)
    a = np.empty((17, 33, 4), np.float32)
    assert vbctx.lib.vb_gbuffer_download(vbctx.h, 0, a.ctypes.data) != 0      # film size changed: no G-buffer of this size yet
    assert vbctx.lib.vb_gbuffer_download(vbctx.h, 9, a.ctypes.data) != 0
    assert vbctx.gbuffer(0)["normal"].shape == (17, 33, 4)

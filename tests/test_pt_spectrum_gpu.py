"""GPU parity of the hero-wavelength spectrum (reference src/render_core/spectrum/hero.cpp; config C4 "dispersion spectral
glass + Disney") against the oracle at identical sampler seeds: wavelength sampling at sampler dimension -1, RGB -> spectrum
uplifting of albedo / illumination / medium coefficients, CIE -> sRGB at every accumulation, dispersive BK7 glass killing the
secondary wavelengths, spectral metal eta / k, principled BSDF tints. Both schedules (wavefront, megakernel)."""
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import scene_py, capi

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

REL_MSE_TOL = 1e-5          # same fp32 operation sequence on both sides: expected bit-identical (see test_pt_materials_gpu.py)
MAX_DIFF_PIXELS = 2e-3


def rel_mse(a, b):
    return float(np.mean((a - b) ** 2 / (b ** 2 + 1e-2)))


@pytest.mark.parametrize("name", ["cbox_dispersion", "cbox_matte", "cbox_fog", "cbox_lights", "cbox_principled", "cbox_lobesets"])
def test_hero_image_matches_oracle(vbctx, oracle, name):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / f"{name}.json", {"resolution": [128, 128], "spp": 8, "spectrum": "hero"})
    assert sc.params.spectrum == capi.VB_SPECTRUM_HERO
    vbctx.upload_scene(sc)
    osc = oracle_lib.OracleScene(oracle, sc, threads=oracle.hw_threads())
    vbctx.film_clear()
    vbctx.render_frames(0, sc.spp, 1, capi.VB_ACCUM_LERP)
    g = vbctx.film_download(capi.VB_FILM_ACCUM)
    o, _, ost = osc.render(0, sc.spp, threads=oracle.hw_threads(), stats=True)
    assert np.isfinite(g).all() and np.isfinite(o).all()
    assert g[..., :3].mean() > 0.005
    err = rel_mse(g[..., :3], o[..., :3])
    diff = np.abs(g[..., :3] - o[..., :3]) > 1e-4 * (np.abs(o[..., :3]) + 1e-3)
    frac = diff.any(axis=-1).mean()
    print(f"{name} (hero): relMSE {err:.3e}, pixels differing {frac:.3e}, bit-exact pixels {(g == o).all(axis=-1).mean():.4f}")
    assert err <= REL_MSE_TOL
    assert frac <= MAX_DIFF_PIXELS
    gst = vbctx.render_stats()
    assert abs(int(gst["any"] + gst["culled"]) - ost["any"]) <= 1e-3 * ost["any"]
    vbctx.set_option("pt.schedule", 1)
    try:
        vbctx.film_clear()
        vbctx.render_frames(0, sc.spp, 1, capi.VB_ACCUM_LERP)
        g2 = vbctx.film_download(capi.VB_FILM_ACCUM)
        assert vbctx.render_stats() == gst
    finally:
        vbctx.set_option("pt.schedule", 0)
    assert np.array_equal(g, g2)


def test_hero_needs_its_tables(vbctx):
    """selecting the hero spectrum without vb_spectrum_set / vb_spds_set fails loudly instead of rendering RGB"""
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", {"resolution": [32, 32], "spp": 1})
    vbctx.upload_scene(sc)
    sc.params.spectrum = capi.VB_SPECTRUM_HERO
    import ctypes as C
    vbctx._check(vbctx.lib.vb_integrator_set(vbctx.h, C.byref(sc.params)), "vb_integrator_set")
    with pytest.raises(capi.VbError):
        vbctx.render_frames(0, 1, 1, capi.VB_ACCUM_LERP)
    sc.params.spectrum = capi.VB_SPECTRUM_SRGB
    vbctx._check(vbctx.lib.vb_integrator_set(vbctx.h, C.byref(sc.params)), "vb_integrator_set")


@pytest.mark.parametrize("name,spectrum", [("cbox_principled", "srgb"), ("cbox_dispersion", "hero")])
def test_split_principled_launches_give_the_same_film(vbctx, name, spectrum):
    """scenes with the principled BSDF shade each bounce with two launches (lean kernel + principled instantiation, filtered by
    the material key of the ordered queue): same film and counters as the single principled launch"""
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / f"{name}.json", {"resolution": [160, 160], "spp": 4, "spectrum": spectrum})
    vbctx.upload_scene(sc)
    films, stats = [], []
    try:
        for split in (1, 0):
            vbctx.set_option("pt.split_principled", split)
            vbctx.film_clear()
            vbctx.render_frames(0, sc.spp, 1, capi.VB_ACCUM_LERP)
            films.append(vbctx.film_download(capi.VB_FILM_ACCUM)); stats.append(vbctx.render_stats())
    finally:
        vbctx.set_option("pt.split_principled", 1)
    assert np.array_equal(films[0], films[1]) and stats[0] == stats[1]


@pytest.mark.parametrize("name", ["cbox_dispersion", "cbox_fog", "cbox_lights"])
def test_cached_constant_slot_coefficients_give_the_same_film(vbctx, name):
    """hero spectrum: the coefficients of constant colour slots (materials, lights, medium sigma) are decoded once per scene by
    k_spec_cache; the per-vertex decode (pt.spec_cache = 0) must give the same bits"""
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / f"{name}.json", {"resolution": [128, 128], "spp": 4, "spectrum": "hero"})
    vbctx.upload_scene(sc)
    films = []
    try:
        for cache in (1, 0):
            vbctx.set_option("pt.spec_cache", cache)
            vbctx.film_clear()
            vbctx.render_frames(0, sc.spp, 1, capi.VB_ACCUM_LERP)
            films.append(vbctx.film_download(capi.VB_FILM_ACCUM))
    finally:
        vbctx.set_option("pt.spec_cache", 1)
    assert np.array_equal(films[0], films[1])


def test_spectrum_api_rejects_bad_input(vbctx):
    import ctypes as C
    capi._bind_pt(vbctx.lib)
    vals = np.arange(6, dtype=np.float32)
    for offs in ([0, 1, 6], [0, 4, 3]):                                   # an SPD with one sample; descending offsets
        o = np.asarray(offs, np.uint32)
        assert vbctx.lib.vb_spds_set(vbctx.h, vals.ctypes.data, o.ctypes.data, 2) != 0
    rec = capi.vb_spectrum_rec()
    rec.type, rec.dimension, rec.rgb2spec_res = capi.VB_SPECTRUM_HERO, 4, 64
    assert vbctx.lib.vb_spectrum_set(vbctx.h, C.byref(rec)) != 0           # dimension 4: not implemented, said so
    rec.dimension = 3
    assert vbctx.lib.vb_spectrum_set(vbctx.h, C.byref(rec)) != 0           # no coefficient table
    rec.type = 7
    assert vbctx.lib.vb_spectrum_set(vbctx.h, C.byref(rec)) != 0           # unknown spectrum

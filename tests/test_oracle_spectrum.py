"""CPU tests of the hero-wavelength spectrum restatement (oracle/vo_spectrum.inl; reference
src/render_core/spectrum/hero.cpp, src/base/color/spd.cpp, cie.h) and of the regenerated RGB -> spectrum table
(tools/rgb2spec_opt.c; the reference's srgb2spec.h is not in its tree). Known answers come from the formulas the
reference states (hero.cpp:16-24) evaluated in double precision, and from the colorimetric identity the table exists for:
the spectrum of an RGB illuminant, integrated against the CIE observer, gives that RGB back."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import scene_py, capi

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def hero_scene(oracle):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", {"resolution": [32, 32], "spp": 1, "spectrum": "hero"})
    return sc, oracle_lib.OracleScene(oracle, sc, threads=2)


def wavelengths(oracle, u24):
    out = (C.c_float * 6)()
    oracle.lib.vo_hero_wavelengths.argtypes = [C.c_uint32, C.c_void_p]
    oracle.lib.vo_hero_wavelengths.restype = None
    oracle.lib.vo_hero_wavelengths(u24, out)
    return np.array(out[:3], np.float64), np.array(out[3:], np.float64)


def decode(oracle, osc, u24, kind, rgb):
    lib = oracle.lib
    lib.vo_hero_decode.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.vo_hero_decode.restype = None
    a = (C.c_float * 3)(*rgb); sp = (C.c_float * 3)(); out = (C.c_float * 3)()
    lib.vo_hero_decode(osc.h, u24, kind, a, sp, out)
    return np.array(sp[:], np.float64), np.array(out[:], np.float64)


def test_wavelength_sampling_matches_the_stated_formulas(oracle):
    for u24 in (0, 1, 1 << 12, 1 << 23, 12345678, (1 << 24) - 1):
        lam, pdf = wavelengths(oracle, u24)
        u = u24 / 2.0 ** 24
        for i in range(3):
            up = np.float32(np.float32(u) + np.float32(i * np.float32(1.0 / 3.0)))
            up = float(up - np.floor(up))
            ref = 538.0 - 138.888889 * np.arctanh(0.85691062 - 1.82750197 * up)          # hero.cpp:16-18
            assert abs(lam[i] - ref) < 2e-3, (u24, i, lam[i], ref)
            assert 359.99 <= lam[i] <= 830.01
            ref_pdf = 0.0039398042 / np.cosh(0.0072 * (lam[i] - 538.0)) ** 2               # hero.cpp:22-24
            assert abs(pdf[i] - ref_pdf) < 1e-5 * ref_pdf + 1e-9


def test_wavelength_pdf_integrates_to_one():
    lam = np.linspace(360.0, 830.0, 200001)
    pdf = 0.0039398042 / np.cosh(0.0072 * (lam - 538.0)) ** 2
    assert abs(np.trapezoid(pdf, lam) - 1.0) < 1e-3


def test_table_layout_and_size():
    t = capi.load_rgb2spec()
    assert t.size == 3 * 64 ** 3 * 4
    t = t.reshape(3, 64, 64, 64, 4)
    assert np.isfinite(t).all() and (t[..., 3] == 0).all()


@pytest.mark.parametrize("rgb", [(0.8, 0.3, 0.1), (0.1, 0.5, 0.9), (0.2, 0.9, 0.3), (0.5, 0.5, 0.5), (3.0, 2.0, 0.5), (0.05, 0.04, 0.01)])
def test_illumination_round_trip(oracle, hero_scene, rgb):
    """E_lambda[ linear_srgb(decode_to_illumination(rgb)) ] == rgb: Monte-Carlo mean over a stratified set of hero draws"""
    _, osc = hero_scene
    acc = np.zeros(3)
    n = 4096
    for k in range(n):
        _, out = decode(oracle, osc, (k << 12) + 2048, 2, rgb)
        acc += out
    acc /= n
    assert np.allclose(acc, rgb, rtol=0.03, atol=0.004 * max(rgb)), (acc, rgb)


def test_albedo_spectrum_is_bounded_and_gray_is_flat(oracle, hero_scene):
    _, osc = hero_scene
    for u24 in (5, 777777, 9999999):
        for rgb in [(0.9, 0.1, 0.1), (0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (0.3, 0.3, 0.3), (0.2, 0.7, 0.7)]:
            sp, _ = decode(oracle, osc, u24, 0, rgb)
            assert (sp >= 0).all() and (sp <= 1).all()
            if rgb[0] == rgb[1] == rgb[2]:
                assert np.allclose(sp, rgb[0], atol=1e-6)


def test_secondary_wavelength_invalidation(oracle):
    lam, pdf = wavelengths(oracle, 4242424 | 0x80000000)
    lam2, pdf2 = wavelengths(oracle, 4242424)
    assert np.array_equal(lam, lam2) and pdf[0] == pdf2[0] and pdf[1] == 0 and pdf[2] == 0


def test_hero_render_agrees_with_srgb_render_in_the_mean(oracle):
    """same scene, both spectra: the images differ per sample (colour noise) but not in their means"""
    import oracle_lib
    means = {}
    for spec in ("srgb", "hero"):
        sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_matte.json", {"resolution": [48, 48], "spp": 16, "spectrum": spec})
        img, _ = oracle_lib.OracleScene(oracle, sc, threads=oracle.hw_threads()).render(0, sc.spp, threads=oracle.hw_threads())
        assert np.isfinite(img).all()
        means[spec] = img[..., :3].reshape(-1, 3).mean(0)
    assert np.allclose(means["hero"], means["srgb"], rtol=0.08), means


def test_dispersion_scene_renders(oracle):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_dispersion.json", {"resolution": [48, 48], "spp": 8})
    assert sc.params.spectrum == capi.VB_SPECTRUM_HERO
    glass = [m for m in sc.materials if m.type == capi.VB_MAT_GLASS][0]
    assert glass.flags & 4 and len(sc.spds[glass.spd[0]]) == 94
    # Schott N-BK7 catalogue: n_F(486.1 nm) = 1.52238, n_d(587.6 nm) = 1.51680, n_C(656.3 nm) = 1.51432
    curve = sc.spds[glass.spd[0]]
    at = lambda lam: np.interp(lam, 360.0 + 5.0 * np.arange(94), curve)
    assert abs(at(486.1) - 1.52238) < 2e-4 and abs(at(587.6) - 1.51680) < 2e-4 and abs(at(656.3) - 1.51432) < 2e-4
    img, _ = oracle_lib.OracleScene(oracle, sc, threads=oracle.hw_threads()).render(0, sc.spp, threads=oracle.hw_threads())
    assert np.isfinite(img).all() and img[..., :3].mean() > 0.05

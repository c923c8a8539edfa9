"""CPU unit tests of the oracle's BSDF restatement (oracle/vo_bsdf.inl): closed-form values, a second
numpy statement of the Fresnel equations, normalisation of the GGX distribution, and white-furnace
energy checks of the lobes with the reference's energy-compensation tables. The reference holds no
golden vectors for this code; its scratch test src/tests/test_bxdf.cpp:26-42 only prints
refract(wo = normalize(1,0,-2), n = (0,0,+-1), eta = 1.5), which is frozen here from the oracle."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import capi, scene_py

ROOT = Path(__file__).resolve().parent.parent
f32 = np.float32


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def lib(oracle):
    l = oracle.lib
    l.vo_probe_fresnel_dielectric.restype = C.c_float
    l.vo_probe_fresnel_dielectric.argtypes = [C.c_float, C.c_float]
    l.vo_probe_fresnel_complex.restype = C.c_float
    l.vo_probe_fresnel_complex.argtypes = [C.c_float] * 3
    l.vo_probe_refract.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
    l.vo_probe_ggx.argtypes = [C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    l.vo_probe_ggx_sample_wh.argtypes = [C.c_float, C.c_float, C.c_void_p, C.c_float, C.c_float, C.c_void_p]
    l.vo_probe_bsdf_eval.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
    l.vo_probe_bsdf_sample.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
    l.vo_vexp.restype = C.c_float; l.vo_vexp.argtypes = [C.c_float]
    l.vo_vlog.restype = C.c_float; l.vo_vlog.argtypes = [C.c_float]
    return l


def test_vexp_vlog_match_libm(lib):
    xs = np.concatenate([np.linspace(-80, 5, 2001), [-1e-6, 0.0, 1e-6]]).astype(f32)
    got = np.array([lib.vo_vexp(float(x)) for x in xs])
    np.testing.assert_allclose(got, np.exp(xs.astype(np.float64)), rtol=4e-7)
    ys = np.concatenate([np.logspace(-7, 3, 2001), [1.0, 0.5, 2.0 ** -24]]).astype(f32)
    got = np.array([lib.vo_vlog(float(y)) for y in ys])
    np.testing.assert_allclose(got, np.log(ys.astype(np.float64)), rtol=4e-7, atol=1e-7)
    assert lib.vo_vexp(-1000.0) == 0.0 and lib.vo_vlog(0.0) == -np.inf


def test_fresnel_dielectric_closed_forms(lib):
    assert abs(lib.vo_probe_fresnel_dielectric(1.0, 1.5) - 0.04) < 1e-7            # ((n-1)/(n+1))^2
    assert abs(lib.vo_probe_fresnel_dielectric(1.0, 1 / 1.5) - 0.04) < 1e-6
    assert lib.vo_probe_fresnel_dielectric(0.0, 1.5) == 1.0
    assert lib.vo_probe_fresnel_dielectric(0.5, 1 / 1.5) == 1.0                     # beyond the critical angle (41.8 deg)
    b = float(np.cos(np.arctan(1.5)))                                               # Brewster: r_parallel = 0
    rs = ((b - 1.5 * np.sqrt(1 - (1 - b * b) / 2.25)) / (b + 1.5 * np.sqrt(1 - (1 - b * b) / 2.25))) ** 2
    assert abs(lib.vo_probe_fresnel_dielectric(b, 1.5) - 0.5 * rs) < 1e-6


def test_fresnel_conductor_vs_numpy_complex(lib):
    rng = np.random.default_rng(1)
    for _ in range(200):
        c, n, k = rng.uniform(0.02, 1), rng.uniform(0.1, 3), rng.uniform(0.0, 8)
        eta = complex(n, k)
        s2 = (1 - c * c) / eta ** 2
        ct = np.sqrt(1 - s2)
        rp = (eta * c - ct) / (eta * c + ct); rs = (c - eta * ct) / (c + eta * ct)
        ref = 0.5 * (abs(rp) ** 2 + abs(rs) ** 2)
        assert abs(lib.vo_probe_fresnel_complex(c, n, k) - ref) < 2e-5


def test_refract_reference_scratch_inputs(lib):
    """inputs of reference src/tests/test_bxdf.cpp:28-40; Snell's law checked, values frozen from the oracle"""
    wo = (np.array([1, 0, -2], f32) / f32(np.sqrt(5))).astype(f32)
    for n in ([0, 0, 1], [0, 0, -1]):
        nn = np.array(n, f32); wt = np.zeros(3, f32)
        ok = lib.vo_probe_refract(_p(wo), _p(nn), 1.5, _p(wt))
        assert ok == 1
        sin_i = np.sqrt(1 - float(wo @ nn) ** 2); sin_t = np.sqrt(wt[0] ** 2 + wt[1] ** 2)
        assert abs(sin_t - sin_i / 1.5) < 1e-6 and abs(np.linalg.norm(wt) - 1) < 1e-6
        assert wt[0] < 0                                      # transmitted direction continues away from wo's side
    wt2 = np.zeros(3, f32)
    lib.vo_probe_refract(_p(wo), _p(np.array([0, 0, 1], f32)), 1.5, _p(wt2))
    np.testing.assert_allclose(wt2, [-0.29814240, 0.0, -0.95452213], atol=1e-6)   # n on the far side of wo: formula as written, optics.h:27-39


def test_ggx_distribution_normalised_and_lambda(lib):
    th = (np.arange(2000) + 0.5) / 2000 * (np.pi / 2)
    ph = (np.arange(256) + 0.5) / 256 * 2 * np.pi
    out = np.zeros(4, f32)
    wo = np.array([0.3, 0.1, 0.9486833], f32)
    for ax, ay in ((0.5, 0.5), (0.2, 0.6), (0.05, 0.05)):
        tot = 0.0
        for t in th[::8]:
            for p in ph[::8]:
                wh = np.array([np.sin(t) * np.cos(p), np.sin(t) * np.sin(p), np.cos(t)], f32)
                lib.vo_probe_ggx(ax, ay, _p(wo), _p(wo), _p(wh), _p(out))
                tot += out[0] * np.cos(t) * np.sin(t)
        tot *= (np.pi / 2 / 250) * (2 * np.pi / 32)
        assert abs(tot - 1) < 2e-2, (ax, ay, tot)
    lib.vo_probe_ggx(0.5, 0.5, _p(wo), _p(wo), _p(np.array([0, 0, 1], f32)), _p(out))
    tan2 = (wo[0] ** 2 + wo[1] ** 2) / wo[2] ** 2
    assert abs(out[3] - 0.5 * (np.sqrt(1 + 0.25 * tan2) - 1)) < 1e-6
    assert abs(out[0] - 1 / (np.pi * 0.25)) < 1e-5               # D(n) = 1 / (pi ax ay)


def _material_scene(oracle):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_materials.json", {"resolution": [8, 8], "spp": 1})
    return sc, oracle_lib.OracleScene(oracle, sc, threads=1)


def test_sampled_direction_pdf_matches_evaluate_and_white_furnace(lib, oracle):
    """for every lobe: (i) sample().eval equals evaluate(wo, wi) at the sampled direction, (ii) the
    Monte-Carlo albedo E[f cos / pdf] is <= 1 and, for the lossless lobes (glass kt=1 with the
    compensation LUT, mirror kr=1), close to 1"""
    sc, osc = _material_scene(oracle)
    types = [m.type for m in sc.materials]
    n = 20000
    for mat_type, lo, hi in ((capi.VB_MAT_GLASS, 0.9, 1.1), (capi.VB_MAT_MIRROR, 0.93, 1.07), (capi.VB_MAT_METAL, 0.3, 1.02),
                             (capi.VB_MAT_SUBSTRATE, 0.3, 1.02), (capi.VB_MAT_DIFFUSE, 0.0, 1.0)):
        mid = types.index(mat_type)
        for wo in ([0, 0, 1], [0.6, 0, 0.8], [0.3, -0.9, 0.3162278]):
            wo = np.array(wo, f32)
            out = np.zeros((n, 8), f32)
            lib.vo_probe_bsdf_sample(osc.h, mid, _p(wo), 12345, n, _p(out))
            ok = out[:, 6] > 0
            assert ok.mean() > 0.5
            w = out[ok, 3:6] / out[ok, 6:7]
            albedo = w.sum(axis=0) / n
            assert np.isfinite(albedo).all() and (albedo <= hi).all(), (mat_type, wo, albedo)
            if mat_type == capi.VB_MAT_MIRROR:
                assert (albedo >= lo).all(), (wo, albedo)
            if mat_type == capi.VB_MAT_GLASS:     # radiance transport scales transmission by 1/eta^2 (microfacet.cpp:169)
                ior = scene_py.glass_ior("BK7")
                F = lib.vo_probe_fresnel_dielectric(float(wo[2]), ior)
                expect = F + (1 - F) / ior ** 2
                assert abs(albedo[0] / expect - 1) < 0.1, (wo, albedo, expect)
            ev = np.zeros(5, f32)
            for k in np.flatnonzero(ok)[:50]:
                lib.vo_probe_bsdf_eval(osc.h, mid, _p(wo), _p(out[k, 0:3].copy()), _p(ev))
                np.testing.assert_allclose(ev[:3], out[k, 3:6], rtol=2e-3, atol=1e-6)
                if mat_type != capi.VB_MAT_GLASS:            # glass: sample() flips wh toward wo, evaluate() does not (lobe.cpp:379-399)
                    np.testing.assert_allclose(ev[3], out[k, 6], rtol=2e-3)


def test_glass_eta_and_transmission_flag(lib, oracle):
    sc, osc = _material_scene(oracle)
    mid = [m.type for m in sc.materials].index(capi.VB_MAT_GLASS)
    wo = np.array([0.0, 0.6, 0.8], f32)
    out = np.zeros((4000, 8), f32)
    lib.vo_probe_bsdf_sample(osc.h, mid, _p(wo), 7, 4000, _p(out))
    ok = out[:, 6] > 0
    refl = ok & (out[:, 2] > 0); trans = ok & (out[:, 2] < 0)
    assert refl.any() and trans.any()
    assert np.all(out[refl, 7] == 1.0)
    ior = scene_py.glass_ior("BK7")
    np.testing.assert_allclose(out[trans, 7], ior, rtol=1e-6)          # wo on the outside: eta = ior
    assert 0.02 < refl.sum() / ok.sum() < 0.2                           # ~ Fresnel reflectance at 37 deg


def test_principled_lobe_set_energy_and_consistency(lib, oracle):
    """PrincipledBSDF (cbox_principled.json: coated metal, transmissive, sheen, plastic, mixed): (i) the sampled
    direction's eval equals evaluate() at that direction (LobeSet evaluates ALL lobes at the sampled wi), (ii) the
    Monte-Carlo albedo E[f cos / pdf] stays <= 1 + noise, (iii) the six sample weights behave: a pure-metal material
    never samples below the surface, the transmissive one does, and eta is 1 on reflection and the corrected ior
    on transmission"""
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_principled.json", {"resolution": [8, 8], "spp": 1})
    osc = oracle_lib.OracleScene(oracle, sc, threads=1)
    mids = [i for i, m in enumerate(sc.materials) if m.type == capi.VB_MAT_PRINCIPLED]
    assert len(mids) == 5
    n = 20000
    for mid in mids:
        P = sc.materials[mid].params
        for wo in ([0, 0, 1], [0.6, 0, 0.8], [0.3, -0.9, 0.3162278]):
            wo = np.array(wo, f32)
            out = np.zeros((n, 8), f32)
            lib.vo_probe_bsdf_sample(osc.h, mid, _p(wo), 4242, n, _p(out))
            assert np.isfinite(out).all()
            ok = out[:, 6] > 0
            assert ok.mean() > 0.5, (mid, wo, ok.mean())
            albedo = (out[ok, 3:6] / out[ok, 6:7]).sum(axis=0) / n
            assert (albedo <= 1.08).all() and (albedo >= 0).all(), (mid, wo, albedo)
            trans = ok & (out[:, 2] < 0)
            if P[13] == 0.0:
                assert not trans.any()                                  # no transmission lobe weight: nothing goes through
            else:
                assert trans.any()
                ior = float(sc.materials[mid].ior.value[0])
                np.testing.assert_allclose(out[trans, 7], ior, rtol=1e-6)
            assert np.all(out[ok & (out[:, 2] > 0), 7] == 1.0)
            ev = np.zeros(5, f32)
            for k in np.flatnonzero(ok)[:40]:
                lib.vo_probe_bsdf_eval(osc.h, mid, _p(wo), _p(out[k, 0:3].copy()), _p(ev))
                np.testing.assert_allclose(ev[:3], out[k, 3:6], rtol=2e-3, atol=1e-6)
                np.testing.assert_allclose(ev[3], out[k, 6], rtol=2e-3)
    # sheen cloth (third principled material): brighter at grazing view than at normal incidence in the sheen tint
    sheen = mids[2]
    assert sc.materials[sheen].params[2] == 1.0


# ---- plastic, metallic, and the LobeSet materials mix / add ------------------------------------------------------
def _lobeset_scene(oracle, tmp_path, materials):
    """cbox_matte with `materials` appended and assigned to successive shapes (so none is dropped as unused); returns name -> record index"""
    import json
    import shutil
    import oracle_lib
    d = scene_py.load_json_with_comments(ROOT / "scenes" / "cbox" / "cbox_matte.json")
    d["materials"] += materials
    for m, s in zip(materials, d["shapes"]):
        s["param"]["material"] = m["name"]
    p = tmp_path / "s.json"
    p.write_text(json.dumps(d))
    shutil.copy(ROOT / "scenes" / "cbox" / "checker.jpg", tmp_path / "checker.jpg")
    sc = scene_py.build_scene(p, {"resolution": [8, 8], "spp": 1})
    order = [s["param"]["material"] for s in d["shapes"]]
    listed = [m["name"] for m in d["materials"]]
    used = [n for n in listed if n in order]          # remove_unused_materials keeps the file's list order
    return sc, oracle_lib.OracleScene(oracle, sc, threads=1), {n: i for i, n in enumerate(used)}


def _eval(lib, osc, mid, wo, wi):
    ev = np.zeros(5, f32)
    lib.vo_probe_bsdf_eval(osc.h, mid, _p(np.asarray(wo, f32)), _p(np.asarray(wi, f32)), _p(ev))
    return ev


def test_plastic_and_metallic_known_answers(lib, oracle, tmp_path):
    mats = [{"type": "plastic", "name": "pl", "param": {"color": [0.2, 0.4, 0.8], "spec": [1, 1, 1], "ior": 1.5, "roughness": 0.3}},
            {"type": "metallic", "name": "me", "param": {"color": [0.9, 0.6, 0.5], "edge_tint": [1, 1, 1], "roughness": 0.25}}]
    sc, osc, idx = _lobeset_scene(oracle, tmp_path, mats)
    assert sc.materials[idx["pl"]].type == capi.VB_MAT_PLASTIC and sc.materials[idx["me"]].type == capi.VB_MAT_METALLIC
    # plastic at normal incidence, wi = wo = n: wh = n, F = F0 = ((1.5 - 1) / 2.5)^2 = 0.04 (spec = 1); alpha = clamp((r^2)^2) = 0.0081:
    # f cos = [Rd / pi (1 - F) + F D G / 4] with D = 1 / (pi alpha^2), G = 1; pdf = lerp(F, 1 / pi, D G1 / 4) (plastic.cpp:37-50)
    n = [0, 0, 1]
    ev = _eval(lib, osc, idx["pl"], n, n)
    a = (0.3 ** 2) ** 2
    D = 1 / (np.pi * a * a)
    F = 0.04
    want = np.array([0.2, 0.4, 0.8]) / np.pi * (1 - F) + F * D / 4
    np.testing.assert_allclose(ev[:3], want, rtol=2e-5)
    np.testing.assert_allclose(ev[3], (1 - F) / np.pi + F * D / 4, rtol=2e-5)
    # metallic (F82 tint, edge_tint = 1 -> B = 0): at normal incidence F = F0 = color, f cos = color * F0 * D / 4 / compensation
    ev = _eval(lib, osc, idx["me"], n, n)
    assert ev[0] > ev[1] > ev[2] > 0 and abs(ev[1] / ev[0] - (0.6 / 0.9) ** 2) < 1e-3            # colour enters twice (kr and F0)
    # sampled directions agree with evaluate, and the Monte-Carlo albedo stays <= 1
    for name, hi in (("pl", 1.02), ("me", 1.02)):
        for wo in ([0, 0, 1], [0.6, 0, 0.8]):
            wo = np.array(wo, f32)
            out = np.zeros((20000, 8), f32)
            lib.vo_probe_bsdf_sample(osc.h, idx[name], _p(wo), 99, len(out), _p(out))
            ok = out[:, 6] > 0
            assert ok.mean() > 0.5
            albedo = (out[ok, 3:6] / out[ok, 6:7]).sum(axis=0) / len(out)
            assert np.isfinite(albedo).all() and (albedo <= hi).all(), (name, wo, albedo)
            for k in np.flatnonzero(ok)[:40]:
                ev = _eval(lib, osc, idx[name], wo, out[k, 0:3].copy())
                np.testing.assert_allclose(ev[:3], out[k, 3:6], rtol=2e-3, atol=1e-6)
                np.testing.assert_allclose(ev[3], out[k, 6], rtol=2e-3)


def test_mix_and_add_are_the_weighted_lobe_sets_of_their_children(lib, oracle, tmp_path):
    """LobeSet::create_mix / create_add + flatten (lobe.cpp:495-562): evaluate = sum_i weight_i f_i, pdf = sum_i sample_weight_i pdf_i over
    the children's lobes; mix weights (1 - frac, frac) for both; add: weights 1, sample weights 1/2; a principled child's sub-lobes keep
    their own weights times the parent's sample weight"""
    diffuse = {"type": "diffuse", "name": "d", "param": {"color": [0.5, 0.3, 0.2]}}
    mirror = {"type": "mirror", "name": "m", "param": {"color": [0.9, 0.9, 0.9], "roughness": 0.4}}
    pr = {"type": "principled_bsdf", "name": "pr", "param": {"color": [0.8, 0.5, 0.3], "roughness": 0.3, "metallic": 0.4, "coat_weight": 0.5, "sheen_weight": 0.3}}
    mats = [dict(diffuse, name="d_alone"), dict(mirror, name="m_alone"), dict(pr, name="pr_alone"),
            {"type": "mix", "name": "mix", "param": {"frac": 0.25, "mat0": diffuse, "mat1": mirror}},
            {"type": "add", "name": "add", "param": {"mat0": diffuse, "mat1": mirror}},
            {"type": "mix", "name": "mixp", "param": {"frac": 0.7, "mat0": diffuse, "mat1": pr}}]
    sc, osc, idx = _lobeset_scene(oracle, tmp_path, mats)
    assert sc.materials[idx["mix"]].type == capi.VB_MAT_MIX and sc.materials[idx["add"]].type == capi.VB_MAT_ADD
    assert len(sc.materials) == len(idx) + 6                      # six child records appended
    for wo, wi in (([0, 0, 1], [0.3, 0.2, 0.9327379]), ([0.6, 0, 0.8], [-0.5, 0.1, 0.8602325]), ([0.1, -0.7, 0.7071068], [0.2, 0.6, 0.7745967])):
        d, m, p = (_eval(lib, osc, idx[k], wo, wi) for k in ("d_alone", "m_alone", "pr_alone"))
        mix, add, mixp = (_eval(lib, osc, idx[k], wo, wi) for k in ("mix", "add", "mixp"))
        np.testing.assert_allclose(mix[:4], 0.75 * d[:4] + 0.25 * m[:4], rtol=1e-5)
        np.testing.assert_allclose(add[:3], d[:3] + m[:3], rtol=1e-5)
        np.testing.assert_allclose(add[3], 0.5 * (d[3] + m[3]), rtol=1e-5)
        np.testing.assert_allclose(mixp[:4], 0.3 * d[:4] + 0.7 * p[:4], rtol=1e-5)
    # sampling: pdf of the sampled direction is the set's pdf there; both children get chosen; albedo bounded
    wo = np.array([0.4, 0.1, 0.9110434], f32)
    for name in ("mix", "add", "mixp"):
        out = np.zeros((20000, 8), f32)
        lib.vo_probe_bsdf_sample(osc.h, idx[name], _p(wo), 321, len(out), _p(out))
        ok = out[:, 6] > 0
        assert ok.mean() > 0.6
        albedo = (out[ok, 3:6] / out[ok, 6:7]).sum(axis=0) / len(out)
        assert np.isfinite(albedo).all() and (albedo <= (2.02 if name == "add" else 1.02)).all(), (name, albedo)
        for k in np.flatnonzero(ok)[:40]:
            ev = _eval(lib, osc, idx[name], wo, out[k, 0:3].copy())
            np.testing.assert_allclose(ev[:3], out[k, 3:6], rtol=2e-3, atol=1e-6)
            np.testing.assert_allclose(ev[3], out[k, 6], rtol=2e-3)

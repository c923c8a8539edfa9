"""CPU: LightSampler's env_separate / env_prob rule (reference src/base/illumination/lightsampler.cpp:11-13,28-33,159-197 over
uniform.cpp:12-33 and power.cpp:16-53) and the directional light (render_core/light/environments/directional.cpp:20-80), as the
oracle states them: known answers worked out by hand from the reference's formulas, and C++ host == Python twin records."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import capi, host_api, scene_py

ROOT = Path(__file__).resolve().parent.parent
f32 = np.float32


def _probes(oracle):
    lib = oracle.lib
    lib.vo_probe_light_select.restype = C.c_uint32
    lib.vo_probe_light_select.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
    lib.vo_probe_light_pmf.restype = C.c_float
    lib.vo_probe_light_pmf.argtypes = [C.c_void_p, C.c_uint32]
    return lib


def _select(lib, osc, u):
    pmf = C.c_float()
    i = lib.vo_probe_light_select(osc.h, f32(u), C.byref(pmf))
    return int(i), float(pmf.value)


def _scene(name, oracle):
    import oracle_lib
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / name, {"resolution": [32, 32]})
    return sc, oracle_lib.OracleScene(oracle, sc, threads=2)


def test_uniform_env_separate_known_answers(oracle):
    lib = _probes(oracle)
    sc, osc = _scene("cbox_env_separate.json", oracle)
    types = [l.type for l in sc.lights]
    assert types == [capi.VB_LIGHT_POINT, capi.VB_LIGHT_SPOT, capi.VB_LIGHT_ENV, capi.VB_LIGHT_AREA]      # first-seen type order
    assert sc.light_sampler_env == (1, float(f32(0.3)))
    env, p = 2, float(f32(0.3))
    # u < env_prob: the environment light with PMF env_prob
    for u in (0.0, 0.1, 0.29):
        assert _select(lib, osc, u) == (env, p)
    # otherwise u' = (u - p) / (1 - p) over the 3 other lights, indices at or past the environment's shifted by one
    others = [0, 1, 3]
    for u in (0.31, 0.5, 0.6, 0.77, 0.9, 0.999):
        up = (f32(u) - f32(p)) / (f32(1) - f32(p))
        k = int(min(f32(up) * f32(3), f32(2)))
        i, pmf = _select(lib, osc, u)
        assert i == others[k], u
        assert pmf == pytest.approx(0.7 / 3, rel=1e-6)
    total = sum(lib.vo_probe_light_pmf(osc.h, i) for i in range(4))
    assert total == pytest.approx(1.0, rel=1e-6)
    # empirical frequencies follow the PMF
    us = (np.arange(20000) + 0.5) / 20000
    counts = np.bincount([_select(lib, osc, u)[0] for u in us], minlength=4) / len(us)
    np.testing.assert_allclose(counts, [0.7 / 3, 0.7 / 3, 0.3, 0.7 / 3], atol=2e-3)


def test_power_env_separate_known_answers(oracle):
    lib = _probes(oracle)
    sc, osc = _scene("cbox_env_separate_power.json", oracle)
    kind, off, n, integral = sc.light_sampler
    assert kind == capi.VB_LIGHT_SAMPLER_POWER and n == len(sc.lights) == 5
    env = [l.type for l in sc.lights].index(capi.VB_LIGHT_ENV)
    w = np.array(sc.alias_func[off:off + n], np.float64)
    assert w[env] == 0.0 and (np.delete(w, env) > 0).all()           # power.cpp:37-45: the infinite light gets weight 0
    # same other weights as without env_separate
    plain = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_lights_power.json", {"resolution": [32, 32]})
    wp = np.array(plain.alias_func[plain.light_sampler[1]:plain.light_sampler[1] + n], np.float64)
    assert np.array_equal(np.delete(w, env), np.delete(wp, env)) and wp[env] > 0
    p = float(f32(0.7))
    want = (1 - p) * w / w.sum()
    want[env] = p
    got = np.array([lib.vo_probe_light_pmf(osc.h, i) for i in range(n)])
    np.testing.assert_allclose(got, want, rtol=1e-5)
    assert got.sum() == pytest.approx(1.0, rel=1e-5)
    us = (np.arange(40000) + 0.5) / 40000
    sel = [_select(lib, osc, u) for u in us]
    assert all(pm == pytest.approx(want[i], rel=1e-5) for i, pm in sel[::97])
    counts = np.bincount([i for i, _ in sel], minlength=n) / len(us)
    np.testing.assert_allclose(counts, want, atol=3e-3)
    assert counts[env] == pytest.approx(0.7, abs=1e-4)


def test_env_prob_is_clamped_like_the_reference_constructor(tmp_path):
    src = scene_py.load_json_with_comments(ROOT / "scenes" / "cbox" / "cbox_env_separate.json")
    for given, want in ((0.0, 0.01), (1.0, 0.99), (0.25, 0.25)):
        src["light_sampler"]["param"]["env_prob"] = given
        p = ROOT / "scenes" / "cbox" / f"_tmp_env_prob_{given}.json"
        try:
            p.write_text(__import__("json").dumps(src))
            sc = scene_py.build_scene(p, {"resolution": [16, 16]})
            assert sc.light_sampler_env == (1, float(f32(want)))
            hp = host_api.HostPipeline(p)
            hp.prepare_records()
            assert hp.records().light_sampler_env == (1, float(f32(want)))
            hp.close()
        finally:
            p.unlink(missing_ok=True)


def test_directional_light_record_and_power():
    """directional.cpp:31-46: direction kept as written, world radius = diameter / 2, power = average * Pi * radius^2"""
    import math
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_directional.json", {"resolution": [16, 16]})
    d = next(l for l in sc.lights if l.type == capi.VB_LIGHT_DIRECTIONAL)
    assert [f32(x) for x in d.direction[:3]] == [f32(-0.35), f32(-0.6), f32(-1)]
    assert d.params[1] * 2 == d.params[0] and d.params[0] >= 20.0                      # min_world_radius 10
    # colour [17, 12, 4] x 0.1 normalised into the scale (number.cpp:40-49)
    np.testing.assert_allclose(list(d.color.value[:3]), [1.0, 12 / 17, 4 / 17], rtol=1e-6)
    assert d.scale == pytest.approx(1.7, rel=1e-6)
    lum = 0.212671 * 1.7 + 0.715160 * 1.2 + 0.072169 * 0.4
    assert scene_py._light_power_luminance(d, sc) == pytest.approx(lum * math.pi * d.params[1] ** 2, rel=1e-5)


def test_directional_is_the_environment_light(oracle):
    """a delta Infinite light: sampled with pdf -1 toward -direction; a path that misses evaluates it too (lightsampler.cpp:290-300)"""
    lib = _probes(oracle)
    sc, osc = _scene("cbox_env_separate_directional.json", oracle)
    env = [l.type for l in sc.lights].index(capi.VB_LIGHT_DIRECTIONAL)
    assert sc.light_sampler_env == (1, 0.5)
    assert lib.vo_probe_light_pmf(osc.h, env) == 0.5
    assert _select(lib, osc, 0.49) == (env, 0.5)
    i, pmf = _select(lib, osc, 0.51)
    assert i != env and pmf == pytest.approx(0.5 / (len(sc.lights) - 1))
    film, _ = osc.render(0, 4, threads=2)[:2]
    assert np.isfinite(film).all() and film[..., :3].mean() > 0.01


@pytest.mark.parametrize("name", ["cbox_directional", "cbox_env_separate", "cbox_env_separate_power", "cbox_env_separate_directional"])
def test_host_records_equal_python_twin(name):
    path = ROOT / "scenes" / "cbox" / f"{name}.json"
    hp = host_api.HostPipeline(path)
    hp.prepare_records()
    a = hp.records()
    hp.close()
    b = scene_py.build_scene(path)
    assert [bytes(x) for x in a.lights] == [bytes(x) for x in b.lights]
    assert a.alias_table == [(float(f32(p)), i) for p, i in b.alias_table] and a.alias_func == b.alias_func
    assert tuple(a.light_sampler) == tuple(b.light_sampler)
    assert tuple(a.light_sampler_env) == tuple(b.light_sampler_env)
    assert "vision-light-directional#constructor" in host_api.plugin_names()

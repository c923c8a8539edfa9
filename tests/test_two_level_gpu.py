"""Two-level acceleration structure ("build.two_level" = 1; the reference's own shape: a GAS per mesh + an IAS over the instances,
src/base/mgr/geometry.cpp:36-53): one bottom-level tree per UNIQUE mesh in object space + a top-level tree over the instances. Its results must
equal the flattened structure's and the brute-force oracle's bit for bit - hit ids, u, v, t, occlusion, films - because every triangle is
still tested in world space through the build's own transform chain."""
from pathlib import Path

import numpy as np
import pytest

from vision_b200 import scene_py, capi, host_api, synth

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture
def two_level(vbctx):
    vbctx.set_option("build.two_level", 1)
    yield vbctx
    vbctx.set_option("build.two_level", 0)


def _blob(n_tri, seed):
    """a small closed-ish cloud of triangles around the origin (indexed mesh with shared vertices)"""
    rng = np.random.default_rng(seed)
    pos = (rng.random((n_tri + 2, 3), np.float32) - 0.5).astype(np.float32)
    tri = np.stack([np.arange(n_tri), np.arange(n_tri) + 1, np.arange(n_tri) + 2], 1).astype(np.uint32)
    return pos, tri


def _transform(rng, scale_range=(0.05, 0.4)):
    a, b, c = rng.random(3) * 2 * np.pi
    rx = np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])
    ry = np.array([[np.cos(b), 0, np.sin(b)], [0, 1, 0], [-np.sin(b), 0, np.cos(b)]])
    rz = np.array([[np.cos(c), -np.sin(c), 0], [np.sin(c), np.cos(c), 0], [0, 0, 1]])
    s = np.diag(rng.uniform(*scale_range, 3))                      # non-uniform scale
    m = np.eye(4)
    m[:3, :3] = rx @ ry @ rz @ s
    m[:3, 3] = rng.uniform(-0.9, 0.9, 3)
    return np.ascontiguousarray(m.T, np.float32).reshape(-1)       # column-major


def test_many_instances_of_few_meshes_vs_oracle_and_flattened(two_level, oracle):
    """400 instances of 3 meshes (rotated, non-uniformly scaled, translated; two of them coincident: ties): two-level == flattened == oracle"""
    ctx = two_level
    rng = np.random.default_rng(7)
    meshes = [_blob(60, 1), _blob(200, 2), _blob(17, 3)]
    ctx.meshes_clear()
    mids = [ctx.mesh_upload(p, t) for p, t in meshes]
    inst = []
    for i in range(400):
        k = i % 3
        inst.append((k, _transform(rng)))
    inst.append((1, inst[1][1]))                                    # an exact duplicate of instance 1: every hit on it ties
    ctx.instances_set([{"mesh_id": mids[k], "o2w": m} for k, m in inst])
    st2 = ctx.accel_build()
    assert st2["n_triangles"] == 60 + 200 + 17                      # unique triangles only
    rays = oracle.synth_rays(200_000, seed=41, seed_any=42)
    g = ctx.trace_closest(rays)
    ga = ctx.trace_any(dict(rays, tmax=rays["tmax_any"]))
    _, nn, nt = ctx.trace_closest(rays, counted=True)
    assert nn > 0 and nt > 0
    ctx.set_option("build.two_level", 0)
    st1 = ctx.accel_build()
    assert st1["n_triangles"] == sum(len(meshes[k][1]) for k, _ in inst)
    f = ctx.trace_closest(rays)
    fa = ctx.trace_any(dict(rays, tmax=rays["tmax_any"]))
    for k in ("inst", "prim", "u", "v", "t"):
        assert g[k].tobytes() == f[k].tobytes(), k
    assert np.array_equal(ga, fa)
    acc = oracle.accel()
    for i, (k, m) in enumerate(inst):
        acc.add_instance(meshes[k][0], meshes[k][1], m, inst_id=i)
    acc.build(2)
    o = acc.trace_closest(rays, brute=True)
    oa, _ = acc.trace_any(dict(rays, tmax=rays["tmax_any"]), brute=True)
    for k in ("inst", "prim", "u", "v", "t"):
        assert g[k].tobytes() == o[k].tobytes(), k
    assert np.array_equal(ga, oa)
    hit = o["inst"] != 0xFFFFFFFF
    assert 0.2 < hit.mean() < 1.0 and (o["inst"][hit] == 400).sum() == 0 and (o["inst"][hit] == 1).sum() > 0     # the duplicate always loses the tie
    assert st2["node_bytes"] + st2["tri_bytes"] < 0.2 * (st1["node_bytes"] + st1["tri_bytes"])


def test_soup_as_one_instance_and_device_rays(two_level, oracle):
    """one 200 k-triangle soup, identity transform, device-resident rays through the record form and the SoA form"""
    import torch
    ctx = two_level
    mesh = synth.synth_mesh(200_000, seed=5)
    rays = synth.synth_rays(1 << 18, seed=6, seed_any=7)
    ctx.meshes_clear()
    mid = ctx.mesh_upload_soup(mesh); ctx.instances_set([{"mesh_id": mid}]); ctx.accel_build()
    d = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
    g = ctx.trace_closest(d)
    rec = torch.from_numpy(capi.rays_to_records(rays)).cuda()
    gr = ctx.trace_closest_records(rec)
    ctx.synchronize()                                      # device-buffer calls are asynchronous on the context's stream
    ctx.set_option("build.two_level", 0)
    ctx.accel_build()
    f = ctx.trace_closest(d)
    ctx.synchronize()
    for k in ("inst", "prim", "u", "v", "t"):
        assert torch.equal(g[k], f[k]), k
    st = ctx.accel_build(two_level=True)                  # the per-build flag (VB_BUILD_TWO_LEVEL) instead of the context option
    g2 = ctx.trace_closest(d)
    ctx.synchronize()
    assert all(torch.equal(g2[k], f[k]) for k in f)
    bad = capi.vb_build_opts(8)
    assert ctx.lib.vb_accel_build(ctx.h, __import__("ctypes").byref(bad), None) != 0      # unknown flag: refused
    ctx.accel_build()
    assert torch.equal(gr[:, 1].cpu(), f["prim"].cpu().view(torch.int32))


@pytest.mark.parametrize("name", ["cbox_matte", "cbox_fog", "cbox_lights", "cbox_principled"])
def test_path_tracer_films_identical(two_level, name):
    """both schedules, G-buffer and the path dump over the two-level structure == the flattened structure"""
    ctx = two_level
    sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / f"{name}.json", {"resolution": [96, 96], "spp": 2})
    out = {}
    for tl in (1, 0):
        ctx.set_option("build.two_level", tl)
        ctx.upload_scene(sc)
        for sched in (0, 1):
            ctx.set_option("pt.schedule", sched)
            ctx.film_clear(); ctx.render_frames(0, 2, 1, capi.VB_ACCUM_LERP)
            out[(tl, sched)] = ctx.film_download(capi.VB_FILM_ACCUM)
        ctx.set_option("pt.schedule", 0)
        out[(tl, "g")] = ctx.gbuffer(1)
        out[(tl, "d")] = ctx.dump_path_trace(40, 50, 1)
    assert np.array_equal(out[(1, 0)], out[(0, 0)]) and np.array_equal(out[(1, 1)], out[(0, 1)]) and np.array_equal(out[(1, 0)], out[(1, 1)])
    for k in out[(1, "g")]:
        assert np.array_equal(out[(1, "g")][k], out[(0, "g")][k]), k
    assert out[(1, "d")] == out[(0, "d")]


def test_classroom_through_the_host_layer(two_level):
    """79 models of the classroom: two-level film == flattened film (2 frames at 480 x 270)"""
    films = {}
    for tl in (1, 0):
        hp = host_api.HostPipeline(ROOT / "scenes/classroom/vision_scene.json", {"resolution": [480, 270]})
        hp.prepare(0)
        c = hp.context()
        c.set_option("build.two_level", tl)
        st = c.accel_build()
        hp.render(2)
        films[tl] = (hp.accum(), st)
        hp.close()
    assert np.array_equal(films[1][0], films[0][0])
    assert films[1][1]["n_triangles"] == films[0][1]["n_triangles"] == 103832       # no mesh is instanced twice in this scene


def test_vision_cli_two_level_flag(tmp_path):
    """vision-cli --two-level writes the same linear film (.pfm) as the default build"""
    import subprocess
    cli = str(ROOT / "vision_b200" / "vision-cli")
    scene = str(ROOT / "scenes" / "cbox" / "cbox_materials.json")
    outs = []
    for k, extra in enumerate(([], ["--two-level"])):
        pfm = tmp_path / f"f{k}.pfm"
        r = subprocess.run([cli, "-s", scene, "-o", str(tmp_path / f"o{k}.png"), "--pfm", str(pfm), "--spp", "2"] + extra, capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stderr[-500:]
        outs.append(pfm.read_bytes())
    assert outs[0] == outs[1] and len(outs[0]) > 256 * 256 * 12

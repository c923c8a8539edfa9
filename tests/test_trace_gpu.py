"""GPU parity: libvision_b200.so (through the C ABI) vs the oracle on the same seeded inputs.
Bar (BASELINE.json north_star): inst_id / prim_id bit-exact; t and barycentrics <= 1e-5 relative -
both sides implement one normative fp32 operation order, so we assert bit equality of t,u,v too."""
import numpy as np
import pytest

from vision_b200 import synth

pytestmark = pytest.mark.gpu

MISS = 0xFFFFFFFF


def _build_soup(vbctx, mesh):
    mid = vbctx.mesh_upload_soup(mesh)
    vbctx.instances_set([{"mesh_id": mid}])
    return vbctx.accel_build()


def _compare_closest(g, o):
    assert np.array_equal(g["inst"], o["inst"]), f"inst mismatches: {(g['inst'] != o['inst']).sum()}"
    assert np.array_equal(g["prim"], o["prim"]), f"prim mismatches: {(g['prim'] != o['prim']).sum()}"
    hit = o["inst"] != MISS
    for f in ("t", "u", "v"):
        np.testing.assert_allclose(g[f][hit], o[f][hit], rtol=1e-5, atol=0)
        assert g[f].tobytes() == o[f].tobytes(), f"{f} not bit-exact"


@pytest.mark.parametrize("ntri", [1, 2, 3, 4, 9, 100, 20000])
def test_closest_and_any_vs_bruteforce_oracle(vbctx, oracle, ntri):
    mesh = oracle.synth_mesh(ntri, seed=21)
    rays = oracle.synth_rays(20000, seed=22, seed_any=23)
    st = _build_soup(vbctx, mesh)
    assert st["n_triangles"] == ntri
    acc = oracle.accel()
    acc.add_soup(mesh)
    acc.build(4)
    brute = ntri <= 100
    o = acc.trace_closest(rays, threads=8, brute=brute)
    g = vbctx.trace_closest(rays)
    _compare_closest(g, o)
    any_rays = dict(rays, tmax=rays["tmax_any"])
    oa, _ = acc.trace_any(any_rays, brute=brute)
    ga = vbctx.trace_any(any_rays)
    assert np.array_equal(ga, oa)


def test_million_triangle_soup_vs_oracle_bvh(vbctx, oracle):
    mesh = synth.synth_mesh(1_000_000, seed=1)
    rays = synth.synth_rays(1 << 18, seed=2, seed_any=3)
    st = _build_soup(vbctx, mesh)
    assert st["n_triangles"] == 1_000_000
    acc = oracle.accel()
    acc.add_soup(mesh)
    acc.build(oracle.hw_threads())
    o = acc.trace_closest(rays, threads=oracle.hw_threads())
    g, nodes, tris = vbctx.trace_closest(rays, counted=True)
    _compare_closest(g, o)
    assert nodes > 0 and tris > 0
    assert (o["inst"] != MISS).mean() > 0.5
    any_rays = dict(rays, tmax=rays["tmax_any"])
    oa, _ = acc.trace_any(any_rays, threads=oracle.hw_threads())
    ga = vbctx.trace_any(any_rays)
    assert np.array_equal(ga, oa)
    assert 0.05 < oa.mean() < 0.95


def test_instanced_indexed_meshes_and_ties(vbctx, oracle):
    """two instances of one indexed quad, the second translated onto the first: equal-t ties
    resolve to the smallest (inst, prim); a third instance is rotated+scaled"""
    pos = np.array([[-1, -1, 0], [1, -1, 0], [-1, 1, 0], [1, 1, 0]], np.float32)
    tri = np.array([[0, 1, 2], [2, 1, 3]], np.uint32)      # reference quad order (quad.cpp:44)
    ident = np.eye(4, dtype=np.float32).T.reshape(-1)
    m3 = np.eye(4, dtype=np.float32)
    c, s = np.cos(0.7), np.sin(0.7)
    m3[:3, :3] = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], np.float32) * 0.5
    m3[:3, 3] = [0.1, 0.2, -0.8]
    o2w3 = np.ascontiguousarray(m3.T).reshape(-1)           # column-major
    mid = vbctx.mesh_upload(pos, tri)
    vbctx.instances_set([{"mesh_id": mid, "o2w": ident}, {"mesh_id": mid, "o2w": ident}, {"mesh_id": mid, "o2w": o2w3}])
    st = vbctx.accel_build()
    assert st["n_triangles"] == 6
    acc = oracle.accel()
    for i, m in enumerate((ident, ident, o2w3)):
        acc.add_instance(pos, tri, m, inst_id=i)
    acc.build(1)
    rays = oracle.synth_rays(50000, seed=31, seed_any=32)
    o = acc.trace_closest(rays, brute=True)
    g = vbctx.trace_closest(rays)
    _compare_closest(g, o)
    hit0 = o["inst"] == 0
    assert hit0.sum() > 1000 and (o["inst"] == 1).sum() == 0     # inst 1 always loses the tie
    assert (o["inst"] == 2).sum() > 100


def test_ragged_and_degenerate_rays(vbctx, oracle):
    mesh = oracle.synth_mesh(3000, seed=41)
    _build_soup(vbctx, mesh)
    acc = oracle.accel()
    acc.add_soup(mesh)
    acc.build(2)
    rays = oracle.synth_rays(1000, seed=42, seed_any=43)
    rays["dx"][0:10] = 0.0; rays["dy"][0:10] = 0.0; rays["dz"][0:10] = 1.0      # axis-aligned
    rays["dx"][10:20] = 0.0; rays["dy"][10:20] = 0.0; rays["dz"][10:20] = 0.0   # zero direction
    rays["tmax"][20:30] = 0.0                                                    # empty interval
    rays["tmin"][30:40] = 0.3                                                    # t_min honoured
    rays["dx"][40:50] *= 7.5; rays["dy"][40:50] *= 7.5; rays["dz"][40:50] *= 7.5  # un-normalised dirs
    rays["ox"][50:55] = np.nan                                                   # NaN rays miss
    for n in (0, 1, 31, 33, 1000):
        sub = {k: np.ascontiguousarray(v[:n]) for k, v in rays.items()}
        if n == 0:
            g = vbctx.trace_closest(sub)
            assert g["inst"].shape == (0,)
            continue
        o = acc.trace_closest(sub, brute=True)
        g = vbctx.trace_closest(sub)
        _compare_closest(g, o)
        oa, _ = acc.trace_any(sub, brute=True)
        assert np.array_equal(vbctx.trace_any(sub), oa)


def test_device_resident_rays(vbctx, oracle):
    import torch
    mesh = oracle.synth_mesh(5000, seed=51)
    _build_soup(vbctx, mesh)
    rays = oracle.synth_rays(10000, seed=52, seed_any=53)
    host = vbctx.trace_closest(rays)
    drays = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
    dev = vbctx.trace_closest(drays)
    vbctx.synchronize()
    torch.cuda.synchronize()
    assert np.array_equal(dev["prim"].cpu().numpy().view(np.uint32), host["prim"])
    assert dev["t"].cpu().numpy().tobytes() == host["t"].tobytes()


def test_host_buffers_pipelined_in_ranges(vbctx, oracle):
    """host-buffer calls are cut into ray ranges (copy-in / trace / copy-out overlapped): same results as one range
    and as device-resident rays, ragged last range included"""
    import torch
    mesh = synth.synth_mesh(20000, seed=61)
    _build_soup(vbctx, mesh)
    n = 300_001
    rays = synth.synth_rays(n, seed=62, seed_any=63)
    any_rays = dict(rays, tmax=rays["tmax_any"])
    vbctx.set_option("trace.host_chunks", 1)
    one = vbctx.trace_closest(rays); one_any = vbctx.trace_any(any_rays)
    try:
        vbctx.set_option("trace.host_chunk_min_rays", 1024)
        for chunks in (3, 16):
            vbctx.set_option("trace.host_chunks", chunks)
            got = vbctx.trace_closest(rays)
            for f in ("inst", "prim", "t", "u", "v"):
                assert got[f].tobytes() == one[f].tobytes(), (chunks, f)
            assert np.array_equal(vbctx.trace_any(any_rays), one_any)
        # device rays -> host results and host rays -> device results take the same path
        drays = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
        dev = vbctx.trace_closest(drays)
        vbctx.synchronize()
        assert dev["prim"].cpu().numpy().view(np.uint32).tobytes() == one["prim"].tobytes()
    finally:
        vbctx.set_option("trace.host_chunks", 4)
        vbctx.set_option("trace.host_chunk_min_rays", 1 << 20)
    acc = oracle.accel()
    acc.add_soup(mesh)
    acc.build(oracle.hw_threads())
    _compare_closest(one, acc.trace_closest(rays, threads=oracle.hw_threads()))


def test_errors_are_reported(vbctx):
    from vision_b200 import VbError
    with pytest.raises(VbError):
        vbctx.instances_set([{"mesh_id": 10_000}])
    with pytest.raises(VbError):
        vbctx.mesh_upload(np.zeros((3, 3), np.float32), np.array([[0, 1, 7]], np.uint32))


def test_full_size_properties_16Mi_rays_10M_triangles(vbctx, oracle):
    """BASELINE.json configs[1] at its full size, through size-independent properties (the oracle would need minutes here):
    (1) closest / any consistency: no triangle lies in front of the reported closest hit (any-hit with t_max just below t is
        free) and the hit triangle itself occludes (any-hit with t_max just above t is occluded); rays that miss are free;
    (2) order independence: the answer of a ray does not depend on where it sits in the batch (reversed batch -> reversed answers);
    (3) a 1/64 sample of the rays agrees bit for bit with the oracle's own BVH over the same 10 M triangles."""
    n_tri, n_rays = 10_000_000, 1 << 24
    mesh = oracle.synth_mesh(n_tri, seed=1, threads=oracle.hw_threads())
    rays = oracle.synth_rays(n_rays, seed=2, seed_any=3, threads=oracle.hw_threads())
    st = _build_soup(vbctx, mesh)
    assert st["n_triangles"] == n_tri
    g = vbctx.trace_closest(rays)
    hit = g["inst"] != MISS
    assert 0.85 < hit.mean() < 0.97
    t = g["t"]
    assert np.all(t[hit] > 0) and np.all(np.isfinite(t[hit])) and np.all(g["u"][hit] >= 0) and np.all(g["u"][hit] + g["v"][hit] <= 1.0 + 1e-6)
    below = dict(rays, tmax=np.where(hit, t * np.float32(1 - 1e-4), rays["tmax"]).astype(np.float32))
    above = dict(rays, tmax=np.where(hit, t * np.float32(1 + 1e-4), rays["tmax"]).astype(np.float32))
    occ_below, occ_above = vbctx.trace_any(below), vbctx.trace_any(above)
    assert not occ_below.any()                                   # nothing in front of the closest hit; misses stay free
    assert np.array_equal(occ_above.astype(bool), hit)           # the hit triangle occludes; misses stay free
    rev = {k: np.ascontiguousarray(v[::-1]) for k, v in rays.items()}
    gr = vbctx.trace_closest(rev)
    for f in ("inst", "prim", "u", "v", "t"):
        assert np.array_equal(gr[f][::-1], g[f]), f
    sub = {k: np.ascontiguousarray(v[::64]) for k, v in rays.items()}
    acc = oracle.accel()
    acc.add_soup(mesh)
    acc.build(oracle.hw_threads())
    o = acc.trace_closest(sub, threads=oracle.hw_threads())
    for f in ("inst", "prim", "u", "v", "t"):
        assert np.array_equal(o[f], g[f][::64]), f


def test_record_form_equals_soa_form(vbctx):
    """vb_trace_*_records (reference Ray / TriangleHit records): same answers as the SoA form from device buffers,
    from host buffers (one range and pipelined ranges), and with "trace.async" = 1 (consecutive calls overlapping over the two
    staging sets, completed by vb_ctx_synchronize)"""
    import torch
    from vision_b200 import capi
    mesh = synth.synth_mesh(200_000, seed=11)
    n = 3 * (1 << 20) + 1234
    rays = synth.synth_rays(n, seed=12, seed_any=13)
    mid = vbctx.mesh_upload_soup(mesh)
    vbctx.instances_set([{"mesh_id": mid}])
    vbctx.accel_build()
    want = vbctx.trace_closest(rays)
    want_occ = vbctx.trace_any(dict(rays, tmax=rays["tmax_any"]))
    rec = capi.rays_to_records(rays)
    rec_any = capi.rays_to_records(rays, "tmax_any")

    def check(h, o):
        h = np.asarray(h).view(np.uint32)
        assert np.array_equal(h[:, 0], want["inst"]) and np.array_equal(h[:, 1], want["prim"])
        assert h[:, 2].tobytes() == want["u"].tobytes() and h[:, 3].tobytes() == want["v"].tobytes()
        assert np.array_equal(np.asarray(o), want_occ)
    # host buffers, synchronous: one range, then pipelined ranges
    for chunks in (1, 3):
        vbctx.set_option("trace.host_chunks", chunks)
        vbctx.set_option("trace.host_chunk_min_rays", 1 << 19)
        check(vbctx.trace_closest_records(rec), vbctx.trace_any_records(rec_any))
    # device buffers
    d = torch.from_numpy(rec).cuda(); d_any = torch.from_numpy(rec_any).cuda()
    dh, do = vbctx.trace_closest_records(d), vbctx.trace_any_records(d_any)
    vbctx.synchronize()                      # device-buffer calls are stream-ordered on the context's stream, not torch's
    check(dh.cpu().numpy(), do.cpu().numpy())
    # asynchronous host calls: several in flight, different outputs each
    vbctx.set_option("trace.async", 1)
    try:
        pin = lambda a: torch.from_numpy(a).pin_memory().numpy()
        hr, hr_any = pin(rec), pin(rec_any)
        outs = [(pin(np.zeros((n, 4), np.uint32)), pin(np.zeros(n, np.uint8))) for _ in range(3)]
        for h, o in outs:
            vbctx.trace_closest_records(hr, h)
            vbctx.trace_any_records(hr_any, o)
        vbctx.synchronize()
        for h, o in outs:
            check(h, o)
    finally:
        vbctx.set_option("trace.async", 0)
        vbctx.set_option("trace.host_chunks", 4)
        vbctx.set_option("trace.host_chunk_min_rays", 1 << 20)
    # empty and tiny inputs
    assert vbctx.trace_closest_records(rec[:0]).shape == (0, 4)
    check_small = vbctx.trace_closest_records(np.ascontiguousarray(rec[:5])).view(np.uint32)
    assert np.array_equal(check_small[:, 1], want["prim"][:5])


def test_trees_deeper_than_the_traversal_stack_are_refused(vbctx, oracle):
    """The traversal stack holds 64 entries, one per wide level in the worst case; a deeper tree would drop pushes and skip subtrees
    silently. The build therefore checks the depth: a clustering (PLOC) tree that is too deep is rebuilt as a radix tree (bounded
    depth), and a tree that is still too deep is refused. "build.max_levels" lowers the accepted depth so that both steps can be
    exercised (no float geometry makes a 64-level wide tree: an agglomerative chain needs positions growing by 2x per element)."""
    from vision_b200 import capi
    mesh = oracle.synth_mesh(20000, seed=51)
    rays = oracle.synth_rays(20000, seed=52, seed_any=53)
    acc = oracle.accel()
    acc.add_soup(mesh)
    acc.build(4)
    want = acc.trace_closest(rays, threads=8)
    try:
        vbctx.set_option("build.ploc", 1)
        vbctx.set_option("build.max_levels", 3)                      # 8^3 < 20000 leaves: no tree fits
        vbctx.profile_reset()
        with pytest.raises(capi.VbError, match="deeper than the traversal stack"):
            _build_soup(vbctx, mesh)
        assert vbctx.profile_get("bvh_build_radix_fallback")[1] == 1         # clustering first, then the radix tree, then refused
        with pytest.raises(capi.VbError, match="accel not built"):
            vbctx.trace_closest(rays)
        vbctx.set_option("build.max_levels", 64)
        vbctx.profile_reset()
        _build_soup(vbctx, mesh)
        assert vbctx.profile_get("bvh_build_radix_fallback")[1] == 0
        _compare_closest(vbctx.trace_closest(rays), want)
    finally:
        vbctx.set_option("build.ploc", -1)
        vbctx.set_option("build.max_levels", 64)

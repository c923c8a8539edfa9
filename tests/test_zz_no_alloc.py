"""GPU: vb_trace_closest_static, the closest-hit trace that allocates nothing (include/vision_b200.h; no counterpart in the
reference). Called through the C ABI on device buffers and compared with the existing
entry point vb_trace_closest on the same rays: identical ids, bit-identical t / u / v (the tolerance tests/test_trace_gpu.py uses),
on calls in a row that share the static scratch array - a sorted batch, a different and shorter sorted batch (stale keys and
permutation entries beyond it must not matter), a batch below the ray-sort threshold, and the first batch again.
This file is collected last on purpose."""
import numpy as np
import pytest

from vision_b200 import capi, synth

pytestmark = pytest.mark.gpu
HIT_FIELDS = ("inst", "prim", "u", "v", "t")


def _device(arrs):
    import torch
    return {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in arrs.items()}


def _empty_hits(n):
    import torch
    out = {f: torch.empty(n, dtype=torch.int32, device="cuda") for f in ("inst", "prim")}
    out.update({f: torch.empty(n, dtype=torch.float32, device="cuda") for f in ("u", "v", "t")})
    return out


def test_static_scratch_trace_equals_vb_trace_closest(vbctx):
    import torch
    mesh = synth.synth_mesh(300_000, seed=31)
    mid = vbctx.mesh_upload_soup(mesh)
    vbctx.instances_set([{"mesh_id": mid}])
    vbctx.accel_build()
    batches = [synth.synth_rays(capi.VB_TRACE_STATIC_MAX_RAYS, seed=32, seed_any=33),      # the maximum, sorted
               synth.synth_rays(150_001, seed=34, seed_any=35),                            # shorter, other rays, sorted
               synth.synth_rays(5_000, seed=36, seed_any=37)]                              # below the sort threshold
    batches.append(batches[0])
    for rays in batches:
        n = rays["ox"].shape[0]
        d = _device({k: rays[k] for k in ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax")})
        want = vbctx.trace_closest(d)                       # existing entry point, device buffers
        got = _empty_hits(n)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        vbctx.trace_closest_static(d, got)                  # stream-ordered on the context's stream
        vbctx.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before, "the device's free memory changed across a no-allocation call"
        hits = 0
        for f in HIT_FIELDS:
            g, w = got[f].cpu().numpy(), want[f].cpu().numpy()
            assert g.tobytes() == w.tobytes(), f"{f} differs on the {n}-ray batch"
            if f == "inst":
                hits = int((w.view(np.uint32) != 0xFFFFFFFF).sum())
        assert hits > n // 4


def test_static_scratch_trace_rejects_what_it_cannot_hold(vbctx):
    mesh = synth.synth_mesh(1000, seed=41)
    mid = vbctx.mesh_upload_soup(mesh)
    vbctx.instances_set([{"mesh_id": mid}])
    vbctx.accel_build()
    n = capi.VB_TRACE_STATIC_MAX_RAYS + 1
    rays = synth.synth_rays(n, seed=42, seed_any=43)
    d = _device({k: rays[k] for k in ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax")})
    with pytest.raises(capi.VbError, match="exceed the static scratch's maximum of 262144"):
        vbctx.trace_closest_static(d, _empty_hits(n))
    host = {k: rays[k][:100] for k in ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax")}
    host_out = {"inst": np.empty(100, np.uint32), "prim": np.empty(100, np.uint32), "u": np.empty(100, np.float32),
                "v": np.empty(100, np.float32), "t": np.empty(100, np.float32)}
    with pytest.raises(capi.VbError, match="device buffers"):
        vbctx.trace_closest_static(host, host_out)

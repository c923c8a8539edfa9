"""ray-sort granularity: trace.sort_skip_bits x trace.sort_oct_shift on the microbench and on classroom (run under gpurun)"""
import sys, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from vision_b200 import Context, synth, host_api

n = 1 << 24
ctx = Context(0)
mesh = synth.synth_mesh(10_000_000, seed=1); rays = synth.synth_rays(n)
mid = ctx.mesh_upload_soup(mesh); ctx.instances_set([{"mesh_id": mid}]); ctx.accel_build()
d = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}; da = dict(d, tmax=d["tmax_any"])
hits = occ = None
for skip, shift in ((0, 0), (4, 0), (8, 0), (8, 8), (8, 5), (12, 0)):
    ctx.set_option("trace.sort_skip_bits", skip); ctx.set_option("trace.sort_oct_shift", shift)
    for _ in range(2):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    ctx.profile_reset()
    for _ in range(4):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    c, _ = ctx.profile_get("trace_closest"); a, _ = ctx.profile_get("trace_any")
    print(f"soup10m skip={skip} oct_shift={shift}: closest {c/4:.2f} ms any {a/4:.2f} ms total {2*n/((c+a)/4)/1e3:.0f} Mrays/s", flush=True)
ctx.close()
hp = host_api.HostPipeline(ROOT / "scenes/classroom/vision_scene.json", {"resolution": [1920, 1080]}); hp.prepare(0)
c = hp.context()
for skip, shift in ((0, 0), (4, 0), (8, 0), (8, 8), (8, 5), (12, 0)):
    c.set_option("trace.sort_skip_bits", skip); c.set_option("trace.sort_oct_shift", shift)
    hp.invalidate(); hp.render(16); c.synchronize()
    c.profile_reset(); hp.invalidate()
    t0 = time.perf_counter(); hp.render(32); c.synchronize(); dt = time.perf_counter() - t0
    st = {k: round(c.profile_get(k)[0], 1) for k in ("trace_closest", "trace_any", "pt_shade")}
    print(f"classroom skip={skip} oct_shift={shift}: {1920*1080*32/dt/1e6:.1f} Msamples/s {st}", flush=True)

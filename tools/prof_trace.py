"""One closest + one any-hit launch on the bench workload, for ncu (run under gpurun)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200 import Context, synth

tris = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 24
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
ctx = Context(0)
mesh = synth.synth_mesh(tris, seed=1)
rays = synth.synth_rays(n)
mid = ctx.mesh_upload_soup(mesh); ctx.instances_set([{"mesh_id": mid}]); st = ctx.accel_build()
d = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
da = dict(d, tmax=d["tmax_any"])
hits = occ = None
for _ in range(reps):
    hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
ctx.synchronize()
c, k = ctx.profile_get("trace_closest"); a, _ = ctx.profile_get("trace_any")
print(f"closest {c/k:.2f} ms any {a/k:.2f} ms build {st['build_ms']:.2f} ms")

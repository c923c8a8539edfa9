"""Sweep the traversal tunables on the bench workload (run under gpurun): trace.refill_thr x trace.tri_thr x trace.blocks_per_sm."""
import itertools, json, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200 import Context, synth

tris = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 24
ctx = Context(0)
mesh = synth.synth_mesh(tris, seed=1)
rays = synth.synth_rays(n)
mid = ctx.mesh_upload_soup(mesh); ctx.instances_set([{"mesh_id": mid}])
d = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
da = dict(d, tmax=d["tmax_any"])
hits = None; occ = None
st = ctx.accel_build()
def run(tag):
    global hits, occ
    for _ in range(2):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    ctx.profile_reset()
    for _ in range(4):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    c, _ = ctx.profile_get("trace_closest"); a, _ = ctx.profile_get("trace_any")
    print(f"{tag}: closest {c/4:.2f} ms any {a/4:.2f} ms total {2*n/((c+a)/4)/1e3:.0f} Mrays/s", flush=True)
for refill, tri, bps in itertools.product((24, 28, 30), (8, 12, 16), (8, 12)):
    ctx.set_option("trace.refill_thr", refill); ctx.set_option("trace.tri_thr", tri); ctx.set_option("trace.blocks_per_sm", bps)
    run(f"refill={refill} tri={tri} bps={bps}")

"""Sweep the traversal tunables on the bench workload (run under gpurun)."""
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
def run(tag):
    global hits, occ
    for _ in range(2):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    ctx.profile_reset()
    for _ in range(3):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    c, _ = ctx.profile_get("trace_closest"); a, _ = ctx.profile_get("trace_any")
    _, nn, nt = ctx.trace_closest(d, hits, counted=True)
    print(f"{tag}: closest {c/3:.2f} ms ({n/(c/3)/1e3:.0f} Mrays/s)  any {a/3:.2f} ms ({n/(a/3)/1e3:.0f} Mrays/s) nodes/ray {nn/n:.1f} tris/ray {nt/n:.1f}", flush=True)
for ml in (1,):
    ctx.set_option("build.max_leaf", ml)
    st = ctx.accel_build()
    for bps in (8, 10):
        ctx.set_option("trace.blocks_per_sm", bps)
        run(f"max_leaf={ml} bps={bps}")

import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from vision_b200 import Context, synth, capi
ctx = Context(0)
mesh = synth.synth_mesh(200_000, seed=5)
rays = synth.synth_rays(1 << 18, seed=6, seed_any=7)
mid = ctx.mesh_upload_soup(mesh); ctx.instances_set([{"mesh_id": mid}])
d = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
ctx.accel_build()
r0 = ctx.trace_closest(d); ctx.synchronize()
f = {k: v.clone() for k, v in r0.items()}
for rep in range(8):
    how = "flag" if rep % 2 else "option"
    if how == "option":
        ctx.set_option("build.two_level", 1); st = ctx.accel_build(); ctx.set_option("build.two_level", 0)
    else:
        st = ctx.accel_build(two_level=True)
    g = ctx.trace_closest(d); ctx.synchronize()
    bad = {k: int((g[k] != f[k]).sum()) for k in f}
    print(rep, how, st["n_nodes"], bad, flush=True)
    if any(bad.values()):
        i = int(torch.nonzero(g["prim"] != f["prim"])[0]) if bad["prim"] else int(torch.nonzero(g["t"] != f["t"])[0])
        print("   ray", i, {k: (g[k][i].item(), f[k][i].item()) for k in f})
    ctx.accel_build()
    f2 = ctx.trace_closest(d); ctx.synchronize()
    print("   flat again equal:", all(torch.equal(f2[k], f[k]) for k in f), flush=True)

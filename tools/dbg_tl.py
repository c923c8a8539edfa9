import sys, faulthandler
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
faulthandler.dump_traceback_later(40, exit=True)
import numpy as np, torch
from vision_b200 import Context, synth
print("import ok", flush=True)
ctx = Context(0)
mesh = synth.synth_mesh(2000, seed=5)
rays = synth.synth_rays(4096, seed=6, seed_any=7)
mid = ctx.mesh_upload_soup(mesh); ctx.instances_set([{"mesh_id": mid}])
print("flat build", ctx.accel_build(), flush=True)
f = ctx.trace_closest(rays)
print("flat trace ok", int((f["inst"] != 0xFFFFFFFF).sum()), flush=True)
ctx.set_option("build.two_level", 1)
print("tl build...", flush=True)
print(ctx.accel_build(), flush=True)
print("tl trace...", flush=True)
g = ctx.trace_closest(rays)
print("tl trace ok", all(g[k].tobytes() == f[k].tobytes() for k in g), flush=True)

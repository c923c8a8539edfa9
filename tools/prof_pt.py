"""One path-tracing launch on a scene, for ncu / quick timing (run under gpurun).
usage: prof_pt.py <scene.json> <width> <height> <spp> [reps] [process_mediums 0/1]"""
import sys
import time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200 import Context, scene_py, capi

path = sys.argv[1]
w, h, spp = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 1
ov = {"resolution": [w, h], "spp": spp}
if len(sys.argv) > 6:
    ov["process_mediums"] = bool(int(sys.argv[6]))
import os
if "VB_MAX_DEPTH" in os.environ:
    ov["max_depth"] = int(os.environ["VB_MAX_DEPTH"])
sc = scene_py.build_scene(path, ov)
import os
ctx = Context(0)
if "VB_PLOC" in os.environ:
    ctx.set_option("build.ploc", int(os.environ["VB_PLOC"]))
if "VB_SPLIT" in os.environ:
    ctx.set_option("build.split_pct", int(os.environ["VB_SPLIT"]))
if "VB_SIZE_CLASSES" in os.environ:
    ctx.set_option("build.size_classes", int(os.environ["VB_SIZE_CLASSES"]))
bst = ctx.upload_scene(sc)
print("build:", bst)
ctx.set_option("pt.schedule", int(os.environ.get("VB_SCHED", "0")))
if "VB_MAX_PATHS" in os.environ:
    ctx.set_option("pt.max_paths", int(os.environ["VB_MAX_PATHS"]))
if os.environ.get("VB_COUNT"):
    ctx.set_option("pt.count_work", 1)
    ctx.film_clear(); ctx.render_frames(0, 1, 1, capi.VB_ACCUM_SUM); ctx.synchronize()
    st, wk = ctx.render_stats(), ctx.render_work()
    print(f"work: closest {wk['closest_nodes'] / st['closest']:.1f} nodes {wk['closest_tris'] / st['closest']:.2f} tris per ray; "
          f"any {wk['any_nodes'] / max(st['any'], 1):.1f} nodes {wk['any_tris'] / max(st['any'], 1):.2f} tris per ray; build", ctx.build_stats if hasattr(ctx, "build_stats") else "")
    ctx.set_option("pt.count_work", 0)
for r in range(reps):
    ctx.film_clear()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ctx.render_frames(0, spp, 1, capi.VB_ACCUM_SUM)
    ctx.synchronize()
    dt = time.perf_counter() - t0
    st = ctx.render_stats()
    rays = st["closest"] + st["any"]
    prof = {k: ctx.profile_get(k) for k in ("trace_closest", "trace_any", "pt_generate", "pt_shade", "pt_shadow_resolve", "pt_film", "path_trace")}
    ctx.profile_reset()
    print("   ", {k: round(v[0], 2) for k, v in prof.items() if v[1]})
    print(f"rep {r}: {dt * 1e3:.2f} ms, {w * h * spp / dt / 1e6:.2f} Msamples/s, {rays / dt / 1e6:.1f} Mrays/s, {rays / (w * h * spp):.2f} rays/sample")

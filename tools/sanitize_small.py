"""small end-to-end pass over the round-2 kernels for compute-sanitizer (run under gpurun):
compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
from vision_b200 import Context, synth, scene_py, capi

ctx = Context(0)
rays = synth.synth_rays(1 << 16, seed=6, seed_any=7)          # >= trace.sort_min_rays: the sorted path
mesh = synth.synth_mesh(3000, seed=5)
for tl in (0, 1):
    ctx.set_option("build.two_level", tl)
    for ploc in (0, 1):
        ctx.set_option("build.ploc", ploc)
        ctx.meshes_clear()
        mid = ctx.mesh_upload_soup(mesh)
        m = np.eye(4, dtype=np.float32); m[:3, 3] = [0.3, 0.1, -0.2]
        ctx.instances_set([{"mesh_id": mid}, {"mesh_id": mid, "o2w": np.ascontiguousarray(m.T).reshape(-1)}])
        st = ctx.accel_build()
        g = ctx.trace_closest(rays); a = ctx.trace_any(dict(rays, tmax=rays["tmax_any"]))
        print("two_level", tl, "ploc", ploc, st["n_nodes"], int((g["inst"] != 0xFFFFFFFF).sum()), int(a.sum()), flush=True)
ctx.set_option("build.ploc", -1)
for tl in (0, 1):
    ctx.set_option("build.two_level", tl)
    for name in ("cbox_matte", "cbox_dispersion", "cbox_lobesets"):
        sc = scene_py.build_scene(ROOT / "scenes" / "cbox" / f"{name}.json", {"resolution": [48, 40], "spp": 1})
        ctx.upload_scene(sc)
        for sched in (0, 1):
            ctx.set_option("pt.schedule", sched)
            ctx.film_clear(); ctx.render_frames(0, 2, 1, capi.VB_ACCUM_LERP)
        ctx.set_option("pt.schedule", 0)
        gb = ctx.gbuffer(0, gradient=True)
        d = ctx.dump_path_trace(20, 20, 0)
        ctx.pass_tonemapping(capi.VB_TONEMAP_ACES); ctx.pass_gamma()
        v = ctx.film_download(capi.VB_FILM_VIEW)
        print(name, "two_level", tl, float(v.mean()), len(d) // 96, float(gb["gradient"][..., 0].mean()), flush=True)
print("done")

"""GPU experiment: time one path-traced scene with an alternative build of libvision_b200 (VISION_B200_LIB=...), print Msamples/s,
the per-stage device times and a checksum of the film (identical across builds that only differ in code generation).
    VISION_B200_LIB=gpurun_variants/libvb_B.so python tools/sweep_shade.py scenes/cbox/cbox_dispersion.json 1024 64"""
import hashlib
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
from vision_b200 import Context, capi, scene_py

path, res, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
opts = dict(kv.split("=") for kv in sys.argv[4:])
sc = scene_py.build_scene(ROOT / path, {"resolution": [res, res]} if res else {})
ctx = Context(0)
for k, v in opts.items():
    ctx.set_option(k, float(v))
ctx.upload_scene(sc)
ctx.film_clear(); ctx.render_frames(0, min(spp, 4), 1, capi.VB_ACCUM_SUM); ctx.synchronize()
best = 1e30
for rep in range(3):
    ctx.film_clear(); ctx.profile_reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    ctx.render_frames(0, spp, 1, capi.VB_ACCUM_SUM); ctx.synchronize()
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
film = ctx.film_download(capi.VB_FILM_ACCUM)
w, h = sc.resolution
stages = {k: round(ctx.profile_get(k)[0], 2) for k in ("trace_closest", "trace_any", "pt_shade", "pt_generate", "pt_shadow_resolve", "pt_film")}
print(json.dumps({"lib": str(capi.LIB_PATH.name), "scene": path, "opts": opts, "ms": round(best, 2), "msamples_s": round(w * h * spp / best / 1e3, 1),
                  "stage_ms_last": stages, "film_sha": hashlib.sha256(film.tobytes()).hexdigest()[:16]}))

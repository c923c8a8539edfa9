"""BVH-quality sweep (run under gpurun): builder (radix tree / PLOC) x collapse rule (greedy / cost-optimal) on the microbench soups and
on the path-tracer scenes. Per variant: build ms, wide nodes, nodes / triangles per ray (instrumented, never timed), Mrays/s or Msamples/s.
One JSON line per variant on stdout.  usage: sweep_bvh.py [soup1m] [soup10m] [classroom] [dragon] [--opt k=v ...]"""
import json, sys, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from vision_b200 import Context, synth, capi, host_api

args = [a for a in sys.argv[1:] if not a.startswith("--")]
extra = [a for a in sys.argv[1:] if a.startswith("--opt=")]
which = args or ["soup1m", "soup10m", "classroom", "dragon"]
VARIANTS = [(p, c) for p in (0, 1) for c in (0, 1)]


def apply_extra(ctx):
    for e in extra:
        k, v = e[6:].split("=")
        ctx.set_option(k, float(v))


def soup(tris, n=1 << 24):
    ctx = Context(0)
    mesh = synth.synth_mesh(tris, seed=1)
    rays = synth.synth_rays(n)
    mid = ctx.mesh_upload_soup(mesh); ctx.instances_set([{"mesh_id": mid}])
    d = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
    da = dict(d, tmax=d["tmax_any"])
    apply_extra(ctx)
    ref = None
    for ploc, col in VARIANTS:
        ctx.set_option("build.ploc", ploc); ctx.set_option("build.collapse", col)
        st = ctx.accel_build()
        hits = occ = None
        for _ in range(2):
            hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
        ctx.profile_reset()
        for _ in range(3):
            hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
        c, _ = ctx.profile_get("trace_closest"); a, _ = ctx.profile_get("trace_any")
        _, nn, nt = ctx.trace_closest(d, hits, counted=True)
        _, an, at = ctx.trace_any(da, occ, counted=True)
        ctx.synchronize()
        sig = (hits["prim"].cpu().numpy().tobytes(), hits["t"].cpu().numpy().tobytes(), occ.cpu().numpy().tobytes())
        if ref is None:
            ref = sig
        print(json.dumps({"scene": f"soup{tris}", "ploc": ploc, "collapse": col, "build_ms": st["build_ms"], "nodes": st["n_nodes"],
                          "closest_ms": c / 3, "any_ms": a / 3, "mrays_s": 2 * n / ((c + a) / 3) / 1e3,
                          "closest_mrays_s": n / (c / 3) / 1e3, "any_mrays_s": n / (a / 3) / 1e3,
                          "closest_nodes_per_ray": nn / n, "closest_tris_per_ray": nt / n, "any_nodes_per_ray": an / n, "any_tris_per_ray": at / n,
                          "same_results_as_first": sig == ref}), flush=True)
    ctx.close()


def scene(name, path, ov, spp):
    hp = host_api.HostPipeline(ROOT / path, ov)
    hp.prepare(0)
    ctx = hp.context()
    apply_extra(ctx)
    info = hp.info()
    w, h = info["width"], info["height"]
    ref = None
    for ploc, col in VARIANTS:
        ctx.set_option("build.ploc", ploc); ctx.set_option("build.collapse", col)
        st = ctx.accel_build()
        hp.invalidate(); hp.render(2); ctx.synchronize()
        ctx.profile_reset()
        hp.invalidate()
        t0 = time.perf_counter(); hp.render(spp); ctx.synchronize(); dt = time.perf_counter() - t0
        stage = {k: ctx.profile_get(k)[0] for k in ("trace_closest", "trace_any", "pt_shade")}
        film = hp.accum().tobytes()
        if ref is None:
            ref = film
        ctx.set_option("pt.count_work", 1)
        ctx.film_clear(); ctx.render_frames(0, 1, 1, capi.VB_ACCUM_SUM); ctx.synchronize()
        s1, wk = ctx.render_stats(), ctx.render_work()
        ctx.set_option("pt.count_work", 0)
        print(json.dumps({"scene": name, "ploc": ploc, "collapse": col, "build_ms": st["build_ms"], "nodes": st["n_nodes"],
                          "msamples_s": w * h * spp / dt / 1e6, "stage_ms": stage,
                          "closest_nodes_per_ray": wk["closest_nodes"] / max(s1["closest"], 1), "closest_tris_per_ray": wk["closest_tris"] / max(s1["closest"], 1),
                          "any_nodes_per_ray": wk["any_nodes"] / max(s1["any"], 1), "any_tris_per_ray": wk["any_tris"] / max(s1["any"], 1),
                          "same_film_as_first": film == ref}), flush=True)
    hp.close()


for wname in which:
    if wname == "soup1m":
        soup(1_000_000)
    elif wname == "soup10m":
        soup(10_000_000)
    elif wname == "classroom":
        scene("classroom", "scenes/classroom/vision_scene.json", {"resolution": [1920, 1080]}, 16)
    elif wname == "dragon":
        scene("dragon", "scenes/cbox_dragon/vision_scene.json", {}, 16)

"""Two-level vs flattened structure (run under gpurun): (a) M(1M) as one instance - the pure cost of the simpler two-level kernel;
(b) 4096 instances of a 20 k-triangle mesh (82 M instanced triangles) - memory, build time, Mrays/s."""
import sys, time
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200 import Context, synth

n = 1 << 22
rays = synth.synth_rays(n)
d = {k: torch.from_numpy(v).cuda() for k, v in rays.items()}
da = dict(d, tmax=d["tmax_any"])
def measure(ctx, tag):
    hits = occ = None
    for _ in range(2):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    ctx.profile_reset()
    for _ in range(3):
        hits = ctx.trace_closest(d, hits); occ = ctx.trace_any(da, occ)
    c, _ = ctx.profile_get("trace_closest"); a, _ = ctx.profile_get("trace_any")
    ctx.synchronize()
    print(f"  {tag}: closest {n/(c/3)/1e3:.0f} any {n/(a/3)/1e3:.0f} Mrays/s", flush=True)
    return hits
ctx = Context(0)
mid = ctx.mesh_upload_soup(synth.synth_mesh(1_000_000, seed=1)); ctx.instances_set([{"mesh_id": mid}])
for tl in (0, 1):
    ctx.set_option("build.two_level", tl); st = ctx.accel_build()
    print(f"M(1M) one instance, two_level={tl}: build {st['build_ms']:.1f} ms, {(st['node_bytes'] + st['tri_bytes'])/1e6:.0f} MB")
    measure(ctx, "trace")
ctx.meshes_clear()
rng = np.random.default_rng(3)
blob = synth.synth_mesh(20_000, seed=9) * np.float32(0.03)
mid = ctx.mesh_upload_soup(np.ascontiguousarray(blob))
inst = []
for i in range(4096):
    a = rng.random() * 2 * np.pi
    m = np.eye(4); m[:3, :3] = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]]) * rng.uniform(0.5, 1.5)
    m[:3, 3] = rng.uniform(-0.95, 0.95, 3)
    inst.append({"mesh_id": mid, "o2w": np.ascontiguousarray(m.T, np.float32).reshape(-1)})
ctx.instances_set(inst)
res = {}
for tl in (1, 0):
    ctx.set_option("build.two_level", tl)
    torch.cuda.synchronize(); t0 = time.perf_counter(); st = ctx.accel_build(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"4096 x 20 k triangles, two_level={tl}: build {st['build_ms']:.1f} ms (wall {1e3*dt:.0f}), {st['n_triangles']} stored triangles, {(st['node_bytes'] + st['tri_bytes'])/1e6:.1f} MB")
    res[tl] = measure(ctx, "trace")
print("identical results:", all(torch.equal(res[0][k], res[1][k]) for k in res[0]))

"""pt.max_paths / trace.blocks_per_sm on the path-tracer configs (run under gpurun)"""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from vision_b200 import host_api
for scene, ov, spp in (("scenes/classroom/vision_scene.json", {"resolution": [1920, 1080]}, 64), ("scenes/cbox/cbox_dispersion.json", {"resolution": [1024, 1024]}, 128)):
    hp = host_api.HostPipeline(ROOT / scene, ov); hp.prepare(0)
    c = hp.context(); i = hp.info()
    for mp, bps in ((1 << 24, 8), (1 << 25, 8), (1 << 26, 8), (1 << 25, 6), (1 << 25, 12), (1 << 25, 16)):
        c.set_option("pt.max_paths", mp); c.set_option("trace.blocks_per_sm", bps)
        hp.invalidate(); hp.render(spp); c.synchronize()
        c.profile_reset(); hp.invalidate()
        t0 = time.perf_counter(); hp.render(spp); c.synchronize(); dt = time.perf_counter() - t0
        st = {k: round(c.profile_get(k)[0], 1) for k in ("trace_closest", "trace_any", "pt_shade")}
        print(f"{Path(scene).parent.name}/{Path(scene).stem} max_paths=2^{mp.bit_length()-1} bps={bps}: {i['width']*i['height']*spp/dt/1e6:.1f} Msamples/s {st}", flush=True)
    hp.close()

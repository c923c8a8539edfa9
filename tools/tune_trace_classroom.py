"""trace.refill_thr x trace.tri_thr on classroom 1920x1080 (run under gpurun)"""
import sys, time, itertools
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from vision_b200 import host_api
hp = host_api.HostPipeline(ROOT / "scenes/classroom/vision_scene.json", {"resolution": [1920, 1080]}); hp.prepare(0)
c = hp.context()
for refill, tri in itertools.product((20, 24, 28), (6, 8, 12, 16)):
    c.set_option("trace.refill_thr", refill); c.set_option("trace.tri_thr", tri)
    hp.invalidate(); hp.render(16); c.synchronize()
    c.profile_reset(); hp.invalidate()
    t0 = time.perf_counter(); hp.render(32); c.synchronize(); dt = time.perf_counter() - t0
    st = {k: round(c.profile_get(k)[0], 1) for k in ("trace_closest", "trace_any", "pt_shade")}
    print(f"classroom refill={refill} tri={tri}: {1920*1080*32/dt/1e6:.1f} Msamples/s {st}", flush=True)

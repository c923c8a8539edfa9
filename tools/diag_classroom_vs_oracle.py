"""classroom 1920x1080 x 2 spp through the C++ host vs the oracle over the same records: differing pixels and, for a few of them, the
first differing vb_dump_path_trace record (run under gpurun)."""
import ctypes as C, sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from vision_b200 import capi, host_api
import oracle_lib

res = [int(sys.argv[1]), int(sys.argv[2])] if len(sys.argv) > 2 else [1920, 1080]
hp = host_api.HostPipeline(ROOT / "scenes/classroom/vision_scene.json", {"resolution": res}); hp.prepare(0)
hp.render(2)
film = hp.accum()
oracle = oracle_lib.load()
osc = oracle_lib.OracleScene(oracle, hp.records(), threads=oracle.hw_threads())
ofilm, _, _ = osc.render(0, 2, threads=oracle.hw_threads(), stats=True)
diff = np.argwhere((film != ofilm).any(-1))
print(f"{len(diff)} of {film.shape[0] * film.shape[1]} pixels differ, max abs {np.abs(film - ofilm).max():.3e}")
ctx = hp.context(); ctx._res = (res[0], res[1])
n = C.sizeof(capi.vb_path_vertex)
shown = 0
for (py, px) in diff[:: max(1, len(diff) // 12)]:
    for f in range(2):
        g = ctx.dump_path_trace(int(px), int(py), f, 1024); t = oracle_lib.oracle_dump_path_trace(osc, int(px), int(py), f, 1024)
        if g == t:
            continue
        ra = (capi.vb_path_vertex * (len(g) // n)).from_buffer_copy(g); rb = (capi.vb_path_vertex * (len(t) // n)).from_buffer_copy(t)
        print(f" pixel ({px},{py}) frame {f}: {len(ra)} vs {len(rb)} records")
        for i, (x, y) in enumerate(zip(ra, rb)):
            if bytes(x) != bytes(y):
                print(f"   record {i} differs (inst {y.inst} prim {y.prim} bounces {y.bounces}; previous record: inst {rb[i-1].inst} prim {rb[i-1].prim} medium {rb[i-1].medium})")
                for name, _ in capi.vb_path_vertex._fields_:
                    vx, vy = getattr(x, name), getattr(y, name)
                    vx = list(vx) if hasattr(vx, "__len__") else vx; vy = list(vy) if hasattr(vy, "__len__") else vy
                    if vx != vy: print(f"      {name}: gpu {vx}  oracle {vy}")
                break
        shown += 1
        break
    if shown >= 8: break

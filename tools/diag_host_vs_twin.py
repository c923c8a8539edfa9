"""Where does a film rendered through the C++ host layer differ from the Python twin's? (run under gpurun)
Renders one frame both ways, counts differing pixels and, for the first one, prints the first differing vb_dump_path_trace record."""
import ctypes as C, sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from vision_b200 import Context, scene_py, capi, host_api

scene = ROOT / (sys.argv[1] if len(sys.argv) > 1 else "scenes/cbox/cbox_matte.json")
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 1
hp = host_api.HostPipeline(scene); hp.prepare(0); hp.render(frames)
a = hp.accum()
sc = scene_py.build_scene(scene)
ctx = Context(0); ctx.upload_scene(sc); ctx.film_clear(); ctx.render_frames(0, frames, 1, capi.VB_ACCUM_LERP)
b = ctx.film_download(capi.VB_FILM_ACCUM)
diff = np.argwhere((a != b).any(-1))
print(f"{scene.name}: {len(diff)} of {a.shape[0] * a.shape[1]} pixels differ, max abs {np.abs(a - b).max():.3e}")
hctx = hp.context()
hctx._res = (a.shape[1], a.shape[0])
for (py, px) in diff[:3]:
    for f in range(frames):
        g = hctx.dump_path_trace(int(px), int(py), f); t = ctx.dump_path_trace(int(px), int(py), f)
        if g == t:
            continue
        n = C.sizeof(capi.vb_path_vertex)
        ra = (capi.vb_path_vertex * (len(g) // n)).from_buffer_copy(g); rb = (capi.vb_path_vertex * (len(t) // n)).from_buffer_copy(t)
        for i, (x, y) in enumerate(zip(ra, rb)):
            if bytes(x) != bytes(y):
                print(f" pixel ({px},{py}) frame {f}: record {i} differs")
                for name, _ in capi.vb_path_vertex._fields_:
                    vx, vy = getattr(x, name), getattr(y, name)
                    vx = list(vx) if hasattr(vx, "__len__") else vx; vy = list(vy) if hasattr(vy, "__len__") else vy
                    if vx != vy: print(f"    {name}: host {vx}  twin {vy}")
                break
        break

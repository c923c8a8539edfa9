"""Copies the reference's STOCK Cornell box (res/render_scene/cbox/vision_scene.json: glass short box, spot + point + spherical +
projector lights besides the ceiling quad, triangle filter, ACES tone mapper, depth 16, 1024x1024; the file also DEFINES an unused
"mix" material) into scenes/cbox_stock/ with its two images (checker.jpg, Painting3.jpg). The only edit: the environment map
textures/spruit_sunrise_2k.hdr is stripped from the reference tree, so light_sampler.lights[2].param.color.fn points at
scenes/cbox/env_small.pfm-style stand-in written here (textures/env_standin.pfm, 128x64 analytic gradient + bright patch).
Run once in the build container: python tools/import_cbox_stock.py [/root/reference]"""
import json
import shutil
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200.scene_py import load_json_with_comments

ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference") / "res" / "render_scene" / "cbox"
dst = Path(__file__).resolve().parent.parent / "scenes" / "cbox_stock"
(dst / "textures").mkdir(parents=True, exist_ok=True)
for f in ("checker.jpg", "Painting3.jpg", "LICENSE.txt"):
    if (ref / f).exists():
        shutil.copyfile(ref / f, dst / f)
w, h = 128, 64
v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
img = np.stack([0.5 + 0.4 * u, 0.5 + 0.3 * (1 - v), 0.9 - 0.4 * u * v], axis=-1).astype(np.float32)
img[10:16, 40:52] = [40.0, 32.0, 22.0]
with open(dst / "textures" / "env_standin.pfm", "wb") as f:
    f.write(b"PF\n%d %d\n-1.0\n" % (w, h))
    f.write(img[::-1].astype("<f4").tobytes())
scene = load_json_with_comments(ref / "vision_scene.json")
n = 0
for l in scene["light_sampler"]["param"]["lights"]:
    c = l["param"].get("color")
    node = c.get("node") if isinstance(c, dict) else None
    tgt = node if isinstance(node, dict) else c
    if isinstance(tgt, dict) and str(tgt.get("fn", "")).endswith(".hdr"):
        tgt["fn"] = "textures/env_standin.pfm"; n += 1
assert n == 1, n
(dst / "vision_scene.json").write_text(json.dumps(scene, indent=1))
print("cbox_stock imported; materials:", [m["type"] for m in scene["materials"]], "lights:", [l["type"] for l in scene["light_sampler"]["param"]["lights"]])

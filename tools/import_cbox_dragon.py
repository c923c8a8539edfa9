"""Writes config C3 into scenes/cbox_dragon/ from the reference's res/cbox-dragon.json:
  * vision_scene.json       : the file as shipped (glass model filled with the "fog" medium = Skin1 x 20, thin-lens camera,
                              box filter, max depth 16, 1024 x 1024); its "dragon.obj" resolves to the stand-in that
                              tools/make_dragon_standin.py writes next to it (the Stanford dragon is not in the reference tree)
  * vision_scene_small.json : the same scene over dragon_small.obj (4.5 k triangles) at 96 x 96 - the parity-test size
Run once in the build container: python tools/import_cbox_dragon.py [/root/reference]"""
import copy
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200.scene_py import load_json_with_comments

ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference") / "res" / "cbox-dragon.json"
dst = Path(__file__).resolve().parent.parent / "scenes" / "cbox_dragon"
dst.mkdir(parents=True, exist_ok=True)
scene = load_json_with_comments(ref)
(dst / "vision_scene.json").write_text(json.dumps(scene, indent=1))
small = copy.deepcopy(scene)
for s in small["shapes"]:
    if s["type"] == "model":
        s["param"]["fn"] = "dragon_small.obj"
small["pipeline"]["param"]["frame_buffer"]["param"]["resolution"] = [96, 96]
small["camera"]["param"]["radiance_collector"]["param"]["resolution"] = [96, 96]
small["output"]["spp"] = 4
(dst / "vision_scene_small.json").write_text(json.dumps(small, indent=1))
print("wrote", sorted(p.name for p in dst.glob("*.json")))

"""Reference-held parity pins: the renders the reference ships in its own tree are the only outputs of the reference's
implementation available offline (SURVEY.md F8, section 8c "look-alike images"). This tool

  1. copies them BYTE FOR BYTE into tests/golden/reference_renders/ (they are data, like the scene files under scenes/) and records
     their sha256 in tests/golden/reference_renders/MANIFEST.json:
        res/render_scene/cbox/srgb.png                   -> cbox_srgb.png        (1024x1024, stock Cornell box, srgb spectrum)
        res/render_scene/classroom/output-1024spp.png    -> classroom_1024spp.png (1280x720, the pipeline block's resolution)
  2. writes scenes/cbox_stock/lookalike.json: the stock scene file with the state it evidently had when srgb.png was rendered
     (the image shows ONLY the ceiling quad light, a plain back wall, a glass tall box and a mirror short box, and its wall
     radiances match `1 - exp(-L)` + sRGB OETF with NO tone curve): free lights removed, BackWall = the other walls' white,
     shape_5 -> "Mirror2", shape_6 -> "Glass", tone mapper "linear". Nothing else is touched (camera, geometry, light, depth).

tests/test_reference_pins.py (oracle, CPU) and tests/test_reference_pins_gpu.py (CUDA path through the C++ host) compare against
these files; nothing reads /root/reference at test time.
Run once in the build container: python tools/import_reference_renders.py [/root/reference]"""
import hashlib
import json
import shutil
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from vision_b200.scene_py import load_json_with_comments

RENDERS = {
    "cbox_srgb.png": "res/render_scene/cbox/srgb.png",
    "classroom_1024spp.png": "res/render_scene/classroom/output-1024spp.png",
}


def main():
    ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    dst = ROOT / "tests" / "golden" / "reference_renders"
    dst.mkdir(parents=True, exist_ok=True)
    manifest = {}
    for name, rel in RENDERS.items():
        shutil.copyfile(ref / rel, dst / name)
        manifest[name] = {"reference_path": rel, "sha256": hashlib.sha256((dst / name).read_bytes()).hexdigest()}
    for lic in ("res/render_scene/cbox/LICENSE.txt", "res/render_scene/classroom/LICENSE.txt"):
        shutil.copyfile(ref / lic, dst / ("LICENSE_" + lic.split("/")[2] + ".txt"))
    (dst / "MANIFEST.json").write_text(json.dumps(manifest, indent=1) + "\n")

    stock = ROOT / "scenes" / "cbox_stock" / "vision_scene.json"
    d = load_json_with_comments(stock)
    d["light_sampler"]["param"]["lights"] = []
    for m in d["materials"]:
        if m.get("name") == "BackWall":
            m["param"]["color"] = [0.725, 0.71, 0.68]
    for s in d["shapes"]:
        if s.get("name") == "shape_5":
            s["param"]["material"] = "Mirror2"
        if s.get("name") == "shape_6":
            s["param"]["material"] = "Glass"

    def set_linear(node):
        if isinstance(node, dict):
            tm = node.get("tone_mapper")
            if isinstance(tm, dict):
                tm["type"] = "linear"
            for v in node.values():
                set_linear(v)
        elif isinstance(node, list):
            for v in node:
                set_linear(v)
    set_linear(d)
    (ROOT / "scenes" / "cbox_stock" / "lookalike.json").write_text(json.dumps(d, indent=1) + "\n")
    print("imported", list(manifest), "and wrote scenes/cbox_stock/lookalike.json")


if __name__ == "__main__":
    main()

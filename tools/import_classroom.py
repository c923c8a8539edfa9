"""Copies the classroom scene data (Bitterli 'Japanese Classroom', CC-BY, see LICENSE.txt) from the reference
checkout into scenes/classroom/ and writes the derived scene file used as config C5:
  * vision_scene.json   : the reference's res/render_scene/classroom/vision_scene.json with
      - light_sampler.lights[0].param.color.fn -> textures/sky_standin.pfm (the reference's
        spaichingen_hill_2k.exr is stripped from the tree: SURVEY.md F9 / A3),
      - pipeline.param.frame_buffer.param.resolution -> [1920, 1080] (config C5; the file ships 1280x720)
  * textures/sky_standin.pfm : analytic sun-and-sky lat-long map, 512x256, fixed formula below
Run once in the build container: python tools/import_classroom.py [/root/reference]"""
import json
import shutil
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200.scene_py import load_json_with_comments

ref = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference") / "res" / "render_scene" / "classroom"
dst = Path(__file__).resolve().parent.parent / "scenes" / "classroom"
(dst / "models").mkdir(parents=True, exist_ok=True)
(dst / "textures").mkdir(parents=True, exist_ok=True)
for f in sorted((ref / "models").glob("*.obj")):
    shutil.copyfile(f, dst / "models" / f.name)
for f in sorted((ref / "textures").glob("*.jpg")):
    shutil.copyfile(f, dst / "textures" / f.name)
shutil.copyfile(ref / "LICENSE.txt", dst / "LICENSE.txt")

# ---- stand-in environment map: lat-long, row = theta (0 at +z of the light's local frame), col = phi ----
W, H = 512, 256
v = (np.arange(H) + 0.5) / H
u = (np.arange(W) + 0.5) / W
theta = (v * np.pi)[:, None]
phi = (u * 2 * np.pi)[None, :]
d = np.stack([np.sin(theta) * np.cos(phi), np.sin(theta) * np.sin(phi), np.cos(theta) * np.ones_like(phi)], -1)
up = d[..., 2]                                           # local +z is "up" after SphericalMap's rotation_x(-90)
sky = np.array([0.35, 0.55, 1.0]) * (0.25 + 0.75 * np.clip(up, 0, 1)[..., None] ** 0.5) * 0.02
horizon = np.array([1.0, 0.9, 0.75]) * (np.exp(-8 * np.abs(up))[..., None]) * 0.01
ground = np.array([0.25, 0.22, 0.2]) * (np.clip(-up, 0, 1)[..., None]) * 0.004
sun_dir = np.array([np.cos(np.radians(35)) * np.cos(np.radians(200)), np.cos(np.radians(35)) * np.sin(np.radians(200)), np.sin(np.radians(35))])
cosang = np.clip((d * sun_dir).sum(-1), -1, 1)
sun = np.array([1.0, 0.93, 0.82]) * (np.exp((cosang - 1) * 900.0)[..., None]) * 6.0
env = (sky + horizon + ground + sun).astype(np.float32)
with open(dst / "textures" / "sky_standin.pfm", "wb") as f:
    f.write(f"PF\n{W} {H}\n-1.0\n".encode())
    f.write(np.ascontiguousarray(env[::-1].astype("<f4")).tobytes())

scene = load_json_with_comments(ref / "vision_scene.json")
scene["light_sampler"]["param"]["lights"][0]["param"]["color"]["fn"] = "textures/sky_standin.pfm"
scene["pipeline"]["param"]["frame_buffer"]["param"]["resolution"] = [1920, 1080]
(dst / "vision_scene.json").write_text(json.dumps(scene, indent=1))
print("classroom imported:", len(list((dst / "models").glob("*.obj"))), "meshes; sky mean", env.mean(axis=(0, 1)), "max", env.max())

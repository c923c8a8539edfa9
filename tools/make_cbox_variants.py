"""Writes the derived cbox test scenes from scenes/cbox/cbox_matte.json (itself the matte variant of the
reference's res/render_scene/cbox/vision_scene.json). Material parameter blocks are the ones the
reference scene files use (cbox/vision_scene.json materials Glass2, Mirror, Metal, substrate;
cbox-dragon.json medium "fog").
  cbox_materials.json : short box = glass (BK7), tall box = mirror (roughness 0.3), floor = substrate,
                        right wall = metal (Cu)
  cbox_fog.json       : short box = glass boundary filled with the Skin1 medium (scale 20), global thin fog,
                        mediums.process = true
  cbox_dispersion.json: hero-wavelength spectrum (hero.cpp) + dispersive BK7 glass + principled + spectral metal (config C4)
  cbox_projector.json : adds the projector light of the reference's cbox scene (image = checker.jpg), area light at 20 %
  cbox_lights.json    : the ceiling quad keeps its emission; adds a point light, a spot light (the reference's
                        commented examples in point.cpp:11-20 / spot.cpp:10-18) and a spherical environment
                        map (spherical.cpp:12-25) over scenes/cbox/env_small.pfm (64x32 analytic gradient with a
                        bright patch, written here); the back wall is removed so that camera rays can miss
"""
import copy
import json
from pathlib import Path
import sys
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vision_b200.scene_py import load_json_with_comments

root = Path(__file__).resolve().parent.parent / "scenes" / "cbox"
base = load_json_with_comments(root / "cbox_matte.json")

extra = [
    {"type": "glass", "name": "Glass2", "param": {"material_name": "BK7", "color": {"channels": "xyz", "node": [1, 1, 1]}, "ior": 1.5, "roughness": [0.001, 0.001]}},
    {"type": "mirror", "name": "Mirror", "param": {"color": {"channels": "xyz", "node": [1, 1, 1]}, "roughness": 0.3}},
    {"type": "metal", "name": "Metal", "param": {"material_name": "Cu", "roughness": [0.001, 0.001]}},
    {"type": "substrate", "name": "substrate", "param": {"roughness": 0.001, "spec": [0.04, 0.04, 0.04], "color": [0.725, 0.71, 0.68]}},
]

def assign(d, mapping):
    for s in d["shapes"]:
        if s["name"] in mapping:
            s["param"]["material"] = mapping[s["name"]]

a = copy.deepcopy(base)
a["materials"] += extra
assign(a, {"shape_5": "Glass2", "shape_6": "Mirror", "shape_0": "substrate", "shape_4": "Metal"})
a["integrator"]["param"]["max_depth"] = 8
(root / "cbox_materials.json").write_text(json.dumps(a, indent=1))

b = copy.deepcopy(base)
b["materials"] += extra[:1]
assign(b, {"shape_5": "Glass2"})
b["mediums"] = {"process": True, "global": "air", "list": [
    {"type": "homogeneous", "name": "fog", "param": {"g": 0, "medium_name": "Skin1", "scale": 20}},
    {"type": "homogeneous", "name": "air", "param": {"g": 0.3, "scale": 0.25, "sigma_s": [1, 1, 1], "sigma_a": [0, 0, 0]}}]}
for s in b["shapes"]:
    if s["name"] == "shape_5":
        s["param"]["medium"] = {"inside": "fog", "outside": "air"}
b["integrator"]["param"]["max_depth"] = 8
(root / "cbox_fog.json").write_text(json.dumps(b, indent=1))

import numpy as np
c = copy.deepcopy(base)
c["shapes"] = [s for s in c["shapes"] if s["name"] != "shape_2"]
c["light_sampler"]["param"]["lights"] = [
    {"type": "point", "param": {"color": {"channels": "xyz", "node": [1, 0.8, 0.6]}, "position": [-0.5, 1.8, 0], "scale": 0.2}},
    {"type": "spot", "param": {"angle": 20, "falloff": 2, "direction": [0.2, -1, 0], "color": [0.4, 0.6, 2], "position": [-0.5, 1.9, 0], "scale": 1}},
    {"type": "spherical", "param": {"flip_u": True, "color": {"fn": "env_small.pfm", "color_space": "linear"},
                                     "o2w": {"type": "Euler", "param": {"yaw": -30}}, "scale": 0.5}},
]
w, h = 64, 32
v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
img = np.stack([0.3 + 0.5 * u, 0.4 + 0.4 * (1 - v), 0.8 - 0.5 * u * v], axis=-1).astype(np.float32)
img[6:10, 20:26] = [30.0, 25.0, 18.0]
with open(root / "env_small.pfm", "wb") as f:
    f.write(b"PF\n%d %d\n-1.0\n" % (w, h))
    f.write(img[::-1].astype("<f4").tobytes())
(root / "cbox_lights.json").write_text(json.dumps(c, indent=1))

# cbox_sphere.json: the short box becomes a glass sphere (sphere.cpp tessellation, sub_div 24), gaussian pixel filter
e = copy.deepcopy(base)
e["materials"] += extra[:1]
for i, s_ in enumerate(e["shapes"]):
    if s_["name"] == "shape_5":
        e["shapes"][i] = {"type": "sphere", "name": "shape_5", "param": {"radius": 0.3, "sub_div": 24, "material": "Glass2",
                          "transform": {"type": "matrix4x4", "param": {"matrix4x4": [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0.33, 0.3, 0.35, 1]]}}}}
e["camera"]["param"]["filter"] = {"type": "gaussian", "name": "gaussianFilter", "param": {"radius": 1.5, "sigma": 0.5}}
e["integrator"]["param"]["max_depth"] = 8
(root / "cbox_sphere.json").write_text(json.dumps(e, indent=1))

# cbox_principled.json: PrincipledBSDF (principled_bsdf.cpp) on every lobe combination: coated metal (tall box),
# transmissive (short box), sheen cloth (floor), plain dielectric plastic with a textured colour (back wall), rough
# tinted specular (right wall)
g = copy.deepcopy(base)
back_color = next(m for m in g["materials"] if m["name"] == next(s_["param"]["material"] for s_ in g["shapes"] if s_["name"] == "shape_2"))["param"]["color"]
g["materials"] += [
    {"type": "principled_bsdf", "name": "P_metal", "param": {"color": [0.95, 0.64, 0.54], "metallic": 1.0, "roughness": 0.35, "spec_tint": [0.9, 0.8, 0.7],
                                                              "coat_weight": 0.8, "coat_roughness": 0.1, "coat_ior": 1.6, "coat_tint": [1, 0.9, 0.8], "anisotropic": 0.5}},
    {"type": "principled_bsdf", "name": "P_trans", "param": {"color": [0.8, 0.95, 0.9], "transmission_weight": 0.9, "roughness": 0.15, "ior": 1.45}},
    {"type": "principled_bsdf", "name": "P_sheen", "param": {"color": [0.3, 0.1, 0.4], "sheen_weight": 1.0, "sheen_roughness": 0.4, "sheen_tint": [1, 0.8, 0.9], "roughness": 0.8}},
    {"type": "principled_bsdf", "name": "P_plastic", "param": {"color": back_color, "roughness": 0.3, "ior": 1.5}},
    {"type": "principled_bsdf", "name": "P_mix", "param": {"color": [0.1, 0.5, 0.1], "metallic": 0.5, "roughness": 0.6, "spec_tint": [1, 0.7, 0.7], "coat_weight": 0.3,
                                                            "sheen_weight": 0.5, "transmission_weight": 0.2}},
]
assign(g, {"shape_6": "P_metal", "shape_5": "P_trans", "shape_0": "P_sheen", "shape_2": "P_plastic", "shape_4": "P_mix"})
g["integrator"]["param"]["max_depth"] = 8
(root / "cbox_principled.json").write_text(json.dumps(g, indent=1))
# cbox_projector.json: the reference's projector light (projector.cpp:11-31 example block, as in cbox/vision_scene.json) over
# the checker image, dimmed area light so that the projected image dominates the back wall
pj = copy.deepcopy(base)
pj["light_sampler"]["param"]["lights"] = [
    {"type": "projector", "param": {"angle": 6, "color": {"channels": "xyz", "node": {"fn": "checker.jpg", "color_space": "srgb"}}, "ratio": 1,
                                    "o2w": {"type": "look_at", "param": {"position": [0, 1, 6.8], "up": [0, 1, 0], "target_pos": [0, 1, 0]}}, "scale": 300}}]
for s_ in pj["shapes"]:
    if "emission" in s_["param"]:
        s_["param"]["emission"]["param"]["scale"] = 0.2
(root / "cbox_projector.json").write_text(json.dumps(pj, indent=1))
# cbox_lights_power.json: cbox_lights + the projector under the "power" light sampler (lightsampler/power.cpp): every light type's power()
pw = copy.deepcopy(c)
pw["light_sampler"]["type"] = "power"
pw["light_sampler"]["param"]["lights"] = pw["light_sampler"]["param"]["lights"] + copy.deepcopy(pj["light_sampler"]["param"]["lights"])
(root / "cbox_lights_power.json").write_text(json.dumps(pw, indent=1))
# cbox_dispersion.json (config C4 "cbox-glass / dispersion spectral glass + Disney"): hero-wavelength spectrum, the short box is
# BK7 glass (material_name -> dispersive "spd" ior slot, glass.cpp:216-234), the tall box the coated principled metal, metal (Cu,
# spectral eta / k) right wall, substrate floor
ds = copy.deepcopy(a)
ds["materials"] += [m for m in g["materials"] if m["name"] == "P_metal"]
assign(ds, {"shape_6": "P_metal"})
ds["spectrum"] = {"type": "hero", "param": {"dimension": 3}}
ds["integrator"]["param"]["max_depth"] = 16
(root / "cbox_dispersion.json").write_text(json.dumps(ds, indent=1))
print("wrote", root / "cbox_dispersion.json", root / "cbox_projector.json", root / "cbox_materials.json", root / "cbox_fog.json", root / "cbox_lights.json", root / "cbox_sphere.json", root / "cbox_principled.json")

# cbox_lobesets.json: plastic and metallic single-lobe materials, and the two LobeSet materials over child
# materials - mix (diffuse + mirror; glass + principled with every lobe switched on) and add (diffuse + metallic). Parameter blocks
# follow the reference's res/render_scene/cbox/cbox.json ("plastic", "metallic", "mix", "mix2").
g = copy.deepcopy(base)
principled = {"type": "principled_bsdf", "name": "pr", "param": {"color": [0.8, 0.5, 0.3], "ior": 1.5, "roughness": 0.3, "metallic": 0.4, "spec_tint": [1, 0.9, 0.8],
              "sheen_weight": 0.5, "sheen_roughness": 0.4, "sheen_tint": [0.9, 0.9, 1], "coat_weight": 0.6, "coat_roughness": 0.1, "coat_ior": 1.6,
              "coat_tint": [1, 1, 1], "transmission_weight": 0.3}}
g["materials"] += [
    {"type": "plastic", "name": "plastic", "param": {"color": [0.2, 0.4, 0.8], "spec": [0.9, 0.9, 0.9], "ior": 1.5, "roughness": 0.3}},
    {"type": "metallic", "name": "metallic", "param": {"color": [0.95, 0.64, 0.54], "edge_tint": [0.98, 0.82, 0.76], "roughness": 0.25, "anisotropic": 0.3}},
    {"type": "mix", "name": "mix_dm", "param": {"frac": 0.35,
        "mat0": {"type": "diffuse", "name": "d0", "param": {"color": {"channels": "xyz", "node": {"type": "image", "param": {"fn": "checker.jpg", "color_space": "srgb"}}}}},
        "mat1": {"type": "mirror", "name": "m1", "param": {"color": [0.9, 0.9, 0.9], "roughness": 0.1}}}},
    {"type": "mix", "name": "mix_gp", "param": {"frac": 0.6,
        "mat0": {"type": "glass", "name": "g0", "param": {"material_name": "BK7", "color": [1, 1, 1], "roughness": 0.05}},
        "mat1": principled}},
    {"type": "add", "name": "add_dm", "param": {
        "mat0": {"type": "diffuse", "name": "d2", "param": {"color": [0.3, 0.3, 0.3]}},
        "mat1": {"type": "metallic", "name": "m2", "param": {"color": [0.6, 0.6, 0.2], "edge_tint": [1, 1, 1], "roughness": 0.4}}}},
]
assign(g, {"shape_5": "mix_gp", "shape_6": "mix_dm", "shape_0": "plastic", "shape_4": "metallic", "shape_2": "add_dm"})
g["integrator"]["param"]["max_depth"] = 8
(root / "cbox_lobesets.json").write_text(json.dumps(g, indent=1))

# cbox_directional.json: the directional light (environments/directional.cpp:12-19 example block) shining in through the
# open front; the back wall is removed (as in cbox_lights) so that paths also MISS into it - it is the scene's environment light.
# cbox_env_separate*.json: LightSampler's env_separate / env_prob (lightsampler.cpp:11-13,159-197) over the uniform and the power rule,
# with the spherical map, and over the directional light.
dr = copy.deepcopy(c)
dr["light_sampler"]["param"]["lights"] = [
    {"type": "point", "param": {"color": {"channels": "xyz", "node": [1, 0.8, 0.6]}, "position": [-0.5, 1.8, 0], "scale": 0.2}},
    {"type": "directional", "param": {"color": {"channels": "xyz", "node": [17, 12, 4]}, "scale": 0.1, "direction": [-0.35, -0.6, -1]}},
]
(root / "cbox_directional.json").write_text(json.dumps(dr, indent=1))
es = copy.deepcopy(c)
es["light_sampler"]["param"].update({"env_separate": True, "env_prob": 0.3})
(root / "cbox_env_separate.json").write_text(json.dumps(es, indent=1))
esp = copy.deepcopy(pw)
esp["light_sampler"]["param"].update({"env_separate": True, "env_prob": 0.7})
(root / "cbox_env_separate_power.json").write_text(json.dumps(esp, indent=1))
esd = copy.deepcopy(dr)
esd["light_sampler"]["param"].update({"env_separate": True})          # env_prob default 0.5
(root / "cbox_env_separate_directional.json").write_text(json.dumps(esd, indent=1))

# cbox_customized.json: the matte box rendered by the `customized` pipeline (render graph integrate -> accumulate ->
# tonemapping -> gamma, src/pipelines/customized/pipeline.cpp:28-44) at 128x128
cu = copy.deepcopy(base)
cu["pipeline"]["type"] = "customized"
cu["pipeline"]["param"]["frame_buffer"]["param"]["resolution"] = [128, 128]
(root / "cbox_customized.json").write_text(json.dumps(cu, indent=1))

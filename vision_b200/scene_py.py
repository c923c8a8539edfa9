"""Python twin of the host-side plugin logic for tests: builds vision_b200.capi.SceneRecords for the
cbox family (quad / cube shapes, diffuse materials, area light, thin-lens camera) directly from a
vision_scene.json. The production host layer is C++ (vision_b200/host); tests compare the two.

Reference logic restated: quad.cpp:21-50, cube.cpp:20-71, alias.h:86-122, sensor.cpp:58-78,153-162,
frame_buffer.cpp:93-102, light.cpp:19-24 + number.cpp:40-49, scene.cpp:118-151, node_desc.cpp:31-58."""
from __future__ import annotations

import json
import math
import re
from pathlib import Path

import numpy as np

from . import capi
from .capi import SceneRecords, make_slot, VB_INVALID_ID

f32 = np.float32


def load_json_with_comments(path) -> dict:
    s = Path(path).read_text()
    out, i, n, in_str = [], 0, len(s), False
    while i < n:                       # strips // and /* */ outside strings (reference json_util.h:12-94)
        c = s[i]
        if in_str:
            out.append(c)
            if c == "\\" and i + 1 < n:
                out.append(s[i + 1]); i += 1
            elif c == '"':
                in_str = False
        elif c == '"':
            in_str = True; out.append(c)
        elif c == "/" and i + 1 < n and s[i + 1] == "/":
            while i < n and s[i] != "\n":
                i += 1
            continue
        elif c == "/" and i + 1 < n and s[i + 1] == "*":
            i = s.index("*/", i + 2) + 1
        else:
            out.append(c)
        i += 1
    return json.loads("".join(out))


def quad_mesh(width=1.0, height=1.0):
    w, h = f32(width) / f32(2), f32(height) / f32(2)
    P = np.array([[w, 0, h], [w, 0, -h], [-w, 0, h], [-w, 0, -h]], f32)
    UV = np.array([[1, 1], [1, 0], [0, 1], [0, 0]], f32)
    ng = np.cross(P[0] - P[2], P[1] - P[2]).astype(f32)
    N = np.tile(ng, (4, 1)).astype(f32)
    T = np.array([[0, 1, 2], [2, 1, 3]], np.uint32)
    return {"pos": P, "nrm": N, "uv": UV, "tri": T}


def cube_mesh(x=1.0, y=1.0, z=1.0):
    y = x if y == 0 else y
    z = y if z == 0 else z
    x, y, z = f32(x) / f32(2), f32(y) / f32(2), f32(z) / f32(2)
    P = np.array([[-x, -y, z], [x, -y, z], [-x, y, z], [x, y, z],
                  [-x, y, -z], [x, y, -z], [-x, -y, -z], [x, -y, -z],
                  [-x, y, z], [x, y, z], [-x, y, -z], [x, y, -z],
                  [-x, -y, z], [x, -y, z], [-x, -y, -z], [x, -y, -z],
                  [x, -y, z], [x, y, z], [x, y, -z], [x, -y, -z],
                  [-x, -y, z], [-x, y, z], [-x, y, -z], [-x, -y, -z]], f32)
    N = np.repeat(np.array([[0, 0, 1], [0, 0, -1], [0, 1, 0], [0, -1, 0], [1, 0, 0], [-1, 0, 0]], f32), 4, axis=0)
    UV = np.array([[0, 0], [1, 0], [0, 1], [1, 1], [0, 1], [1, 1], [0, 0], [1, 0], [0, 1], [1, 1], [0, 0], [1, 0],
                   [0, 1], [1, 1], [0, 0], [1, 0], [0, 1], [1, 1], [1, 0], [0, 0], [0, 1], [1, 1], [1, 0], [0, 0]], f32)
    T = np.array([[0, 1, 3], [0, 3, 2], [6, 5, 7], [4, 5, 6], [10, 9, 11], [8, 9, 10],
                  [13, 14, 15], [13, 12, 14], [18, 17, 19], [17, 16, 19], [21, 22, 23], [20, 21, 23]], np.uint32)
    return {"pos": P, "nrm": N, "uv": UV, "tri": T}


def alias_build(weights):
    """AliasTable::build (alias.h:86-122): sums in double, probabilities in float, LIFO stacks"""
    w = [float(f32(x)) for x in weights]
    n = len(w)
    s = sum(w)
    ratio = n / s
    prob = [f32(x * ratio) for x in w]
    alias = list(range(n))
    over = [i for i in range(n) if prob[i] > f32(1.0)]
    under = [i for i in range(n) if not (prob[i] > f32(1.0))]
    # the reference pushes in index order into two stacks and pops from the back
    while over and under:
        o, u = over.pop(), under.pop()
        prob[o] = f32(prob[o] - f32(f32(1.0) - prob[u]))
        alias[u] = o
        if prob[o] > f32(1.0):
            over.append(o)
        elif prob[o] < f32(1.0):
            under.append(o)
    for i in over:
        prob[i], alias[i] = f32(1.0), i
    for i in under:
        prob[i], alias[i] = f32(1.0), i
    integral = f32(s / n)
    return [(float(prob[i]), alias[i]) for i in range(n)], [float(f32(x)) for x in w], float(integral)


def triangle_areas(mesh):
    P, T = mesh["pos"].astype(f32), mesh["tri"]
    out = []
    for t in T:
        a, b, c = P[t[0]], P[t[1]], P[t[2]]
        cr = np.cross((b - a).astype(f32), (c - a).astype(f32)).astype(f32)
        out.append(f32(0.5) * f32(np.sqrt(f32(np.dot(cr, cr)))))
    return out


def look_at(pos, target, up):
    pos, target, up = (np.asarray(v, np.float64) for v in (pos, target, up))
    fwd = target - pos; fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, up); right /= np.linalg.norm(right)
    up2 = np.cross(right, fwd)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = right, up2, fwd, pos     # columns
    return m


def camera_to_world(m):
    """Sensor::update_mat + camera_to_world (sensor.cpp:73-78,153-162): yaw/pitch from the matrix, roll dropped"""
    pitch = math.degrees(math.atan2(m[2, 1], m[1, 1]))     # m[1][2] in column-major [c][r] = row 2, col 1
    yaw = math.degrees(math.atan2(m[0, 2], m[0, 0]))       # m[2][0] = row 0, col 2
    y, p = math.radians(yaw), math.radians(-pitch)
    ry = np.array([[math.cos(y), 0, math.sin(y), 0], [0, 1, 0, 0], [-math.sin(y), 0, math.cos(y), 0], [0, 0, 0, 1]])
    rx = np.array([[1, 0, 0, 0], [0, math.cos(p), -math.sin(p), 0], [0, math.sin(p), math.cos(p), 0], [0, 0, 0, 1]])
    sc = np.diag([1.0, 1.0, -1.0, 1.0])
    tr = np.eye(4); tr[:3, 3] = m[:3, 3]
    return tr @ sc @ ry @ rx


def _color_slot(v, textures, base_dir):
    """a colour attribute: [r,g,b] | {"channels","node":[..]} | {"node":{"type":"image","param":{"fn"}}} | {"type":"number"}"""
    if isinstance(v, dict):
        if "node" in v:
            return _color_slot(v["node"], textures, base_dir)
        if v.get("type") == "number":
            return _color_slot(v["param"]["value"], textures, base_dir)
        if v.get("type") == "image" or "fn" in v:
            prm = v.get("param", v)
            textures.append(load_image(Path(base_dir) / prm["fn"], prm.get("color_space", "srgb")))
            return make_slot([1, 1, 1, 1], len(textures) - 1)
    if isinstance(v, (int, float)):
        v = [v, v, v]
    return make_slot([float(x) for x in v])


def load_image(path, color_space):
    from PIL import Image
    im = np.asarray(Image.open(path).convert("RGBA"), np.float32) / f32(255.0)
    if color_space == "srgb" and Path(path).suffix.lower() not in (".exr", ".hdr"):
        c = im[..., :3]
        lin = np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4).astype(f32)
        im = np.concatenate([lin, im[..., 3:]], axis=-1)
    return np.ascontiguousarray(im, f32)


def _number(v, default):
    """a Number attribute read with as_scalar(): x | [x, y] | {"node": ...} | {"type":"number","param":{"value":..}}"""
    if v is None:
        return float(default)
    if isinstance(v, dict):
        if "node" in v:
            return _number(v["node"], default)
        if "param" in v:
            return _number(v["param"].get("value"), default)
        return float(default)
    if isinstance(v, (list, tuple)):
        return float(v[0])
    return float(v)


_DATA = Path(__file__).resolve().parent / "data"


def glass_ior(name, lam=602.785):
    """IORCurve::eta for BK7 / LASF9 (glass.cpp:104-134) at rgb_spectrum_peak_wavelengths.x (non-spectral spectrum)"""
    coef = {"BK7": ((1.03961212, 0.00600069867), (0.231792344, 0.0200179144), (1.01046945, 103.560653)),
            "LASF9": ((2.00029547, 0.0121426017), (0.298926886, 0.0538736236), (1.80691843, 156.530829))}
    c = coef.get(name, coef["BK7"])
    l = f32(lam) / f32(1000.0)
    l2 = f32(l * l)
    f = f32(0)
    acc = None
    for b, cc in c:
        term = f32(f32(f32(b) * l2) / f32(l2 - f32(cc)))
        acc = term if acc is None else f32(acc + term)
    return float(f32(np.sqrt(f32(acc + f32(1)))))


def _material_record(m, textures, base, spectrum_complete=False):
    """plugin -> POD record: diffuse.cpp:36-42, mirror.cpp:35-41, glass.cpp:183-238, metal.cpp:59-133, substrate.cpp:113-124"""
    prm = m.get("param", {})
    rec = capi.vb_material_rec()
    for f in ("color", "roughness", "spec", "ior", "sigma"):           # unused slots: zero constant, no texture
        setattr(rec, f, make_slot([0.0, 0.0, 0.0, 0.0]))
    t = m["type"]
    remap = bool(prm.get("remapping_roughness", True))
    rec.flags = 2 if remap else 0
    rec.params[0] = _number(prm.get("anisotropic"), 0.0)
    if t == "diffuse":
        rec.type = capi.VB_MAT_DIFFUSE
        rec.flags = 0
        rec.color = _color_slot(prm.get("color", [0.5, 0.5, 0.5]), textures, base)
        if "sigma" in prm:
            rec.flags |= 1
            rec.sigma = make_slot([_number(prm["sigma"], 0.5)] * 3)
    elif t == "mirror":
        rec.type = capi.VB_MAT_MIRROR
        rec.color = _color_slot(prm.get("color", [1, 1, 1]), textures, base)
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.001)] * 3)
    elif t == "glass":
        rec.type = capi.VB_MAT_GLASS
        rec.color = _color_slot(prm.get("color", [1, 1, 1]), textures, base)
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.5)] * 3)
        name = prm.get("material_name", "")
        ior = _number(prm.get("ior"), 1.5) if not name else glass_ior(name)
        rec.ior = make_slot([ior] * 3)
    elif t == "metal":
        rec.type = capi.VB_MAT_METAL
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.01)] * 3)
        table = json.loads((_DATA / "metal_ior_srgb.json").read_text())
        name = prm.get("material_name", "")
        names = sorted(table)                                  # std::map order; unknown name -> first key (metal.cpp:106-108)
        e = table[name if name in table else names[0]]
        rec.eta[:3] = e["eta"]; rec.k[:3] = e["k"]
    elif t == "substrate":
        rec.type = capi.VB_MAT_SUBSTRATE
        rec.color = _color_slot(prm.get("color", [1, 1, 1]), textures, base)
        rec.spec = _color_slot(prm.get("spec", [0.05, 0.05, 0.05]), textures, base)
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.5)] * 3)
    else:
        return None
    return rec


def _medium_records(d):
    """MediumsDesc (scene_desc.cpp:24-34) + MediumDesc::init (node_desc.cpp:182-197) + HomogeneousMedium ctor"""
    md = d.get("mediums")
    if not isinstance(md, dict):               # missing key or a bare list: nothing loaded, "process" keeps its default (true)
        return [], {}, True, ""
    lst, process, glob = md.get("list", []), bool(md.get("process", True)), md.get("global", "")
    table = json.loads((_DATA / "medium_scatter.json").read_text())
    recs, index = [], {}
    for m in lst:
        prm = m.get("param", {})
        r = capi.vb_medium_rec()
        r.type = 0
        name = prm.get("medium_name", "")
        if name:
            e = table.get(name) or next(iter(table.values()))
            ss, sa = e["sigma_s"], e["sigma_a"]
        else:
            ss, sa = prm.get("sigma_s", [0, 0, 0]), prm.get("sigma_a", [0, 0, 0])
        r.sigma_s[:3] = [float(f32(x)) for x in ss]; r.sigma_a[:3] = [float(f32(x)) for x in sa]
        r.scale = float(f32(_number(prm.get("scale"), 0.0)))
        r.g = float(f32(min(max(_number(prm.get("g"), 0.0), -0.99), 0.99)))
        index[m.get("name", "")] = len(recs)
        recs.append(r)
    return recs, index, process, glob


def build_scene(path, overrides=None) -> SceneRecords:
    """overrides: dict with optional keys resolution, max_depth, min_depth, spp, shape_material {name: mat}, drop_lights"""
    ov = overrides or {}
    d = load_json_with_comments(path)
    base = Path(path).parent
    sc = SceneRecords()
    # materials (Scene::load_materials)
    mat_index = {}
    for m in d.get("materials", []):
        rec = _material_record(m, sc.textures, base)
        if rec is None:
            mat_index[m["name"]] = None          # present but unsupported (principled_bsdf, mix, ...)
            continue
        mat_index.setdefault(m["name"], len(sc.materials))
        sc.materials.append(rec)
    if any(r.type in (capi.VB_MAT_MIRROR, capi.VB_MAT_GLASS, capi.VB_MAT_METAL) for r in sc.materials) and ov.get("luts", True):
        sc.luts = capi.load_luts()
    sc.mediums, med_index, process_mediums, global_medium = _medium_records(d)
    if "process_mediums" in ov:
        process_mediums = ov["process_mediums"]
    gm = med_index.get(global_medium, VB_INVALID_ID)
    # shapes -> meshes, instances, area lights (Scene::load_shapes / add_shape)
    for sh in d["shapes"]:
        prm = sh["param"]
        if sh["type"] == "quad":
            mesh = quad_mesh(prm.get("width", 1.0), prm.get("height", 1.0))
        elif sh["type"] == "cube":
            mesh = cube_mesh(prm.get("x", 1.0), prm.get("y", 1.0), prm.get("z", 1.0))
        else:
            raise NotImplementedError(sh["type"])
        mat_name = ov.get("shape_material", {}).get(sh["name"], prm.get("material"))
        if mat_name in mat_index and mat_index[mat_name] is None:
            raise NotImplementedError(f"material {mat_name} of {sh['name']}")
        mid = mat_index.get(mat_name, VB_INVALID_ID)                # unknown / empty name: no material (medium boundary)
        o2w = np.asarray(prm["transform"]["param"]["matrix4x4"], f32).reshape(-1)   # JSON rows are columns
        inst = {"mesh_id": len(sc.meshes), "material_id": mid, "light_id": VB_INVALID_ID, "o2w": o2w}
        if "medium" in prm:                                       # ShapeGroup::post_init, shape.cpp:245-270
            inst["inside_medium"] = med_index.get(prm["medium"].get("inside", ""), VB_INVALID_ID)
            inst["outside_medium"] = med_index.get(prm["medium"].get("outside", ""), VB_INVALID_ID)
        else:
            inst["inside_medium"] = inst["outside_medium"] = gm
        sc.meshes.append(mesh)
        em = prm.get("emission")
        if em:
            ep = em["param"]
            l = capi.vb_light_rec()
            l.type = capi.VB_LIGHT_AREA
            l.inst_id = len(sc.instances)
            l.two_sided = 1 if ep.get("two_sided", False) else 0
            col = ep["color"]["node"] if isinstance(ep["color"], dict) else ep["color"]
            col = [float(f32(c)) for c in col]
            scale = ep.get("scale", 1.0)
            if isinstance(scale, dict):
                scale = scale["node"]["param"]["value"] if isinstance(scale["node"], dict) else scale["node"]
            scale = f32(scale)
            mx = max(col)                                   # Light::initialize_slots -> color_->normalize()
            if mx >= 1.0:
                col = [float(f32(f32(c) / f32(mx))) for c in col]
                scale = f32(scale * f32(mx))
            l.color = make_slot(col)
            l.scale = float(scale)
            table, func, integral = alias_build(triangle_areas(mesh))
            l.alias_offset, l.alias_size, l.alias_integral = len(sc.alias_table), len(table), integral
            sc.alias_table += table; sc.alias_func += func
            inst["light_id"] = len(sc.lights)
            sc.lights.append(l)
        sc.instances.append(inst)
    # Scene::remove_unused_materials (scene.cpp:79-91): drop materials no instance refers to, re-index the rest
    used = sorted({i["material_id"] for i in sc.instances if i["material_id"] != VB_INVALID_ID})
    remap = {old: new for new, old in enumerate(used)}
    sc.materials = [sc.materials[i] for i in used]
    for i in sc.instances:
        if i["material_id"] != VB_INVALID_ID:
            i["material_id"] = remap[i["material_id"]]
    if not any(r.type in (capi.VB_MAT_MIRROR, capi.VB_MAT_GLASS, capi.VB_MAT_METAL) for r in sc.materials):
        sc.luts = {}
    if not ov.get("drop_lights", True) and d.get("light_sampler", {}).get("param", {}).get("lights"):
        raise NotImplementedError("non-area lights")
    # camera
    cam = d["camera"]["param"]
    res = ov.get("resolution") or d.get("pipeline", {}).get("param", {}).get("frame_buffer", {}).get("param", {}).get("resolution", [1280, 720])
    sc.resolution = (int(res[0]), int(res[1]))
    tp = cam["transform"]["param"]
    m = look_at(tp["position"], tp["target_pos"], tp.get("up", [0, 1, 0]))
    c2w = camera_to_world(m)
    sc.camera.c2w[:] = [float(f32(x)) for x in c2w.T.reshape(-1)]          # column-major
    sc.camera.tan_fov_y_over_2 = float(f32(math.tan(math.radians(float(f32(cam.get("fov_y", 20.0)))) * 0.5)))
    sc.camera.lens_radius = float(cam.get("lens_radius", 0.0))
    sc.camera.focal_distance = float(cam.get("focal_distance", 5.0))
    # Photosensory ctor (photosensory.cpp:15-33): only when mediums are processed; named medium or the global one
    sc.camera.medium_id = (med_index.get(cam["medium"], VB_INVALID_ID) if "medium" in cam else gm) if process_mediums else VB_INVALID_ID
    # the flag only matters when some medium exists (all ids are invalid otherwise): keep the kernel's fast path
    sc.params.process_mediums = 1 if (process_mediums and sc.mediums) else 0
    ratio = f32(sc.resolution[0]) * f32(1.0) / f32(sc.resolution[1])
    lo, hi = [-1.0, -1.0], [1.0, 1.0]
    if ratio > 1.0:
        lo[0], hi[0] = -float(ratio), float(ratio)
    else:
        lo[1], hi[1] = -float(f32(1.0) / ratio), float(f32(1.0) / ratio)
    sc.camera.screen_lo[:] = lo; sc.camera.screen_hi[:] = hi
    flt = cam.get("filter", {"type": "gaussian", "param": {}})
    sc.camera.filter_type = capi.VB_FILTER_TRIANGLE if flt["type"] == "triangle" else capi.VB_FILTER_BOX
    r = flt.get("param", {}).get("radius", 0.5)
    r = r if isinstance(r, (list, tuple)) else [r, r]
    sc.camera.filter_radius[:] = [float(r[0]), float(r[0])]        # Filter ctor reads as_float: x only (filter.h:43)
    # integrator
    ip = d.get("integrator", {}).get("param", {})
    sc.params.max_depth = ov.get("max_depth", ip.get("max_depth", 16))
    sc.params.min_depth = ov.get("min_depth", ip.get("min_depth", 5))
    sc.params.rr_threshold = float(ip.get("rr_threshold", 1.0))
    sc.params.mis_mode = int(ip.get("mis_mode", 0))
    sc.params.ray_offset_factor = float(d.get("render_setting", {}).get("ray_offset_factor", 1.0))
    sc.spp = ov.get("spp", d.get("output", {}).get("spp", 1)) or 1
    fb = d.get("pipeline", {}).get("param", {}).get("frame_buffer", {}).get("param", {})
    sc.exposure = float(fb.get("exposure", 1.0))
    tm = fb.get("tone_mapper", {}).get("type", "linear")
    sc.tonemap = {"linear": capi.VB_TONEMAP_LINEAR, "aces": capi.VB_TONEMAP_ACES, "reinhard": capi.VB_TONEMAP_REINHARD}[tm]
    return sc


CBOX_MATTE_OVERRIDES = {"resolution": [256, 256], "max_depth": 5, "spp": 16,
                        "shape_material": {"shape_5": "ShortBox"}, "drop_lights": True}

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


_TEX_KEYS = {}          # id(texture list) -> {(path, colour space): index}; reset per build_scene


def _color_slot(v, textures, base_dir):
    """a colour attribute: [r,g,b] | {"channels","node":[..]} | {"node":{"type":"image","param":{"fn"}}} | {"type":"number"}"""
    if isinstance(v, dict):
        if "node" in v:
            return _color_slot(v["node"], textures, base_dir)
        if v.get("type") == "number":
            return _color_slot(v["param"]["value"], textures, base_dir)
        if v.get("type") == "image" or "fn" in v:
            prm = v.get("param", v)
            # ImagePool::obtain_image (image_pool.cpp:13-35): one texture per (file, colour space)
            keys = _TEX_KEYS.setdefault(id(textures), {})
            key = (str(Path(base_dir) / prm["fn"]), prm.get("color_space", "srgb"))
            if key not in keys or keys[key] >= len(textures):
                textures.append(load_image(Path(base_dir) / prm["fn"], prm.get("color_space", "srgb")))
                keys[key] = len(textures) - 1
            return make_slot([1, 1, 1, 1], keys[key])
    if isinstance(v, (int, float)):
        v = [v, v, v]
    return make_slot([float(x) for x in v])


def read_pfm(path):
    with open(path, "rb") as f:
        magic = f.readline().strip()
        w, h = (int(x) for x in f.readline().split())
        scale = float(f.readline())
        ch = 3 if magic == b"PF" else 1
        a = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4").reshape(h, w, ch)[::-1]
    if ch == 1:
        a = np.repeat(a, 3, axis=-1)
    return np.ascontiguousarray(np.concatenate([a, np.ones((h, w, 1), np.float32)], axis=-1), np.float32)


def model_meshes(path, prm):
    """Model shape (model.cpp:27-36): the meshes of one model file, in the importer's order. With the reference's importer library
    built into the host layer (tools/build_assimp.py) they come from it - the twin does not restate assimp's post-processing, only
    the in-tree OBJ loader below ("importer": "obj", or no assimp: file-order primitive ids)."""
    if prm.get("importer", "assimp") != "obj":
        from . import host_api
        if host_api.assimp_available():
            return host_api.load_model(path, "assimp", bool(prm.get("swap_handed", False)), bool(prm.get("smooth", False)),
                                       bool(prm.get("flip_uv", True)))
    return [obj_mesh(path, smooth=bool(prm.get("smooth", False)), flip_uv=bool(prm.get("flip_uv", True)))]


def obj_mesh(path, smooth=False, flip_uv=True):
    """Model shape (model.cpp:27-36). The reference imports through assimp with JoinIdenticalVertices |
    GenNormals | ImproveCacheLocality | FindDegenerates | OptimizeMeshes | Triangulate | FlipUVs
    (assimp_parser.cpp:309-341), whose cache-locality pass re-orders triangles; this loader keeps FILE order
    (prim_id = order of the f statements): one mesh per file, vertices de-duplicated by
    their (v, vt, vn) index triple in first-use order, polygons fan-triangulated, v -> 1 - v when flip_uv,
    faces with a repeated position index dropped, missing normals = face normals (or their sum per
    position when smooth)."""
    V, VT, VN, faces = [], [], [], []
    with open(path) as f:
        for line in f:
            t = line.split()
            if not t:
                continue
            if t[0] == "v":
                V.append([float(x) for x in t[1:4]])
            elif t[0] == "vt":
                VT.append([float(t[1]), float(t[2]) if len(t) > 2 else 0.0])
            elif t[0] == "vn":
                VN.append([float(x) for x in t[1:4]])
            elif t[0] == "f":
                idx = []
                for c in t[1:]:
                    q = (c.split("/") + ["", ""])[:3]
                    a = [int(x) if x else 0 for x in q]
                    a = [(x - 1 if x > 0 else (len(L) + x if x < 0 else -1)) for x, L in zip(a, (V, VT, VN))]
                    idx.append(tuple(a))
                for k in range(1, len(idx) - 1):
                    tri = (idx[0], idx[k], idx[k + 1])
                    if len({tri[0][0], tri[1][0], tri[2][0]}) == 3:
                        faces.append(tri)
    V = np.asarray(V, f32).reshape(-1, 3); VT = np.asarray(VT, f32).reshape(-1, 2); VN = np.asarray(VN, f32).reshape(-1, 3)
    key_to_index, pos, nrm, uv, tris = {}, [], [], [], []
    def cross3(a, b, t):
        return np.array([t(t(a[1] * b[2]) - t(a[2] * b[1])), t(t(a[2] * b[0]) - t(a[0] * b[2])), t(t(a[0] * b[1]) - t(a[1] * b[0]))], t)
    acc = None
    if smooth and not len(VN):
        acc = np.zeros((len(V), 3), np.float64)
        for tri in faces:
            p = V[[c[0] for c in tri]].astype(np.float64)
            acc[[c[0] for c in tri]] += cross3(p[1] - p[0], p[2] - p[0], np.float64)
    for tri in faces:
        fn = None
        if not len(VN) and acc is None:
            p = V[[c[0] for c in tri]]
            n = cross3((p[1] - p[0]).astype(f32), (p[2] - p[0]).astype(f32), f32)
            l = f32(np.sqrt(f32(f32(f32(n[0] * n[0]) + f32(n[1] * n[1])) + f32(n[2] * n[2]))))
            fn = tuple(float(x) for x in (n / l).astype(f32)) if l > 0 else (0.0, 0.0, 0.0)
        out = []
        for (iv, it, inn) in tri:
            key = (iv, it, inn, fn)
            j = key_to_index.get(key)
            if j is None:
                j = len(pos)
                key_to_index[key] = j
                pos.append(V[iv])
                if inn >= 0:
                    nrm.append(VN[inn])
                elif acc is not None:
                    a = acc[iv]; l = math.sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2])
                    nrm.append((a / l if l > 0 else a).astype(f32))
                else:
                    nrm.append(np.asarray(fn, f32))
                t = VT[it] if it >= 0 else np.zeros(2, f32)
                uv.append(np.array([t[0], f32(1.0) - t[1] if flip_uv else t[1]], f32))
            out.append(j)
        tris.append(out)
    return {"pos": np.asarray(pos, f32).reshape(-1, 3), "nrm": np.asarray(nrm, f32).reshape(-1, 3),
            "uv": np.asarray(uv, f32).reshape(-1, 2), "tri": np.asarray(tris, np.uint32).reshape(-1, 3)}


K_DEG = 3.14159265358979323846 / 180.0


def mat4_mul_f32(a, b):
    """vh::mat4_mul: float operands, products summed in double in k order, one rounding to float"""
    a = np.asarray(a, f32).astype(np.float64); b = np.asarray(b, f32).astype(np.float64)
    r = np.zeros((4, 4), f32)
    for i in range(4):
        for j in range(4):
            acc = 0.0
            for k in range(4):
                acc += float(a[i, k]) * float(b[k, j])
            r[i, j] = f32(acc)
    return r


def rotation_f32(axis, deg):
    """vh::rotation_x/y/z: float degrees, double trigonometry, float entries (row-major here)"""
    ang = float(f32(deg)) * K_DEG
    c, s_ = math.cos(ang), math.sin(ang)
    m = np.eye(4, dtype=f32)
    i, j = {"x": (1, 2), "y": (2, 0), "z": (0, 1)}[axis]
    m[i, i] = f32(c); m[j, j] = f32(c); m[i, j] = f32(-s_); m[j, i] = f32(s_)
    return m


def euler_matrix(prm):
    """TransformDesc 'Euler' (node_desc.cpp:43-48): translation * pitch(x) * roll(z) * yaw(y), degrees"""
    t = np.eye(4, dtype=f32); t[:3, 3] = np.asarray(prm.get("position", [0, 0, 0]), f32)
    m = mat4_mul_f32(t, rotation_f32("x", prm.get("pitch", 0.0)))
    m = mat4_mul_f32(m, rotation_f32("z", prm.get("roll", 0.0)))
    return mat4_mul_f32(m, rotation_f32("y", prm.get("yaw", 0.0)))


def trs_matrix(prm):
    """TransformDesc 'trs' (node_desc.cpp:48-52): translation * axis-angle rotation (x, y, z, degrees) * scale;
    arithmetic as vh_desc.cpp (float products of the axis components, double trigonometry, float entries)"""
    T = np.eye(4, dtype=f32); T[:3, 3] = np.asarray(prm.get("t", [0, 0, 0]), f32)
    S = np.eye(4, dtype=f32)
    sv = np.asarray(prm.get("s", [1, 1, 1]), f32)
    S[0, 0], S[1, 1], S[2, 2] = sv[0], sv[1], sv[2]
    R = np.eye(4, dtype=f32)
    r = [f32(x) for x in prm.get("r", [1, 0, 0, 0])]
    if len(r) == 4 and r[3] != 0:
        l = math.sqrt(float(r[0]) * float(r[0]) + float(r[1]) * float(r[1]) + float(r[2]) * float(r[2]))
        ax, ay, az = (f32(float(r[i]) / l) for i in range(3))
        ang = float(r[3]) * K_DEG
        c, sn = math.cos(ang), math.sin(ang)
        k = 1 - c
        p = lambda u, v: float(f32(u * v))
        # R[row, col]; the host's column-major R[4 * col + row]
        R[0, 0] = f32(c + p(ax, ax) * k); R[1, 0] = f32(p(ay, ax) * k + float(az) * sn); R[2, 0] = f32(p(az, ax) * k - float(ay) * sn)
        R[0, 1] = f32(p(ax, ay) * k - float(az) * sn); R[1, 1] = f32(c + p(ay, ay) * k); R[2, 1] = f32(p(az, ay) * k + float(ax) * sn)
        R[0, 2] = f32(p(ax, az) * k + float(ay) * sn); R[1, 2] = f32(p(ay, az) * k - float(ax) * sn); R[2, 2] = f32(c + p(az, az) * k)
    return mat4_mul_f32(mat4_mul_f32(T, R), S)


def transform_matrix(tr):
    """row-major numpy 4x4 from a transform block (matrix4x4 rows are columns; look_at; Euler)"""
    if not tr:
        return np.eye(4)
    t, prm = tr.get("type", "matrix4x4"), tr.get("param", {})
    if t == "matrix4x4":
        return np.asarray(prm["matrix4x4"], np.float64).T
    if t == "look_at":
        return look_at(prm["position"], prm["target_pos"], prm.get("up", [0, 1, 0]))
    if t == "Euler":
        return euler_matrix(prm).astype(np.float64)
    if t == "trs":
        return trs_matrix(prm).astype(np.float64)
    raise NotImplementedError(t)


def invert4(m):
    """Gauss-Jordan with partial pivoting in double (vh invert4); m row-major"""
    a = [[float(m[r][c]) for c in range(4)] + [1.0 if r == c else 0.0 for c in range(4)] for r in range(4)]
    for i in range(4):
        p = i
        for r in range(i + 1, 4):
            if abs(a[r][i]) > abs(a[p][i]):
                p = r
        if a[p][i] == 0.0:
            raise ValueError("singular matrix")
        if p != i:
            a[i], a[p] = a[p], a[i]
        d = a[i][i]
        a[i] = [x / d for x in a[i]]
        for r in range(4):
            if r != i:
                f = a[r][i]
                if f != 0.0:
                    a[r] = [x - f * y for x, y in zip(a[r], a[i])]
    return np.array([row[4:] for row in a], np.float64)


def env_o2w(prm):
    """SphericalMap ctor (spherical.cpp:36-45): desc.o2w * rotation_x(-90) * scale(1, flip_u ? 1 : -1, 1); the two
    products in double from the float o2w, one rounding to float; w2o = inverse of the rounded matrix"""
    A = transform_matrix(prm.get("o2w")).astype(f32).astype(np.float64)
    ang = -90.0 * 3.14159265358979323846 / 180.0
    c, s_ = math.cos(ang), math.sin(ang)
    RX = np.array([[1, 0, 0, 0], [0, c, -s_, 0], [0, s_, c, 0], [0, 0, 0, 1]], np.float64)
    SC = np.diag([1.0, 1.0 if prm.get("flip_u", False) else -1.0, 1.0, 1.0])
    def mul(x, y):
        o = np.zeros((4, 4))
        for i in range(4):
            for j in range(4):
                v = 0.0
                for k in range(4):
                    v += float(x[i, k]) * float(y[k, j])
                o[i, j] = v
        return o
    o2w = mul(mul(A, RX), SC).astype(f32)
    w2o = invert4(o2w.astype(np.float64)).astype(f32)
    return o2w, w2o


def alias2d_build(weights, rx, ry):
    """AliasTable2D::build (alias2d.cpp:47-82): one table per row, marginal over the row integrals"""
    cond_table, cond_func, marg = [], [], []
    for v in range(ry):
        t, fn, integral = alias_build(weights[v * rx:(v + 1) * rx]) if any(weights[v * rx:(v + 1) * rx]) else \
            ([(1.0, i) for i in range(rx)], [0.0] * rx, 0.0)
        cond_table += t; cond_func += fn; marg.append(integral)
    mt, mf, mi = alias_build(marg)
    return mt, mf, mi, cond_table, cond_func


def read_hdr(path):
    """Radiance RGBE (.hdr): flat and new-style run-length scanlines -> (h, w, 4) float32, linear"""
    data = Path(path).read_bytes()
    pos = data.index(b"\n\n") + 2
    assert data.startswith(b"#?") and b"FORMAT=32-bit_rle_rgbe" in data[:pos], path
    end = data.index(b"\n", pos)
    tok = data[pos:end].split()
    assert tok[0] == b"-Y" and tok[2] == b"+X", tok
    h, w = int(tok[1]), int(tok[3])
    pos = end + 1
    out = np.zeros((h, w, 4), np.uint8)
    for y in range(h):
        head = data[pos:pos + 4]
        if 8 <= w < 32768 and head[0] == 2 and head[1] == 2 and (head[2] << 8 | head[3]) == w:
            pos += 4
            for c in range(4):
                x = 0
                while x < w:
                    n = data[pos]; pos += 1
                    if n > 128:
                        out[y, x:x + n - 128, c] = data[pos]; pos += 1; x += n - 128
                    else:
                        out[y, x:x + n, c] = np.frombuffer(data, np.uint8, n, pos); pos += n; x += n
        else:
            out[y] = np.frombuffer(data, np.uint8, 4 * w, pos).reshape(w, 4); pos += 4 * w
    e = out[..., 3].astype(np.int32)
    scale = np.where(e == 0, f32(0), np.ldexp(f32(1.0), e - 136)).astype(f32)
    rgb = out[..., :3].astype(f32) * scale[..., None]
    return np.ascontiguousarray(np.concatenate([rgb, np.ones((h, w, 1), f32)], axis=-1), f32)


def load_image(path, color_space):
    if Path(path).suffix.lower() == ".pfm":
        return read_pfm(path)
    if Path(path).suffix.lower() == ".hdr":
        return read_hdr(path)
    from PIL import Image
    im = np.asarray(Image.open(path).convert("RGBA"), np.float32) / f32(255.0)
    if color_space == "srgb" and Path(path).suffix.lower() not in (".exr", ".hdr"):
        c = im[..., :3].astype(np.float64)            # double precision, rounded once: the C++ host layer does the same (bit-identical textures)
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


def setup_hero_spectrum(sc, dimension=3):
    """HeroWavelengthSpectrum ctor (hero.cpp:235-245): 5 nm SPDs of the CIE observer and D65 (SPD::create_cie_*, spd.cpp:24-30,89-103:
    every 5th of the 1 nm samples, ceil(470 / 5) = 94 of them), SPD::cie_y_integral (spd.cpp:11-21,105-108) and the coefficient table"""
    if dimension != 3:
        raise NotImplementedError("hero spectrum: dimension 3 only")
    cie = capi.load_cie_1nm()
    n = int(np.ceil(f32(470.0) / f32(5.0)))
    sc.spds = [np.ascontiguousarray(cie[k][0:n * 5:5], np.float32) for k in range(4)]
    s = f32(0.0)
    for i in range(int(f32(470.0) / f32(5.0))):
        s = f32(s + f32(f32(0.5) * f32(cie[1][i * 5] + cie[1][(i + 1) * 5])))
    sp = sc.spectrum
    sp.type, sp.dimension = capi.VB_SPECTRUM_HERO, 3
    sp.spd_x, sp.spd_y, sp.spd_z, sp.spd_d65 = 0, 1, 2, 3
    sp.cie_y_integral = float(f32(s * f32(5.0)))
    sp.rgb2spec_res = capi.RGB2SPEC_RES
    sc.rgb2spec = capi.load_rgb2spec()
    sc.params.spectrum = capi.VB_SPECTRUM_HERO


def _material_record(m, textures, base, spectrum_complete=False, spds=None):
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
        if name and spectrum_complete:                             # glass.cpp:221-224: SPD::to_list(curve): 360, 365, ... < 830 nm
            rec.flags |= 4
            rec.spd[0] = len(spds)
            lst, lam = [], f32(360.0)
            while lam < f32(830.0):
                lst.append(glass_ior(name, lam)); lam = f32(lam + f32(5.0))
            spds.append(np.asarray(lst, np.float32))
    elif t == "metal":
        rec.type = capi.VB_MAT_METAL
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.01)] * 3)
        table = json.loads((_DATA / "metal_ior_srgb.json").read_text())
        name = prm.get("material_name", "")
        names = sorted(table)                                  # std::map order; unknown name -> first key (metal.cpp:106-108)
        e = table[name if name in table else names[0]]
        rec.eta[:3] = e["eta"]; rec.k[:3] = e["k"]
        if spectrum_complete:                                      # metal.cpp:117-121: the full measured tables as "spd" slots
            full = json.loads((_DATA / "metal_ior_spd.json").read_text())[name if name in table else names[0]]
            rec.flags |= 4
            rec.spd[0] = len(spds); spds.append(np.asarray(full["eta"], np.float32))
            rec.spd[1] = len(spds); spds.append(np.asarray(full["k"], np.float32))
    elif t == "substrate":
        rec.type = capi.VB_MAT_SUBSTRATE
        rec.color = _color_slot(prm.get("color", [1, 1, 1]), textures, base)
        rec.spec = _color_slot(prm.get("spec", [0.05, 0.05, 0.05]), textures, base)
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.5)] * 3)
    elif t == "principled_bsdf":                                   # principled_bsdf.cpp:264-290 (slot defaults)
        rec.type = capi.VB_MAT_PRINCIPLED
        rec.flags = 0
        rec.color = _color_slot(prm.get("color", [1, 1, 1]), textures, base)
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.5)] * 3)
        rec.spec = _color_slot(prm.get("spec_tint", [1, 1, 1]), textures, base)
        rec.ior = make_slot([_number(prm.get("ior"), 1.5)] * 3)
        def const3(key, default):
            s = _color_slot(prm.get(key, default), textures, base)
            if s.tex_id != VB_INVALID_ID:
                raise NotImplementedError(f"principled_bsdf: textured {key}")
            return [float(s.value[0]), float(s.value[1]), float(s.value[2])]
        P = rec.params
        P[1] = _number(prm.get("metallic"), 0.0)
        P[2] = _number(prm.get("sheen_weight"), 0.0); P[3] = _number(prm.get("sheen_roughness"), 0.5)
        P[4], P[5], P[6] = const3("sheen_tint", [1, 1, 1])
        P[7] = _number(prm.get("coat_weight"), 0.0); P[8] = _number(prm.get("coat_roughness"), 0.2); P[9] = _number(prm.get("coat_ior"), 1.5)
        P[10], P[11], P[12] = const3("coat_tint", [1, 1, 1])
        P[13] = _number(prm.get("transmission_weight"), 0.0)
        P[14] = 1.0                                                # sheen_mode_ = Approximate (principled_bsdf.cpp:260)
        P[15] = 63.0                                               # all lobe switches on (principled_bsdf.cpp:261-265)
    elif t == "plastic":                                           # plastic.cpp:90-99
        rec.type = capi.VB_MAT_PLASTIC
        rec.color = _color_slot(prm.get("color", [1, 1, 1]), textures, base)
        rec.spec = _color_slot(prm.get("spec", [0.05, 0.05, 0.05]), textures, base)
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.5)] * 3)
        rec.ior = make_slot([_number(prm.get("ior"), 1.3)] * 3)
    elif t == "metallic":                                          # metallic.cpp:28-36; remapping_roughness_ is not a desc attribute
        rec.type = capi.VB_MAT_METALLIC
        rec.flags = 2
        rec.color = _color_slot(prm.get("color", [1, 1, 1]), textures, base)
        rec.spec = _color_slot(prm.get("edge_tint", [1, 1, 1]), textures, base)
        rec.roughness = make_slot([_number(prm.get("roughness"), 0.5)] * 3)
    elif t in ("mix", "add"):                                      # mix.cpp:29-36, add.cpp:25-31: children are filled in by build_scene
        rec.type = capi.VB_MAT_MIX if t == "mix" else capi.VB_MAT_ADD
        rec.flags = 0
        rec.params[0] = 0.0
        if t == "mix":
            rec.sigma = make_slot([_number(prm.get("frac"), 0.5) if "frac" in prm else 0.5] * 3)
    else:
        return None
    return rec


_libm = None


def _cfun(name, x):
    """glibc sinf / cosf: the C++ host calls std::sin / std::cos on floats; numpy's float32 kernels may differ by an ulp"""
    global _libm
    if _libm is None:
        import ctypes
        _libm = ctypes.CDLL("libm.so.6")
        for n in ("sinf", "cosf", "tanf", "acosf", "sqrtf"):
            getattr(_libm, n).restype = ctypes.c_float
            getattr(_libm, n).argtypes = [ctypes.c_float]
    return f32(getattr(_libm, name)(float(f32(x))))


def _light_common(l, prm, textures, base, default_color=(1, 1, 1)):
    """Light ctor + initialize_slots (light.cpp:10-24): colour slot, scale, colour normalised into the scale"""
    slot = _color_slot(prm.get("color", list(default_color)), textures, base)
    scale = f32(_number(prm.get("scale", 1.0), 1.0))
    if slot.tex_id == VB_INVALID_ID:
        col = [f32(c) for c in slot.value[:3]]
        mx = max(col)
        if mx >= 1.0:
            slot = make_slot([float(f32(c / mx)) for c in col])
            scale = f32(scale * mx)
    l.color = slot
    l.scale = float(scale)


def env_weights(tex):
    """SphericalMap::calculate_weights (spherical.cpp:170-196) incl. its index quirk: v = idx / res.y + 0.5,
    theta = v / res.x (integer division by the HEIGHT, normalised by the WIDTH); luminance = Rec.709"""
    h, w = tex.shape[:2]
    px = tex.reshape(-1, 4).astype(f32)
    lum = (f32(0.212671) * px[:, 0] + f32(0.715160) * px[:, 1]) + f32(0.072169) * px[:, 2]
    idx = np.arange(w * h, dtype=np.int64)
    theta = ((idx // h).astype(f32) + f32(0.5)) / f32(w)
    ang = (f32(math.pi) * theta).astype(f32)
    uniq = np.unique(ang)
    table = {float(a): _cfun("sinf", a) for a in uniq}
    sin_t = np.array([table[float(a)] for a in ang], f32)
    return (lum * sin_t).astype(f32)


PI_F = f32(3.14159265358979323846)


def _light_power_luminance(l, sc, shape=None):
    """luminance(light->power()), the PowerLightSampler weight (power.cpp:35-50): area.cpp:87-89, point.cpp:32-34, spot.cpp:44-46,
    projector.cpp:58-91, spherical.cpp:56-59; Light::average (light.h:80-83) = slot average x scale; a texture's average
    (ocarina Image::average_vector, absent) = per-channel mean, summed in double in pixel order; luminance = Rec.709.
    `shape` carries what the record does not keep: spot (angle, falloff) in radians."""
    if l.color.tex_id == VB_INVALID_ID:
        avg = [f32(l.color.value[i]) for i in range(3)]
    else:
        t = np.asarray(sc.textures[l.color.tex_id], np.float32)
        px = t.reshape(-1, t.shape[-1])[:, :3].astype(np.float64)
        avg = [f32(np.add.accumulate(px[:, c])[-1] / len(px)) for c in range(3)]
    avg = [f32(a * f32(l.scale)) for a in avg]
    if l.type == capi.VB_LIGHT_AREA:
        area = f32(0.0)
        for w in sc.alias_func[l.alias_offset:l.alias_offset + l.alias_size]:
            area = f32(area + f32(w))
        k = f32(2.0 if l.two_sided else 1.0)
        p = [f32(f32(f32(k * a) * area) * PI_F) for a in avg]
    elif l.type == capi.VB_LIGHT_POINT:
        p = [f32(f32(f32(4.0) * PI_F) * a) for a in avg]
    elif l.type == capi.VB_LIGHT_SPOT:
        angle, falloff = shape
        s = f32(f32(1.0) - f32(f32(0.5) * f32(f32(angle * f32(2.0)) + falloff)))
        p = [f32(f32(f32(f32(2.0) * PI_F) * a) * s) for a in avg]
    elif l.type == capi.VB_LIGHT_ENV:
        r = f32(f32(l.params[0]) / f32(2.0))
        p = [f32(f32(PI_F * f32(r * r)) * a) for a in avg]
    elif l.type == capi.VB_LIGHT_DIRECTIONAL:                          # directional.cpp:44-46
        r = f32(f32(l.params[0]) / f32(2.0))
        p = [f32(f32(a * PI_F) * f32(r * r)) for a in avg]
    else:                                                              # projector: Girard's theorem on the frustum corner directions
        ratio = f32(l.params[1])
        y = _cfun("sqrtf", f32(f32(1.0) / f32(f32(ratio * ratio) + f32(1.0))))
        x = f32(ratio * y)
        z = _cfun("sqrtf", f32(f32(x * x) + f32(y * y)))
        P0, P1, P2 = (x, y, z), (x, f32(-y), z), (f32(-x), f32(-y), z)

        def cross(a, b):
            return (f32(f32(a[1] * b[2]) - f32(a[2] * b[1])), f32(f32(a[2] * b[0]) - f32(a[0] * b[2])), f32(f32(a[0] * b[1]) - f32(a[1] * b[0])))

        def dot(a, b):
            return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2]))

        def norm(a):
            l2 = dot(a, a)
            if not l2 > 0:
                return a
            inv = f32(f32(1.0) / _cfun("sqrtf", l2))
            return tuple(f32(c * inv) for c in a)

        c01, c12, c20 = norm(cross(P0, P1)), norm(cross(P1, P2)), norm(cross(P2, P0))
        neg = lambda a: tuple(f32(-c) for c in a)
        ac = lambda a, b: _cfun("acosf", min(max(dot(a, b), f32(-1.0)), f32(1.0)))
        solid = f32(abs(f32(f32(f32(ac(c01, neg(c12)) + ac(c12, neg(c20))) + ac(c20, neg(c01))) - PI_F)))
        k = f32(f32(f32(2.0) * solid) / f32(f32(4.0) * PI_F))
        p = [f32(k * a) for a in avg]
    return f32(f32(f32(f32(0.212671) * p[0]) + f32(f32(0.715160) * p[1])) + f32(f32(0.072169) * p[2]))


def _free_light_record(ld, sc, base, world_diameter):
    """lights listed under light_sampler.param.lights: point.cpp:24-27, spot.cpp:29-35, spherical.cpp:36-45,199-211"""
    t, prm = ld["type"], ld.get("param", {})
    l = capi.vb_light_rec()
    l.inst_id = VB_INVALID_ID
    _light_common(l, prm, sc.textures, base)
    if t in ("point", "spot"):
        l.type = capi.VB_LIGHT_POINT if t == "point" else capi.VB_LIGHT_SPOT
        l.position[:3] = [float(f32(x)) for x in prm.get("position", [0, 0, 0])]
        if t == "spot":
            angle = f32(math.radians(float(np.clip(f32(prm.get("angle", 45.0)), f32(1.0), f32(89.0)))))
            falloff = f32(math.radians(float(np.clip(f32(prm.get("falloff", 10.0)), f32(0.0), angle))))
            dv = np.asarray(prm.get("direction", [0, 0, 1]), f32)
            inv = f32(1.0) / f32(np.sqrt(f32(f32(f32(dv[0] * dv[0]) + f32(dv[1] * dv[1])) + f32(dv[2] * dv[2]))))
            l.direction[:3] = [float(f32(x * inv)) for x in dv]
            l.params[0] = float(_cfun("cosf", angle))
            l.params[1] = float(_cfun("cosf", max(f32(0.0), f32(angle - falloff))))
            l._shape = (angle, falloff)
    elif t == "spherical":
        l.type = capi.VB_LIGHT_ENV
        o2w, w2o = env_o2w(prm)
        l.o2w[:] = [float(x) for x in o2w.T.reshape(-1)]
        l.w2o[:] = [float(x) for x in w2o.T.reshape(-1)]
        if l.color.tex_id != VB_INVALID_ID:
            tex = sc.textures[l.color.tex_id]
            ry, rx = tex.shape[:2]
            weights = env_weights(tex)
        else:
            rx = ry = 1
            weights = np.ones(1, f32)
        mt, mf, mi, ct, cf = alias2d_build([float(x) for x in weights], rx, ry)
        l.alias_offset, l.alias_size, l.alias_integral = len(sc.alias_table), ry, mi
        sc.alias_table += mt; sc.alias_func += mf
        l.cond_offset = len(sc.alias_table)
        sc.alias_table += ct; sc.alias_func += cf
        l.env_res[:] = [rx, ry]
        l.params[0] = world_diameter
    elif t == "directional":
        # directional.cpp:20-38: "direction" stored as written; world radius = diameter / 2 (scene.h:108-109)
        l.type = capi.VB_LIGHT_DIRECTIONAL
        l.direction[:3] = [float(f32(x)) for x in prm.get("direction", [-1, -1, -1])]
        l.params[0] = world_diameter
        l.params[1] = float(f32(f32(world_diameter) * f32(0.5)))
    elif t == "projector":
        # projector.cpp:33-52: point light at o2w[3] projecting its image along local +z
        l.type = capi.VB_LIGHT_PROJECTOR
        o2w = transform_matrix(prm.get("o2w")).astype(f32)
        w2o = invert4(o2w.astype(np.float64)).astype(f32)
        l.o2w[:] = [float(x) for x in o2w.T.reshape(-1)]
        l.w2o[:] = [float(x) for x in w2o.T.reshape(-1)]
        l.position[:3] = [float(x) for x in o2w[:3, 3]]
        l.direction[:3] = [float(x) for x in o2w[:3, 2]]
        ratio = f32(_number(prm.get("ratio", 1.0), 1.0))
        if ratio == 0 and l.color.tex_id != VB_INVALID_ID:
            th, tw = sc.textures[l.color.tex_id].shape[:2]
            ratio = f32(f32(tw) / f32(th))
        angle = f32(math.radians(float(np.clip(f32(prm.get("angle", 45.0)), f32(1.0), f32(89.0)))))
        l.params[0] = float(_cfun("tanf", angle))
        l.params[1] = float(ratio)
    else:
        raise NotImplementedError(f"light {t}")
    return l


def world_diameter(sc, min_world_radius):
    """Scene::world_diameter (scene.h:108-109) over the instance boxes (shape.cpp:73-85), accumulated in double"""
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    for inst in sc.instances:
        m = np.asarray(inst["o2w"], np.float64).reshape(4, 4).T          # row-major
        mesh = sc.meshes[inst["mesh_id"]]
        p = mesh["pos"][np.unique(mesh["tri"])].astype(np.float64)
        q = np.stack([p[:, 0] * m[r, 0] + p[:, 1] * m[r, 1] + p[:, 2] * m[r, 2] + m[r, 3] for r in range(3)], axis=1)
        if len(q):
            lo, hi = np.minimum(lo, q.min(0)), np.maximum(hi, q.max(0))
    d = hi - lo
    radius = 0.5 * math.sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]) if len(sc.instances) and np.all(np.isfinite(d)) else 0.0
    return float(f32(2.0 * max(radius, float(min_world_radius))))


def sphere_mesh(radius=1.0, sub_div=60):
    """Sphere shape (sphere.cpp:20-88): theta_div = sub_div rings, phi_div = 2 * sub_div segments; vertex and
    triangle order as the reference emits them; trigonometry in float via glibc cosf / sinf like the C++ host"""
    radius = f32(radius); theta_div = int(sub_div); phi_div = 2 * theta_div
    pos, nrm, uv, tri = [[0.0, float(radius), 0.0]], [[0.0, 1.0, 0.0]], [[0.0, 0.0]], []
    def unit(p):
        p = np.asarray(p, f32)
        l = f32(np.sqrt(f32(f32(f32(p[0] * p[0]) + f32(p[1] * p[1])) + f32(p[2] * p[2]))))
        return [float(f32(x / l)) for x in p]
    two_pi = f32(6.283185307179586)
    for i in range(1, theta_div):
        v = f32(f32(i) / f32(theta_div))
        theta = f32(f32(math.pi) * v)
        y = f32(radius * _cfun("cosf", theta)); r = f32(radius * _cfun("sinf", theta))
        p0 = [float(r), float(y), 0.0]
        pos.append(p0); nrm.append(unit(p0)); uv.append([0.0, float(v)])
        for j in range(1, phi_div):
            u = f32(f32(j) / f32(phi_div))
            phi = f32(u * two_pi)
            p = [float(f32(_cfun("cosf", phi) * r)), float(y), float(f32(_cfun("sinf", phi) * r))]
            pos.append(p); nrm.append(unit(p)); uv.append([float(u), float(v)])
    pos.append([0.0, float(-radius), 0.0]); nrm.append([0.0, -1.0, 0.0]); uv.append([0.0, 1.0])
    for i in range(phi_div):
        tri.append([0, (i + 1) % phi_div + 1, i + 1])
    for i in range(theta_div - 2):
        vs = 1 + i * phi_div
        for j in range(phi_div):
            if j != phi_div - 1:
                tri.append([vs, vs + 1, vs + phi_div]); tri.append([vs + 1, vs + phi_div + 1, vs + phi_div])
            else:
                tri.append([vs, vs + 1 - phi_div, vs + phi_div]); tri.append([vs + 1 - phi_div, vs + 1, vs + phi_div])
            vs += 1
    ve = len(pos) - 1
    for i in range(phi_div):
        tri.append([ve, ve - ((1 + i) % phi_div + 1), ve - (i + 1)])
    return {"pos": np.asarray(pos, f32), "nrm": np.asarray(nrm, f32), "uv": np.asarray(uv, f32), "tri": np.asarray(tri, np.uint32)}


def _expf(x):
    global _libm
    _cfun("sinf", 0.0)
    import ctypes
    _libm.expf.restype = ctypes.c_float; _libm.expf.argtypes = [ctypes.c_float]
    return f32(_libm.expf(float(f32(x))))


def fitted_filter_weights(flt):
    """FilterSampler::build (fitted_curve.h:36-58): |f| at the 20x20 cell centres of the positive quadrant.
    gaussian.cpp:49-53, mitchell.cpp:33-52, sinc.cpp:27-30 + math/util.h:37-66; float arithmetic as in the C++ host"""
    t, prm = flt["type"], flt.get("param", {})
    r = prm.get("radius", 0.5); r = f32(r[0] if isinstance(r, (list, tuple)) else r)
    N = capi.VB_FILTER_TABLE_SIZE
    PI = f32(math.pi)
    if t == "gaussian":
        sigma = f32(prm.get("sigma", 1.0))
        def g(x):
            a = f32(f32(1.0) / f32(np.sqrt(f32(f32(f32(2.0) * PI) * f32(sigma * sigma)))))
            return f32(a * _expf(f32(f32(-f32(x * x)) / f32(f32(2.0) * f32(sigma * sigma)))))
        ex = g(r)
        f = lambda x, y: f32(max(f32(0.0), f32(g(x) - ex)) * max(f32(0.0), f32(g(y) - ex)))
    elif t == "mitchell":
        b, c = f32(prm.get("b", f32(1.0) / f32(3.0))), f32(prm.get("c", f32(1.0) / f32(3.0)))
        def m1(x):
            x = f32(abs(x)); x2 = f32(x * x); x3 = f32(x2 * x); s6 = f32(f32(1.0) / f32(6.0))
            if x <= 1:
                return f32(f32(f32(f32(f32(f32(12) - f32(9 * b)) - f32(6 * c)) * x3) + f32(f32(f32(f32(-18) + f32(12 * b)) + f32(6 * c)) * x2) + f32(f32(6) - f32(2 * b))) * s6)
            if x <= 2:
                return f32(f32(f32(f32(f32(f32(-b) - f32(6 * c)) * x3) + f32(f32(f32(6 * b) + f32(30 * c)) * x2)) + f32(f32(f32(-12 * b) - f32(48 * c)) * x) + f32(f32(8 * b) + f32(24 * c))) * s6)
            return f32(0.0)
        f = lambda x, y: f32(m1(f32(f32(2 * x) / r)) * m1(f32(f32(2 * y) / r)))
    elif t == "sinc":
        tau = f32(prm.get("tau", 3.0))
        def sinc(x):
            x = f32(PI * x)
            return f32(1.0) if f32(f32(1.0) + f32(x * x)) == 1 else f32(_cfun("sinf", x) / x)
        ws = lambda x: f32(0.0) if abs(x) > r else f32(sinc(x) * sinc(f32(x / tau)))
        f = lambda x, y: f32(f32(ws(x) * ws(y)) * f32(4.0))
    else:
        raise NotImplementedError(f"filter {t}")
    out = []
    for i in range(N * N):
        x, y = i % N, i // N
        px, py = f32(f32(f32(x + f32(0.5)) / f32(N)) * r), f32(f32(f32(y + f32(0.5)) / f32(N)) * r)
        out.append(float(abs(f(px, py))))
    return out


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
    _TEX_KEYS.pop(id(sc.textures), None)
    # materials (Scene::load_materials): records are made further down, for the materials in use only
    mat_descs, mat_index = [], {}
    for m in d.get("materials", []):
        mat_index.setdefault(m["name"], len(mat_descs))
        mat_descs.append(m)
    sc.mediums, med_index, process_mediums, global_medium = _medium_records(d)
    spec_type = ov.get("spectrum", d.get("spectrum", {}).get("type", "srgb"))       # SpectrumDesc, scene_desc.cpp:39
    hero = spec_type == "hero"
    if hero:
        setup_hero_spectrum(sc, int(d.get("spectrum", {}).get("param", {}).get("dimension", 3)))
    elif spec_type != "srgb":
        raise NotImplementedError(f"spectrum {spec_type}")
    if "process_mediums" in ov:
        process_mediums = ov["process_mediums"]
    gm = med_index.get(global_medium, VB_INVALID_ID)
    # shapes -> meshes, instances, area lights (Scene::load_shapes / add_shape)
    area_lights = []                                                # (emission block, instance index)
    for sh in d["shapes"]:
        prm = sh["param"]
        if sh["type"] == "quad":
            mesh = quad_mesh(prm.get("width", 1.0), prm.get("height", 1.0))
        elif sh["type"] == "cube":
            mesh = cube_mesh(prm.get("x", 1.0), prm.get("y", 1.0), prm.get("z", 1.0))
        elif sh["type"] == "sphere":
            mesh = sphere_mesh(prm.get("radius", 1.0), prm.get("sub_div", 60))
        elif sh["type"] == "model":
            mesh = model_meshes(base / prm["fn"], prm)
        else:
            raise NotImplementedError(sh["type"])
        mat_name = ov.get("shape_material", {}).get(sh["name"], prm.get("material"))
        mid = mat_index.get(mat_name, VB_INVALID_ID)                # unknown / empty name: no material (medium boundary)
        o2w = transform_matrix(prm.get("transform")).astype(f32).T.reshape(-1)       # column-major
        for one in (mesh if isinstance(mesh, list) else [mesh]):   # a model yields one instance per imported mesh (model.cpp:34-35)
            inst = {"mesh_id": len(sc.meshes), "material_id": mid, "light_id": VB_INVALID_ID, "o2w": o2w}
            if "medium" in prm:                                       # ShapeGroup::post_init, shape.cpp:245-270
                inst["inside_medium"] = med_index.get(prm["medium"].get("inside", ""), VB_INVALID_ID)
                inst["outside_medium"] = med_index.get(prm["medium"].get("outside", ""), VB_INVALID_ID)
            else:
                inst["inside_medium"] = inst["outside_medium"] = gm
            sc.meshes.append(one)
            if prm.get("emission"):
                area_lights.append((prm["emission"], len(sc.instances)))
            sc.instances.append(inst)
    # Scene::remove_unused_materials (scene.cpp:79-91): drop materials no instance refers to, re-index the rest;
    # records (and their textures) are made in the new index order
    used = sorted({i["material_id"] for i in sc.instances if i["material_id"] != VB_INVALID_ID})
    remap = {old: new for new, old in enumerate(used)}
    for old in used:
        rec = _material_record(mat_descs[old], sc.textures, base, hero, sc.spds)
        if rec is None:
            raise NotImplementedError(f"material {mat_descs[old]['name']} ({mat_descs[old]['type']})")
        sc.materials.append(rec)
    # children of mix / add materials: records appended after the scene's own, in the order mat0, mat1 (the host layer's order)
    for k in range(len(used)):
        md = mat_descs[used[k]]
        if md["type"] not in ("mix", "add"):
            continue
        for c, key in enumerate(("mat0", "mat1")):
            cd = md.get("param", {}).get(key)
            if cd is None:
                raise ValueError(f"material {md.get('name')}: \"{key}\" is missing")
            if cd.get("type", "diffuse") in ("mix", "add"):
                raise NotImplementedError(f"material {md.get('name')}: a mix / add material nested in another one is not supported")
            cd = dict(cd); cd.setdefault("type", "diffuse")
            crec = _material_record(cd, sc.textures, base, hero, sc.spds)
            if crec is None:
                raise NotImplementedError(f"material {md.get('name')}.{key} ({cd['type']})")
            sc.materials[k].spd[c] = len(sc.materials)
            sc.materials.append(crec)
    for i in sc.instances:
        if i["material_id"] != VB_INVALID_ID:
            i["material_id"] = remap[i["material_id"]]
    lut_users = (capi.VB_MAT_MIRROR, capi.VB_MAT_GLASS, capi.VB_MAT_METAL, capi.VB_MAT_PRINCIPLED, capi.VB_MAT_METALLIC)
    if any(r.type in lut_users for r in sc.materials) and ov.get("luts", True):
        sc.luts = capi.load_luts(principled=any(r.type == capi.VB_MAT_PRINCIPLED for r in sc.materials))
    # LightSampler ctor (lightsampler.cpp:11-27): listed lights come first, the shapes' emissions follow; tidy_up
    # (lightsampler.cpp:67-77) groups them by type in first-seen order; records are made in that order
    lsp = d.get("light_sampler", {}).get("param", {})                # lightsampler.cpp:11-13: env_separate, env_prob in [0.01, 0.99]
    env_separate = bool(lsp.get("env_separate", False))
    sc.light_sampler_env = (1 if env_separate else 0, float(min(max(f32(lsp.get("env_prob", 0.5)), f32(0.01)), f32(0.99))))
    ls_type = d.get("light_sampler", {}).get("type", "uniform")
    if ls_type not in ("uniform", "power"):
        raise NotImplementedError("light sampler " + ls_type)
    free = [] if ov.get("drop_lights", False) else d.get("light_sampler", {}).get("param", {}).get("lights", [])
    type_of = {"point": capi.VB_LIGHT_POINT, "spot": capi.VB_LIGHT_SPOT, "spherical": capi.VB_LIGHT_ENV, "projector": capi.VB_LIGHT_PROJECTOR,
               "directional": capi.VB_LIGHT_DIRECTIONAL}
    pending = [(type_of.get(ld["type"], -1), ld, None) for ld in free] + [(capi.VB_LIGHT_AREA, em, ii) for em, ii in area_lights]
    first = {}
    for t, _, _ in pending:
        first.setdefault(t, len(first))
    pending.sort(key=lambda x: first[x[0]])                          # stable
    wd = world_diameter(sc, d.get("render_setting", {}).get("min_world_radius", 10.0)) if free else 0.0
    for t, ld, ii in pending:
        if t != capi.VB_LIGHT_AREA:
            sc.lights.append(_free_light_record(ld, sc, base, wd))
            continue
        ep = ld["param"]
        l = capi.vb_light_rec()
        l.type = capi.VB_LIGHT_AREA
        l.inst_id = ii
        l.two_sided = 1 if ep.get("two_sided", False) else 0
        _light_common(l, ep, sc.textures, base, default_color=[0.5, 0.5, 0.5])
        table, func, integral = alias_build(triangle_areas(sc.meshes[sc.instances[ii]["mesh_id"]]))
        l.alias_offset, l.alias_size, l.alias_integral = len(sc.alias_table), len(table), integral
        sc.alias_table += table; sc.alias_func += func
        sc.instances[ii]["light_id"] = len(sc.lights)
        sc.lights.append(l)
    if ls_type == "power" and sc.lights:                             # PowerLightSampler::prepare (power.cpp:33-53)
        infinite = (capi.VB_LIGHT_ENV, capi.VB_LIGHT_DIRECTIONAL)       # power.cpp:37-45: weight 0 when sampled separately
        weights = [f32(0.0) if (env_separate and l.type in infinite) else _light_power_luminance(l, sc, getattr(l, "_shape", None))
                   for l in sc.lights]
        table, func, integral = alias_build(weights)
        sc.light_sampler = (capi.VB_LIGHT_SAMPLER_POWER, len(sc.alias_table), len(table), integral)
        sc.alias_table += table; sc.alias_func += func
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
    sc.camera.filter_type = {"triangle": capi.VB_FILTER_TRIANGLE, "box": capi.VB_FILTER_BOX}.get(flt["type"], capi.VB_FILTER_FITTED)
    if sc.camera.filter_type == capi.VB_FILTER_FITTED:
        N = capi.VB_FILTER_TABLE_SIZE
        mt, mf, _, ct, cf = alias2d_build(fitted_filter_weights(flt), N, N)
        sc.camera.filter_table_offset = len(sc.alias_table)
        sc.alias_table += mt + ct; sc.alias_func += mf + cf
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

"""ctypes binding of oracle/liboracle.so - the CPU checker. Test infrastructure only."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "oracle" / "liboracle.so"

RAY_FIELDS = ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax")
_lib = None


def build():
    subprocess.run(["make", "-C", str(ROOT / "oracle")], check=True, capture_output=True)


def load() -> "Oracle":
    global _lib
    if _lib is None:
        if not LIB.exists():
            build()
        _lib = Oracle(C.CDLL(str(LIB)))
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        vp = C.c_void_p
        lib.vo_hw_threads.restype = C.c_int
        lib.vo_synth_mesh.argtypes = [C.c_uint64, C.c_uint64, vp, C.c_int]
        lib.vo_synth_mesh.restype = None
        lib.vo_synth_rays.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64] + [vp] * 9 + [C.c_int]
        lib.vo_synth_rays.restype = None
        lib.vo_accel_create.restype = vp
        lib.vo_accel_destroy.argtypes = [vp]
        lib.vo_accel_add_instance.argtypes = [vp, vp, C.c_uint32, vp, C.c_uint32, vp, C.c_uint32]
        lib.vo_accel_add_instance.restype = None
        lib.vo_accel_add_soup.argtypes = [vp, vp, C.c_uint64, C.c_uint32]
        lib.vo_accel_add_soup.restype = None
        lib.vo_accel_build.argtypes = [vp, C.c_int]
        lib.vo_accel_build.restype = C.c_double
        lib.vo_accel_num_nodes.argtypes = [vp]
        lib.vo_accel_num_nodes.restype = C.c_uint64
        lib.vo_trace_closest.argtypes = [vp, C.c_uint64] + [vp] * 8 + [vp] * 5 + [C.c_int, C.c_int, vp]
        lib.vo_trace_closest.restype = C.c_double
        lib.vo_trace_any.argtypes = [vp, C.c_uint64] + [vp] * 8 + [vp, C.c_int, C.c_int, vp]
        lib.vo_trace_any.restype = C.c_double

    def hw_threads(self) -> int:
        return self.lib.vo_hw_threads()

    def synth_mesh(self, n: int, seed: int = 1, threads: int = 8) -> np.ndarray:
        out = np.empty((n, 9), np.float32)
        self.lib.vo_synth_mesh(seed, n, _p(out), threads)
        return out

    def synth_rays(self, n: int, seed: int = 2, seed_any: int = 3, threads: int = 8) -> dict:
        r = {f: np.empty(n, np.float32) for f in RAY_FIELDS + ("tmax_any",)}
        self.lib.vo_synth_rays(seed, seed_any, n, *[_p(r[f]) for f in RAY_FIELDS], _p(r["tmax_any"]), threads)
        return r

    def accel(self) -> "OracleAccel":
        return OracleAccel(self)


class OracleAccel:
    def __init__(self, o: Oracle):
        self.o = o
        self.h = C.c_void_p(o.lib.vo_accel_create())
        self.build_s = 0.0

    def __del__(self):
        if self.h:
            self.o.lib.vo_accel_destroy(self.h)
            self.h = None

    def add_instance(self, pos, tri, o2w=None, inst_id=0):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        tri = np.ascontiguousarray(tri, np.uint32).reshape(-1, 3)
        m = None if o2w is None else np.ascontiguousarray(o2w, np.float32).reshape(16)
        self.o.lib.vo_accel_add_instance(self.h, _p(pos), len(pos), _p(tri), len(tri), _p(m), inst_id)

    def add_soup(self, pos9, inst_id=0):
        pos9 = np.ascontiguousarray(pos9, np.float32)
        self.o.lib.vo_accel_add_soup(self.h, _p(pos9), pos9.shape[0], inst_id)

    def build(self, threads=8):
        self.build_s = self.o.lib.vo_accel_build(self.h, threads)
        return self.build_s

    def trace_closest(self, rays, threads=8, brute=False, stats=False):
        n = rays["ox"].shape[0]
        out = {"inst": np.empty(n, np.uint32), "prim": np.empty(n, np.uint32), "u": np.empty(n, np.float32),
               "v": np.empty(n, np.float32), "t": np.empty(n, np.float32)}
        st = np.zeros(2, np.uint64) if stats else None
        dt = self.o.lib.vo_trace_closest(self.h, n, *[_p(np.ascontiguousarray(rays[f])) for f in RAY_FIELDS],
                                         *[_p(out[f]) for f in ("inst", "prim", "u", "v", "t")],
                                         threads, 1 if brute else 0, _p(st))
        out["seconds"] = dt
        if stats:
            out["stats"] = st
        return out

    def trace_any(self, rays, threads=8, brute=False):
        n = rays["ox"].shape[0]
        occ = np.empty(n, np.uint8)
        dt = self.o.lib.vo_trace_any(self.h, n, *[_p(np.ascontiguousarray(rays[f])) for f in RAY_FIELDS], _p(occ),
                                     threads, 1 if brute else 0, None)
        return occ, dt


# ---------------------------------------------------------------------------------------------
# oracle path tracer
# ---------------------------------------------------------------------------------------------
class OracleScene:
    """uploads a vision_b200.capi.SceneRecords into the oracle and renders it on the CPU"""

    def __init__(self, o: Oracle, sc, brute=False, threads=8):
        from vision_b200 import capi
        self.o, self.sc = o, sc
        lib = o.lib
        vp = C.c_void_p
        lib.vo_scene_create.restype = vp
        lib.vo_scene_destroy.argtypes = [vp]
        lib.vo_scene_add_mesh.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, C.c_uint32]
        lib.vo_scene_add_mesh.restype = C.c_uint32
        lib.vo_scene_set_instances.argtypes = [vp, vp, C.c_uint32]
        lib.vo_scene_set_materials.argtypes = [vp, vp, C.c_uint32]
        lib.vo_scene_set_lights.argtypes = [vp, vp, C.c_uint32, vp, vp, C.c_uint32]
        lib.vo_scene_set_textures.argtypes = [vp, vp, C.c_uint32]
        lib.vo_scene_set_camera.argtypes = [vp, vp]
        lib.vo_scene_set_lut.argtypes = [vp, C.c_int, vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        lib.vo_scene_set_lut.restype = None
        lib.vo_scene_set_mediums.argtypes = [vp, vp, C.c_uint32]
        lib.vo_scene_set_mediums.restype = None
        lib.vo_scene_set_integrator.argtypes = [vp, vp]
        lib.vo_scene_build.argtypes = [vp, C.c_int, C.c_int]
        lib.vo_render.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, vp, C.c_int, vp]
        lib.vo_render.restype = C.c_double
        lib.vo_primary_hits.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp, vp, vp, vp, C.c_int]
        for f in ("vo_scene_destroy", "vo_scene_set_instances", "vo_scene_set_materials", "vo_scene_set_lights",
                  "vo_scene_set_textures", "vo_scene_set_camera", "vo_scene_set_integrator", "vo_scene_build", "vo_primary_hits"):
            getattr(lib, f).restype = None
        self.h = C.c_void_p(lib.vo_scene_create())
        for m in sc.meshes:
            pos = np.ascontiguousarray(m["pos"], np.float32).reshape(-1, 3)
            tri = np.ascontiguousarray(m["tri"], np.uint32).reshape(-1, 3)
            nrm = None if m.get("nrm") is None else np.ascontiguousarray(m["nrm"], np.float32)
            uv = None if m.get("uv") is None else np.ascontiguousarray(m["uv"], np.float32)
            lib.vo_scene_add_mesh(self.h, _p(pos), _p(nrm), _p(uv), len(pos), _p(tri), len(tri))
        inst = (capi.vb_instance * len(sc.instances))()
        for i, d in enumerate(sc.instances):
            inst[i].mesh_id = d["mesh_id"]
            for k in ("material_id", "light_id", "inside_medium", "outside_medium", "lightmap_id"):
                setattr(inst[i], k, d.get(k, 0xFFFFFFFF))
            inst[i].o2w[:] = [float(x) for x in np.asarray(d["o2w"], np.float32).reshape(-1)]
        lib.vo_scene_set_instances(self.h, C.cast(inst, vp), len(sc.instances))
        mats = (capi.vb_material_rec * max(1, len(sc.materials)))(*sc.materials)
        lib.vo_scene_set_materials(self.h, C.cast(mats, vp), len(sc.materials))
        lights = (capi.vb_light_rec * max(1, len(sc.lights)))(*sc.lights)
        na = len(sc.alias_table)
        tab = (capi.vb_alias_entry * max(1, na))(*[capi.vb_alias_entry(p, a) for p, a in sc.alias_table])
        func = np.ascontiguousarray(sc.alias_func if na else [0.0], np.float32)
        lib.vo_scene_set_lights(self.h, C.cast(lights, vp), len(sc.lights), C.cast(tab, vp), _p(func), na)
        lib.vo_scene_set_light_sampler.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float]
        lib.vo_scene_set_light_sampler.restype = None
        lib.vo_scene_set_light_sampler(self.h, *getattr(sc, "light_sampler", (0, 0, 0, 0.0)))
        lib.vo_scene_set_light_sampler_env.argtypes = [vp, C.c_uint32, C.c_float]
        lib.vo_scene_set_light_sampler_env.restype = None
        lib.vo_scene_set_light_sampler_env(self.h, *getattr(sc, "light_sampler_env", (0, 0.0)))
        texs = (capi.vb_texture * max(1, len(sc.textures)))()
        keep = []
        for i, t in enumerate(sc.textures):
            a = np.ascontiguousarray(t, np.float32); keep.append(a)
            texs[i].texels = a.ctypes.data; texs[i].height, texs[i].width = a.shape[0], a.shape[1]
        lib.vo_scene_set_textures(self.h, C.cast(texs, vp), len(sc.textures))
        for which, (a, (nx, ny, nz, ch)) in getattr(sc, "luts", {}).items():
            a = np.ascontiguousarray(a, np.float32)
            lib.vo_scene_set_lut(self.h, which, _p(a), nx, ny, nz, ch)
        meds = getattr(sc, "mediums", [])
        if meds:
            arr = (capi.vb_medium_rec * len(meds))(*meds)
            lib.vo_scene_set_mediums(self.h, C.cast(arr, vp), len(meds))
        if getattr(sc, "spds", None):
            vals, offs = capi.pack_spds(sc.spds)
            lib.vo_scene_set_spds.argtypes = [vp, vp, vp, C.c_uint32]; lib.vo_scene_set_spds.restype = None
            lib.vo_scene_set_spds(self.h, _p(vals), _p(offs), len(sc.spds))
        if getattr(sc, "rgb2spec", None) is not None:
            spec = capi.vb_spectrum_rec.from_buffer_copy(sc.spectrum)
            t = np.ascontiguousarray(sc.rgb2spec, np.float32)
            spec.rgb2spec = t.ctypes.data
            lib.vo_scene_set_spectrum.argtypes = [vp, vp]; lib.vo_scene_set_spectrum.restype = None
            lib.vo_scene_set_spectrum(self.h, C.addressof(spec))
        lib.vo_scene_set_camera(self.h, C.addressof(sc.camera))
        lib.vo_scene_set_integrator(self.h, C.addressof(sc.params))
        lib.vo_scene_build(self.h, threads, 1 if brute else 0)

    def __del__(self):
        if getattr(self, "h", None):
            self.o.lib.vo_scene_destroy(self.h)
            self.h = None

    def render(self, first, n, stride=1, mode=0, frames_done=0, accum=None, threads=8, stats=False):
        w, h = self.sc.resolution
        if accum is None:
            accum = np.zeros((h, w, 4), np.float32)
        st = np.zeros(5, np.uint64) if stats else None
        dt = self.o.lib.vo_render(self.h, w, h, first, n, stride, mode, frames_done, _p(accum), threads, _p(st))
        if stats:
            return accum, dt, {"closest": int(st[0]), "any": int(st[1]), "vertices": int(st[2]), "nodes": int(st[3]), "tris": int(st[4])}
        return accum, dt

    def primary_hits(self, frame, threads=8):
        w, h = self.sc.resolution
        n = w * h
        out = {"inst": np.empty(n, np.uint32), "prim": np.empty(n, np.uint32), "u": np.empty(n, np.float32),
               "v": np.empty(n, np.float32), "t": np.empty(n, np.float32)}
        self.o.lib.vo_primary_hits(self.h, w, h, frame, *[_p(out[f]) for f in ("inst", "prim", "u", "v", "t")], threads)
        return out


def oracle_gbuffer(osc, frame, threads=8, gradient=False):
    """rt_geom's G-buffer planes from the oracle: {"normal", "albedo", "emission", "geom", "motion"[, "gradient"]} -> (h, w, 4) float32;
    gradient=True adds rt_gradient's plane (normal_fwidth, depth_gradient, 0, 0)"""
    w, h = osc.sc.resolution
    out = np.empty((5, h, w, 4), np.float32)
    lib = osc.o.lib
    lib.vo_gbuffer.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]
    lib.vo_gbuffer.restype = None
    lib.vo_gbuffer(osc.h, w, h, frame, _p(out), threads)
    d = dict(zip(("normal", "albedo", "emission", "geom", "motion"), out))
    if gradient:
        g = np.empty((h, w, 4), np.float32)
        lib.vo_gbuffer_gradient.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        lib.vo_gbuffer_gradient.restype = None
        lib.vo_gbuffer_gradient(_p(np.ascontiguousarray(d["normal"])), _p(np.ascontiguousarray(d["geom"])), w, h, _p(g))
        d["gradient"] = g
    return d


def oracle_dump_path_trace(osc, px, py, frame, max_vertices=256) -> bytes:
    """raw vb_path_vertex records of one pixel-sample from the oracle (same layout as capi.Context.dump_path_trace)"""
    from vision_b200 import capi
    w, h = osc.sc.resolution
    buf = (capi.vb_path_vertex * max_vertices)()
    lib = osc.o.lib
    lib.vo_dump_path_trace.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32]
    lib.vo_dump_path_trace.restype = C.c_uint32
    n = lib.vo_dump_path_trace(osc.h, w, h, px, py, frame, buf, max_vertices)
    assert n <= max_vertices
    return bytes(buf)[:n * C.sizeof(capi.vb_path_vertex)]


def oracle_camera_prev_set(osc, pc):
    """pc: capi.vb_camera_prev or None (None = back to the static camera)"""
    lib = osc.o.lib
    lib.vo_scene_set_camera_prev.argtypes = [C.c_void_p, C.c_void_p]
    lib.vo_scene_set_camera_prev.restype = None
    lib.vo_scene_set_camera_prev(osc.h, C.byref(pc) if pc is not None else None)


def oracle_pass(o, name, src, dst=None, arg=0):
    """render-graph passes of the customized pipeline on (h, w, 4) float32 pictures: accumulate (dst updated in place, arg = frame index),
    tonemapping (arg = tone mapper), gamma"""
    lib = o.lib
    src = np.ascontiguousarray(src, np.float32)
    n = src.size // 4
    if name == "accumulate":
        lib.vo_pass_accumulate.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]; lib.vo_pass_accumulate.restype = None
        lib.vo_pass_accumulate(_p(src), _p(dst), n, arg)
        return dst
    out = np.empty_like(src)
    if name == "tonemapping":
        lib.vo_pass_tonemapping.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int]; lib.vo_pass_tonemapping.restype = None
        lib.vo_pass_tonemapping(_p(src), _p(out), n, arg)
    elif name == "gamma":
        lib.vo_pass_gamma.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]; lib.vo_pass_gamma.restype = None
        lib.vo_pass_gamma(_p(src), _p(out), n)
    else:
        raise ValueError(name)
    return out

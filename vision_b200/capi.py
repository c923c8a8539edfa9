"""ctypes binding of libvision_b200.so (the C ABI declared in include/vision_b200.h).

This is the thin Python harness over the product library: it owns no algorithm. It fails loudly
when the CUDA library is missing or no sm_100a device is present - there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libvision_b200.so"
if os.environ.get("VISION_B200_LIB"):       # experiments only (tools/sweep_*.py): an alternative build of the same library
    LIB_PATH = Path(os.environ["VISION_B200_LIB"])

VB_MEM_HOST, VB_MEM_DEVICE = 0, 1
VB_INVALID_ID = 0xFFFFFFFF

_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)
_u8p = C.POINTER(C.c_uint8)


class VbError(RuntimeError):
    pass


class vb_instance(C.Structure):
    _fields_ = [("mesh_id", C.c_uint32), ("material_id", C.c_uint32), ("light_id", C.c_uint32),
                ("inside_medium", C.c_uint32), ("outside_medium", C.c_uint32), ("lightmap_id", C.c_uint32),
                ("o2w", C.c_float * 16)]


class vb_build_opts(C.Structure):
    _fields_ = [("flags", C.c_uint32)]


class vb_build_stats(C.Structure):
    _fields_ = [("n_triangles", C.c_uint64), ("n_nodes", C.c_uint64), ("node_bytes", C.c_uint64),
                ("tri_bytes", C.c_uint64), ("build_ms", C.c_double),
                ("scene_lo", C.c_float * 3), ("scene_hi", C.c_float * 3), ("n_references", C.c_uint64)]


class vb_rays_soa(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax")] + [("mem", C.c_int)]


class vb_hits_soa(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("inst", "prim", "u", "v", "t")] + [("mem", C.c_int)]


_lib = None


def load_library() -> C.CDLL:
    """dlopen libvision_b200.so; raises VbError when it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise VbError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(vision_b200 has no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH))
    vp = C.c_void_p
    sigs = {
        "vb_ctx_create": [C.c_int, C.POINTER(vp)],
        "vb_ctx_destroy": [vp],
        "vb_ctx_set_stream": [vp, vp],
        "vb_ctx_synchronize": [vp],
        "vb_ctx_set_option": [vp, C.c_char_p, C.c_double],
        "vb_profile_reset": [vp],
        "vb_probe_read_bandwidth": [vp, C.c_uint64, C.c_int, C.POINTER(C.c_double)],
        "vb_profile_get": [vp, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_uint64)],
        "vb_mesh_upload": [vp, vp, vp, vp, C.c_uint32, vp, C.c_uint32, _u32p],
        "vb_mesh_upload_soup": [vp, vp, C.c_int, C.c_uint64, _u32p],
        "vb_instances_set": [vp, C.POINTER(vb_instance), C.c_uint32],
        "vb_meshes_clear": [vp],
        "vb_accel_build": [vp, C.POINTER(vb_build_opts), C.POINTER(vb_build_stats)],
        "vb_accel_download": [vp, vp, vp],
        "vb_trace_closest": [vp, C.POINTER(vb_rays_soa), C.c_uint64, C.POINTER(vb_hits_soa)],
        "vb_trace_any": [vp, C.POINTER(vb_rays_soa), C.c_uint64, vp, C.c_int],
        "vb_trace_closest_static": [vp, C.POINTER(vb_rays_soa), C.c_uint64, C.POINTER(vb_hits_soa)],
        "vb_trace_closest_records": [vp, vp, C.c_uint64, vp, C.c_int],
        "vb_trace_any_records": [vp, vp, C.c_uint64, vp, C.c_int],
        "vb_trace_closest_counted": [vp, C.POINTER(vb_rays_soa), C.c_uint64, C.POINTER(vb_hits_soa),
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)],
        "vb_trace_any_counted": [vp, C.POINTER(vb_rays_soa), C.c_uint64, vp, C.c_int,
                                 C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    lib.vb_last_error.argtypes = [vp]
    lib.vb_last_error.restype = C.c_char_p
    _bind_pt(lib)                       # path-tracer / film / multi-GPU signatures (defined below)
    _lib = lib
    return lib


def _ptr(a) -> int:
    """address of a numpy array (host) or torch tensor (host or device)"""
    if a is None:
        return 0
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    return a.data_ptr()   # torch tensor


def _is_device(a) -> bool:
    return (not isinstance(a, np.ndarray)) and a is not None and a.is_cuda


RAY_FIELDS = ("ox", "oy", "oz", "tmin", "dx", "dy", "dz", "tmax")
HIT_FIELDS = ("inst", "prim", "u", "v", "t")


class Context:
    """One vb_ctx (one GPU). Methods map 1:1 onto the C ABI."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.vb_ctx_create(device, C.byref(h))
        if rc != 0:
            raise VbError(f"vb_ctx_create failed ({rc}): {self.lib.vb_last_error(None).decode()}")
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.vb_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, what: str):
        if rc != 0:
            raise VbError(f"{what} failed ({rc}): {self.lib.vb_last_error(self.h).decode()}")

    # -- stream / profiling -------------------------------------------------------------
    def set_stream(self, cuda_stream_ptr: int | None):
        self._check(self.lib.vb_ctx_set_stream(self.h, C.c_void_p(cuda_stream_ptr or 0)), "vb_ctx_set_stream")

    def synchronize(self):
        self._check(self.lib.vb_ctx_synchronize(self.h), "vb_ctx_synchronize")

    def set_option(self, key: str, value: float):
        self._check(self.lib.vb_ctx_set_option(self.h, key.encode(), float(value)), "vb_ctx_set_option")

    def profile_reset(self):
        self._check(self.lib.vb_profile_reset(self.h), "vb_profile_reset")

    def probe_read_bandwidth(self, nbytes: int, reps: int = 20) -> float:
        g = C.c_double()
        self._check(self.lib.vb_probe_read_bandwidth(self.h, nbytes, reps, C.byref(g)), "vb_probe_read_bandwidth")
        return g.value

    def profile_get(self, name: str):
        ms, n = C.c_double(), C.c_uint64()
        self._check(self.lib.vb_profile_get(self.h, name.encode(), C.byref(ms), C.byref(n)), "vb_profile_get")
        return ms.value, n.value

    # -- geometry ---------------------------------------------------------------------
    def meshes_clear(self):
        self._check(self.lib.vb_meshes_clear(self.h), "vb_meshes_clear")

    def mesh_upload(self, pos, tri, nrm=None, uv=None) -> int:
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        tri = np.ascontiguousarray(tri, np.uint32).reshape(-1, 3)
        nrm = None if nrm is None else np.ascontiguousarray(nrm, np.float32).reshape(-1, 3)
        uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
        mid = C.c_uint32()
        self._check(self.lib.vb_mesh_upload(self.h, _ptr(pos), _ptr(nrm), _ptr(uv), len(pos), _ptr(tri), len(tri),
                                            C.byref(mid)), "vb_mesh_upload")
        return mid.value

    def mesh_upload_soup(self, pos9) -> int:
        """pos9: (nt, 9) float32 numpy array or CUDA torch tensor"""
        nt = pos9.shape[0]
        if isinstance(pos9, np.ndarray):
            pos9 = np.ascontiguousarray(pos9, np.float32)
        mid = C.c_uint32()
        self._check(self.lib.vb_mesh_upload_soup(self.h, _ptr(pos9), VB_MEM_DEVICE if _is_device(pos9) else VB_MEM_HOST,
                                                 nt, C.byref(mid)), "vb_mesh_upload_soup")
        return mid.value

    def instances_set(self, instances):
        """instances: list of dicts(mesh_id, o2w (16 floats column-major), material_id, light_id, ...)"""
        arr = (vb_instance * len(instances))()
        for i, d in enumerate(instances):
            arr[i].mesh_id = d["mesh_id"]
            arr[i].material_id = d.get("material_id", VB_INVALID_ID)
            arr[i].light_id = d.get("light_id", VB_INVALID_ID)
            arr[i].inside_medium = d.get("inside_medium", VB_INVALID_ID)
            arr[i].outside_medium = d.get("outside_medium", VB_INVALID_ID)
            arr[i].lightmap_id = d.get("lightmap_id", VB_INVALID_ID)
            o2w = d.get("o2w")
            if o2w is None:
                o2w = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]
            arr[i].o2w[:] = [float(x) for x in np.asarray(o2w, np.float32).reshape(-1)]
        self._check(self.lib.vb_instances_set(self.h, arr, len(instances)), "vb_instances_set")

    def accel_build(self, two_level: bool = False) -> dict:
        """two_level=True: VB_BUILD_TWO_LEVEL for this build (otherwise the context option build.two_level decides)"""
        st = vb_build_stats()
        opts = vb_build_opts(1 if two_level else 0)
        self._check(self.lib.vb_accel_build(self.h, C.byref(opts), C.byref(st)), "vb_accel_build")
        return {"n_triangles": st.n_triangles, "n_nodes": st.n_nodes, "node_bytes": st.node_bytes,
                "tri_bytes": st.tri_bytes, "build_ms": st.build_ms, "n_references": st.n_references,
                "scene_lo": list(st.scene_lo), "scene_hi": list(st.scene_hi)}

    def accel_download(self, n_nodes: int, n_tris: int):
        nodes = np.empty((n_nodes, 96), np.uint8)
        tris = np.empty((n_tris, 12), np.float32)
        self._check(self.lib.vb_accel_download(self.h, _ptr(nodes), _ptr(tris)), "vb_accel_download")
        return nodes, tris

    # -- trace ------------------------------------------------------------------------
    @staticmethod
    def _rays_struct(rays: dict) -> vb_rays_soa:
        r = vb_rays_soa()
        dev = _is_device(rays["ox"])
        for f in RAY_FIELDS:
            a = rays[f]
            assert _is_device(a) == dev, "all ray arrays must live in the same memory space"
            setattr(r, f, _ptr(a))
        r.mem = VB_MEM_DEVICE if dev else VB_MEM_HOST
        return r

    def trace_closest(self, rays: dict, out: dict | None = None, counted: bool = False):
        """rays/out: dicts of numpy arrays (host) or CUDA torch tensors (device), SoA."""
        n = rays["ox"].shape[0]
        if out is None:
            if _is_device(rays["ox"]):
                import torch
                dev = rays["ox"].device
                out = {"inst": torch.empty(n, dtype=torch.int32, device=dev), "prim": torch.empty(n, dtype=torch.int32, device=dev),
                       "u": torch.empty(n, dtype=torch.float32, device=dev), "v": torch.empty(n, dtype=torch.float32, device=dev),
                       "t": torch.empty(n, dtype=torch.float32, device=dev)}
            else:
                out = {"inst": np.empty(n, np.uint32), "prim": np.empty(n, np.uint32), "u": np.empty(n, np.float32),
                       "v": np.empty(n, np.float32), "t": np.empty(n, np.float32)}
        r = self._rays_struct(rays)
        h = vb_hits_soa()
        for f in HIT_FIELDS:
            setattr(h, f, _ptr(out[f]))
        h.mem = VB_MEM_DEVICE if _is_device(out["inst"]) else VB_MEM_HOST
        if counted:
            a, b = C.c_uint64(), C.c_uint64()
            self._check(self.lib.vb_trace_closest_counted(self.h, C.byref(r), n, C.byref(h), C.byref(a), C.byref(b)),
                        "vb_trace_closest_counted")
            return out, a.value, b.value
        self._check(self.lib.vb_trace_closest(self.h, C.byref(r), n, C.byref(h)), "vb_trace_closest")
        return out

    def trace_closest_static(self, rays: dict, out: dict):
        """vb_trace_closest_static: device buffers only, n <= VB_TRACE_STATIC_MAX_RAYS, allocates nothing (the caller passes `out`)"""
        n = rays["ox"].shape[0]
        r = self._rays_struct(rays)
        h = vb_hits_soa()
        for f in HIT_FIELDS:
            setattr(h, f, _ptr(out[f]))
        h.mem = VB_MEM_DEVICE if _is_device(out["inst"]) else VB_MEM_HOST
        self._check(self.lib.vb_trace_closest_static(self.h, C.byref(r), n, C.byref(h)), "vb_trace_closest_static")
        return out

    def trace_any(self, rays: dict, out=None, counted: bool = False):
        n = rays["ox"].shape[0]
        if out is None:
            if _is_device(rays["ox"]):
                import torch
                out = torch.empty(n, dtype=torch.uint8, device=rays["ox"].device)
            else:
                out = np.empty(n, np.uint8)
        r = self._rays_struct(rays)
        mem = VB_MEM_DEVICE if _is_device(out) else VB_MEM_HOST
        if counted:
            a, b = C.c_uint64(), C.c_uint64()
            self._check(self.lib.vb_trace_any_counted(self.h, C.byref(r), n, _ptr(out), mem, C.byref(a), C.byref(b)),
                        "vb_trace_any_counted")
            return out, a.value, b.value
        self._check(self.lib.vb_trace_any(self.h, C.byref(r), n, _ptr(out), mem), "vb_trace_any")
        return out


    # -- record form: rays (n, 8) float32 = vb_ray, hits (n, 4) uint32 = vb_hit {inst, prim, u bits, v bits} ---------------
    def trace_closest_records(self, rays, hits=None):
        """rays: (n, 8) float32 numpy array (host, ideally pinned) or CUDA torch tensor; returns hits (n, 4) uint32 / int32"""
        n = rays.shape[0]
        dev = _is_device(rays)
        if hits is None:
            if dev:
                import torch
                hits = torch.empty((n, 4), dtype=torch.int32, device=rays.device)
            else:
                hits = np.empty((n, 4), np.uint32)
        assert _is_device(hits) == dev, "rays and hits must live in the same memory space"
        self._check(self.lib.vb_trace_closest_records(self.h, _ptr(rays), n, _ptr(hits), VB_MEM_DEVICE if dev else VB_MEM_HOST),
                    "vb_trace_closest_records")
        return hits

    def trace_any_records(self, rays, occ=None):
        n = rays.shape[0]
        dev = _is_device(rays)
        if occ is None:
            if dev:
                import torch
                occ = torch.empty(n, dtype=torch.uint8, device=rays.device)
            else:
                occ = np.empty(n, np.uint8)
        assert _is_device(occ) == dev, "rays and occlusion bytes must live in the same memory space"
        self._check(self.lib.vb_trace_any_records(self.h, _ptr(rays), n, _ptr(occ), VB_MEM_DEVICE if dev else VB_MEM_HOST),
                    "vb_trace_any_records")
        return occ


def rays_to_records(rays: dict, tmax_key="tmax") -> np.ndarray:
    """SoA dict -> (n, 8) float32 vb_ray records"""
    return np.ascontiguousarray(np.stack([rays[k] for k in ("ox", "oy", "oz", "tmin", "dx", "dy", "dz")] + [rays[tmax_key]], axis=1), np.float32)


# ---------------------------------------------------------------------------------------------
# path-tracing records (include/vision_b200_pt.h)
# ---------------------------------------------------------------------------------------------
class vb_slot(C.Structure):
    _fields_ = [("value", C.c_float * 4), ("tex_id", C.c_uint32), ("pad", C.c_uint32 * 3)]


class vb_texture(C.Structure):
    _fields_ = [("texels", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32)]


class vb_material_rec(C.Structure):
    _fields_ = [("type", C.c_uint32), ("flags", C.c_uint32), ("spd", C.c_uint32 * 2),
                ("color", vb_slot), ("roughness", vb_slot), ("spec", vb_slot), ("ior", vb_slot), ("sigma", vb_slot),
                ("eta", C.c_float * 4), ("k", C.c_float * 4), ("params", C.c_float * 16)]


class vb_medium_rec(C.Structure):
    _fields_ = [("type", C.c_uint32), ("g", C.c_float), ("scale", C.c_float), ("pad", C.c_float),
                ("sigma_a", C.c_float * 4), ("sigma_s", C.c_float * 4)]


VB_LUT_PURE_REFLECTION, VB_LUT_DIELECTRIC, VB_LUT_DIELECTRIC_INV, VB_LUT_SPECULAR, VB_LUT_COAT, VB_LUT_SHEEN_VOLUME, VB_LUT_SHEEN_APPROX = range(7)
VB_LUT_COUNT = 7
LUT_FILES = {VB_LUT_PURE_REFLECTION: ("lut_pure_reflection_32x32x1x1.f32", (32, 32, 1, 1)),
             VB_LUT_DIELECTRIC: ("lut_dielectric_32x32x32x2.f32", (32, 32, 32, 2)),
             VB_LUT_DIELECTRIC_INV: ("lut_dielectric_inv_32x32x32x2.f32", (32, 32, 32, 2)),
             VB_LUT_SPECULAR: ("lut_specular_32x32x32x1.f32", (32, 32, 32, 1)),
             VB_LUT_COAT: ("lut_coat_32x32x32x1.f32", (32, 32, 32, 1)),
             VB_LUT_SHEEN_VOLUME: ("lut_sheen_volume_32x32x1x4.f32", (32, 32, 1, 4)),
             VB_LUT_SHEEN_APPROX: ("lut_sheen_approx_32x32x1x4.f32", (32, 32, 1, 4))}


def load_luts(principled=False) -> dict:
    """the energy-compensation tables shipped under vision_b200/data (tools/extract_luts.py): the three microfacet
    tables, plus the principled BSDF's specular / coat / sheen tables when asked"""
    from pathlib import Path
    d = Path(__file__).resolve().parent / "data"
    return {k: (np.fromfile(d / fn, np.float32), dims) for k, (fn, dims) in LUT_FILES.items() if principled or k <= VB_LUT_DIELECTRIC_INV}


class vb_alias_entry(C.Structure):
    _fields_ = [("prob", C.c_float), ("alias", C.c_uint32)]


class vb_light_rec(C.Structure):
    _fields_ = [("type", C.c_uint32), ("inst_id", C.c_uint32), ("two_sided", C.c_uint32), ("alias_offset", C.c_uint32),
                ("alias_size", C.c_uint32), ("alias_integral", C.c_float), ("scale", C.c_float), ("pad0", C.c_float),
                ("color", vb_slot), ("position", C.c_float * 4), ("direction", C.c_float * 4), ("params", C.c_float * 8),
                ("env_res", C.c_uint32 * 2), ("cond_offset", C.c_uint32), ("pad1", C.c_uint32),
                ("w2o", C.c_float * 16), ("o2w", C.c_float * 16)]


class vb_camera(C.Structure):
    _fields_ = [("c2w", C.c_float * 16), ("tan_fov_y_over_2", C.c_float), ("lens_radius", C.c_float),
                ("focal_distance", C.c_float), ("medium_id", C.c_uint32), ("screen_lo", C.c_float * 2),
                ("screen_hi", C.c_float * 2), ("filter_type", C.c_uint32), ("filter_radius", C.c_float * 2),
                ("filter_table_offset", C.c_uint32)]


class vb_camera_prev(C.Structure):
    _fields_ = [("w2c", C.c_float * 16), ("tan_fov_y_over_2", C.c_float), ("screen_lo", C.c_float * 2), ("screen_hi", C.c_float * 2),
                ("pad", C.c_float * 3)]


class vb_path_vertex(C.Structure):
    _fields_ = [("ray_o", C.c_float * 3), ("ray_tmin", C.c_float), ("ray_d", C.c_float * 3), ("ray_tmax", C.c_float),
                ("inst", C.c_uint32), ("prim", C.c_uint32), ("u", C.c_float), ("v", C.c_float),
                ("t", C.c_float), ("bounces", C.c_uint32), ("sampler_state", C.c_uint32), ("medium", C.c_uint32),
                ("throughput", C.c_float * 3), ("scatter_pdf", C.c_float), ("radiance", C.c_float * 3), ("eta_scale", C.c_float)]


VB_PATH_END = 0xFFFFFFFE


class vb_pt_params(C.Structure):
    _fields_ = [("max_depth", C.c_uint32), ("min_depth", C.c_uint32), ("rr_threshold", C.c_float),
                ("mis_mode", C.c_uint32), ("spectrum", C.c_uint32), ("sampler", C.c_uint32),
                ("process_mediums", C.c_uint32), ("ray_offset_factor", C.c_float)]


class vb_spectrum_rec(C.Structure):
    _fields_ = [("type", C.c_uint32), ("dimension", C.c_uint32), ("spd_x", C.c_uint32), ("spd_y", C.c_uint32),
                ("spd_z", C.c_uint32), ("spd_d65", C.c_uint32), ("cie_y_integral", C.c_float), ("rgb2spec_res", C.c_uint32),
                ("rgb2spec", C.c_void_p)]


VB_SPECTRUM_SRGB, VB_SPECTRUM_HERO = 0, 1
RGB2SPEC_FILE, RGB2SPEC_RES = "rgb2spec_srgb_3x64x64x64x4.f32", 64


def load_rgb2spec() -> np.ndarray:
    """the sRGB -> sigmoid-polynomial coefficient table of the hero spectrum, float[3][64][64][64][4]; written by
    build() (tools/rgb2spec_opt.c - the reference's srgb2spec.h is not in its tree)"""
    from pathlib import Path
    f = Path(__file__).resolve().parent / "data" / RGB2SPEC_FILE
    if not f.exists():
        raise FileNotFoundError(f"{f} is missing: run __graft_entry__.build() (tools/rgb2spec_opt.c) first")
    return np.fromfile(f, np.float32)


def load_cie_1nm() -> np.ndarray:
    """CIE 1931 X, Y, Z and D65 at 1 nm, 360..830 nm (4, 471) float32 (tools/extract_cie.py)"""
    from pathlib import Path
    return np.fromfile(Path(__file__).resolve().parent / "data" / "cie_1nm_4x471.f32", np.float32).reshape(4, 471)


VB_MAT_DIFFUSE, VB_MAT_MIRROR, VB_MAT_GLASS, VB_MAT_METAL, VB_MAT_SUBSTRATE, VB_MAT_PRINCIPLED, VB_MAT_BLACK_BODY = range(7)
VB_MAT_PLASTIC, VB_MAT_METALLIC, VB_MAT_MIX, VB_MAT_ADD = 7, 8, 9, 10
VB_TRACE_STATIC_MAX_RAYS = 1 << 18
VB_LIGHT_AREA, VB_LIGHT_POINT, VB_LIGHT_SPOT, VB_LIGHT_ENV, VB_LIGHT_PROJECTOR, VB_LIGHT_DIRECTIONAL = range(6)
VB_LIGHT_SAMPLER_UNIFORM, VB_LIGHT_SAMPLER_POWER = 0, 1
VB_FILTER_BOX, VB_FILTER_TRIANGLE, VB_FILTER_FITTED = 0, 1, 2
VB_FILTER_TABLE_SIZE = 20
VB_ACCUM_LERP, VB_ACCUM_SUM = 0, 1
VB_FILM_RT, VB_FILM_ACCUM, VB_FILM_OUTPUT, VB_FILM_VIEW = 0, 1, 2, 3
VB_TONEMAP_LINEAR, VB_TONEMAP_ACES, VB_TONEMAP_REINHARD = 0, 1, 2


def make_slot(value, tex_id=VB_INVALID_ID) -> vb_slot:
    s = vb_slot()
    v = list(value) + [1.0] * (4 - len(value))
    s.value[:] = [float(x) for x in v[:4]]
    s.tex_id = tex_id
    return s


def _bind_pt(lib):
    vp = C.c_void_p
    sigs = {
        "vb_textures_set": [vp, C.POINTER(vb_texture), C.c_uint32],
        "vb_materials_set": [vp, C.POINTER(vb_material_rec), C.c_uint32],
        "vb_luts_set": [vp, C.c_int, vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32],
        "vb_mediums_set": [vp, C.POINTER(vb_medium_rec), C.c_uint32],
        "vb_spds_set": [vp, vp, vp, C.c_uint32],
        "vb_spectrum_set": [vp, C.POINTER(vb_spectrum_rec)],
        "vb_lights_set": [vp, C.POINTER(vb_light_rec), C.c_uint32, C.POINTER(vb_alias_entry), vp, C.c_uint32],
        "vb_light_sampler_set": [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float],
        "vb_light_sampler_env_set": [vp, C.c_uint32, C.c_float],
        "vb_camera_set": [vp, C.POINTER(vb_camera)],
        "vb_integrator_set": [vp, C.POINTER(vb_pt_params)],
        "vb_film_create": [vp, C.c_uint32, C.c_uint32],
        "vb_film_clear": [vp],
        "vb_render_frames": [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int],
        "vb_film_resolve": [vp, C.c_int, C.c_float, C.c_int],
        "vb_film_download": [vp, vp, C.c_int],
        "vb_film_device_ptr": [vp, C.c_int, C.POINTER(vp)],
        "vb_dump_primary_hits": [vp, C.c_uint32, C.POINTER(vb_hits_soa)],
        "vb_render_stats": [vp, C.POINTER(C.c_uint64)],
        "vb_dump_path_trace": [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, C.c_uint32, C.POINTER(C.c_uint32)],
        "vb_gbuffer_render": [vp, C.c_uint32],
        "vb_gbuffer_download": [vp, C.c_int, vp],
        "vb_gbuffer_gradient": [vp],
        "vb_camera_prev_set": [vp, C.POINTER(vb_camera_prev)],
        "vb_pass_tonemapping": [vp, C.c_int],
        "vb_pass_gamma": [vp],
        "vb_render_work": [vp, C.POINTER(C.c_uint64)],
        "vb_nccl_unique_id": [vp],
        "vb_nccl_comm_create": [vp, vp, C.c_int, C.c_int, C.POINTER(vp)],
        "vb_nccl_comm_destroy": [vp, vp],
        "vb_film_allreduce": [vp, vp],
        "vb_film_reduce_peers": [C.POINTER(vp), C.c_uint32, C.c_uint32, C.c_float, C.c_int],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int


class SceneRecords:
    """Plain container of the POD records of one scene, uploadable to a Context (GPU) or to the oracle."""

    def __init__(self):
        self.meshes = []        # dicts: pos (nv,3), nrm (nv,3), uv (nv,2), tri (nt,3)
        self.instances = []     # dicts as Context.instances_set
        self.materials = []     # vb_material_rec
        self.lights = []        # vb_light_rec
        self.alias_table = []   # (prob, alias)
        self.alias_func = []    # float
        self.light_sampler = (0, 0, 0, 0.0)   # (kind, alias_offset, alias_size, alias_integral): vb_light_sampler_set
        self.light_sampler_env = (0, 0.0)     # (env_separate, env_prob): vb_light_sampler_env_set
        self.textures = []      # (h, w, 4) float32 arrays
        self.mediums = []       # vb_medium_rec
        self.spds = []          # SPD pool: float32 arrays (hero spectrum: CIE x, y, z, D65 first, then material curves)
        self.spectrum = vb_spectrum_rec()   # rgb2spec is filled at upload time from `rgb2spec` below
        self.rgb2spec = None    # float32 array 3 * res^3 * 4 (hero spectrum)
        self.luts = {}          # which -> (float32 array, (nx, ny, nz, channels)); filled when a material needs them
        self.camera = vb_camera()
        self.params = vb_pt_params()
        self.resolution = (256, 256)
        self.exposure = 1.0
        self.tonemap = VB_TONEMAP_LINEAR
        self.spp = 1


def pack_spds(spds):
    """SPD pool -> (values float32, offsets uint32[n + 1]) as vb_spds_set takes them"""
    offs = np.zeros(len(spds) + 1, np.uint32)
    for i, s in enumerate(spds):
        offs[i + 1] = offs[i] + len(s)
    vals = np.ascontiguousarray(np.concatenate([np.asarray(s, np.float32) for s in spds]) if spds else np.zeros(1, np.float32), np.float32)
    return vals, offs


def _pt_methods():
    def upload_scene(self, sc: SceneRecords):
        """meshes + instances + BVH + tables + camera + film"""
        _bind_pt(self.lib)
        self.meshes_clear()                  # a reused context drops the previous scene's meshes (mesh ids restart at 0)
        ids = [self.mesh_upload(m["pos"], m["tri"], m.get("nrm"), m.get("uv")) for m in sc.meshes]
        inst = [dict(d, mesh_id=ids[d["mesh_id"]]) for d in sc.instances]
        self.instances_set(inst)
        st = self.accel_build()
        texs = (vb_texture * max(1, len(sc.textures)))()
        keep = []
        for i, t in enumerate(sc.textures):
            a = np.ascontiguousarray(t, np.float32)
            keep.append(a)
            texs[i].texels = a.ctypes.data; texs[i].height, texs[i].width = a.shape[0], a.shape[1]
        self._check(self.lib.vb_textures_set(self.h, texs, len(sc.textures)), "vb_textures_set")
        mats = (vb_material_rec * max(1, len(sc.materials)))(*sc.materials)
        self._check(self.lib.vb_materials_set(self.h, mats, len(sc.materials)), "vb_materials_set")
        for which in range(VB_LUT_COUNT):
            if which in sc.luts:
                a, (nx, ny, nz, ch) = sc.luts[which]
                a = np.ascontiguousarray(a, np.float32)
                self._check(self.lib.vb_luts_set(self.h, which, a.ctypes.data, nx, ny, nz, ch), "vb_luts_set")
            else:
                self._check(self.lib.vb_luts_set(self.h, which, None, 0, 0, 0, 0), "vb_luts_set")
        meds = (vb_medium_rec * max(1, len(sc.mediums)))(*sc.mediums)
        self._check(self.lib.vb_mediums_set(self.h, meds, len(sc.mediums)), "vb_mediums_set")
        vals, offs = pack_spds(sc.spds)
        self._check(self.lib.vb_spds_set(self.h, vals.ctypes.data, offs.ctypes.data, len(sc.spds)), "vb_spds_set")
        spec = vb_spectrum_rec.from_buffer_copy(sc.spectrum)
        t = None
        if sc.rgb2spec is not None:
            t = np.ascontiguousarray(sc.rgb2spec, np.float32)
            spec.rgb2spec = t.ctypes.data
        self._check(self.lib.vb_spectrum_set(self.h, C.byref(spec)), "vb_spectrum_set")
        lights = (vb_light_rec * max(1, len(sc.lights)))(*sc.lights)
        na = len(sc.alias_table)
        tab = (vb_alias_entry * max(1, na))(*[vb_alias_entry(p, a) for p, a in sc.alias_table])
        func = np.ascontiguousarray(sc.alias_func if na else [0.0], np.float32)
        self._check(self.lib.vb_lights_set(self.h, lights, len(sc.lights), tab, func.ctypes.data, na), "vb_lights_set")
        k, off, size, integral = sc.light_sampler
        self._check(self.lib.vb_light_sampler_set(self.h, k, off, size, integral), "vb_light_sampler_set")
        sep, prob = getattr(sc, "light_sampler_env", (0, 0.0))
        self._check(self.lib.vb_light_sampler_env_set(self.h, sep, prob), "vb_light_sampler_env_set")
        self._check(self.lib.vb_camera_set(self.h, C.byref(sc.camera)), "vb_camera_set")
        self._check(self.lib.vb_camera_prev_set(self.h, None), "vb_camera_prev_set")          # a new scene starts with a static camera
        self._check(self.lib.vb_integrator_set(self.h, C.byref(sc.params)), "vb_integrator_set")
        self._check(self.lib.vb_film_create(self.h, sc.resolution[0], sc.resolution[1]), "vb_film_create")
        self._res = sc.resolution
        return st

    def render_frames(self, first, n, stride=1, mode=VB_ACCUM_LERP):
        self._check(self.lib.vb_render_frames(self.h, first, n, stride, mode), "vb_render_frames")

    def film_clear(self):
        self._check(self.lib.vb_film_clear(self.h), "vb_film_clear")

    def film_resolve(self, mode=VB_ACCUM_LERP, exposure=1.0, tonemap=VB_TONEMAP_LINEAR):
        self._check(self.lib.vb_film_resolve(self.h, mode, exposure, tonemap), "vb_film_resolve")

    def film_download(self, which=VB_FILM_ACCUM):
        w, h = self._res
        out = np.empty((h, w, 4), np.float32)
        self._check(self.lib.vb_film_download(self.h, out.ctypes.data, which), "vb_film_download")
        return out

    def film_device_ptr(self, which=VB_FILM_ACCUM) -> int:
        p = C.c_void_p()
        self._check(self.lib.vb_film_device_ptr(self.h, which, C.byref(p)), "vb_film_device_ptr")
        return p.value

    def gbuffer(self, frame, gradient=False) -> dict:
        """rt_geom's G-buffer planes of one frame index: {"normal", "albedo", "emission", "geom", "motion"[, "gradient"]} -> (h, w, 4) float32;
        gradient=True also runs rt_gradient (vb_gbuffer_gradient)"""
        self._check(self.lib.vb_gbuffer_render(self.h, frame), "vb_gbuffer_render")
        if gradient:
            self._check(self.lib.vb_gbuffer_gradient(self.h), "vb_gbuffer_gradient")
        w, h = self._res
        out = {}
        for which, name in enumerate(("normal", "albedo", "emission", "geom", "motion") + (("gradient",) if gradient else ())):
            a = np.empty((h, w, 4), np.float32)
            self._check(self.lib.vb_gbuffer_download(self.h, which, a.ctypes.data), "vb_gbuffer_download")
            out[name] = a
        return out

    def camera_prev_set(self, w2c, tan_fov_y_over_2, screen_lo, screen_hi):
        """Sensor::store_prev_data: the previous frame's camera (w2c = inverse of its camera_to_world, 16 floats column-major)"""
        pc = vb_camera_prev()
        pc.w2c[:] = [float(x) for x in np.asarray(w2c, np.float32).reshape(16)]
        pc.tan_fov_y_over_2 = float(tan_fov_y_over_2)
        pc.screen_lo[:] = [float(x) for x in screen_lo]
        pc.screen_hi[:] = [float(x) for x in screen_hi]
        self._check(self.lib.vb_camera_prev_set(self.h, C.byref(pc)), "vb_camera_prev_set")
        return pc

    def pass_tonemapping(self, tonemap):
        self._check(self.lib.vb_pass_tonemapping(self.h, tonemap), "vb_pass_tonemapping")

    def pass_gamma(self):
        self._check(self.lib.vb_pass_gamma(self.h), "vb_pass_gamma")

    def dump_path_trace(self, px, py, frame, max_vertices=256) -> bytes:
        """the path of one pixel-sample as raw vb_path_vertex records (96 bytes each, the VB_PATH_END record last)"""
        buf = (vb_path_vertex * max_vertices)()
        n = C.c_uint32()
        self._check(self.lib.vb_dump_path_trace(self.h, px, py, frame, buf, max_vertices, C.byref(n)), "vb_dump_path_trace")
        if n.value > max_vertices:
            raise VbError(f"vb_dump_path_trace: the path has {n.value} vertices, room for {max_vertices}")
        return bytes(buf)[:n.value * C.sizeof(vb_path_vertex)]

    def dump_primary_hits(self, frame):
        w, h = self._res
        n = w * h
        out = {"inst": np.empty(n, np.uint32), "prim": np.empty(n, np.uint32), "u": np.empty(n, np.float32),
               "v": np.empty(n, np.float32), "t": np.empty(n, np.float32)}
        hs = vb_hits_soa()
        for f in HIT_FIELDS:
            setattr(hs, f, _ptr(out[f]))
        hs.mem = VB_MEM_HOST
        self._check(self.lib.vb_dump_primary_hits(self.h, frame, C.byref(hs)), "vb_dump_primary_hits")
        return out

    def nccl_comm_create(self, unique_id: bytes, n_ranks: int, rank: int) -> int:
        comm = C.c_void_p()
        buf = (C.c_ubyte * 128).from_buffer_copy(unique_id)
        self._check(self.lib.vb_nccl_comm_create(self.h, buf, n_ranks, rank, C.byref(comm)), "vb_nccl_comm_create")
        return comm.value

    def nccl_comm_destroy(self, comm: int):
        self._check(self.lib.vb_nccl_comm_destroy(self.h, C.c_void_p(comm)), "vb_nccl_comm_destroy")

    def film_allreduce(self, comm: int):
        self._check(self.lib.vb_film_allreduce(self.h, C.c_void_p(comm)), "vb_film_allreduce")

    def render_stats(self):
        st = (C.c_uint64 * 4)()
        self._check(self.lib.vb_render_stats(self.h, st), "vb_render_stats")
        return {"closest": st[0], "any": st[1], "vertices": st[2], "culled": st[3]}

    def render_work(self):
        w = (C.c_uint64 * 4)()
        self._check(self.lib.vb_render_work(self.h, w), "vb_render_work")
        return {"closest_nodes": w[0], "closest_tris": w[1], "any_nodes": w[2], "any_tris": w[3]}

    for f in (render_work, upload_scene, render_frames, film_clear, film_resolve, film_download, film_device_ptr,
              dump_primary_hits, dump_path_trace, render_stats, gbuffer, camera_prev_set, pass_tonemapping, pass_gamma, nccl_comm_create, nccl_comm_destroy, film_allreduce):
        setattr(Context, f.__name__, f)


_pt_methods()


def nccl_unique_id() -> bytes:
    """ncclGetUniqueId through the library's run-time NCCL binding (rank 0 calls it and ships the 128 bytes to every rank)"""
    lib = load_library()
    buf = (C.c_ubyte * 128)()
    if lib.vb_nccl_unique_id(buf) != 0:
        raise VbError("vb_nccl_unique_id: NCCL unavailable")
    return bytes(buf)


def film_reduce_peers(ctxs, root=0, exposure=1.0, tonemap=VB_TONEMAP_LINEAR):
    """one process, several contexts on different GPUs: fused peer-memory reduce + resolve into ctxs[root]'s films"""
    lib = load_library()
    arr = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    rc = lib.vb_film_reduce_peers(arr, len(ctxs), root, exposure, tonemap)
    if rc != 0:
        raise VbError(f"vb_film_reduce_peers failed ({rc}): {lib.vb_last_error(ctxs[root].h).decode()}")

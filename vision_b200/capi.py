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
                ("scene_lo", C.c_float * 3), ("scene_hi", C.c_float * 3)]


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
        "vb_profile_get": [vp, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_uint64)],
        "vb_mesh_upload": [vp, vp, vp, vp, C.c_uint32, vp, C.c_uint32, _u32p],
        "vb_mesh_upload_soup": [vp, vp, C.c_int, C.c_uint64, _u32p],
        "vb_instances_set": [vp, C.POINTER(vb_instance), C.c_uint32],
        "vb_accel_build": [vp, C.POINTER(vb_build_opts), C.POINTER(vb_build_stats)],
        "vb_accel_download": [vp, vp, vp],
        "vb_trace_closest": [vp, C.POINTER(vb_rays_soa), C.c_uint64, C.POINTER(vb_hits_soa)],
        "vb_trace_any": [vp, C.POINTER(vb_rays_soa), C.c_uint64, vp, C.c_int],
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

    def profile_get(self, name: str):
        ms, n = C.c_double(), C.c_uint64()
        self._check(self.lib.vb_profile_get(self.h, name.encode(), C.byref(ms), C.byref(n)), "vb_profile_get")
        return ms.value, n.value

    # -- geometry ---------------------------------------------------------------------
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

    def accel_build(self) -> dict:
        st = vb_build_stats()
        opts = vb_build_opts(0)
        self._check(self.lib.vb_accel_build(self.h, C.byref(opts), C.byref(st)), "vb_accel_build")
        return {"n_triangles": st.n_triangles, "n_nodes": st.n_nodes, "node_bytes": st.node_bytes,
                "tri_bytes": st.tri_bytes, "build_ms": st.build_ms,
                "scene_lo": list(st.scene_lo), "scene_hi": list(st.scene_hi)}

    def accel_download(self, n_nodes: int, n_tris: int):
        nodes = np.empty((n_nodes, 80), np.uint8)
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

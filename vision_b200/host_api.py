"""ctypes binding of libvision_host.so - the C++ host layer (scene JSON -> desc -> plugins -> C ABI)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from . import capi

LIB_PATH = Path(__file__).resolve().parent / "libvision_host.so"
_lib = None


def load():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise capi.VbError(f"{LIB_PATH} missing: run __graft_entry__.build()")
        capi.load_library()
        lib = C.CDLL(str(LIB_PATH))
        vp = C.c_void_p
        lib.vh_last_error.restype = C.c_char_p
        lib.vh_import_scene.argtypes = [C.c_char_p]; lib.vh_import_scene.restype = vp
        lib.vh_destroy.argtypes = [vp]; lib.vh_destroy.restype = None
        lib.vh_override.argtypes = [vp, C.c_char_p, C.c_double, C.c_double]
        lib.vh_prepare_records.argtypes = [vp]
        lib.vh_prepare.argtypes = [vp, C.c_int]
        lib.vh_render.argtypes = [vp, C.c_uint32]
        lib.vh_final_picture.argtypes = [vp, vp]
        lib.vh_download_accum.argtypes = [vp, vp]
        lib.vh_info.argtypes = [vp, vp, vp]; lib.vh_info.restype = None
        lib.vh_records_blob.argtypes = [vp, C.c_int, vp, C.c_uint64]; lib.vh_records_blob.restype = C.c_uint64
        lib.vh_save_image.argtypes = [C.c_char_p, vp, C.c_uint32, C.c_uint32]; lib.vh_save_image.restype = C.c_int
        lib.vh_load_image.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), vp, C.c_uint64]; lib.vh_load_image.restype = C.c_int
        lib.vh_rgbe_encode.argtypes = [vp, vp]; lib.vh_rgbe_encode.restype = None
        lib.vh_rgbe_decode.argtypes = [vp, vp]; lib.vh_rgbe_decode.restype = None
        lib.vh_mesh_info.argtypes = [vp, C.c_uint32, vp, vp]; lib.vh_mesh_info.restype = None
        lib.vh_mesh_copy.argtypes = [vp, C.c_uint32, vp, vp, vp, vp]; lib.vh_mesh_copy.restype = None
        lib.vh_texture_info.argtypes = [vp, C.c_uint32, vp, vp]; lib.vh_texture_info.restype = None
        lib.vh_texture_copy.argtypes = [vp, C.c_uint32, vp]; lib.vh_texture_copy.restype = None
        lib.vh_plugin_names.argtypes = [vp, C.c_uint64]; lib.vh_plugin_names.restype = C.c_uint64
        lib.vh_scene_log.argtypes = [vp, vp, C.c_uint64]; lib.vh_scene_log.restype = C.c_uint64
        lib.vh_ctx.argtypes = [vp]; lib.vh_ctx.restype = vp
        lib.vh_set_partition.argtypes = [vp, C.c_uint32, C.c_uint32]
        lib.vh_nccl_init.argtypes = [vp, vp]
        lib.vh_film_allreduce.argtypes = [vp]
        lib.vh_synchronize.argtypes = [vp]
        lib.vh_invalidate.argtypes = [vp]
        lib.vh_build_stats.argtypes = [vp, vp]; lib.vh_build_stats.restype = None
        lib.vh_assimp_available.restype = C.c_int
        lib.vh_model_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int]; lib.vh_model_load.restype = vp
        lib.vh_model_count.argtypes = [vp]; lib.vh_model_count.restype = C.c_uint32
        lib.vh_model_mesh_info.argtypes = [vp, C.c_uint32, vp, vp]; lib.vh_model_mesh_info.restype = None
        lib.vh_model_mesh_copy.argtypes = [vp, C.c_uint32, vp, vp, vp, vp]; lib.vh_model_mesh_copy.restype = None
        lib.vh_model_free.argtypes = [vp]; lib.vh_model_free.restype = None
        lib.vh_render_on_gpus.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, vp, C.c_uint64, vp]; lib.vh_render_on_gpus.restype = vp
        _lib = lib
    return _lib


def plugin_names():
    lib = load()
    n = lib.vh_plugin_names(None, 0)
    buf = C.create_string_buffer(n + 1)
    lib.vh_plugin_names(buf, n + 1)
    return [s for s in buf.value.decode().split("\n") if s]


class _LightSamplerRec(C.Structure):       # SceneRecords::LightSamplerRec (vh_scene.h)
    _fields_ = [("kind", C.c_uint32), ("alias_offset", C.c_uint32), ("alias_size", C.c_uint32), ("alias_integral", C.c_float),
                ("env_separate", C.c_uint32), ("env_prob", C.c_float)]


class HostPipeline:
    """Importer::import_scene + Pipeline::prepare/render/final_picture of the reference, through libvision_host"""

    def __init__(self, scene_path, overrides=None):
        self.lib = load()
        self.h = C.c_void_p(self.lib.vh_import_scene(str(scene_path).encode()))
        if not self.h:
            raise capi.VbError("vh_import_scene: " + self.lib.vh_last_error().decode())
        for k, v in (overrides or {}).items():
            a, b = (v if isinstance(v, (list, tuple)) else (v, 0))
            self._check(self.lib.vh_override(self.h, k.encode(), float(a), float(b)), "vh_override")

    def _check(self, rc, what):
        if rc:
            raise capi.VbError(f"{what}: {self.lib.vh_last_error().decode()}")

    def close(self):
        if self.h:
            self.lib.vh_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def prepare_records(self):
        self._check(self.lib.vh_prepare_records(self.h), "vh_prepare_records")

    def prepare(self, device=0):
        self._check(self.lib.vh_prepare(self.h, device), "vh_prepare")

    def info(self):
        v = (C.c_uint32 * 10)()
        e = C.c_float()
        self.lib.vh_info(self.h, v, C.byref(e))
        keys = ("width", "height", "spp", "n_meshes", "n_instances", "n_materials", "n_lights", "n_alias", "n_textures", "tonemap")
        d = dict(zip(keys, list(v)))
        d["exposure"] = e.value
        return d

    def render(self, n_frames):
        self._check(self.lib.vh_render(self.h, n_frames), "vh_render")

    def accum(self):
        i = self.info()
        out = np.empty((i["height"], i["width"], 4), np.float32)
        self._check(self.lib.vh_download_accum(self.h, out.ctypes.data), "vh_download_accum")
        return out

    def final_picture(self):
        i = self.info()
        out = np.empty((i["height"], i["width"], 4), np.float32)
        self._check(self.lib.vh_final_picture(self.h, out.ctypes.data), "vh_final_picture")
        return out

    # -- multi-GPU (one process per GPU): partition before prepare(), communicator after it ------------------------------
    def set_partition(self, rank, world):
        self._check(self.lib.vh_set_partition(self.h, rank, world), "vh_set_partition")

    def nccl_init(self, unique_id: bytes):
        """every rank calls this with the 128 bytes rank 0 got from capi.nccl_unique_id(); blocks until all ranks have"""
        buf = (C.c_ubyte * 128).from_buffer_copy(unique_id)
        self._check(self.lib.vh_nccl_init(self.h, buf), "vh_nccl_init")

    def film_allreduce(self):
        self._check(self.lib.vh_film_allreduce(self.h), "vh_film_allreduce")

    def synchronize(self):
        self._check(self.lib.vh_synchronize(self.h), "vh_synchronize")

    def build_stats(self) -> dict:
        st = capi.vb_build_stats()
        self.lib.vh_build_stats(self.h, C.byref(st))
        return {"n_triangles": st.n_triangles, "n_nodes": st.n_nodes, "node_bytes": st.node_bytes, "tri_bytes": st.tri_bytes, "build_ms": st.build_ms}

    def invalidate(self):
        """Pipeline::invalidate: empty film, frame index 0"""
        self._check(self.lib.vh_invalidate(self.h), "vh_invalidate")

    def context(self) -> "capi.Context":
        """the pipeline's own vb_ctx (after prepare()), BORROWED: valid while this HostPipeline lives, never closed by the view"""
        h = self.lib.vh_ctx(self.h)
        if not h:
            raise capi.VbError("HostPipeline.context(): prepare() has not run")
        view = capi.Context.__new__(capi.Context)
        view.lib = capi.load_library()
        view.h = C.c_void_p(h)
        view.device = 0
        i = self.info()
        view._res = (i["width"], i["height"])
        view.close = lambda: None
        return view

    def log(self):
        n = self.lib.vh_scene_log(self.h, None, 0)
        buf = C.create_string_buffer(n + 1)
        self.lib.vh_scene_log(self.h, buf, n + 1)
        return [s for s in buf.value.decode().split("\n") if s]

    def _blob(self, which, ctype):
        n = self.lib.vh_records_blob(self.h, which, None, 0)
        cnt = n // C.sizeof(ctype)
        arr = (ctype * max(1, cnt))()
        self.lib.vh_records_blob(self.h, which, arr, n)
        return list(arr)[:cnt]

    def records(self) -> capi.SceneRecords:
        """the POD records the C++ host layer produced, as a capi.SceneRecords (feedable to the oracle)"""
        i = self.info()
        sc = capi.SceneRecords()
        for m in range(i["n_meshes"]):
            nv, nt = C.c_uint32(), C.c_uint32()
            self.lib.vh_mesh_info(self.h, m, C.byref(nv), C.byref(nt))
            pos = np.empty((nv.value, 3), np.float32); nrm = np.empty((nv.value, 3), np.float32)
            uv = np.empty((nv.value, 2), np.float32); tri = np.empty((nt.value, 3), np.uint32)
            self.lib.vh_mesh_copy(self.h, m, pos.ctypes.data, nrm.ctypes.data, uv.ctypes.data, tri.ctypes.data)
            sc.meshes.append({"pos": pos, "nrm": nrm, "uv": uv, "tri": tri})
        for r in self._blob(0, capi.vb_instance):
            sc.instances.append({"mesh_id": r.mesh_id, "material_id": r.material_id, "light_id": r.light_id,
                                 "inside_medium": r.inside_medium, "outside_medium": r.outside_medium,
                                 "lightmap_id": r.lightmap_id, "o2w": np.array(list(r.o2w), np.float32)})
        sc.materials = self._blob(1, capi.vb_material_rec)
        sc.lights = self._blob(2, capi.vb_light_rec)
        sc.alias_table = [(e.prob, e.alias) for e in self._blob(3, capi.vb_alias_entry)]
        sc.alias_func = [float(x) for x in self._blob(4, C.c_float)]
        ls = self._blob(102, _LightSamplerRec)[0]
        sc.light_sampler = (ls.kind, ls.alias_offset, ls.alias_size, ls.alias_integral)
        sc.light_sampler_env = (ls.env_separate, ls.env_prob)
        sc.camera = self._blob(5, capi.vb_camera)[0]
        sc.params = self._blob(6, capi.vb_pt_params)[0]
        sc.mediums = self._blob(7, capi.vb_medium_rec)
        for which, (_, dims) in capi.LUT_FILES.items():
            vals = self._blob(8 + which, C.c_float)
            if vals:
                sc.luts[which] = (np.array(vals, np.float32), dims)
        sc.spectrum = self._blob(100, capi.vb_spectrum_rec)[0]
        sc.spectrum.rgb2spec = None
        sc.spds = [np.array(self._blob(200 + k, C.c_float), np.float32) for k in range(self._blob(101, C.c_uint64)[0])]
        if sc.spectrum.type == capi.VB_SPECTRUM_HERO:
            sc.rgb2spec = capi.load_rgb2spec()                      # the same data file the host layer reads
        for t in range(i["n_textures"]):
            w, h = C.c_uint32(), C.c_uint32()
            self.lib.vh_texture_info(self.h, t, C.byref(w), C.byref(h))
            a = np.empty((h.value, w.value, 4), np.float32)
            self.lib.vh_texture_copy(self.h, t, a.ctypes.data)
            sc.textures.append(a)
        sc.resolution = (i["width"], i["height"]); sc.spp = i["spp"]; sc.exposure = i["exposure"]; sc.tonemap = i["tonemap"]
        return sc


def render_on_gpus(scene_path, n_gpus, spp=0, use_nccl=False):
    """one process, n GPUs (vh::render_on_gpus): returns (rank 0's HostPipeline, final picture (h, w, 4), times in ms)"""
    lib = load()
    probe = HostPipeline(scene_path)
    probe.prepare_records()
    i = probe.info()
    probe.close()
    pic = np.empty((i["height"], i["width"], 4), np.float32)
    times = (C.c_double * 3)()
    h = lib.vh_render_on_gpus(str(scene_path).encode(), n_gpus, spp, int(use_nccl), pic.ctypes.data, pic.size, times)
    if not h:
        raise capi.VbError("vh_render_on_gpus: " + lib.vh_last_error().decode())
    hp = HostPipeline.__new__(HostPipeline)
    hp.lib, hp.h = lib, C.c_void_p(h)
    return hp, pic, {"prepare_ms": times[0], "render_ms": times[1], "reduce_ms": times[2]}


def assimp_available() -> bool:
    return bool(load().vh_assimp_available())


def load_model(path, importer="assimp", swap_handed=False, smooth=False, flip_uv=True):
    """the `model` plugin's loader by itself: list of {"pos", "nrm", "uv", "tri"} meshes. importer = "assimp" (the reference's
    importer library: reference vertex / triangle order) or "obj" (in-tree loader: file order)"""
    lib = load()
    h = lib.vh_model_load(str(path).encode(), int(importer == "assimp"), int(swap_handed), int(smooth), int(flip_uv))
    if not h:
        raise capi.VbError("vh_model_load: " + lib.vh_last_error().decode())
    out = []
    try:
        for i in range(lib.vh_model_count(h)):
            nv, nt = C.c_uint32(), C.c_uint32()
            lib.vh_model_mesh_info(h, i, C.byref(nv), C.byref(nt))
            pos = np.empty((nv.value, 3), np.float32); nrm = np.empty((nv.value, 3), np.float32)
            uv = np.empty((nv.value, 2), np.float32); tri = np.empty((nt.value, 3), np.uint32)
            lib.vh_model_mesh_copy(h, i, pos.ctypes.data, nrm.ctypes.data, uv.ctypes.data, tri.ctypes.data)
            out.append({"pos": pos, "nrm": nrm, "uv": uv, "tri": tri})
    finally:
        lib.vh_model_free(h)
    return out


def save_image(path, rgba) -> None:
    """host layer's output writer: .png / .ppm get the display gamma, .hdr / .pfm stay linear (pipeline.cpp:337-354)"""
    a = np.ascontiguousarray(rgba, np.float32)
    if load().vh_save_image(str(path).encode(), a.ctypes.data, a.shape[1], a.shape[0]) != 0:
        raise capi.VbError(load().vh_last_error().decode())


def load_image(path, srgb_to_linear=False) -> np.ndarray:
    lib = load()
    w, h = C.c_uint32(), C.c_uint32()
    if lib.vh_load_image(str(path).encode(), int(srgb_to_linear), C.byref(w), C.byref(h), None, 0) != 0:
        raise capi.VbError(lib.vh_last_error().decode())
    a = np.empty((h.value, w.value, 4), np.float32)
    lib.vh_load_image(str(path).encode(), int(srgb_to_linear), C.byref(w), C.byref(h), a.ctypes.data, a.size)
    return a


def rgbe_encode(rgb) -> bytes:
    a = np.ascontiguousarray(rgb, np.float32); out = (C.c_ubyte * 4)()
    load().vh_rgbe_encode(a.ctypes.data, out)
    return bytes(out)

"""raytracing_course_hw_b200 -- Python harness over the two C-ABI libraries.

The product is C++/CUDA (csrc/); this module only binds it with ctypes so that tests, bench.py
and scripts can drive the same entry points the CLI uses.  Names mirror the reference's
`sceneio` namespace: load_scene(...) / render_scene(...)  (hw8/src/include/sceneio.h:7-11).

There is NO CPU fallback: if libb200rt.so is missing, or no CUDA device is visible, Renderer()
raises.  Nothing here imports or touches oracle/.
"""
import ctypes
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(_PKG, "lib")

MODE_HW1, MODE_HW2, MODE_HW3, MODE_HW4, MODE_HW5, MODE_HW6, MODE_HW7, MODE_HW8 = range(1, 9)
FLAG_COUNTERS = 1
FLAG_MEGAKERNEL = 2   # glTF modes: per-path megakernel instead of the wavefront pipeline
FLAG_REFERENCE_BVH = 8  # glTF modes: traverse the description's BVH instead of the library's rebuilt one
FLAG_NO_FIREFLY_CLAMP = 64  # hw7 / hw8: unbiased estimator instead of the reference's weight clamp at 6
FLAG_KEEP_DARK_RAYS = 16
FLAG_WAVEFRONT = 128  # hw6: the wavefront pipeline even for scenes small enough for the per-pixel kernel
FLAG_EXPLICIT_CAMERA_STATE = 32
FLAG_SERIAL_WAVES = 4  # wavefront: one wave in flight (stage times in stats() are then exclusive kernel times)


class B200rtError(RuntimeError):
    def __init__(self, code, what):
        super().__init__(f"b200rt error {code}: {what}")
        self.code = code


class Options(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int32), ("seed", ctypes.c_uint32), ("flags", ctypes.c_int32)]


class Stats(ctypes.Structure):
    _fields_ = [("kernel_ms", ctypes.c_double), ("paths", ctypes.c_uint64), ("rays", ctypes.c_uint64),
                ("light_queries", ctypes.c_uint64), ("node_tests", ctypes.c_uint64), ("tri_tests", ctypes.c_uint64),
                ("light_node_tests", ctypes.c_uint64), ("light_tri_tests", ctypes.c_uint64),
                ("shaded_hits", ctypes.c_uint64), ("tex_lookups", ctypes.c_uint64), ("launches", ctypes.c_int32),
                ("extend_launches", ctypes.c_int32), ("extend_ms", ctypes.c_double), ("shade_ms", ctypes.c_double),
                ("light_ms", ctypes.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class BvhLayout(ctypes.Structure):
    """b200rt_bvh_layout (include/b200rt.h)."""
    _fields_ = [("n_records", ctypes.c_uint64), ("n_tris", ctypes.c_uint64), ("root", ctypes.c_uint32),
                ("quant_bits", ctypes.c_int32), ("guard", ctypes.c_float),
                ("base", ctypes.c_float * 3), ("scale", ctypes.c_float * 3), ("inv_scale", ctypes.c_float * 3)]


class SceneHeader(ctypes.Structure):
    """Leading scalar fields of b200rt_scene_desc (include/b200rt.h); pointers are not bound here."""
    _fields_ = [("mode", ctypes.c_int32), ("width", ctypes.c_int32), ("height", ctypes.c_int32),
                ("samples", ctypes.c_int32), ("ray_depth", ctypes.c_int32)]


class Image(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int32), ("height", ctypes.c_int32), ("rgb", ctypes.POINTER(ctypes.c_uint8))]


class SceneDesc(ctypes.Structure):
    """Full b200rt_scene_desc (include/b200rt.h), for tests and tools that inspect loaded scenes."""
    _fields_ = [("mode", ctypes.c_int32), ("width", ctypes.c_int32), ("height", ctypes.c_int32),
                ("samples", ctypes.c_int32), ("ray_depth", ctypes.c_int32),
                ("bg_color", ctypes.c_float * 3), ("ambient", ctypes.c_float * 3),
                ("cam_pos", ctypes.c_float * 3), ("cam_right", ctypes.c_float * 3),
                ("cam_up", ctypes.c_float * 3), ("cam_forward", ctypes.c_float * 3), ("cam_fov", ctypes.c_float),
                ("prims", ctypes.c_void_p), ("n_prims", ctypes.c_int32),
                ("lights", ctypes.c_void_p), ("n_lights", ctypes.c_int32),
                ("tris", ctypes.c_void_p), ("n_tris", ctypes.c_int32),
                ("nodes", ctypes.c_void_p), ("n_nodes", ctypes.c_int32), ("root", ctypes.c_int32),
                ("light_tris", ctypes.c_void_p), ("n_light_tris", ctypes.c_int32),
                ("light_nodes", ctypes.c_void_p), ("n_light_nodes", ctypes.c_int32),
                ("materials", ctypes.c_void_p), ("n_materials", ctypes.c_int32),
                ("images", ctypes.POINTER(Image)), ("n_images", ctypes.c_int32),
                ("env_map", ctypes.POINTER(Image)),
                ("light_prims", ctypes.c_void_p), ("n_light_prims", ctypes.c_int32), ("n_bvh_prims", ctypes.c_int32)]


_host = None
_dev = None


def host_lib():
    global _host
    if _host is None:
        path = os.path.join(LIB_DIR, "libb200rt_host.so")
        if not os.path.exists(path):
            raise ImportError(f"{path} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = ctypes.CDLL(path)
        L.b200rt_host_desc.restype = ctypes.c_void_p
        L.b200rt_host_desc.argtypes = [ctypes.c_void_p]
        L.b200rt_host_strerror.restype = ctypes.c_char_p
        L.b200rt_host_free.argtypes = [ctypes.c_void_p]
        L.b200rt_host_free.restype = None
        L.b200rt_host_set_frame.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32]
        L.b200rt_host_set_env_map.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        L.b200rt_host_timings.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
        L.b200rt_host_write_ppm.argtypes = [ctypes.c_char_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_host_write_png.argtypes = [ctypes.c_char_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_host_write_pfm.argtypes = [ctypes.c_char_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32]
        _host = L
    return _host


def device_lib():
    """The CUDA library.  Fails loudly when it was not built (no fallback of any kind)."""
    global _dev
    if _dev is None:
        path = os.path.join(LIB_DIR, "libb200rt.so")
        if not os.path.exists(path):
            raise ImportError(f"{path} missing: the CUDA extension is required (no CPU fallback); "
                              "run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = ctypes.CDLL(path)
        L.b200rt_strerror.restype = ctypes.c_char_p
        L.b200rt_last_error.restype = ctypes.c_char_p
        L.b200rt_last_error.argtypes = [ctypes.c_void_p]
        L.b200rt_destroy.argtypes = [ctypes.c_void_p]
        L.b200rt_destroy.restype = None
        L.b200rt_upload_scene.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_render.argtypes = [ctypes.c_void_p] + [ctypes.c_int32] * 5
        L.b200rt_clear.argtypes = [ctypes.c_void_p]
        L.b200rt_resolve.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_download_accum.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_upload_accum.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_accum_device_ptr.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]
        L.b200rt_export_scene.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint64)]
        L.b200rt_import_scene_begin.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
        L.b200rt_import_scene_end.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_replicate_scene.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_primary_ids.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_synchronize.argtypes = [ctypes.c_void_p]
        L.b200rt_get_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(Stats)]
        L.b200rt_stream.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
        L.b200rt_trace_rays.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p,
                                        ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        L.b200rt_scene_bvh_layout.argtypes = [ctypes.c_void_p, ctypes.POINTER(BvhLayout)]
        L.b200rt_scene_bvh_download.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_resolve_noalloc.argtypes = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                             ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_resolve_peers.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_void_p), ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_reduce_peers.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.c_int32]
        L.b200rt_enable_peer.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        L.b200rt_accum_ipc_export.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_ipc_open.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
        L.b200rt_ipc_close.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_group_create.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int32), ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_group_destroy.argtypes = [ctypes.c_void_p]
        L.b200rt_group_destroy.restype = None
        L.b200rt_group_size.argtypes = [ctypes.c_void_p]
        L.b200rt_group_member.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        L.b200rt_group_member.restype = ctypes.c_void_p
        L.b200rt_group_upload_scene.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_group_render.argtypes = [ctypes.c_void_p] + [ctypes.c_int32] * 5
        L.b200rt_group_clear.argtypes = [ctypes.c_void_p]
        L.b200rt_group_resolve.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]
        L.b200rt_group_download_accum.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.b200rt_group_get_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(Stats)]
        L.b200rt_group_last_error.argtypes = [ctypes.c_void_p]
        L.b200rt_group_last_error.restype = ctypes.c_char_p
        _dev = L
    return _dev


DEVICE_SYMBOLS = ["b200rt_create", "b200rt_destroy", "b200rt_upload_scene", "b200rt_render", "b200rt_clear",
                  "b200rt_resolve", "b200rt_download_accum", "b200rt_upload_accum", "b200rt_accum_device_ptr",
                  "b200rt_export_scene", "b200rt_import_scene_begin", "b200rt_import_scene_end", "b200rt_replicate_scene",
                  "b200rt_primary_ids", "b200rt_synchronize", "b200rt_get_stats", "b200rt_stream",
                  "b200rt_strerror", "b200rt_last_error", "b200rt_device_count",
                  "b200rt_trace_rays", "b200rt_scene_bvh_layout", "b200rt_scene_bvh_download",
                  "b200rt_resolve_noalloc", "b200rt_resolve_peers", "b200rt_reduce_peers", "b200rt_enable_peer", "b200rt_accum_ipc_export",
                  "b200rt_ipc_open", "b200rt_ipc_close", "b200rt_group_create", "b200rt_group_destroy", "b200rt_group_size",
                  "b200rt_group_member", "b200rt_group_upload_scene", "b200rt_group_render", "b200rt_group_clear",
                  "b200rt_group_resolve", "b200rt_group_download_accum", "b200rt_group_get_stats", "b200rt_group_last_error"]
HOST_SYMBOLS = ["b200rt_host_load_text", "b200rt_host_load_gltf", "b200rt_host_load_gltf_ex", "b200rt_host_set_env_map",
                "b200rt_host_desc", "b200rt_host_set_frame", "b200rt_host_free", "b200rt_host_write_ppm",
                "b200rt_host_write_png", "b200rt_host_write_pfm", "b200rt_host_timings", "b200rt_host_strerror"]


class Scene:
    """A loaded scene (host arrays + reference-shaped b200rt_scene_desc)."""

    def __init__(self, handle):
        self._h = handle

    @property
    def desc_ptr(self):
        return host_lib().b200rt_host_desc(self._h)

    @property
    def header(self):
        return SceneHeader.from_address(self.desc_ptr)

    @property
    def desc(self):
        return SceneDesc.from_address(self.desc_ptr)

    mode = property(lambda s: s.header.mode)
    width = property(lambda s: s.header.width)
    height = property(lambda s: s.header.height)
    samples = property(lambda s: s.header.samples)
    ray_depth = property(lambda s: s.header.ray_depth)

    def set_frame(self, width=0, height=0, samples=0):
        host_lib().b200rt_host_set_frame(self._h, width, height, samples)

    def set_env_map(self, path):
        rc = host_lib().b200rt_host_set_env_map(self._h, os.fsencode(path))
        if rc:
            raise B200rtError(rc, host_lib().b200rt_host_strerror(rc).decode())

    def timings(self):
        a, b = ctypes.c_double(), ctypes.c_double()
        host_lib().b200rt_host_timings(self._h, ctypes.byref(a), ctypes.byref(b))
        return {"parse_ms": a.value, "bvh_ms": b.value}

    def close(self):
        if self._h:
            host_lib().b200rt_host_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


HOST_FLAG_REFERENCE_BVH = 1


def load_scene(path, width=None, height=None, samples=None, mode=0, env_map=None, reference_bvh=False):
    """sceneio::loadScene: text scenes take everything from the file; glTF scenes take
    width/height/samples from the caller (argv in the reference, hw8/src/main.cpp:9-11).
    reference_bvh: glTF only -- also run the reference's SAH build (triangles in its order + both trees); the
    default description carries no trees, b200rt_upload_scene builds its own."""
    L = host_lib()
    h = ctypes.c_void_p()
    p = os.fsencode(path)
    if str(path).lower().endswith((".gltf",)):
        if not (width and height and samples):
            raise ValueError("glTF scenes need width, height and samples")
        rc = L.b200rt_host_load_gltf_ex(p, mode, width, height, samples, HOST_FLAG_REFERENCE_BVH if reference_bvh else 0, ctypes.byref(h))
    else:
        rc = L.b200rt_host_load_text(p, mode, ctypes.byref(h))
    if rc:
        raise B200rtError(rc, L.b200rt_host_strerror(rc).decode())
    sc = Scene(h)
    if not str(path).lower().endswith(".gltf") and (width or height or samples):
        sc.set_frame(width or 0, height or 0, samples or 0)
    if env_map:
        sc.set_env_map(env_map)
    return sc


class Renderer:
    """One b200rt context = one GPU."""

    def __init__(self, device=0, seed=0, flags=0):
        self._L = device_lib()
        self._ctx = ctypes.c_void_p()
        opt = Options(device, seed, flags)
        rc = self._L.b200rt_create(ctypes.byref(self._ctx), ctypes.byref(opt))
        if rc:
            raise B200rtError(rc, self._L.b200rt_strerror(rc).decode())
        self.width = self.height = 0

    def _ck(self, rc):
        if rc:
            raise B200rtError(rc, self._L.b200rt_strerror(rc).decode() + " / " + self._L.b200rt_last_error(self._ctx).decode())

    def upload(self, scene):
        self._ck(self._L.b200rt_upload_scene(self._ctx, scene.desc_ptr))
        self.width, self.height, self.samples, self.mode = scene.width, scene.height, scene.samples, scene.mode

    def render(self, spp_begin=0, spp_end=None, width=0, height=0, ray_depth=0):
        if spp_end is None:
            spp_end = self.samples
        if width > 0:
            self.width = width
        if height > 0:
            self.height = height
        self._ck(self._L.b200rt_render(self._ctx, width, height, spp_begin, spp_end, ray_depth))

    def clear(self):
        self._ck(self._L.b200rt_clear(self._ctx))

    def synchronize(self):
        self._ck(self._L.b200rt_synchronize(self._ctx))

    def resolve(self, total_spp=None, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 3), np.uint8)
        self._ck(self._L.b200rt_resolve(self._ctx, total_spp or self.samples, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def download_accum(self):
        out = np.empty((self.height, self.width, 3), np.float32)
        self._ck(self._L.b200rt_download_accum(self._ctx, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def upload_accum(self, arr):
        arr = np.ascontiguousarray(arr, np.float32)
        self._ck(self._L.b200rt_upload_accum(self._ctx, arr.ctypes.data_as(ctypes.c_void_p)))

    def accum_device_ptr(self):
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        self._ck(self._L.b200rt_accum_device_ptr(self._ctx, ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    # scene replication (include/b200rt.h: b200rt_export_scene / b200rt_import_scene_*)
    SCENE_IMAGE_BYTES = 2048

    def export_scene(self):
        """(header bytes, device pointer, byte count) of the uploaded glTF-mode scene."""
        hdr = ctypes.create_string_buffer(self.SCENE_IMAGE_BYTES)
        p, n = ctypes.c_void_p(), ctypes.c_uint64()
        self._ck(self._L.b200rt_export_scene(self._ctx, hdr, ctypes.byref(p), ctypes.byref(n)))
        return hdr.raw, p.value, n.value

    def import_scene_begin(self, header):
        """Sizes this context's arena for the exported scene; returns the device pointer the blob must be copied to."""
        hdr = ctypes.create_string_buffer(bytes(header), self.SCENE_IMAGE_BYTES)
        p = ctypes.c_void_p()
        self._ck(self._L.b200rt_import_scene_begin(self._ctx, hdr, ctypes.byref(p)))
        return p.value

    def import_scene_end(self, header, scene=None, frame=None):
        """Frame metadata comes from `scene` (a description this rank loaded too) or from `frame` =
        (width, height, samples, mode) when only the uploading rank read the file."""
        hdr = ctypes.create_string_buffer(bytes(header), self.SCENE_IMAGE_BYTES)
        self._ck(self._L.b200rt_import_scene_end(self._ctx, hdr))
        if scene is not None:
            frame = (scene.width, scene.height, scene.samples, scene.mode)
        self.width, self.height, self.samples, self.mode = frame

    def replicate_from(self, other, scene=None, frame=None):
        """Device-to-device copy of `other`'s uploaded scene into this context (same process, any two devices);
        frame metadata from `scene` or `frame` = (width, height, samples, mode)."""
        self._ck(self._L.b200rt_replicate_scene(other._ctx, self._ctx))
        if scene is not None:
            frame = (scene.width, scene.height, scene.samples, scene.mode)
        self.width, self.height, self.samples, self.mode = frame

    # several GPUs, peer level (include/b200rt.h): fused reduce + resolve over peer-readable buffers
    @staticmethod
    def _ptr_array(ptrs):
        return (ctypes.c_void_p * max(len(ptrs), 1))(*ptrs)

    def resolve_peers(self, total_spp, peer_ptrs, out=None):
        """RGB8 of (own accumulation buffer + the peers'), summed in the order given, in ONE kernel."""
        if out is None:
            out = np.empty((self.height, self.width, 3), np.uint8)
        self._ck(self._L.b200rt_resolve_peers(self._ctx, total_spp, self._ptr_array(peer_ptrs), len(peer_ptrs), out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def reduce_peers(self, peer_ptrs):
        self._ck(self._L.b200rt_reduce_peers(self._ctx, self._ptr_array(peer_ptrs), len(peer_ptrs)))

    def enable_peer(self, device):
        self._ck(self._L.b200rt_enable_peer(self._ctx, device))

    def accum_ipc_export(self):
        h = ctypes.create_string_buffer(64)
        self._ck(self._L.b200rt_accum_ipc_export(self._ctx, h))
        return h.raw

    def ipc_open(self, handle):
        p = ctypes.c_void_p()
        self._ck(self._L.b200rt_ipc_open(self._ctx, ctypes.create_string_buffer(bytes(handle), 64), ctypes.byref(p)))
        return p.value

    def ipc_close(self, ptr):
        self._ck(self._L.b200rt_ipc_close(self._ctx, ctypes.c_void_p(ptr)))

    # measurement cross-check: the production traversal kernel on a caller's rays + the flattened tree
    def trace_rays(self, rays):
        """rays: (n, 6) float32 (origin, direction).  Returns (hits (n, 4) float32 {t, u, v, bits}, boxes tested,
        triangles tested) from wf_extend's counting build."""
        rays = np.ascontiguousarray(rays, np.float32)
        hits = np.empty((rays.shape[0], 4), np.float32)
        nb, nt = ctypes.c_uint64(), ctypes.c_uint64()
        self._ck(self._L.b200rt_trace_rays(self._ctx, rays.ctypes.data_as(ctypes.c_void_p), rays.shape[0],
                                           hits.ctypes.data_as(ctypes.c_void_p), ctypes.byref(nb), ctypes.byref(nt)))
        return hits, nb.value, nt.value

    def scene_bvh(self):
        """(BvhLayout, records (n, 8) uint32, triangle records (n, 12) float32) exactly as they sit in device memory."""
        lay = BvhLayout()
        self._ck(self._L.b200rt_scene_bvh_layout(self._ctx, ctypes.byref(lay)))
        rec = np.empty((lay.n_records, 8), np.uint32)
        tri = np.empty((lay.n_tris, 12), np.float32)
        self._ck(self._L.b200rt_scene_bvh_download(self._ctx, rec.ctypes.data_as(ctypes.c_void_p), tri.ctypes.data_as(ctypes.c_void_p)))
        return lay, rec, tri

    def primary_ids(self):
        out = np.empty((self.height, self.width), np.int32)
        self._ck(self._L.b200rt_primary_ids(self._ctx, self.width, self.height, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def stats(self):
        s = Stats()
        self._ck(self._L.b200rt_get_stats(self._ctx, ctypes.byref(s)))
        return s.as_dict()

    def stream(self):
        p = ctypes.c_void_p()
        self._ck(self._L.b200rt_stream(self._ctx, ctypes.byref(p)))
        return p.value or 0

    def close(self):
        if self._ctx:
            self._L.b200rt_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


ERR_TOO_LARGE = -7
NOALLOC_MAX_PIXELS = 2048 * 1024


def resolve_noalloc(device, accum_device_ptr, width, height, total_spp, mode, stream=0, out=None):
    """b200rt_resolve_noalloc: RGB8 of a device-resident radiance SUM buffer through the library's static scratch
    (no context, no allocation); frames above NOALLOC_MAX_PIXELS raise B200rtError(code ERR_TOO_LARGE)."""
    L = device_lib()
    if out is None:
        out = np.empty((height, width, 3), np.uint8)
    rc = L.b200rt_resolve_noalloc(device, ctypes.c_void_p(stream), ctypes.c_void_p(accum_device_ptr), width, height,
                                  total_spp, mode, out.ctypes.data_as(ctypes.c_void_p))
    if rc:
        raise B200rtError(rc, L.b200rt_strerror(rc).decode())
    return out


class Group:
    """Several contexts driven by one process (b200rt_group_*): one upload + device-to-device replication, samples
    split contiguously over the members, ONE fused peer reduce + resolve kernel on member 0.  `devices` may repeat."""

    def __init__(self, devices, seed=0, flags=0):
        self._L = device_lib()
        self._g = ctypes.c_void_p()
        opt = Options(0, seed, flags)
        dv = (ctypes.c_int32 * len(devices))(*devices)
        rc = self._L.b200rt_group_create(ctypes.byref(self._g), dv, len(devices), ctypes.byref(opt))
        if rc:
            raise B200rtError(rc, self._L.b200rt_strerror(rc).decode())
        self.width = self.height = 0

    def _ck(self, rc):
        if rc:
            raise B200rtError(rc, self._L.b200rt_strerror(rc).decode() + " / " + self._L.b200rt_group_last_error(self._g).decode())

    def upload(self, scene):
        self._ck(self._L.b200rt_group_upload_scene(self._g, scene.desc_ptr))
        self.width, self.height, self.samples, self.mode = scene.width, scene.height, scene.samples, scene.mode

    def render(self, spp_begin=0, spp_end=None, width=0, height=0, ray_depth=0):
        if spp_end is None:
            spp_end = self.samples
        if width > 0:
            self.width = width
        if height > 0:
            self.height = height
        self._ck(self._L.b200rt_group_render(self._g, width, height, spp_begin, spp_end, ray_depth))

    def clear(self):
        self._ck(self._L.b200rt_group_clear(self._g))

    def resolve(self, total_spp=None, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 3), np.uint8)
        self._ck(self._L.b200rt_group_resolve(self._g, total_spp or self.samples, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def download_accum(self):
        out = np.empty((self.height, self.width, 3), np.float32)
        self._ck(self._L.b200rt_group_download_accum(self._g, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def member_accum(self, i):
        """Radiance sum buffer of member i alone (tests)."""
        out = np.empty((self.height, self.width, 3), np.float32)
        rc = self._L.b200rt_download_accum(self._L.b200rt_group_member(self._g, i), out.ctypes.data_as(ctypes.c_void_p))
        self._ck(rc)
        return out

    def stats(self):
        s = Stats()
        self._ck(self._L.b200rt_group_get_stats(self._g, ctypes.byref(s)))
        return s.as_dict()

    def __len__(self):
        return self._L.b200rt_group_size(self._g)

    def close(self):
        if self._g:
            self._L.b200rt_group_destroy(self._g)
            self._g = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def write_ppm(path, rgb):
    rgb = np.ascontiguousarray(rgb, np.uint8)
    rc = host_lib().b200rt_host_write_ppm(os.fsencode(path), rgb.shape[1], rgb.shape[0], rgb.ctypes.data_as(ctypes.c_void_p))
    if rc:
        raise B200rtError(rc, host_lib().b200rt_host_strerror(rc).decode())


def write_png(path, rgb):
    rgb = np.ascontiguousarray(rgb, np.uint8)
    rc = host_lib().b200rt_host_write_png(os.fsencode(path), rgb.shape[1], rgb.shape[0], rgb.ctypes.data_as(ctypes.c_void_p))
    if rc:
        raise B200rtError(rc, host_lib().b200rt_host_strerror(rc).decode())


def write_pfm(path, accum_sum, total_spp):
    a = np.ascontiguousarray(accum_sum, np.float32)
    rc = host_lib().b200rt_host_write_pfm(os.fsencode(path), a.shape[1], a.shape[0], a.ctypes.data_as(ctypes.c_void_p), total_spp)
    if rc:
        raise B200rtError(rc, host_lib().b200rt_host_strerror(rc).decode())


def render_scene(scene, out_path=None, device=0, seed=0):
    """sceneio::renderScene (hw8/src/sceneio.cpp:381-402): render all samples, tonemap, write the PPM."""
    r = Renderer(device=device, seed=seed)
    try:
        r.upload(scene)
        r.render(0, scene.samples)
        img = r.resolve(scene.samples)
    finally:
        r.close()
    if out_path:
        write_ppm(out_path, img)
    return img

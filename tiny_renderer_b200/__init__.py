"""tiny_renderer_b200 — B200-native (sm_100a) rasterization backend for litzendraht/tiny_renderer.

Python is only a thin ctypes binding over the C ABI in include/tinyr.h (libtinyr.so, built in-tree from
tiny_renderer_b200/csrc).  `Scene` mirrors the reference's `Scene` (src/scene.rs:47-151): same method names,
argument meaning and error behaviour (panics become exceptions).  There is no CPU fallback: constructing a
Scene without the CUDA library or without a GPU raises.
"""
import ctypes
import os
import subprocess

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_PKG, "csrc")
LIB_PATH = os.environ.get("TINYR_LIB") or os.path.join(_PKG, "libtinyr.so")     # (TINYR_LIB: another build of the same ABI, for A/B runs)
CLI_PATH = os.path.join(_PKG, "tiny_renderer")
NO_TRIANGLE = 0xFFFFFFFF
PIPELINES = ("default", "phong", "normal_map", "specular", "darboux", "shadow", "occlusion")
KERNEL_SLOTS = ("K1_vertex_setup", "K2_bin", "K3_raster_visibility", "K4_shadow_depth", "K5_shade", "K6_ao_accumulate",
                "K7_clear_pack", "K8_checksum")
DIAG_SLOTS = ("shadow_index_oob", "shadow_x_wrap", "texel_oob", "nan_depth", "w_zero", "singular_basis", "binned", "reserved")

TR_ERRORS = {-1: "TR_ERR_INVALID_ARGUMENT", -2: "TR_ERR_UNKNOWN_PIPELINE", -3: "TR_ERR_BAD_POLYGON", -4: "TR_ERR_IO",
             -5: "TR_ERR_FORMAT", -6: "TR_ERR_CUDA", -7: "TR_ERR_NO_DEVICE", -8: "TR_ERR_REFERENCE_UNDEFINED",
             -9: "TR_ERR_RESERVED_9", -10: "TR_ERR_UNSUPPORTED"}


class TinyrError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("%s: %s" % (TR_ERRORS.get(status, status), message))
        self.status = status
        self.name = TR_ERRORS.get(status, str(status))


class _Image(ctypes.Structure):
    _fields_ = [("rgb8", ctypes.c_void_p), ("width", ctypes.c_uint32), ("height", ctypes.c_uint32)]


class _SceneDesc(ctypes.Structure):
    _fields_ = [("width", ctypes.c_uint32), ("height", ctypes.c_uint32),
                ("positions", ctypes.c_void_p), ("n_positions", ctypes.c_uint32),
                ("uvs", ctypes.c_void_p), ("n_uvs", ctypes.c_uint32),
                ("normals", ctypes.c_void_p), ("n_normals", ctypes.c_uint32),
                ("tri_idx", ctypes.c_void_p), ("n_triangles", ctypes.c_uint32),
                ("texture", _Image), ("normal_map", _Image), ("normal_map_tangent", _Image), ("specular_map", _Image),
                ("pipeline_name", ctypes.c_char_p), ("device", ctypes.c_int32), ("ring", ctypes.c_uint32)]


class _Mesh(ctypes.Structure):
    _fields_ = [("positions", ctypes.POINTER(ctypes.c_float)), ("n_positions", ctypes.c_uint32),
                ("uvs", ctypes.POINTER(ctypes.c_float)), ("n_uvs", ctypes.c_uint32),
                ("normals", ctypes.POINTER(ctypes.c_float)), ("n_normals", ctypes.c_uint32),
                ("tri_idx", ctypes.POINTER(ctypes.c_uint32)), ("n_triangles", ctypes.c_uint32)]


class _Pixels(ctypes.Structure):
    _fields_ = [("rgb8", ctypes.POINTER(ctypes.c_uint8)), ("width", ctypes.c_uint32), ("height", ctypes.c_uint32)]


class TrView(ctypes.Structure):
    _fields_ = [("look_from", ctypes.c_float * 3), ("look_at", ctypes.c_float * 3), ("up", ctypes.c_float * 3),
                ("light_direction", ctypes.c_float * 3)]


def build(verbose=False):
    """Compile libtinyr.so and the tiny_renderer CLI in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", _CSRC, "all"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("building libtinyr.so failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)
    return LIB_PATH


_LIB = None


def lib():
    """The loaded C ABI.  Fails loudly when the CUDA library has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("tiny_renderer_b200: %s is missing — run __graft_entry__.build() (no CPU fallback exists)" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        vp, u32, i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int
        sig = {
            "tr_scene_create": [ctypes.POINTER(vp), ctypes.POINTER(_SceneDesc)],
            "tr_scene_clear": [vp], "tr_scene_render": [vp], "tr_scene_synchronize": [vp],
            "tr_scene_set_light_direction": [vp, vp], "tr_scene_set_camera": [vp, vp, vp, vp],
            "tr_scene_get_frame_buffer": [vp, vp], "tr_scene_get_z_buffer": [vp, vp], "tr_scene_get_shadow_buffer": [vp, vp],
            "tr_scene_read_frame_raw": [vp, vp], "tr_scene_read_depth": [vp, vp], "tr_scene_read_shadow": [vp, vp],
            "tr_scene_read_visibility": [vp, vp], "tr_scene_set_pipeline": [vp, ctypes.c_char_p],
            "tr_scene_diagnostics": [vp, vp], "tr_scene_get_uniforms": [vp, vp], "tr_scene_get_ao_uniforms": [vp, vp],
            "tr_scene_render_views": [vp, vp, u32, vp, vp], "tr_scene_render_ao": [vp, vp, u32, i32, vp],
            "tr_scene_ao_device_buffer": [vp, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)], "tr_scene_read_ao": [vp, vp],
            "tr_scene_ao_ipc_handle": [vp, vp], "tr_scene_ao_open_peers": [vp, vp, u32, u32],
            "tr_scene_ao_reduce_peers": [vp], "tr_scene_ao_close_peers": [vp],
            "tr_device_count": [ctypes.POINTER(i32)], "tr_scene_set_stream": [vp, vp],
            "tr_scene_device_frame": [vp, ctypes.POINTER(vp)], "tr_scene_timer_begin": [vp],
            "tr_scene_timer_end": [vp, ctypes.POINTER(ctypes.c_float)], "tr_scene_set_profiling": [vp, i32],
            "tr_scene_kernel_times": [vp, vp, vp],
            "tr_load_obj": [ctypes.c_char_p, ctypes.POINTER(_Mesh)], "tr_load_tga": [ctypes.c_char_p, ctypes.POINTER(_Pixels)],
            "tr_write_tga": [ctypes.c_char_p, vp, u32, u32, i32],
            "tr_scene_create_from_dir": [ctypes.POINTER(vp), ctypes.c_char_p, ctypes.c_char_p, u32, u32, ctypes.c_int32, u32],
            "tr_selftest_shared_divisor": [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)],
            "tr_scene_set_debug_flags": [vp, u32],
            "tr_selftest_sqrt": [ctypes.c_int, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)],
            "tr_selftest_round_u32": [ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)],
            "tr_bench_atomic_min": [ctypes.c_int, u32, u32, u32, u32, ctypes.POINTER(ctypes.c_double)],
            "tr_selftest_exact_quotient": [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)],
            "tr_multi_create": [ctypes.POINTER(vp), ctypes.POINTER(_SceneDesc), vp, u32],
            "tr_multi_create_from_dir": [ctypes.POINTER(vp), ctypes.c_char_p, ctypes.c_char_p, u32, u32, vp, u32, u32],
            "tr_multi_set_camera": [vp, vp, vp, vp], "tr_multi_render_views": [vp, vp, u32, vp, vp],
            "tr_multi_render_ao": [vp, vp, u32, vp], "tr_multi_time_exchange": [vp, u32, ctypes.POINTER(ctypes.c_double)],
            "tr_host_invalidate": [vp, ctypes.c_size_t], "tr_scene_transfer_stats": [vp, vp, i32],
            "tr_ao_generate_lights": [vp, u32, vp, vp, ctypes.c_uint64, u32, ctypes.c_float, vp, ctypes.POINTER(ctypes.c_uint64)],
        }
        for name, argtypes in sig.items():
            fn = getattr(L, name)
            fn.argtypes = argtypes
            fn.restype = ctypes.c_int
        L.tr_scene_destroy.argtypes = [vp]; L.tr_scene_destroy.restype = None
        L.tr_multi_destroy.argtypes = [vp]; L.tr_multi_destroy.restype = None
        L.tr_multi_device_count.argtypes = [vp]; L.tr_multi_device_count.restype = ctypes.c_uint32
        L.tr_multi_scene.argtypes = [vp, u32]; L.tr_multi_scene.restype = vp
        L.tr_multi_exchange_kind.argtypes = [vp]; L.tr_multi_exchange_kind.restype = ctypes.c_char_p
        L.tr_mesh_free.argtypes = [ctypes.POINTER(_Mesh)]; L.tr_mesh_free.restype = None
        L.tr_pixels_free.argtypes = [ctypes.POINTER(_Pixels)]; L.tr_pixels_free.restype = None
        L.tr_last_error.restype = ctypes.c_char_p
        L.tr_version.restype = ctypes.c_char_p
        L.tr_checksum_rgb8.argtypes = [vp, ctypes.c_size_t]; L.tr_checksum_rgb8.restype = ctypes.c_uint64
        L.tr_checksum_rgb8_static.argtypes = [ctypes.c_int, vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]
        L.tr_checksum_rgb8_static.restype = ctypes.c_int
        L.tr_host_alloc.argtypes = [ctypes.c_size_t]; L.tr_host_alloc.restype = vp
        L.tr_host_free.argtypes = [vp]; L.tr_host_free.restype = None
        _LIB = L
    return _LIB


def _check(rc):
    if rc != 0:
        raise TinyrError(rc, lib().tr_last_error().decode(errors="replace"))


def _f3(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float32).reshape(3))


from .scene import (Scene, Model, load_obj, load_tga, write_tga, load_model_dir, checksum_rgb8, checksum_rgb8_static, device_count, pinned_empty,  # noqa: E402,F401
                    generate_ao_lights, AO_LIGHT_SEED, AO_NDC_MARGIN, bench_atomic_min, host_invalidate)

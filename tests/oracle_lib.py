"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The method names mirror the reference `Scene` (src/scene.rs:47-151).
"""
import ctypes
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ORACLE_DIR = os.path.join(_ROOT, "oracle")
_LIB = None

NO_TRIANGLE = 0xFFFFFFFF


class _Image(ctypes.Structure):
    _fields_ = [("rgb8", ctypes.c_void_p), ("w", ctypes.c_uint32), ("h", ctypes.c_uint32)]


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", _ORACLE_DIR, "liboracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        # TINYR_ORACLE_LIB: an alternative build of the oracle (tools/oracle_sensitivity.py)
        path = os.environ.get("TINYR_ORACLE_LIB") or os.path.join(_ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        L = ctypes.CDLL(path)
        L.orc_scene_new.restype = ctypes.c_void_p
        L.orc_scene_new.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                    ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32,
                                    ctypes.POINTER(_Image), ctypes.c_char_p]
        for name in ("orc_scene_free", "orc_scene_clear"):
            getattr(L, name).argtypes = [ctypes.c_void_p]
            getattr(L, name).restype = None
        L.orc_scene_render.argtypes = [ctypes.c_void_p]
        L.orc_scene_render.restype = ctypes.c_int
        L.orc_scene_panic.argtypes = [ctypes.c_void_p]
        L.orc_scene_panic.restype = ctypes.c_char_p
        L.orc_scene_set_light_direction.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.orc_scene_set_camera.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p] * 3
        for name in ("orc_scene_get_frame_buffer", "orc_scene_get_z_buffer", "orc_scene_get_shadow_buffer",
                     "orc_scene_raw_frame", "orc_scene_raw_z", "orc_scene_raw_shadow", "orc_scene_raw_visibility",
                     "orc_scene_diagnostics", "orc_scene_uniforms"):
            getattr(L, name).argtypes = [ctypes.c_void_p, ctypes.c_void_p]
            getattr(L, name).restype = None
        L.orc_scene_ao_uniforms.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.orc_scene_ao_uniforms.restype = ctypes.c_int
        L.orc_scene_render_ao.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p]
        L.orc_scene_render_ao.restype = ctypes.c_int
        _LIB = L
    return _LIB


def _f3(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float32).reshape(3))


class OracleScene:
    """CPU oracle with the reference Scene's surface.  `model` is a tests.assets.Model."""

    def __init__(self, width, height, model, pipeline_name):
        L = lib()
        self.width, self.height = int(width), int(height)
        self._keep = [np.ascontiguousarray(model.positions, np.float32), np.ascontiguousarray(model.uvs, np.float32),
                      np.ascontiguousarray(model.normals, np.float32), np.ascontiguousarray(model.tri_idx, np.uint32)]
        maps = (_Image * 4)()
        for k, img in enumerate(model.maps):
            img = np.ascontiguousarray(img, np.uint8)
            self._keep.append(img)
            maps[k].rgb8 = img.ctypes.data
            maps[k].h, maps[k].w = img.shape[0], img.shape[1]
        p, t, n, idx = self._keep[:4]
        self._h = L.orc_scene_new(self.width, self.height, p.ctypes.data, len(p), t.ctypes.data, len(t), n.ctypes.data,
                                  len(n), idx.ctypes.data, len(idx), maps, pipeline_name.encode())
        if not self._h:
            raise ValueError("oracle: unsupported pipeline name or bad indices: %r" % pipeline_name)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_scene_free(self._h)
            self._h = None

    def clear(self):
        lib().orc_scene_clear(self._h)

    def set_light_direction(self, d):
        lib().orc_scene_set_light_direction(self._h, _f3(d).ctypes.data)

    def set_camera(self, look_from, look_at, up):
        a, b, c = _f3(look_from), _f3(look_at), _f3(up)
        lib().orc_scene_set_camera(self._h, a.ctypes.data, b.ctypes.data, c.ctypes.data)

    def render(self, allow_panic=False):
        rc = lib().orc_scene_render(self._h)
        if rc and not allow_panic:
            raise RuntimeError("reference would panic: " + lib().orc_scene_panic(self._h).decode())
        return rc

    def _out(self, fn, shape, dtype):
        a = np.empty(shape, dtype)
        getattr(lib(), fn)(self._h, a.ctypes.data)
        return a

    def get_frame_buffer(self):
        return self._out("orc_scene_get_frame_buffer", (self.height, self.width, 3), np.uint8)

    def get_z_buffer(self):
        return self._out("orc_scene_get_z_buffer", (self.height, self.width, 3), np.uint8)

    def get_shadow_buffer(self):
        return self._out("orc_scene_get_shadow_buffer", (self.height, self.width, 3), np.uint8)

    # raw bottom-up state (row j = y, pixel (0,0) bottom-left)
    def raw_frame(self):
        return self._out("orc_scene_raw_frame", (self.height, self.width, 3), np.uint8)

    def raw_z(self):
        return self._out("orc_scene_raw_z", (self.height, self.width), np.float32)

    def raw_shadow(self):
        return self._out("orc_scene_raw_shadow", (self.height, self.width), np.float32)

    def raw_visibility(self):
        return self._out("orc_scene_raw_visibility", (self.height, self.width), np.uint32)

    def diagnostics(self):
        d = self._out("orc_scene_diagnostics", (6,), np.uint64)
        return dict(zip(("shadow_index_oob", "shadow_x_wrap", "texel_oob", "nan_depth", "bbox_tests", "fragments"),
                        (int(v) for v in d)))

    def uniforms(self):
        u = self._out("orc_scene_uniforms", (102,), np.float32)
        names = ("vpmv", "i_vpmv", "m", "i_m", "it_m", "shadow")
        out = {n: u[16 * k:16 * k + 16].reshape(4, 4).T.copy() for k, n in enumerate(names)}   # column-major -> [r][c]
        out["camera_direction"] = u[96:99].copy()
        out["t_light_direction"] = u[99:102].copy()
        return out

    def ao_uniforms(self):
        """Per-pass constants of the occlusion fragment: 16 sample offsets (16, 3) and shadow * i_vpmv ([r][c])."""
        u = np.empty(64, np.float32)
        if lib().orc_scene_ao_uniforms(self._h, u.ctypes.data):
            raise RuntimeError("reference would panic in the occlusion fragment's set-up")
        return {"ao_step": u[:48].reshape(16, 3).copy(), "shadow_from_screen": u[48:].reshape(4, 4).T.copy()}

    def render_ao(self, light_dirs, accum=None):
        lights = np.ascontiguousarray(light_dirs, np.float32).reshape(-1, 3)
        if accum is None:
            accum = np.zeros((self.height, self.width), np.uint32)
        rc = lib().orc_scene_render_ao(self._h, lights.ctypes.data, len(lights), accum.ctypes.data)
        if rc:
            raise RuntimeError("reference would panic: " + lib().orc_scene_panic(self._h).decode())
        return accum

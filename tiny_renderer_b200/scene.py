"""Host-side mirror of the reference's Scene surface (src/scene.rs:47-151) over the C ABI (include/tinyr.h)."""
import ctypes

import numpy as np

from . import (DIAG_SLOTS, KERNEL_SLOTS, TrView, _check, _f3, _Image, _Mesh, _Pixels, _SceneDesc, lib)


class Model:
    """Flattened obj::raw::RawObj + the four RgbImages of util.rs:16-22."""

    def __init__(self, positions, uvs, normals, tri_idx, maps, name="model"):
        self.positions = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        self.uvs = np.ascontiguousarray(uvs, np.float32).reshape(-1, 2)
        self.normals = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        self.tri_idx = np.ascontiguousarray(tri_idx, np.uint32).reshape(-1, 9)
        self.maps = tuple(np.ascontiguousarray(m, np.uint8) for m in maps)   # texture, normal_map, normal_map_tangent, specular_map
        self.name = name


def device_count():
    n = ctypes.c_int(0)
    lib().tr_device_count(ctypes.byref(n))
    return n.value


def bench_atomic_min(width, height, rounds=4, repeats=5, device=0):
    """Global 64-bit atomicMin rate (atomics/s) on a width x height key plane (SURVEY.md 8(d) micro-benchmark)."""
    rate = ctypes.c_double(0.0)
    _check(lib().tr_bench_atomic_min(device, width, height, rounds, repeats, ctypes.byref(rate)))
    return rate.value


def load_obj(path):
    """obj::raw::parse_obj as the reference uses it (app.rs:94): returns positions, uvs, normals, tri_idx."""
    m = _Mesh()
    _check(lib().tr_load_obj(str(path).encode(), ctypes.byref(m)))
    try:
        p = np.ctypeslib.as_array(m.positions, (m.n_positions, 3)).copy() if m.n_positions else np.zeros((0, 3), np.float32)
        t = np.ctypeslib.as_array(m.uvs, (m.n_uvs, 2)).copy() if m.n_uvs else np.zeros((0, 2), np.float32)
        n = np.ctypeslib.as_array(m.normals, (m.n_normals, 3)).copy() if m.n_normals else np.zeros((0, 3), np.float32)
        i = np.ctypeslib.as_array(m.tri_idx, (m.n_triangles, 9)).copy() if m.n_triangles else np.zeros((0, 9), np.uint32)
    finally:
        lib().tr_mesh_free(ctypes.byref(m))
    return p, t, n, i


def load_tga(path):
    """image::open(path)?.into_rgb8() (app.rs:99): (H, W, 3) u8, row 0 = top."""
    px = _Pixels()
    _check(lib().tr_load_tga(str(path).encode(), ctypes.byref(px)))
    try:
        return np.ctypeslib.as_array(px.rgb8, (px.height, px.width, 3)).copy()
    finally:
        lib().tr_pixels_free(ctypes.byref(px))


def write_tga(path, rgb8_topdown, rle=True):
    img = np.ascontiguousarray(rgb8_topdown, np.uint8)
    _check(lib().tr_write_tga(str(path).encode(), img.ctypes.data, img.shape[1], img.shape[0], 1 if rle else 0))


def load_model_dir(asset_dir, name=None):
    """The five fixed file names of app.rs:87-91."""
    import os
    p, t, n, i = load_obj(os.path.join(asset_dir, "model.obj"))
    maps = tuple(load_tga(os.path.join(asset_dir, f)) for f in
                 ("texture.tga", "normal_map.tga", "normal_map_tangent.tga", "specular_map.tga"))
    return Model(p, t, n, i, maps, name or os.path.basename(os.path.normpath(asset_dir)))


def checksum_rgb8(img):
    a = np.ascontiguousarray(img, np.uint8)
    return int(lib().tr_checksum_rgb8(a.ctypes.data, a.size))


STATIC_CHECKSUM_MAX_BYTES = 8 << 20     # TR_STATIC_CHECKSUM_MAX_BYTES (include/tinyr.h)


def checksum_rgb8_static(img, device=0):
    """tr_checksum_rgb8_static: the device checksum of a host frame with its scratch in a static device array (no
    allocation on the device; at most STATIC_CHECKSUM_MAX_BYTES per call)."""
    a = np.ascontiguousarray(img, np.uint8)
    out = ctypes.c_uint64(0)
    _check(lib().tr_checksum_rgb8_static(int(device), a.ctypes.data, a.size, ctypes.byref(out)))
    return int(out.value)


AO_LIGHT_SEED = 0x5EED0A0
AO_NDC_MARGIN = 0.97


def generate_ao_lights(positions, n_lights, seed=AO_LIGHT_SEED, look_at=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), ndc_margin=AO_NDC_MARGIN):
    """Deterministic light directions for Scene.render_ao (tr_ao_generate_lights): (n, 3) f32 and the draw count."""
    pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
    out = np.zeros((int(n_lights), 3), np.float32)
    drawn = ctypes.c_uint64(0)
    at, upv = _f3(look_at), _f3(up)          # keep the arrays alive across the call
    _check(lib().tr_ao_generate_lights(pos.ctypes.data, len(pos), at.ctypes.data, upv.ctypes.data, int(seed),
                                       int(n_lights), float(ndc_margin), out.ctypes.data, ctypes.byref(drawn)))
    return out, int(drawn.value)


def host_invalidate(array):
    """Tell the library that the caller wrote into a pinned_empty() array (tr_host_invalidate): the next frames delivered
    into it are rewritten whole instead of sparsely."""
    _check(lib().tr_host_invalidate(array.ctypes.data, array.nbytes))


def pinned_empty(shape, dtype):
    """numpy array over tr_host_alloc memory: pinned, device-mapped, zeroed; frames delivered into it travel sparsely
    (include/tinyr.h "Host frames").  Freed with the array."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = lib().tr_host_alloc(max(n, 1))
    if not p:
        raise MemoryError("tr_host_alloc failed")

    class _Owner:
        def __init__(self, ptr):
            self.ptr = ptr

        def __del__(self, _lib=lib):
            try:
                _lib().tr_host_free(self.ptr)
            except Exception:           # interpreter shutdown: the module globals may already be gone
                pass

    buf = (ctypes.c_uint8 * max(n, 1)).from_address(p)
    buf._owner = _Owner(p)
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


def _scene_desc(width, height, model, pipeline_name, device, ring):
    """tr_scene_desc for a Model; returns (desc, arrays to keep alive across the call)."""
    d = _SceneDesc()
    d.width, d.height = width, height
    keep = [np.ascontiguousarray(model.positions, np.float32), np.ascontiguousarray(model.uvs, np.float32),
            np.ascontiguousarray(model.normals, np.float32), np.ascontiguousarray(model.tri_idx, np.uint32)]
    p, t, n, idx = keep
    d.positions, d.n_positions = p.ctypes.data, len(p)
    d.uvs, d.n_uvs = t.ctypes.data, len(t)
    d.normals, d.n_normals = n.ctypes.data, len(n)
    d.tri_idx, d.n_triangles = idx.ctypes.data, len(idx)
    for field, img in zip(("texture", "normal_map", "normal_map_tangent", "specular_map"), model.maps):
        img = np.ascontiguousarray(img, np.uint8)
        keep.append(img)
        setattr(d, field, _Image(img.ctypes.data, img.shape[1], img.shape[0]))
    name = pipeline_name.encode()
    keep.append(name)
    d.pipeline_name = name
    d.device, d.ring = int(device), int(ring)
    return d, keep


class Scene:
    """Scene::new(width, height, obj, texture, normal_map, normal_map_tangent, specular_map, pipeline_name)
    (scene.rs:47-56) with the model flattened into a `Model`."""

    def __init__(self, width, height, model, pipeline_name, device=0, ring=1):
        self.width, self.height = int(width), int(height)
        self.pipeline_name = pipeline_name
        d, self._keep = _scene_desc(self.width, self.height, model, pipeline_name, device, ring)
        h = ctypes.c_void_p()
        self._h = None
        _check(lib().tr_scene_create(ctypes.byref(h), ctypes.byref(d)))
        self._h = h

    @classmethod
    def from_dir(cls, asset_dir, pipeline_name, width=800, height=800, device=0, ring=1):
        self = cls.__new__(cls)
        self.width, self.height, self.pipeline_name, self._keep = int(width), int(height), pipeline_name, []
        h = ctypes.c_void_p()
        self._h = None
        _check(lib().tr_scene_create_from_dir(ctypes.byref(h), str(asset_dir).encode(), pipeline_name.encode(), self.width,
                                              self.height, int(device), int(ring)))
        self._h = h
        return self

    def close(self, _lib=lib):
        if getattr(self, "_h", None):
            try:
                _lib().tr_scene_destroy(self._h)
            except Exception:           # interpreter shutdown: the module globals may already be gone
                pass
            self._h = None

    __del__ = close

    def set_debug_flags(self, flags):
        """Test hooks (tr_scene_set_debug_flags): bit 0 = occlusion samples always use the range-tested division, bit 1 =
        single-frame renders use the batch CTA shape of the tile kernel."""
        _check(lib().tr_scene_set_debug_flags(self._h, int(flags)))

    # ---- reference surface ----
    def clear(self):
        _check(lib().tr_scene_clear(self._h))

    def set_light_direction(self, light_direction):
        d = _f3(light_direction)
        _check(lib().tr_scene_set_light_direction(self._h, d.ctypes.data))

    def set_camera(self, look_from, look_at, up):
        a, b, c = _f3(look_from), _f3(look_at), _f3(up)
        _check(lib().tr_scene_set_camera(self._h, a.ctypes.data, b.ctypes.data, c.ctypes.data))

    def render(self):
        _check(lib().tr_scene_render(self._h))

    def _read(self, fn, shape, dtype, out=None):
        a = out if out is not None else np.empty(shape, dtype)
        _check(getattr(lib(), fn)(self._h, a.ctypes.data))
        return a

    def get_frame_buffer(self, out=None):
        return self._read("tr_scene_get_frame_buffer", (self.height, self.width, 3), np.uint8, out)

    def get_z_buffer(self):
        return self._read("tr_scene_get_z_buffer", (self.height, self.width, 3), np.uint8)

    def get_shadow_buffer(self):
        return self._read("tr_scene_get_shadow_buffer", (self.height, self.width, 3), np.uint8)

    # ---- additions ----
    def raw_frame(self):
        return self._read("tr_scene_read_frame_raw", (self.height, self.width, 3), np.uint8)

    def raw_z(self):
        return self._read("tr_scene_read_depth", (self.height, self.width), np.float32)

    def raw_shadow(self):
        return self._read("tr_scene_read_shadow", (self.height, self.width), np.float32)

    def raw_visibility(self):
        return self._read("tr_scene_read_visibility", (self.height, self.width), np.uint32)

    def set_pipeline(self, name):
        _check(lib().tr_scene_set_pipeline(self._h, name.encode()))
        self.pipeline_name = name

    def synchronize(self):
        _check(lib().tr_scene_synchronize(self._h))

    def diagnostics(self):
        d = self._read("tr_scene_diagnostics", (8,), np.uint64)
        return dict(zip(DIAG_SLOTS, (int(v) for v in d)))

    def uniforms(self):
        u = self._read("tr_scene_get_uniforms", (102,), np.float32)
        names = ("vpmv", "i_vpmv", "m", "i_m", "it_m", "shadow")
        out = {n: u[16 * k:16 * k + 16].reshape(4, 4).copy() for k, n in enumerate(names)}
        out["camera_direction"] = u[96:99].copy()
        out["t_light_direction"] = u[99:102].copy()
        return out

    def ao_uniforms(self):
        """Hoisted constants of the shadow / occlusion fragments (tr_scene_get_ao_uniforms)."""
        u = self._read("tr_scene_get_ao_uniforms", (64,), np.float32)
        return {"ao_step": u[:48].reshape(16, 3).copy(), "shadow_from_screen": u[48:].reshape(4, 4).copy()}

    def render_views(self, views, want_checksums=True, frames_out=None):
        """views: (n, 12) f32 rows look_from, look_at, up, light_direction.  Returns u64 checksums (or None)."""
        v = np.ascontiguousarray(views, np.float32).reshape(-1, 12)
        cs = np.zeros(len(v), np.uint64) if want_checksums else None
        _check(lib().tr_scene_render_views(self._h, v.ctypes.data, len(v), cs.ctypes.data if cs is not None else None,
                                           frames_out.ctypes.data if frames_out is not None else None))
        return cs

    def render_ao(self, light_dirs, reset=True, read=True):
        lights = np.ascontiguousarray(light_dirs, np.float32).reshape(-1, 3)
        out = np.empty((self.height, self.width), np.uint32) if read else None
        _check(lib().tr_scene_render_ao(self._h, lights.ctypes.data, len(lights), 1 if reset else 0,
                                        out.ctypes.data if out is not None else None))
        return out

    def ao_device_buffer(self):
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        _check(lib().tr_scene_ao_device_buffer(self._h, ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def read_ao(self):
        return self._read("tr_scene_read_ao", (self.height, self.width), np.uint32)

    # Multi-GPU sum of the AO accumulators over NVLink peer memory (include/tinyr.h; one process per GPU).
    def ao_ipc_handle(self):
        """64-byte CUDA IPC handle of this scene's accumulator, for the other ranks' ao_open_peers."""
        h = ctypes.create_string_buffer(64)
        _check(lib().tr_scene_ao_ipc_handle(self._h, h))
        return h.raw

    def ao_open_peers(self, handles, rank):
        """handles: one 64-byte handle per rank, in rank order (this rank's own entry is ignored)."""
        world = len(handles)
        if any(len(h) != 64 for h in handles):
            raise ValueError("every handle is 64 bytes")
        blob = ctypes.create_string_buffer(b"".join(handles), 64 * world)
        _check(lib().tr_scene_ao_open_peers(self._h, blob, world, rank))

    def ao_reduce_peers(self):
        """Collective over the ranks of ao_open_peers; asynchronous on the scene stream."""
        _check(lib().tr_scene_ao_reduce_peers(self._h))

    def ao_close_peers(self):
        _check(lib().tr_scene_ao_close_peers(self._h))

    def transfer_stats(self, reset=False):
        """Host-frame delivery counters (tr_scene_transfer_stats)."""
        o = np.zeros(5, np.uint64)
        _check(lib().tr_scene_transfer_stats(self._h, o.ctypes.data, 1 if reset else 0))
        return {"sparse_bytes": int(o[0]), "sparse_rows": int(o[1]), "frames_sparse": int(o[2]), "frames_copied": int(o[3]),
                "copied_bytes": int(o[4])}

    def set_stream(self, cuda_stream):
        _check(lib().tr_scene_set_stream(self._h, ctypes.c_void_p(cuda_stream)))

    def device_frame(self):
        p = ctypes.c_void_p()
        _check(lib().tr_scene_device_frame(self._h, ctypes.byref(p)))
        return p.value

    def timer_begin(self):
        _check(lib().tr_scene_timer_begin(self._h))

    def timer_end(self):
        ms = ctypes.c_float()
        _check(lib().tr_scene_timer_end(self._h, ctypes.byref(ms)))
        return ms.value

    def set_profiling(self, enabled):
        _check(lib().tr_scene_set_profiling(self._h, 1 if enabled else 0))

    def kernel_times(self):
        ms = np.zeros(8, np.float32)
        n = np.zeros(8, np.uint64)
        _check(lib().tr_scene_kernel_times(self._h, ms.ctypes.data, n.ctypes.data))
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(KERNEL_SLOTS)}


class _BorrowedScene(Scene):
    """A scene owned by a MultiScene: closing (or collecting) the view does not destroy it."""

    def close(self):
        self._h = None

    __del__ = close


class MultiScene:
    """Native multi-GPU driver (tr_multi_*, include/tinyr.h): one host process, one scene per device.  `devices` lists
    CUDA ordinals; listing a device twice puts two scenes on it (a single-GPU box can exercise the sharded paths)."""

    def __init__(self, width, height, model, pipeline_name, devices, ring=1):
        self.width, self.height = int(width), int(height)
        d, self._keep = _scene_desc(self.width, self.height, model, pipeline_name, 0, ring)
        dev = np.ascontiguousarray(devices, np.int32)
        h = ctypes.c_void_p()
        self._h = None
        _check(lib().tr_multi_create(ctypes.byref(h), ctypes.byref(d), dev.ctypes.data, len(dev)))
        self._h = h

    def close(self, _lib=lib):
        if getattr(self, "_h", None):
            try:
                _lib().tr_multi_destroy(self._h)
            except Exception:
                pass
            self._h = None

    __del__ = close

    def device_count(self):
        return int(lib().tr_multi_device_count(self._h))

    def scene(self, i):
        """Borrowed view of the i-th device's scene (owned by the MultiScene)."""
        s = _BorrowedScene.__new__(_BorrowedScene)
        s.width, s.height, s.pipeline_name, s._keep = self.width, self.height, None, []
        s._h = ctypes.c_void_p(lib().tr_multi_scene(self._h, i))
        return s

    def set_camera(self, look_from, look_at, up):
        a, b, c = _f3(look_from), _f3(look_at), _f3(up)
        _check(lib().tr_multi_set_camera(self._h, a.ctypes.data, b.ctypes.data, c.ctypes.data))

    def render_views(self, views, want_checksums=True, frames_out=None):
        v = np.ascontiguousarray(views, np.float32).reshape(-1, 12)
        cs = np.zeros(len(v), np.uint64) if want_checksums else None
        _check(lib().tr_multi_render_views(self._h, v.ctypes.data, len(v), cs.ctypes.data if cs is not None else None,
                                           frames_out.ctypes.data if frames_out is not None else None))
        return cs

    def exchange_kind(self):
        """"nvls" (in-switch multimem reduction) or "peer: <why not nvls>" (tr_multi_exchange_kind)."""
        return lib().tr_multi_exchange_kind(self._h).decode()

    def time_exchange(self, reps=5):
        """us per exchange of the accumulators, slowest device (tr_multi_time_exchange); the sums are garbage afterwards."""
        us = ctypes.c_double(0.0)
        _check(lib().tr_multi_time_exchange(self._h, int(reps), ctypes.byref(us)))
        return us.value

    def render_ao(self, light_dirs):
        lights = np.ascontiguousarray(light_dirs, np.float32).reshape(-1, 3)
        out = np.empty((self.height, self.width), np.uint32)
        _check(lib().tr_multi_render_ao(self._h, lights.ctypes.data, len(lights), out.ctypes.data))
        return out


def make_views(look_from, look_at, up, light):
    """Stack per-view vectors into the (n, 12) layout of tr_view."""
    n = len(look_from)
    v = np.zeros((n, 12), np.float32)
    v[:, 0:3] = look_from; v[:, 3:6] = look_at; v[:, 6:9] = up; v[:, 9:12] = light
    return v

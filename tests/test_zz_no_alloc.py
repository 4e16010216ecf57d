"""-m gpu, collected last: tr_checksum_rgb8_static — the device checksum with its scratch in a static __device__ array
(csrc/tinyr_static.cu; no cudaMalloc) — against the existing entry point's device checksums (tr_scene_render_views),
bit for bit (integer work: tolerance 0), on two calls in a row, and its size limit."""
import ctypes
import math

import numpy as np
import pytest

import assets
from conftest import get_model

pytestmark = pytest.mark.gpu


def _views(n):
    v = np.zeros((n, 12), np.float32)
    for k in range(n):
        lf, la, up, light = assets.view_vectors(2 * math.pi * k / n, 0.7)
        v[k, 0:3], v[k, 3:6], v[k, 6:9], v[k, 9:12] = lf, la, up, light
    return v


def _static_checksum(tr, frame, device=0):
    """Straight through the C ABI."""
    a = np.ascontiguousarray(frame, np.uint8)
    out = ctypes.c_uint64(0xDEADBEEF)
    rc = tr.lib().tr_checksum_rgb8_static(device, a.ctypes.data, a.size, ctypes.byref(out))
    return rc, int(out.value)


def test_static_checksum_equals_render_views_checksums(tinyr):
    from tiny_renderer_b200.scene import Scene
    tr = tinyr
    m = get_model("african_head")
    w, h = 640, 360
    g = Scene(w, h, m, "phong")
    frames = np.zeros((4, h, w, 3), np.uint8)
    checksums = g.render_views(_views(4), frames_out=frames)        # the existing entry point: K8 on cudaMalloc'ed scratch
    g.close()
    assert len({int(c) for c in checksums}) == 4            # four different frames
    for rep in range(2):                                    # two calls in a row per frame: the array keeps the last call's sum
        for k in (0, 1, 2, 3, 3, 0):
            rc, got = _static_checksum(tr, frames[k])
            assert rc == 0, tr.lib().tr_last_error()
            assert got == int(checksums[k]), (rep, k)
            assert got == tr.checksum_rgb8(frames[k])       # and the host twin
    # a shorter input after a longer one: nothing of the longer frame is read; a length that is no multiple of 4
    for n in (0, 4, 1000, 12345, w * 3):
        part = np.ascontiguousarray(frames[1].reshape(-1)[:n])
        rc, got = _static_checksum(tr, part)
        assert rc == 0 and got == tr.checksum_rgb8(part), n
    assert tr.checksum_rgb8_static(frames[2]) == int(checksums[2])


def test_static_checksum_largest_size_and_limit(tinyr):
    tr = tinyr
    from tiny_renderer_b200.scene import STATIC_CHECKSUM_MAX_BYTES
    rng = np.random.default_rng(5)
    full = rng.integers(0, 256, STATIC_CHECKSUM_MAX_BYTES, dtype=np.uint8)
    for rep in range(2):
        rc, got = _static_checksum(tr, full)
        assert rc == 0 and got == tr.checksum_rgb8(full), rep
    too_long = np.zeros(STATIC_CHECKSUM_MAX_BYTES + 4, np.uint8)
    rc, got = _static_checksum(tr, too_long)
    assert rc != 0 and got == 0xDEADBEEF
    msg = tr.lib().tr_last_error().decode()
    assert "exceed" in msg and str(STATIC_CHECKSUM_MAX_BYTES) in msg, msg
    with pytest.raises(tr.TinyrError):
        tr.checksum_rgb8_static(too_long)
    rc, got = _static_checksum(tr, full[:4096])             # still usable after the rejected call
    assert rc == 0 and got == tr.checksum_rgb8(full[:4096])

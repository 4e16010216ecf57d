"""GPU <-> oracle comparison with the tolerances BASELINE.json's north_star states:
visibility (triangle id per pixel) bit-exact, depth within 1 ulp of f32, RGB within 1 LSB per channel on
>= 99.9% of pixels.  Returns a dict of statistics so callers can assert tighter bounds too."""
import numpy as np

import assets


def ulp_diff(a, b):
    """|a-b| in units of f32 ulps via the monotone integer mapping (handles sign, -0 == +0)."""
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, np.int64(-2147483648) - ia, ia)
    ib = np.where(ib < 0, np.int64(-2147483648) - ib, ib)
    return np.abs(ia - ib)


def drive(scene, look_from, look_at, up, light, clear=True):
    """The app loop's per-frame sequence (app.rs:170, 208-210)."""
    if clear:
        scene.clear()
    scene.set_light_direction(light)
    scene.set_camera(look_from, look_at, up)
    scene.render()


def compare(gpu, orc, two_pass=False, rgb_frac=0.999):
    stats = {}
    vg, vo = gpu.raw_visibility(), orc.raw_visibility()
    stats["vis_mismatch"] = int((vg != vo).sum())
    stats["covered"] = int((vo != 0xFFFFFFFF).sum())
    zg, zo = gpu.raw_z(), orc.raw_z()
    stats["depth_max_ulp"] = int(ulp_diff(zg, zo).max())
    stats["depth_bits_differ"] = int((zg.view(np.uint32) != zo.view(np.uint32)).sum())
    if two_pass:
        sg, so = gpu.raw_shadow(), orc.raw_shadow()
        stats["shadow_max_ulp"] = int(ulp_diff(sg, so).max())
    fg, fo = gpu.get_frame_buffer(), orc.get_frame_buffer()
    d = np.abs(fg.astype(np.int16) - fo.astype(np.int16)).max(axis=2)
    stats["rgb_exact_frac"] = float((d == 0).mean())
    stats["rgb_within1_frac"] = float((d <= 1).mean())
    stats["rgb_max_diff"] = int(d.max())
    # raw (bottom-up) frame must be the flip of the top-down one
    stats["raw_is_flip"] = bool(np.array_equal(gpu.raw_frame()[::-1], fg))
    assert stats["vis_mismatch"] == 0, stats
    assert stats["depth_max_ulp"] <= 1, stats
    if two_pass:
        assert stats["shadow_max_ulp"] <= 1, stats
    assert stats["rgb_within1_frac"] >= rgb_frac, stats
    assert stats["raw_is_flip"], stats
    return stats


def both(tr, oracle_lib, model, pipeline, width, height, camera_angle, light_angle, device=0):
    from tiny_renderer_b200.scene import Scene
    g = Scene(width, height, model, pipeline, device=device)
    o = oracle_lib.OracleScene(width, height, model, pipeline)
    lf, la, up, light = assets.view_vectors(camera_angle, light_angle)
    drive(g, lf, la, up, light)
    drive(o, lf, la, up, light)
    return g, o

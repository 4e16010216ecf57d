"""-m gpu: the `tiny_renderer` command line (reference: src/main.rs:12-26 `-p`, `-s`; window replaced by TGA output)."""
import os
import re
import subprocess

import numpy as np
import pytest

import assets
import parity
from conftest import get_model

pytestmark = pytest.mark.gpu


def _run(tinyr, *args):
    out = subprocess.run([tinyr.CLI_PATH, *map(str, args)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    return out.stdout


def test_cli_config1_headless_tga(tinyr, oracle, tmp_path):
    """BASELINE config 1 through the command line: african_head, default, 800x800 -> TGA == the oracle's frame."""
    d = assets.asset_dir("african_head")
    if d is None:
        pytest.skip("reference assets not staged")
    path = tmp_path / "out.tga"
    text = _run(tinyr, "-p", d, "-s", "default", "--out", path)
    assert "loading model from" in text and "default" in text
    frame = tinyr.load_tga(path)
    o = oracle.OracleScene(800, 800, get_model("african_head"), "default")
    lf, la, up, l = assets.view_vectors(0.0, 0.0)
    parity.drive(o, lf, la, up, l)
    assert frame.shape == (800, 800, 3) and np.array_equal(frame, o.get_frame_buffer())


def test_cli_true_normal_alias_and_debug_views(tinyr, tmp_path):
    d = assets.asset_dir("diablo")
    if d is None:
        pytest.skip("reference assets not staged")
    a, b, z, sh = (tmp_path / n for n in ("a.tga", "b.tga", "z.tga", "s.tga"))
    _run(tinyr, "-p", d, "-s", "true_normal", "--width", 320, "--height", 200, "--camera-angle", 0.5, "--out", a)
    _run(tinyr, "-p", d, "-s", "normal_map", "--width", 320, "--height", 200, "--camera-angle", 0.5, "--out", b)
    assert np.array_equal(tinyr.load_tga(a), tinyr.load_tga(b))
    _run(tinyr, "-p", d, "-s", "shadow", "--width", 320, "--height", 200, "--light-angle", 0.7, "--out", a, "--dump-z", z, "--dump-shadow", sh)
    from tiny_renderer_b200.scene import Scene
    g = Scene(320, 200, get_model("diablo"), "shadow")
    lf, la, up, l = assets.view_vectors(0.0, 0.7)
    parity.drive(g, lf, la, up, l)
    assert np.array_equal(tinyr.load_tga(z), g.get_z_buffer())
    assert np.array_equal(tinyr.load_tga(sh), g.get_shadow_buffer())
    out = subprocess.run([tinyr.CLI_PATH, "-p", d, "-s", "true_normals"], capture_output=True, text=True)
    assert out.returncode == 1 and "not supported" in out.stderr            # shader.rs:108


def test_cli_scripted_replay(tinyr, tmp_path):
    """The app loop (app.rs:166-247) replayed from a key script with a fixed frame time: the angles integrate
    SPEED * dt in f32 exactly as the reference does, and the last frame is the render at those angles."""
    d = assets.asset_dir("african_head")
    if d is None:
        pytest.skip("reference assets not staged")
    script = tmp_path / "keys.txt"
    lines = ["d", "d", "dq", "", "a", "e", "e", "d x"]
    script.write_text("\n".join(lines) + "\n")
    path = tmp_path / "last.tga"
    text = _run(tinyr, "-p", d, "-s", "phong", "--width", 256, "--height", 256, "--replay", script, "--dt", 0.016, "--out", path, "--quiet")
    m = re.search(r"replayed (\d+) frames .* camera angle (\S+), light angle (\S+)", text)
    assert m, text
    f32 = np.float32
    cam, light, dt = f32(0), f32(0), f32(0.016)
    frame_time = f32(0)
    pressed = ""
    shown = 0
    for ln in lines:                         # frame k applies the keys read during frame k-1 with frame k-1's dt
        if "d" in pressed: cam = f32(cam + f32(3.0) * frame_time)
        if "a" in pressed: cam = f32(cam - f32(3.0) * frame_time)
        if "e" in pressed: light = f32(light + f32(3.0) * frame_time)
        if "q" in pressed: light = f32(light - f32(3.0) * frame_time)
        shown += 1
        pressed = ln
        frame_time = dt
        if "x" in ln:
            break
    assert int(m.group(1)) == shown == 8
    assert f32(float(m.group(2))) == cam and f32(float(m.group(3))) == light
    # the last frame equals a single-frame render at the final angles (same process, same sinf/cosf)
    single = tmp_path / "single.tga"
    _run(tinyr, "-p", d, "-s", "phong", "--width", 256, "--height", 256, "--camera-angle", m.group(2), "--light-angle", m.group(3),
         "--out", single)
    assert np.array_equal(tinyr.load_tga(path), tinyr.load_tga(single))
    assert tinyr.load_tga(path).any()


def test_cli_batch_modes_views_and_lights(tinyr, oracle, tmp_path):
    """SURVEY.md §5 flags: --views N (config 5 shape) and --lights N --seed S (config 4 shape) at test size through the
    command line and the native multi-GPU driver, on 1 GPU and — when the box has them — on 2; --dump-visibility."""
    import torch
    d = assets.asset_dir("diablo")
    if d is None:
        pytest.skip("reference assets not staged")
    m = get_model("diablo")
    w, h = 320, 240
    gpus = [1] + ([2] if torch.cuda.device_count() >= 2 else [])
    # config 4: the CLI's image is round(sum of gray frames / n) of the oracle's sum over the same seeded lights
    lights, _ = tinyr.generate_ao_lights(m.positions, 6, seed=1234)
    o = oracle.OracleScene(w, h, m, "occlusion")
    lf, la, up, _ = assets.view_vectors(0.0, 0.0)
    o.set_camera(lf, la, up)
    acc = o.render_ao(lights)
    want = ((acc.astype(np.uint64) * 2 + 6) // 12).astype(np.uint8)[::-1]
    for g in gpus:
        path, vis = tmp_path / ("ao%d.tga" % g), tmp_path / ("vis%d.bin" % g)
        text = _run(tinyr, "-p", d, "-s", "occlusion", "--width", w, "--height", h, "--lights", 6, "--seed", 1234, "--gpus", g,
                    "--out", path, "--dump-visibility", vis)
        assert re.search(r"ao: 6 lights .* on %d device" % g, text), text
        img = tinyr.load_tga(path)
        assert np.array_equal(img[..., 0], want) and np.array_equal(img[..., 1], want)
        assert np.array_equal(np.fromfile(vis, np.uint32).reshape(h, w), o.raw_visibility())
    # config 5: per-view checksums equal the host twin's checksum of the oracle-exact frames (phong: bit-exact pipeline)
    n = 5
    sums = []
    og = oracle.OracleScene(w, h, m, "phong")
    import ctypes
    libm = ctypes.CDLL("libm.so.6")             # the CLI's sinf / cosf (numpy's float32 sin is its own SIMD routine)
    for f in (libm.sinf, libm.cosf):
        f.argtypes, f.restype = [ctypes.c_float], ctypes.c_float
    for k in range(n):
        a = np.float32(0.25) + np.float32(2.0) * np.float32(3.14159274101257324) * np.float32(k) / np.float32(n)
        lf = np.array([libm.sinf(float(a)), 0.0, libm.cosf(float(a))], np.float32)
        light = np.array([libm.sinf(0.7), 0.0, libm.cosf(0.7)], np.float32)
        parity.drive(og, lf, np.zeros(3, np.float32), np.array([0, 1, 0], np.float32), light)
        sums.append(tinyr.checksum_rgb8(og.get_frame_buffer()))
    for g in gpus:
        path = tmp_path / ("views%d.tga" % g)
        text = _run(tinyr, "-p", d, "-s", "phong", "--width", w, "--height", h, "--views", n, "--camera-angle", 0.25, "--light-angle", 0.7,
                    "--gpus", g, "--out", path)
        got = [int(x) for x in re.findall(r"view \d+ checksum (\d+)", text)]
        assert got == sums, (g, got, sums)
        assert np.array_equal(tinyr.load_tga(path), og.get_frame_buffer())

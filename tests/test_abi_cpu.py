"""CPU suite (-m "not gpu"), part 2: the C-ABI library and the host-side code that needs no GPU.

* libtinyr.so loads and exports every symbol include/tinyr.h declares;
* the host loaders kept from the reference (OBJ: obj-rs raw::parse_obj as used at app.rs:94, TGA:
  image::open(..).into_rgb8() app.rs:99) agree with independent Python loaders, the TGA writer round-trips;
* without a CUDA device every compute entry point fails loudly (TR_ERR_NO_DEVICE) — there is no CPU fallback;
* the product never touches oracle/.
"""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import assets
from conftest import ROOT, get_model


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_every_declared_symbol_is_exported(tinyr):
    hdr = open(os.path.join(ROOT, "include", "tinyr.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(tr_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 39, names
    L = ctypes.CDLL(tinyr.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert tinyr.lib().tr_version().decode().startswith("tinyr")


def test_library_is_sm100a_only(tinyr):
    out = subprocess.run(["cuobjdump", "-lelf", tinyr.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


@pytest.mark.skipif(_has_gpu(), reason="box has a GPU")
def test_no_cpu_fallback(tinyr):
    from tiny_renderer_b200.scene import Scene
    with pytest.raises(tinyr.TinyrError) as e:
        Scene(64, 64, get_model("african_head"), "default")
    assert e.value.name == "TR_ERR_NO_DEVICE"
    assert "no CPU fallback" in str(e.value)
    assert tinyr.device_count() == 0


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "tiny_renderer_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".inc")) or f == "Makefile":
                text = open(os.path.join(base, f), errors="replace").read()
                assert "oracle" not in text.lower(), os.path.join(base, f)


def test_null_and_argument_errors(tinyr):
    L = tinyr.lib()
    assert L.tr_scene_clear(None) == -1 and b"null scene" in L.tr_last_error()
    assert L.tr_scene_render(None) == -1
    assert L.tr_scene_create(None, None) == -1
    m = tinyr.scene._Mesh() if hasattr(tinyr, "scene") else None
    with pytest.raises(tinyr.TinyrError) as e:
        tinyr.load_obj("/nonexistent/model.obj")
    assert e.value.name == "TR_ERR_IO"
    with pytest.raises(tinyr.TinyrError) as e:
        tinyr.load_tga("/nonexistent/texture.tga")
    assert e.value.name == "TR_ERR_IO"
    del m


@pytest.mark.skipif(assets.asset_dir("diablo") is None, reason="reference assets not staged")
@pytest.mark.parametrize("asset", ["african_head", "diablo"])
def test_native_loaders_match_python_loaders(tinyr, asset):
    d = assets.asset_dir(asset)
    ref = get_model(asset)
    p, t, n, idx = tinyr.load_obj(os.path.join(d, "model.obj"))
    assert np.array_equal(p.view(np.uint32), ref.positions.view(np.uint32))     # correctly rounded decimal -> f32
    assert np.array_equal(t.view(np.uint32), ref.uvs.view(np.uint32))
    assert np.array_equal(n.view(np.uint32), ref.normals.view(np.uint32))
    assert np.array_equal(idx, ref.tri_idx)
    for k, f in enumerate(assets.MAP_FILES):          # RLE true-colour 24/32 bpp and RLE gray, bottom-left origin
        assert np.array_equal(tinyr.load_tga(os.path.join(d, f)), ref.maps[k]), f


def test_obj_parser_edge_cases(tinyr, tmp_path):
    """Negative (relative) indices, comments, extra corners (only the first three are used, scene.rs:224-226),
    4-component positions / 3-component uvs (ignored components), and rejection of non p/t/n polygons
    (scene.rs:216-219)."""
    good = tmp_path / "a.obj"
    good.write_text("# comment\nv 0 0 0 1.0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nvt 0.1 0.2 0.9\nvt 0.3 0.4\nvt 0.5 0.6\n"
                    "vn 0 0 1\nvn 0 1 0\nvn 1 0 0\n\nf 1/1/1 2/2/2 3/3/3 4/1/1\nf -1/-1/-1 -2/-2/-2 -3/-3/-3\n")
    p, t, n, idx = tinyr.load_obj(good)
    assert p.shape == (4, 3) and t.shape == (3, 2) and n.shape == (3, 3)
    assert np.allclose(t[0], [0.1, 0.2])
    assert idx.tolist() == [[0, 0, 0, 1, 1, 1, 2, 2, 2], [3, 2, 2, 2, 1, 1, 1, 0, 0]]
    bad = tmp_path / "b.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    with pytest.raises(tinyr.TinyrError) as e:
        tinyr.load_obj(bad)
    assert e.value.name in ("TR_ERR_BAD_POLYGON", "TR_ERR_FORMAT")
    assert np.array_equal(assets.parse_obj_py(good)[3], idx)


@pytest.mark.parametrize("rle", [False, True])
@pytest.mark.parametrize("shape", [(1, 1), (3, 130), (64, 64), (37, 300)])
def test_tga_write_read_roundtrip(tinyr, tmp_path, shape, rle):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    img = rng.integers(0, 256, shape + (3,)).astype(np.uint8)
    img[:, : shape[1] // 2] = img[:, :1]                     # long runs for the RLE packets (> 128 px)
    path = tmp_path / "x.tga"
    tinyr.write_tga(path, img, rle=rle)
    assert np.array_equal(tinyr.load_tga(path), img)
    assert np.array_equal(assets.load_tga_pil(path), img)     # independent decoder agrees


def test_tga_rejects_garbage(tinyr, tmp_path):
    p = tmp_path / "bad.tga"
    p.write_bytes(b"\x00" * 10)
    with pytest.raises(tinyr.TinyrError) as e:
        tinyr.load_tga(p)
    assert e.value.name in ("TR_ERR_FORMAT", "TR_ERR_IO")


def test_checksum_host_twin(tinyr):
    """tr_checksum_rgb8 = sum_i word_i * (2i+1) mod 2^64 over little-endian u32 words."""
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, 4096 * 3).astype(np.uint8)
    w = a.view("<u4").astype(object)
    expect = sum(int(x) * (2 * i + 1) for i, x in enumerate(w)) % (1 << 64)
    assert tinyr.checksum_rgb8(a) == expect


def test_cli_usage_and_missing_assets(tinyr):
    """The CLI keeps -p / -s (main.rs:12-26); missing assets are an I/O error, not a crash."""
    out = subprocess.run([tinyr.CLI_PATH, "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "-p asset_dir" in out.stdout and "true_normal" in out.stdout
    out = subprocess.run([tinyr.CLI_PATH, "-p", "/nonexistent", "-s", "phong", "--out", "/tmp/none.tga"], capture_output=True, text=True)
    assert out.returncode == 1 and "TR_ERR" not in out.stdout and "model.obj" in (out.stderr + out.stdout)


def test_bench_finds_the_committed_ncu_summaries():
    """bench.py fills roofline.traffic and the issue-rate rooflines from the newest profiles/r*_traffic.json; a kernel
    renamed without a fresh capture would silently turn them into nulls."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    rk, sk = bench.raster_kernel, bench.setup_kernel
    for want in (rk("phong", 256), rk("darboux", 256), rk("phong", 128), rk("darboux", 128), rk("light_depth", 128),
                 "k_ao_accumulate", sk("phong"), sk("darboux")):
        e = bench.ncu_kernel_summary(want)
        assert e is not None and e["warp_instructions"] > 0 and e["dram_bytes_per_launch"] > 0, want
    assert bench.ncu_kernel_summary(rk("darboux", 256))["dram_bytes_per_launch"] > bench.ncu_kernel_summary(rk("phong", 256))["dram_bytes_per_launch"] > 0
    r = bench.issue_roofline([rk("light_depth", 128), "k_ao_accumulate"], 56.6, 148, 1965.0)
    assert 0.5 < r["frac"] < 1.0 and abs(r["peak"] - 148 * 4 * 1.965) < 1e-6
    assert bench.issue_roofline(["no_such_kernel"], 10.0, 148, 1965.0) is None

"""-m gpu: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bars (BASELINE.json north_star): visibility bit-exact, depth <= 1 ulp, RGB within 1 LSB on >= 99.9% of pixels."""
import math

import numpy as np
import pytest

import assets
import parity
from conftest import get_model

pytestmark = pytest.mark.gpu

PIPELINES = ("default", "phong", "normal_map", "specular", "darboux", "shadow", "occlusion")
TWO_PASS = ("shadow", "occlusion")


@pytest.mark.parametrize("pipeline", PIPELINES)
@pytest.mark.parametrize("asset", ["african_head", "diablo"])
def test_all_pipelines_800(tinyr, oracle, asset, pipeline):
    """BASELINE config 1 shape (800x800) for every pipeline x both assets, two views each."""
    model = get_model(asset)
    for cam, light in ((0.0, 0.0), (2.3, 0.7)):
        g, o = parity.both(tinyr, oracle, model, pipeline, 800, 800, cam, light)
        st = parity.compare(g, o, two_pass=pipeline in TWO_PASS)
        assert st["covered"] > 10000
        if pipeline != "specular":          # only powf may differ between libm and CUDA
            assert st["rgb_exact_frac"] == 1.0, st
        gd, od = g.diagnostics(), o.diagnostics()
        for k in ("shadow_index_oob", "shadow_x_wrap", "texel_oob", "nan_depth"):
            assert gd[k] == od[k], (k, gd, od)
        g.close()


def test_config1_african_head_default_frame(tinyr, oracle):
    """BASELINE config 1: african_head, default, 800x800, camera/light angle 0: frame bit-identical."""
    model = get_model("african_head")
    g, o = parity.both(tinyr, oracle, model, "default", 800, 800, 0.0, 0.0)
    assert np.array_equal(g.get_frame_buffer(), o.get_frame_buffer())
    assert np.array_equal(g.get_z_buffer(), o.get_z_buffer())
    assert np.array_equal(g.raw_visibility(), o.raw_visibility())


@pytest.mark.parametrize("pipeline", ["phong", "darboux"])
def test_config2_diablo_4k(tinyr, oracle, pipeline):
    """BASELINE config 2: diablo, phong + darboux, 3840x2160, light 0.7 rad; 2 of the 8 camera angles here
    (the oracle needs ~0.5 s per 4K frame), all 8 in the bench's parity leg."""
    model = get_model("diablo")
    for k in (0, 3):
        g, o = parity.both(tinyr, oracle, model, pipeline, 3840, 2160, 2 * math.pi * k / 8, 0.7)
        st = parity.compare(g, o)
        assert st["rgb_exact_frac"] == 1.0, st
        g.close()


def test_config3_diablo_shadow_4k(tinyr, oracle):
    model = get_model("diablo")
    g, o = parity.both(tinyr, oracle, model, "shadow", 3840, 2160, 2 * math.pi * 1 / 8, 2.1)
    st = parity.compare(g, o, two_pass=True)
    assert st["rgb_exact_frac"] == 1.0, st
    assert g.diagnostics()["shadow_index_oob"] == 0


@pytest.mark.parametrize("size", [(1, 1), (7, 5), (64, 32), (65, 33), (127, 250), (801, 603)])
def test_ragged_sizes(tinyr, oracle, size):
    """Frame sizes that are not multiples of the 64x32 tile or of the 4-pixel shading group."""
    model = get_model("african_head")
    for pipeline in ("phong", "shadow"):
        if pipeline == "shadow" and min(size) == 1:
            # viewport scale (w-1)/2 = 0 makes vpmv singular: the reference panics in shadow_pass_prepare_2
            # (shader.rs:277); the backend reports it instead of rendering.
            with pytest.raises(tinyr.TinyrError) as e:
                parity.both(tinyr, oracle, model, pipeline, size[0], size[1], 0.9, -0.4)
            assert e.value.name == "TR_ERR_REFERENCE_UNDEFINED"
            continue
        g, o = parity.both(tinyr, oracle, model, pipeline, size[0], size[1], 0.9, -0.4)
        parity.compare(g, o, two_pass=pipeline in TWO_PASS, rgb_frac=1.0)
        g.close()


def test_render_without_clear_and_rerender(tinyr, oracle):
    """Buffer::new starts z at 0.0 (shader.rs:46-47): a render before any clear only draws z > 0; a second
    render without clear keeps earlier winners on ties (scene.rs contract: render does not clear)."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("diablo")
    g = Scene(400, 300, model, "phong")
    o = oracle.OracleScene(400, 300, model, "phong")
    for cam, light in ((0.0, 0.0), (0.5, 1.0), (0.5, 2.0)):
        lf, la, up, l = assets.view_vectors(cam, light)
        parity.drive(g, lf, la, up, l, clear=False)
        parity.drive(o, lf, la, up, l, clear=False)
        parity.compare(g, o, rgb_frac=1.0)
    g.clear(); o.clear()
    assert np.array_equal(g.get_frame_buffer(), o.get_frame_buffer())       # cleared frame, materialised lazily
    assert np.array_equal(g.raw_z(), o.raw_z())
    assert np.array_equal(g.raw_shadow(), o.raw_shadow())


def test_empty_model(tinyr):
    from tiny_renderer_b200.scene import Model, Scene
    m = get_model("african_head")
    empty = Model(m.positions, m.uvs, m.normals, np.zeros((0, 9), np.uint32), m.maps)
    g = Scene(100, 80, empty, "default")
    g.clear(); g.render()
    assert not g.get_frame_buffer().any()
    assert (g.raw_visibility() == 0xFFFFFFFF).all()
    assert (g.raw_z() == np.float32(-3.4028235e38)).all()


def test_many_triangles_in_one_tile(tinyr, oracle):
    """More polygons in a single tile than the staged list holds (multi-round path): whole model in 40x30 px."""
    model = get_model("diablo")
    g, o = parity.both(tinyr, oracle, model, "phong", 40, 30, 0.3, 0.3)
    parity.compare(g, o, rgb_frac=1.0)


def test_error_codes(tinyr):
    from tiny_renderer_b200.scene import Model, Scene
    m = get_model("african_head")
    with pytest.raises(tinyr.TinyrError) as e:
        Scene(64, 64, m, "true_normals")                       # shader.rs:108 panic
    assert e.value.name == "TR_ERR_UNKNOWN_PIPELINE"
    Scene(64, 64, m, "true_normal").close()                    # README spelling is accepted as normal_map
    bad = Model(m.positions, m.uvs, m.normals, np.full((1, 9), 10 ** 6, np.uint32), m.maps)
    with pytest.raises(tinyr.TinyrError) as e:
        Scene(64, 64, bad, "default")
    assert e.value.name == "TR_ERR_BAD_POLYGON"
    s = Scene(64, 64, m, "occlusion")
    s.set_light_direction([0, 0, -1])                          # scene.rs:66 default: rotation_between is None
    with pytest.raises(tinyr.TinyrError) as e:
        s.render()
    assert e.value.name == "TR_ERR_REFERENCE_UNDEFINED"


def test_exact_quotient_selftest(tinyr):
    """The 3-instruction division used for the barycentrics (integer operands, |cz| <= 2^21; proof in DESIGN.md §4)
    against IEEE division on 2^31 pseudo-random and adversarial pairs: zero mismatching bit patterns."""
    import ctypes
    bad = ctypes.c_uint64(123)
    for seed in (1, 0xDEADBEEF):
        rc = tinyr.lib().tr_selftest_exact_quotient(0, 1 << 30, seed, ctypes.byref(bad))
        assert rc == 0 and bad.value == 0, (rc, bad.value)


@pytest.mark.parametrize("pipeline", PIPELINES)
@pytest.mark.parametrize("asset", ["african_head", "diablo"])
def test_turntable_sweep_8_angles(tinyr, oracle, asset, pipeline):
    """SURVEY.md §4 test plan: every pipeline x both assets x 8 camera angles (2*pi*k/8) with a moving light, one
    scene object reused across the views like the app loop does (clear between frames)."""
    from tiny_renderer_b200.scene import Scene
    model = get_model(asset)
    w, h = 416, 400
    g = Scene(w, h, model, pipeline)
    o = oracle.OracleScene(w, h, model, pipeline)
    for k in range(8):
        lf, la, up, light = assets.view_vectors(2 * math.pi * k / 8, 0.7 + 0.35 * k)
        if pipeline in TWO_PASS and abs(math.cos(0.7 + 0.35 * k)) > 0.98 and math.cos(0.7 + 0.35 * k) < 0:
            continue                                    # light ~ -z: rotation_between is (nearly) undefined, shader.rs:921
        parity.drive(g, lf, la, up, light)
        parity.drive(o, lf, la, up, light)
        st = parity.compare(g, o, two_pass=pipeline in TWO_PASS)
        if pipeline != "specular":
            assert st["rgb_exact_frac"] == 1.0, (k, st)
        assert st["depth_bits_differ"] == 0, (k, st)
    g.close()


def test_shared_divisor_selftest(tinyr):
    """Several quotients with one divisor (normalisations, 3x3 inverse, homogeneous divides of the fragment stage):
    nvcc's own division sequence with the divisor-only instructions hoisted, against `a / b` on 2^31 pairs."""
    import ctypes
    bad = ctypes.c_uint64(123)
    for seed in (7, 0xC0FFEE):
        rc = tinyr.lib().tr_selftest_shared_divisor(0, 1 << 30, seed, ctypes.byref(bad))
        assert rc == 0 and bad.value == 0, (rc, bad.value)


def test_large_polygons_literal_path(tinyr, oracle):
    """Polygons whose bounding box exceeds kExactExtent (1024 px) take the literal f32 raster path (separate sweep):
    two screen-filling triangles in front of / behind the model's own small ones, 3000x2000."""
    from tiny_renderer_b200.scene import Model
    m = get_model("african_head")
    big = np.array([[-1.1, -1.1, -0.4], [1.1, -1.1, -0.2], [1.1, 1.1, 0.3], [-1.1, 1.1, 0.1]], np.float32)
    pos = np.concatenate([m.positions, big])
    v0 = len(m.positions)
    quad = np.array([[v0, 0, 0, v0 + 1, 1, 1, v0 + 2, 2, 2], [v0, 0, 0, v0 + 2, 2, 2, v0 + 3, 3, 3]], np.uint32)
    tri = np.concatenate([m.tri_idx[:1200], quad, m.tri_idx[1200:]])
    model = Model(pos, m.uvs, m.normals, tri, m.maps, "head+quad")
    for pipeline in ("phong", "shadow"):
        g, o = parity.both(tinyr, oracle, model, pipeline, 3000, 2000, 0.3, 0.5)
        st = parity.compare(g, o, two_pass=pipeline == "shadow", rgb_frac=1.0)
        vis = g.raw_visibility()
        assert ((vis == 1200) | (vis == 1201)).sum() > 10 ** 6          # the big polygons do win a large area
        assert st["depth_bits_differ"] == 0
        g.close()


def test_atomic_min_microbenchmark_runs(tinyr):
    """the global 64-bit atomicMin rate is measured on the box (a yardstick, not on the render path)."""
    rate = tinyr.bench_atomic_min(1024, 512, rounds=2, repeats=2)
    assert rate > 1e8, rate
    with pytest.raises(tinyr.TinyrError):
        tinyr.bench_atomic_min(4, 2)


def test_round_u32_selftest(tinyr):
    """Shadow-lookup rounding (`f32::round() as u32`, shader.rs:774-775) in its two-instruction form against the
    literal one on every f32 bit pattern."""
    import ctypes
    bad = ctypes.c_uint64(123)
    rc = tinyr.lib().tr_selftest_round_u32(0, ctypes.byref(bad))
    assert rc == 0 and bad.value == 0, (rc, bad.value)


def test_sqrt_fast_path_selftest(tinyr):
    """The normalisations' square root (nvcc's IEEE sqrt fast path spelled out, k_tile.cuh) against sqrtf() on every
    f32 bit pattern inside its range test."""
    import ctypes
    bad, n = ctypes.c_uint64(123), ctypes.c_uint64(0)
    rc = tinyr.lib().tr_selftest_sqrt(0, ctypes.byref(bad), ctypes.byref(n))
    assert rc == 0 and bad.value == 0, (rc, bad.value)
    assert n.value == 0x72800000, hex(n.value)           # every pattern the range test accepts was compared


@pytest.mark.parametrize("pipeline", PIPELINES)
def test_fuzz_cameras_partly_off_screen(tinyr, oracle, pipeline):
    """Random cameras, look-at offsets, up vectors and lights (seeded): the model is cut by the frame border, vertices
    land off screen, bounding boxes are clamped, shadow lookups leave the light's frame.  Everything — visibility,
    depth, shadow buffer, colours and the counts of the events on which the reference would panic — must equal the
    oracle's; views the reference cannot prepare must be refused with the same status."""
    from tiny_renderer_b200.scene import Scene
    rng = np.random.default_rng(20260 + PIPELINES.index(pipeline))
    w, h = 336, 250
    done = 0
    for trial in range(10):
        model = get_model("diablo" if trial % 2 else "african_head")
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        lf = (d * rng.uniform(0.45, 2.5)).astype(np.float32)
        la = (rng.uniform(-0.6, 0.6, 3) * (trial % 3 != 0)).astype(np.float32)
        up = np.array([rng.uniform(-0.3, 0.3), 1.0, rng.uniform(-0.3, 0.3)], np.float32)
        l = rng.normal(size=3)
        light = (l / np.linalg.norm(l)).astype(np.float32)
        g = Scene(w, h, model, pipeline)
        o = oracle.OracleScene(w, h, model, pipeline)
        o.clear(); o.set_light_direction(light); o.set_camera(lf, la, up)
        if o.render(allow_panic=True):
            with pytest.raises(tinyr.TinyrError) as e:
                parity.drive(g, lf, la, up, light)
            assert e.value.name == "TR_ERR_REFERENCE_UNDEFINED", (trial, e.value)
            g.close()
            continue
        parity.drive(g, lf, la, up, light)
        st = parity.compare(g, o, two_pass=pipeline in TWO_PASS)
        assert st["depth_bits_differ"] == 0, (trial, st)
        if pipeline != "specular":
            assert st["rgb_exact_frac"] == 1.0, (trial, st)
        if pipeline in TWO_PASS:
            assert np.array_equal(g.raw_shadow().view(np.uint32), o.raw_shadow().view(np.uint32)), trial
        # The oracle, like the reference, runs the fragment closure for every fragment that passes the z-test when it
        # is drawn, the backend only for the final winner of a pixel: its event counts cover the visible fragments,
        # a subset of the oracle's.
        dg, do = g.diagnostics(), o.diagnostics()
        for k in ("shadow_index_oob", "shadow_x_wrap", "texel_oob", "nan_depth"):
            assert dg[k] <= do[k], (trial, k, dg, do)
        done += 1
        g.close()
    assert done >= 5


@pytest.mark.parametrize("pipeline", ["phong", "darboux"])
def test_4k_view_reaches_top_and_right_border(tinyr, oracle, pipeline):
    """3840x2160 has a half-height tile row at the top (2160 = 67.5 x 32): a view whose model runs off the top and the
    right border draws into it (the TMA store of that row has its own tensor map) and into the last tile column."""
    model = get_model("diablo")
    from tiny_renderer_b200.scene import Scene
    g = Scene(3840, 2160, model, pipeline)
    o = oracle.OracleScene(3840, 2160, model, pipeline)
    lf = np.array([-0.6138697862625122, -0.4760490953922272, -0.6958363652229309], np.float32)
    la = np.array([0.6074419021606445, -0.6972569227218628, 0.28669044375419617], np.float32)
    up = np.array([0.22205309569835663, 1.0, -0.1636088788509369], np.float32)
    light = np.array([0.4, 0.3, 0.85], np.float32)
    parity.drive(g, lf, la, up, light)
    parity.drive(o, lf, la, up, light)
    st = parity.compare(g, o, rgb_frac=1.0)
    vis = g.raw_visibility()                                # bottom-up rows
    assert (vis[-1] != 0xFFFFFFFF).any() and (vis[:, -1] != 0xFFFFFFFF).any()     # top row and last column are drawn into
    assert st["depth_bits_differ"] == 0
    g.close()


@pytest.mark.parametrize("pipeline", ["phong", "darboux", "shadow", "occlusion"])
def test_fuzz_sequences_with_and_without_clear(tinyr, oracle, pipeline):
    """One scene driven through a random sequence of views where only some renders are preceded by `clear`
    (scene.rs contract: render does not clear): frames pile up on the old depth buffer, tiles move between the cleared,
    stale and drawn states, and after every render all buffers must equal the oracle's."""
    from tiny_renderer_b200.scene import Scene
    rng = np.random.default_rng(31337 + PIPELINES.index(pipeline))
    model = get_model("diablo")
    w, h = 336, 250
    g = Scene(w, h, model, pipeline)
    o = oracle.OracleScene(w, h, model, pipeline)
    rendered = 0
    for step in range(9):
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        lf = (d * rng.uniform(0.5, 2.5)).astype(np.float32)
        la = (rng.uniform(-0.5, 0.5, 3) * (step % 2)).astype(np.float32)
        up = np.array([rng.uniform(-0.2, 0.2), 1.0, rng.uniform(-0.2, 0.2)], np.float32)
        l = rng.normal(size=3)
        light = (l / np.linalg.norm(l)).astype(np.float32)
        clear = step == 0 or rng.random() < 0.5
        if clear:
            o.clear()
        o.set_light_direction(light); o.set_camera(lf, la, up)
        if o.render(allow_panic=True):
            break               # the reference would have panicked here: nothing left to compare on this scene
        parity.drive(g, lf, la, up, light, clear=clear)
        st = parity.compare(g, o, two_pass=pipeline in TWO_PASS)
        assert st["depth_bits_differ"] == 0 and st["rgb_exact_frac"] == 1.0, (step, clear, st)
        rendered += 1
    assert rendered >= 4
    g.clear(); o.clear()
    assert np.array_equal(g.get_frame_buffer(), o.get_frame_buffer())
    assert np.array_equal(g.raw_z(), o.raw_z())
    g.close()


def _tile_pile_model(n_in_tile, n_elsewhere, seed):
    """Small front-facing triangles: n_in_tile of them strictly inside the bottom-left 64x32 tile of a 128x64 frame
    (camera at +z looking at the origin), n_elsewhere strictly inside the tile to its right — the same binning cell,
    so they are in the tile's source list but fail its bounding-box filter."""
    from tiny_renderer_b200.scene import Model
    m = get_model("african_head")
    rng = np.random.default_rng(seed)
    pos, tri = [], []
    for k in range(n_in_tile + n_elsewhere):
        left = k < n_in_tile
        cx = rng.uniform(-0.9, -0.15) if left else rng.uniform(0.15, 0.9)
        cy = rng.uniform(-0.9, -0.15)
        z = rng.uniform(-0.5, 0.5)
        r = rng.uniform(0.02, 0.08)
        a = rng.uniform(0, 2 * math.pi)
        v0 = len(pos)
        for j in range(3):              # counter-clockwise seen from +z: front-facing
            pos.append([cx + r * math.cos(a + 2.1 * j), cy + r * math.sin(a + 2.1 * j), z + 0.05 * j])
        tri.append([v0, (7 * k) % len(m.uvs), k % len(m.normals), v0 + 1, (7 * k + 1) % len(m.uvs), (k + 1) % len(m.normals),
                    v0 + 2, (7 * k + 2) % len(m.uvs), (k + 2) % len(m.normals)])
    return Model(np.array(pos, np.float32), m.uvs, m.normals, np.array(tri, np.uint32), m.maps, "pile")


@pytest.mark.parametrize("batch_shape", [False, True])
def test_tile_list_exactly_full(tinyr, oracle, batch_shape):
    """A tile whose polygon list fills the staged list EXACTLY (and 1 over, and twice): the round that fills the list
    ends with `more` set, the next one stages nothing — the fragment stage must still run (k_raster.cuh: the empty
    closing round).  Both CTA shapes: 256 threads / 512 entries (lone frame) and 128 threads / 192 entries (batch)."""
    from tiny_renderer_b200.scene import Scene
    cap = 192 if batch_shape else 512
    lf, la, up, light = assets.view_vectors(0.0, 0.3)
    for n_in, n_out in ((cap, 100), (cap + 1, 37), (2 * cap, 0), (cap, 0), (cap - 1, 300)):
        model = _tile_pile_model(n_in, n_out, 1000 + n_in + n_out)
        for pipeline in ("phong", "darboux"):
            o = oracle.OracleScene(128, 64, model, pipeline)
            parity.drive(o, lf, la, up, light)
            vis = o.raw_visibility()
            assert (vis[:32, :64] != 0xFFFFFFFF).sum() > 200 and (vis[32:, :] == 0xFFFFFFFF).all()
            for rep in range(3):        # the staging order (hence which warp sees the list full) varies from run to run
                if batch_shape:
                    g = Scene(128, 64, model, pipeline, ring=2)
                    v = np.zeros((2, 12), np.float32)
                    v[:, 0:3], v[:, 3:6], v[:, 6:9], v[:, 9:12] = lf, la, up, light
                    frames = np.zeros((2, 64, 128, 3), np.uint8)
                    g.render_views(v, want_checksums=False, frames_out=frames)
                    assert np.array_equal(frames[0], o.get_frame_buffer()), (n_in, n_out, pipeline, rep)
                    assert np.array_equal(frames[1], o.get_frame_buffer()), (n_in, n_out, pipeline, rep)
                else:
                    g = Scene(128, 64, model, pipeline)
                    parity.drive(g, lf, la, up, light)
                st = parity.compare(g, o, rgb_frac=1.0)
                assert st["rgb_exact_frac"] == 1.0 and st["depth_bits_differ"] == 0, (n_in, n_out, pipeline, rep, st)
                g.close()


@pytest.mark.parametrize("pipeline", PIPELINES)
def test_k0_uniforms_equal_the_oracle_bit_for_bit(tinyr, oracle, pipeline):
    """K0 (csrc/tr_prepare.h) against the oracle's prepare functions directly, not through colours: vpmv, i_vpmv, m, i_m,
    it_m, the light matrix, camera_direction, t_light_direction — and for the two-pass pipelines the constants K0 hoists
    out of the fragment stage (shadow_matrix * i_vpmv; the 16 occlusion sample offsets, shader.rs:916-929)."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("african_head")
    rng = np.random.default_rng(99)
    for w, h in ((800, 800), (3840, 2160), (333, 201)):
        g = Scene(w, h, model, pipeline)
        o = oracle.OracleScene(w, h, model, pipeline)
        for trial in range(4):
            if trial == 0:
                lf, la, up, light = assets.view_vectors(0.9, 0.7)
            else:
                d = rng.normal(size=3); d /= np.linalg.norm(d)
                lf = (d * rng.uniform(0.5, 2.5)).astype(np.float32)
                la = rng.uniform(-0.3, 0.3, 3).astype(np.float32)
                up = np.array([rng.uniform(-0.2, 0.2), 1.0, rng.uniform(-0.2, 0.2)], np.float32)
                l = rng.normal(size=3)
                light = (l / np.linalg.norm(l)).astype(np.float32)
            o.clear(); o.set_light_direction(light); o.set_camera(lf, la, up)
            if o.render(allow_panic=True):
                continue
            parity.drive(g, lf, la, up, light)
            gu, ou = g.uniforms(), o.uniforms()
            names = ["vpmv", "m", "it_m", "camera_direction", "t_light_direction"]
            if pipeline in TWO_PASS:
                names += ["i_vpmv", "i_m", "shadow"]
            for n in names:
                assert np.array_equal(gu[n].view(np.uint32), ou[n].view(np.uint32)), (w, h, trial, n, gu[n], ou[n])
            if pipeline in TWO_PASS:
                ga, oa = g.ao_uniforms(), o.ao_uniforms()
                assert np.array_equal(ga["shadow_from_screen"].view(np.uint32), oa["shadow_from_screen"].view(np.uint32)), (w, h, trial)
                if pipeline == "occlusion":
                    assert np.array_equal(ga["ao_step"].view(np.uint32), oa["ao_step"].view(np.uint32)), (w, h, trial)
        g.close()


def test_hundreds_of_renders_without_clear(tinyr, oracle):
    """The reference never clears by itself (scene.rs:151) and allows any number of renders on top of each other.  The
    key's epoch field has 8 bits: past 255 renders the stored winners are folded into epoch 0 (k_collapse_epochs) and the
    sequence goes on — all buffers still equal the oracle's after 300 renders of alternating views."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("african_head")
    w, h = 96, 80
    g = Scene(w, h, model, "phong")
    o = oracle.OracleScene(w, h, model, "phong")
    g.clear(); o.clear()
    views = [assets.view_vectors(0.3 * k, 0.5 + 0.2 * k) for k in range(5)]
    for i in range(300):
        lf, la, up, l = views[i % 5]
        parity.drive(g, lf, la, up, l, clear=False)
        parity.drive(o, lf, la, up, l, clear=False)
        if i in (0, 7, 254, 255, 256, 257, 299):
            st = parity.compare(g, o, rgb_frac=1.0)
            assert st["depth_bits_differ"] == 0 and st["rgb_exact_frac"] == 1.0, (i, st)
    g.close()


@pytest.mark.parametrize("pipeline", ["darboux", "shadow"])
def test_lone_4k_frames_in_a_row_launch_order_follows_the_previous_frame(tinyr, oracle, pipeline):
    """Lone frames walk the tiles in the order of what they cost in the previous lone frame (k_setup.cuh sort_tiles,
    k_raster.cuh tile_cursor / tile_cost): the order must stay a permutation of all tiles whatever the previous view was,
    so every frame of a sequence with very different views equals the oracle's, bit for bit."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("diablo")
    w, h = 3840, 2160
    g = Scene(w, h, model, pipeline)
    o = oracle.OracleScene(w, h, model, pipeline)
    two = pipeline in TWO_PASS
    # four frames of one view (from the third on the launch order follows the previous frame's costs), the same view
    # again WITHOUT a clear, a view one degree on, then very different ones
    seq = [(0.4, 0.7, True)] * 4 + [(0.4, 0.7, False), (0.4 + math.pi / 180, 0.7, True), (3.3, 2.1, True), (1.9, 0.2, True), (1.9, 0.2, True), (1.9, 0.2, True)]
    for step, (cam, light, clear) in enumerate(seq):
        lf, la, up, l = assets.view_vectors(cam, light)
        if step == 6:       # a very different frame in between (camera turned away: the model projects mirrored, no clipping)
            away = (lf[0] * 3.0, lf[1] * 3.0, lf[2] * 3.0)
            parity.drive(g, lf, away, up, l)
        parity.drive(g, lf, la, up, l, clear=clear)
        parity.drive(o, lf, la, up, l, clear=clear)
        st = parity.compare(g, o, two_pass=two)
        assert st["rgb_exact_frac"] == 1.0 and st["depth_bits_differ"] == 0, (step, st)
    g.close()


@pytest.mark.parametrize("size", [(2100, 1100), (7680, 4320)])
def test_sorted_launch_order_at_other_tile_counts(tinyr, oracle, size):
    """The sort block of the set-up kernel handles its positions in rounds of 4,096 (k_setup.cuh): 1,155 tiles (a ragged
    last round, a partial top tile row) and 16,200 tiles (four rounds, the 8K frame of config 5) — still views, so frames
    three and four run with the sorted and then the re-used order."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("diablo")
    w, h = size
    g = Scene(w, h, model, "phong")
    o = oracle.OracleScene(w, h, model, "phong")
    lf, la, up, l = assets.view_vectors(0.9, 0.7)
    parity.drive(o, lf, la, up, l)
    for step in range(5):
        parity.drive(g, lf, la, up, l)
        if step >= 2:
            st = parity.compare(g, o)
            assert st["rgb_exact_frac"] == 1.0 and st["depth_bits_differ"] == 0, (step, st)
    lf, la, up, l = assets.view_vectors(0.9 + math.pi / 360, 0.7)        # half a degree on: sorted by the old view's costs
    parity.drive(g, lf, la, up, l)
    parity.drive(o, lf, la, up, l)
    st = parity.compare(g, o)
    assert st["rgb_exact_frac"] == 1.0 and st["depth_bits_differ"] == 0, st
    g.close()


def test_setup_output_is_reused_only_for_identical_inputs(tinyr, oracle):
    """A lone frame of a single-pass pipeline re-uses the set-up kernel's output of the previous frame when camera, light
    and pipeline are bit-identical (tinyr.cu run_setup).  Everything that changes the set-up output must bring the kernel
    back: the light alone (vertex intensities), the pipeline (varyings), a view batch in between (it overwrites the
    slot's records), another view and back."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("diablo")
    w, h = 2560, 1440
    g = Scene(w, h, model, "phong", ring=2)
    oracles = {p: oracle.OracleScene(w, h, model, p) for p in ("phong", "darboux")}
    now = ["phong"]

    def frame(cam, light, tag, clear=True):
        lf, la, up, l = assets.view_vectors(cam, light)
        o = oracles[now[0]]
        parity.drive(g, lf, la, up, l, clear=clear)
        parity.drive(o, lf, la, up, l, clear=clear)
        st = parity.compare(g, o)
        assert st["rgb_exact_frac"] == 1.0 and st["depth_bits_differ"] == 0, (tag, st)

    for i in range(6):
        frame(0.8, 0.7, ("still", i))
    frame(0.8, 1.9, "light moved")
    frame(0.8, 1.9, "light moved, again")
    frame(0.8, 0.7, "light back")
    g.set_pipeline("darboux"); now[0] = "darboux"
    frame(0.8, 0.7, "pipeline changed")
    frame(0.8, 0.7, "pipeline changed, again")
    g.set_pipeline("phong"); now[0] = "phong"
    frame(0.8, 0.7, "pipeline back")
    frame(0.8, 0.7, "pipeline back, again")
    # a view batch on the same scene overwrites the slots' set-up output
    v = np.zeros((3, 12), np.float32)
    for k in range(3):
        lf, la, up, l = assets.view_vectors(2.0 + k, 0.3)
        v[k, 0:3], v[k, 3:6], v[k, 6:9], v[k, 9:12] = lf, la, up, l
    g.render_views(v)
    frame(0.8, 0.7, "after a view batch")
    frame(2.5, 0.7, "another view")
    frame(0.8, 0.7, "and back")
    frame(0.8, 0.7, "and back, no clear", clear=False)
    g.close()


@pytest.mark.parametrize("pipeline", ["shadow", "occlusion"])
def test_light_pass_is_reused_only_while_the_shadow_buffer_stands(tinyr, oracle, pipeline):
    """Lone frames of the two-pass pipelines re-use the shadow buffer of the previous frame while light, target and up are
    bit-identical and nothing else wrote to it (tinyr.cu render_impl): a camera that moves under a fixed light, frames
    without a clear, a getter that materialises a pending clear in between, a light that moves and comes back."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("diablo")
    w, h = 2560, 1440
    g = Scene(w, h, model, pipeline)
    o = oracle.OracleScene(w, h, model, pipeline)

    def frame(cam, light, tag, clear=True):
        lf, la, up, l = assets.view_vectors(cam, light)
        parity.drive(g, lf, la, up, l, clear=clear)
        parity.drive(o, lf, la, up, l, clear=clear)
        st = parity.compare(g, o, two_pass=True)
        assert st["rgb_exact_frac"] == 1.0 and st["depth_bits_differ"] == 0 and st["shadow_max_ulp"] == 0, (tag, st)
        gd, od = g.diagnostics(), o.diagnostics()
        for k in ("shadow_index_oob", "shadow_x_wrap", "texel_oob", "nan_depth"):
            assert gd[k] == od[k], (tag, k, gd, od)

    for i in range(5):
        frame(0.8, 2.1, ("still", i))
    for i in range(3):
        frame(0.8 + 0.3 * (i + 1), 2.1, ("camera moves, light fixed", i))
    frame(1.7, 2.1, "no clear, same light", clear=False)
    frame(1.7, 0.7, "no clear, light moved", clear=False)          # the buffer now holds the maximum of two passes
    frame(1.7, 0.7, "no clear, same light again", clear=False)
    frame(1.7, 0.7, "cleared, same light")
    frame(1.7, 0.7, "cleared, same light, again")
    g.clear(); o.clear()
    assert np.array_equal(g.get_shadow_buffer(), o.get_shadow_buffer())     # materialises the pending clear
    frame(1.7, 0.7, "after a getter between clear and render", clear=False)
    frame(1.7, 2.1, "light back")
    frame(1.7, 2.1, "light back, again")
    g.close()


@pytest.mark.parametrize("pipeline", ["darboux", "shadow"])
def test_debug_bit_2_runs_every_pass_and_changes_nothing(tinyr, pipeline):
    """Debug bit 2 disables the cost-sorted launch order and the re-use of unchanged passes for a scene; frames, depth and
    diagnostics must not depend on it."""
    from tiny_renderer_b200.scene import Scene
    model = get_model("diablo")
    w, h = 2560, 1440
    a, b = Scene(w, h, model, pipeline), Scene(w, h, model, pipeline)
    b.set_debug_flags(4)
    for cam, light in [(0.8, 2.1)] * 5 + [(1.1, 2.1), (1.1, 0.7), (1.1, 0.7)]:
        lf, la, up, l = assets.view_vectors(cam, light)
        parity.drive(a, lf, la, up, l)
        parity.drive(b, lf, la, up, l)
        assert np.array_equal(a.raw_frame(), b.raw_frame()) and np.array_equal(a.raw_z(), b.raw_z())
        assert np.array_equal(a.raw_visibility(), b.raw_visibility())
        if pipeline == "shadow":
            assert np.array_equal(a.raw_shadow(), b.raw_shadow())
        assert a.diagnostics() == b.diagnostics()
    a.close(); b.close()

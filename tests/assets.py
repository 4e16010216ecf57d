"""Test-side asset access: independent Python loaders (OBJ text parse, TGA through PIL), the location of
the staged reference assets, and a procedural synthetic model for boxes where they are absent.

The reference assets (african_head, diablo: OBJ + 4 RLE TGAs each, /root/reference/assets) are NOT committed;
`__graft_entry__.build()` stages a copy under assets/_ref/ (git-ignored, travels with gpurun) when
/root/reference is present.  Nothing here reads /root/reference at test time on the GPU box.
"""
import os
from dataclasses import dataclass

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STAGED = os.path.join(ROOT, "assets", "_ref")
MAP_FILES = ("texture.tga", "normal_map.tga", "normal_map_tangent.tga", "specular_map.tga")   # app.rs:87-91


@dataclass
class Model:
    positions: np.ndarray   # (V,3) f32
    uvs: np.ndarray         # (N,2) f32 (u, v) as in the OBJ (v NOT flipped)
    normals: np.ndarray     # (N,3) f32
    tri_idx: np.ndarray     # (T,9) u32: p0 t0 n0 p1 t1 n1 p2 t2 n2, 0-based
    maps: tuple             # 4 x (H,W,3) u8, top-down rows: texture, normal_map, normal_map_tangent, specular_map
    name: str = "model"


def asset_dir(name):
    """Directory holding model.obj + the 4 TGAs for `name`, or None."""
    for base in (STAGED, "/root/reference/assets"):
        d = os.path.join(base, name)
        if os.path.isfile(os.path.join(d, "model.obj")):
            return d
    return None


def parse_obj_py(path):
    """Independent restatement of obj-rs 0.7.0 raw::parse_obj for the subset the reference accepts
    (v / vt / vn / f with p/t/n corners; 1-based or negative relative indices; first 3 corners used)."""
    pos, tex, nrm, tris = [], [], [], []
    with open(path, "r") as f:
        for line in f:
            line = line.split("#", 1)[0]
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "v":
                pos.append([np.float32(x) for x in tok[1:4]])
            elif tok[0] == "vt":
                vals = [np.float32(x) for x in tok[1:3]]
                while len(vals) < 2:
                    vals.append(np.float32(0.0))
                tex.append(vals)
            elif tok[0] == "vn":
                nrm.append([np.float32(x) for x in tok[1:4]])
            elif tok[0] == "f":
                corners = []
                for c in tok[1:4]:
                    p, t, n = (int(x) for x in c.split("/"))
                    p = p - 1 if p > 0 else len(pos) + p
                    t = t - 1 if t > 0 else len(tex) + t
                    n = n - 1 if n > 0 else len(nrm) + n
                    corners += [p, t, n]
                tris.append(corners)
    return (np.array(pos, np.float32).reshape(-1, 3), np.array(tex, np.float32).reshape(-1, 2),
            np.array(nrm, np.float32).reshape(-1, 3), np.array(tris, np.uint32).reshape(-1, 9))


def load_tga_pil(path):
    """What image::open(path).into_rgb8() yields: RGB8, top-down rows."""
    from PIL import Image
    return np.asarray(Image.open(path).convert("RGB"), dtype=np.uint8).copy()


def load_model_py(name):
    d = asset_dir(name)
    if d is None:
        return None
    p, t, n, idx = parse_obj_py(os.path.join(d, "model.obj"))
    maps = tuple(load_tga_pil(os.path.join(d, f)) for f in MAP_FILES)
    return Model(p, t, n, idx, maps, name)


def synthetic_model(n_lat=36, n_lon=70, tex=256, seed=7, name="synthetic"):
    """Procedural stand-in for diablo (5,022 faces): a bumpy closed blob with 2*n_lon*(n_lat-1) p/t/n triangles
    (36x70 -> 4,900), per-vertex normals, uv strictly inside (0,1), and four procedural RGB8 maps."""
    rng = np.random.default_rng(seed)
    lat = np.linspace(0.0, np.pi, n_lat + 1)[1:-1]                     # interior rings
    lon = np.linspace(0.0, 2 * np.pi, n_lon, endpoint=False)
    la, lo = np.meshgrid(lat, lon, indexing="ij")
    r = 0.75 + 0.12 * np.sin(3 * lo) * np.sin(2 * la) + 0.06 * np.cos(5 * la + 2 * lo)
    x = r * np.sin(la) * np.cos(lo) * 0.8
    y = r * np.cos(la)
    z = r * np.sin(la) * np.sin(lo) * 0.8
    ring = np.stack([x, y, z], -1).reshape(-1, 3)
    top = np.array([[0.0, 0.8, 0.0]]); bot = np.array([[0.0, -0.8, 0.0]])
    pos = np.concatenate([ring, top, bot]).astype(np.float32)
    n_ring = n_lat - 1
    nrm = pos / np.linalg.norm(pos, axis=1, keepdims=True)
    nrm = (nrm + 0.15 * rng.standard_normal(nrm.shape)).astype(np.float32)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    # uv per (ring, lon) with a seam column so u stays in (0,1)
    uu = (np.arange(n_lon + 1) + 0.5) / (n_lon + 2)
    vv = (np.arange(n_ring) + 1.0) / (n_ring + 1)
    uvg = np.stack(np.meshgrid(vv, uu, indexing="ij")[::-1], -1).reshape(-1, 2)
    uvs = np.concatenate([uvg, [[0.5, 0.01], [0.5, 0.99]]]).astype(np.float32)
    vid = lambda i, j: i * n_lon + (j % n_lon)
    tid = lambda i, j: i * (n_lon + 1) + j
    TOP, BOT = n_ring * n_lon, n_ring * n_lon + 1
    TTOP, TBOT = n_ring * (n_lon + 1), n_ring * (n_lon + 1) + 1
    tris = []
    for j in range(n_lon):
        a, b = vid(0, j), vid(0, j + 1)
        tris.append([TOP, TTOP, TOP, b, tid(0, j + 1), b, a, tid(0, j), a])
        a, b = vid(n_ring - 1, j), vid(n_ring - 1, j + 1)
        tris.append([BOT, TBOT, BOT, a, tid(n_ring - 1, j), a, b, tid(n_ring - 1, j + 1), b])
    for i in range(n_ring - 1):
        for j in range(n_lon):
            a, b, c, d = vid(i, j), vid(i, j + 1), vid(i + 1, j), vid(i + 1, j + 1)
            ta, tb, tc, td = tid(i, j), tid(i, j + 1), tid(i + 1, j), tid(i + 1, j + 1)
            tris.append([a, ta, a, b, tb, b, c, tc, c])
            tris.append([b, tb, b, d, td, d, c, tc, c])
    tri_idx = np.array(tris, np.uint32)
    yy, xx = np.mgrid[0:tex, 0:tex]
    checker = (((xx // 16) + (yy // 16)) % 2).astype(np.uint8)
    texture = np.stack([120 + 100 * checker, 90 + (xx * 130 // tex), 60 + (yy * 160 // tex)], -1).astype(np.uint8)
    nm = rng.integers(0, 256, (tex, tex, 3)).astype(np.uint8)
    nm[..., 2] = np.maximum(nm[..., 2], 140)
    nmt = np.stack([128 + 40 * np.sin(xx / 5.0), 128 + 40 * np.cos(yy / 7.0), np.full_like(xx, 230)], -1).astype(np.uint8)
    spec = np.repeat(((xx + yy) % 200)[..., None], 3, -1).astype(np.uint8)
    return Model(pos, uvs, nrm, tri_idx, (texture, nm, nmt, spec), name)


def get_model(name):
    """Reference asset if staged/available, else the synthetic model (callers can check .name)."""
    m = load_model_py(name)
    return m if m is not None else synthetic_model(name="synthetic-for-" + name)


def view_vectors(camera_angle, light_angle):
    """app.rs:200-207 with f32 sin/cos."""
    ca, la = np.float32(camera_angle), np.float32(light_angle)
    look_from = np.array([np.sin(ca), 0.0, np.cos(ca)], np.float32)
    light = np.array([np.sin(la), 0.0, np.cos(la)], np.float32)
    return look_from, np.zeros(3, np.float32), np.array([0, 1, 0], np.float32), light

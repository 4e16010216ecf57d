"""Generates tests/golden/oracle_golden.json: sha256 of the oracle's visibility / depth / frame (and shadow)
buffers for the reference's own assets on fixed views.  Run in the build container (reads the staged copy of
/root/reference/assets).  The reference itself ships no golden vectors and cannot be run here (Rust, windowed),
so these pin the ORACLE's behaviour (regressions, GPU parity on the GPU box) — not the Rust binary's
("parity unpinned", see oracle/oracle.cpp).

    python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import numpy as np  # noqa: E402

import assets  # noqa: E402
import oracle_lib  # noqa: E402

CASES = [
    # asset, pipeline, width, height, camera_angle, light_angle
    ("african_head", "default", 800, 800, 0.0, 0.0),          # BASELINE config 1
    ("african_head", "specular", 800, 800, 0.4, 0.7),
    ("african_head", "normal_map", 800, 800, 5.5, 1.0),
    ("diablo", "default", 800, 800, 0.0, 0.0),
    ("diablo", "phong", 800, 800, 0.7853982, 0.7),
    ("diablo", "darboux", 800, 800, 2.3561945, 0.7),
    ("diablo", "shadow", 800, 800, 0.4, 2.1),
    ("diablo", "occlusion", 800, 800, 0.0, 0.7),
    ("diablo", "phong", 3840, 2160, 0.0, 0.7),                # BASELINE config 2
    ("diablo", "darboux", 3840, 2160, 3.9269908, 0.7),
    ("diablo", "shadow", 3840, 2160, 0.7853982, 0.7),         # BASELINE config 3
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_case(case):
    asset, pipeline, w, h, cam, light = case
    model = assets.load_model_py(asset)
    s = oracle_lib.OracleScene(w, h, model, pipeline)
    lf, la, up, l = assets.view_vectors(cam, light)
    s.clear(); s.set_light_direction(l); s.set_camera(lf, la, up); s.render()
    vis = s.raw_visibility()
    out = {"case": list(case), "covered": int((vis != 0xFFFFFFFF).sum()), "visibility": sha(vis), "depth": sha(s.raw_z()),
           "frame_topdown": sha(s.get_frame_buffer()), "diagnostics": s.diagnostics(),
           # what the reference's own public getters return (scene.rs:101-125): the hashes rust/pin_oracle prints
           "z_gray": sha(s.get_z_buffer()), "shadow_gray": sha(s.get_shadow_buffer())}
    if pipeline in ("shadow", "occlusion"):
        out["shadow"] = sha(s.raw_shadow())
    return out


def write_pin_cases(path):
    """rust/pin_oracle/cases.txt: the golden cases with camera / light vectors as f32 bit patterns."""
    with open(path, "w") as f:
        f.write("# asset pipeline width height look_from(3) look_at(3) up(3) light(3) as f32 bit patterns (hex); written by tests/golden/make_golden.py\n")
        for asset, pipeline, w, h, cam, light in CASES:
            lf, la, up, l = assets.view_vectors(cam, light)
            words = np.concatenate([lf, la, up, l]).astype(np.float32).view(np.uint32)
            f.write("%s %s %d %d %s\n" % (asset, pipeline, w, h, " ".join("%08x" % x for x in words)))


if __name__ == "__main__":
    res = [run_case(c) for c in CASES]
    with open(os.path.join(HERE, "oracle_golden.json"), "w") as f:
        json.dump(res, f, indent=1)
    for r in res:
        print(r["case"], r["covered"], r["frame_topdown"][:12])
    write_pin_cases(os.path.join(os.path.dirname(os.path.dirname(HERE)), "rust", "pin_oracle", "cases.txt"))
    # config 4 light set: the accepted directions are part of the fixture
    import tiny_renderer_b200 as tr  # noqa: E402  (host-only entry point: no GPU needed)
    lights, drawn = tr.generate_ao_lights(assets.load_model_py("diablo").positions, 1024)
    np.save(os.path.join(HERE, "ao_lights_diablo_1024.npy"), lights)
    print("ao lights: 1024 accepted of", drawn)

"""How much can "parity unpinned" cost?  (mitigation 2.)  CPU only; test infrastructure.

The oracle restates nalgebra 0.31.4's operation order from its published algorithms, because the crate's source is
not under /root/reference and the Rust reference cannot be built here.  This tool rebuilds the oracle with ONE
primitive changed to a different, equally plausible order at a time (switches in oracle/na32.h, plus FMA contraction
as the worst case), renders the same views, and counts the pixels whose visibility / depth / colour move.

    python tools/oracle_sensitivity.py            # writes oracle/sensitivity.json and prints the table
"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

VARIANTS = {
    "baseline": [],
    "gemv_pairwise": ["-DNA32_ALT_GEMV_PAIRWISE"],
    "dot3_right": ["-DNA32_ALT_DOT3_RIGHT"],
    "normalize_recip": ["-DNA32_ALT_NORMALIZE_RECIP"],
    "inv4_divide": ["-DNA32_ALT_INV4_DIVIDE"],
    "fma_contraction": ["-ffp-contract=fast", "-mfma"],
}
CASES = [("african_head", "default", 800, 800, 0.0, 0.0), ("diablo", "phong", 1920, 1080, 0.7854, 0.7),
         ("diablo", "darboux", 1920, 1080, 2.3562, 0.7), ("diablo", "shadow", 1920, 1080, 0.7854, 2.1),
         ("african_head", "specular", 1920, 1080, 3.9270, 0.7), ("diablo", "occlusion", 800, 800, 0.0, 0.7)]


def build(name, flags):
    out_dir = os.path.join(ROOT, "oracle", "_alt")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "liboracle_%s.so" % name)
    base = ["g++", "-O2", "-std=c++17", "-fPIC", "-fno-fast-math", "-shared", "-o", so, os.path.join(ROOT, "oracle", "oracle.cpp")]
    contract = [] if any(f.startswith("-ffp-contract") for f in flags) else ["-ffp-contract=off"]
    subprocess.check_call(base + contract + flags)
    return so


def render_child(out_path):
    """Runs in a subprocess with TINYR_ORACLE_LIB pointing at one variant."""
    import assets
    import oracle_lib
    import parity
    res = {}
    for k, (asset, pipe, w, h, cam, light) in enumerate(CASES):
        m = assets.get_model(asset)
        o = oracle_lib.OracleScene(w, h, m, pipe)
        lf, la, up, l = assets.view_vectors(cam, light)
        parity.drive(o, lf, la, up, l)
        res["vis%d" % k] = o.raw_visibility()
        res["z%d" % k] = o.raw_z()
        res["rgb%d" % k] = o.get_frame_buffer()
    np.savez_compressed(out_path, **res)


def main():
    if len(sys.argv) == 3 and sys.argv[1] == "--child":
        render_child(sys.argv[2])
        return
    outs = {}
    for name, flags in VARIANTS.items():
        so = build(name, flags)
        path = "/tmp/oracle_sens_%s.npz" % name
        env = dict(os.environ, TINYR_ORACLE_LIB=so)
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "--child", path], env=env)
        outs[name] = np.load(path)
    base = outs["baseline"]
    report = []
    for name in VARIANTS:
        if name == "baseline":
            continue
        for k, case in enumerate(CASES):
            vis_b, vis = base["vis%d" % k], outs[name]["vis%d" % k]
            z_b, z = base["z%d" % k], outs[name]["z%d" % k]
            d = np.abs(base["rgb%d" % k].astype(np.int16) - outs[name]["rgb%d" % k].astype(np.int16)).max(axis=2)
            n = vis_b.size
            same_vis = vis_b == vis
            report.append({"variant": name, "case": "%s/%s %dx%d" % case[:4], "pixels": int(n),
                           "covered": int((vis_b != 0xFFFFFFFF).sum()),
                           "visibility_differs": int((~same_vis).sum()),
                           "depth_bits_differ": int((z_b.view(np.uint32) != z.view(np.uint32)).sum()),
                           "rgb_differs": int((d > 0).sum()), "rgb_differs_by_more_than_1": int((d > 1).sum()),
                           "rgb_within_1_frac": float((d <= 1).mean())})
    with open(os.path.join(ROOT, "oracle", "sensitivity.json"), "w") as f:
        json.dump(report, f, indent=1)
    print("| variant | case | visibility differs | depth bits differ | rgb differs | rgb off by > 1 | within 1 LSB |\n|---|---|---|---|---|---|---|")
    for r in report:
        print("| %s | %s | %d | %d | %d | %d | %.5f |" % (r["variant"], r["case"], r["visibility_differs"], r["depth_bits_differ"],
                                                          r["rgb_differs"], r["rgb_differs_by_more_than_1"], r["rgb_within_1_frac"]))


if __name__ == "__main__":
    main()

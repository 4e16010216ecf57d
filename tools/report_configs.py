"""Per-configuration report (BASELINE.json configs 1-5): device time per frame through the C ABI
(single-frame sequence clear/set_*/render, CUDA events on the scene stream, median of N) beside the CPU oracle on one
host thread for the same frame.  Run under gpurun; writes gpurun_out/configs.json."""
import json
import math
import os
import statistics
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import assets  # noqa: E402
import oracle_lib  # noqa: E402
import parity  # noqa: E402
import tiny_renderer_b200 as tr  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

CASES = [
    ("config1", "african_head", "default", 800, 800, 0.0, 0.0),
    ("config2a", "diablo", "phong", 3840, 2160, 2 * math.pi / 8, 0.7),
    ("config2b", "diablo", "darboux", 3840, 2160, 2 * math.pi / 8, 0.7),
    ("config3", "diablo", "shadow", 3840, 2160, 2 * math.pi / 8, 2.1),
    ("config4-one-light", "diablo", "occlusion", 3840, 2160, 0.0, 0.7),
    ("config5-head", "african_head", "specular", 7680, 4320, 2 * math.pi * 256 / 2048, 0.7),
    ("config5-diablo", "diablo", "specular", 7680, 4320, 2 * math.pi * 256 / 2048, 0.7),
]


def main():
    rows = []
    for name, asset, pipe, w, h, cam, light in CASES:
        model = assets.get_model(asset)
        lf, la, up, l = assets.view_vectors(cam, light)
        g = Scene(w, h, model, pipe)
        for _ in range(3):
            parity.drive(g, lf, la, up, l)
        times = []
        for _ in range(15):
            g.timer_begin()
            parity.drive(g, lf, la, up, l)
            times.append(g.timer_end())
        gpu_ms = statistics.median(times)
        o = oracle_lib.OracleScene(w, h, model, pipe)
        t0 = time.perf_counter()
        parity.drive(o, lf, la, up, l)
        cpu_ms = (time.perf_counter() - t0) * 1e3
        st = parity.compare(g, o, two_pass=pipe in ("shadow", "occlusion"))
        rows.append({"config": name, "asset": model.name, "pipeline": pipe, "size": [w, h], "gpu_ms_per_frame": gpu_ms,
                     "gpu_mpix_per_s": w * h / gpu_ms / 1e3, "cpu_1thread_ms_per_frame": cpu_ms,
                     "cpu_mpix_per_s": w * h / cpu_ms / 1e3, "speedup": cpu_ms / gpu_ms, "covered": st["covered"],
                     "vis_mismatch": st["vis_mismatch"], "depth_max_ulp": st["depth_max_ulp"], "rgb_exact_frac": st["rgb_exact_frac"],
                     "rgb_within1_frac": st["rgb_within1_frac"]})
        print(rows[-1], flush=True)
        g.close()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "configs.json"), "w") as f:
        json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()

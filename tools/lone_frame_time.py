"""Device time of a lone frame (clear / set_* / render through the C ABI, CUDA events on the scene stream, median of
31) for the single-frame BASELINE configs, without the oracle: quick A/B of launch-path changes, e.g.
`TINYR_NO_PDL=1 python tools/lone_frame_time.py` against `python tools/lone_frame_time.py`.  Run under gpurun."""
import math
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import assets  # noqa: E402
import parity  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

CASES = [("config1", "african_head", "default", 800, 800, 0.0, 0.0),
         ("config2a", "diablo", "phong", 3840, 2160, 2 * math.pi / 8, 0.7),
         ("config2b", "diablo", "darboux", 3840, 2160, 2 * math.pi / 8, 0.7),
         ("config3", "diablo", "shadow", 3840, 2160, 2 * math.pi / 8, 2.1)]
out = []
for name, asset, pipe, w, h, cam, light in CASES:
    model = assets.get_model(asset)
    lf, la, up, l = assets.view_vectors(cam, light)
    g = Scene(w, h, model, pipe)
    for _ in range(5):
        parity.drive(g, lf, la, up, l)
    times = []
    for _ in range(31):
        g.timer_begin()
        parity.drive(g, lf, la, up, l)
        times.append(g.timer_end())
    out.append("%s %.1f us" % (name, statistics.median(times) * 1e3))
    g.close()
print("pdl=%s " % ("off" if os.environ.get("TINYR_NO_PDL") else "on") + " | ".join(out))

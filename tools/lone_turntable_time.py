"""Device time of LONE 4K frames whose view changes from frame to frame (64-angle turntable, 5.6 degrees per frame, the
app loop's clear / set_* / render per frame, CUDA events per frame on the scene stream, median over two turns), beside
the same for a view that does not change.  The launch order of a lone frame is sorted by the previous frame's tile
costs, so this is the case that matters for it:  python tools/lone_turntable_time.py;  TINYR_LONE_SORT=0 python ...
Run under gpurun."""
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

N = int(os.environ.get("TURNTABLE_ANGLES", "64"))
out = []
for pipe, light in (("phong", 0.7), ("darboux", 0.7), ("shadow", 2.1)):
    model = assets.get_model("diablo")
    g = Scene(3840, 2160, model, pipe)
    views = [assets.view_vectors(2 * math.pi * k / N, light) for k in range(N)]
    for lf, la, up, l in views[:4]:
        parity.drive(g, lf, la, up, l)
    moving, still = [], []
    for _ in range(2):
        for lf, la, up, l in views:
            g.timer_begin()
            parity.drive(g, lf, la, up, l)
            moving.append(g.timer_end())
    lf, la, up, l = views[N // 8]
    for _ in range(40):
        g.timer_begin()
        parity.drive(g, lf, la, up, l)
        still.append(g.timer_end())
    out.append("%s moving %.1f us (p90 %.1f) still %.1f us" % (pipe, statistics.median(moving) * 1e3, sorted(moving)[int(len(moving) * 0.9)] * 1e3,
                                                              statistics.median(still[5:]) * 1e3))
    g.close()
print("sort=%s angles=%d " % (os.environ.get("TINYR_LONE_SORT", "1"), N) + " | ".join(out))

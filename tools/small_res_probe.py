"""View-batch throughput at small resolutions (the reference's own 800x800, main.rs:6-7, plus 1280x720 and 1920x1080) with
several views per launch (default) and with TINYR_MULTI_TILES=0 (two launches per view, frames overlapped on the slot
streams), and the lone-frame time: device-timed Gpx/s of tr_scene_render_views on a 64-angle turntable.  Run under gpurun:
    python tools/small_res_probe.py;  TINYR_MULTI_TILES=0 python tools/small_res_probe.py;  TINYR_MULTI_TILES=2000 python ..."""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import assets  # noqa: E402
import bench  # noqa: E402
import parity  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

out = {"multi_tiles": os.environ.get("TINYR_MULTI_TILES", "600 (default)"), "ring": int(os.environ.get("PROBE_RING", "8"))}
views = bench.view_batch()
for asset, pipe, w, h in (("african_head", "default", 800, 800), ("diablo", "phong", 800, 800), ("diablo", "phong", 1280, 720),
                          ("diablo", "darboux", 1920, 1080)):
    model = assets.get_model(asset)
    g = Scene(w, h, model, pipe, ring=out["ring"])
    for _ in range(3):
        g.render_views(views, want_checksums=False)
    g.synchronize()
    reps = 20
    g.timer_begin()
    for _ in range(reps):
        g.render_views(views, want_checksums=False)
    ms = g.timer_end()
    lone = []
    v = views[5]
    for _ in range(21):
        g.timer_begin()
        parity.drive(g, v[0:3], v[3:6], v[6:9], v[9:12])
        lone.append(g.timer_end())
    out["%s %s %dx%d" % (asset, pipe, w, h)] = {"batch_gpx_per_s": reps * len(views) * w * h / (ms * 1e-3) / 1e9, "us_per_view": ms * 1e3 / (reps * len(views)),
                                                 "lone_frame_us": statistics.median(lone) * 1e3, "lone_gpx_per_s": w * h / (statistics.median(lone) * 1e-3) / 1e9}
    g.close()
print(json.dumps(out))

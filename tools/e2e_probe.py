"""e2e leg of bench.py alone (tr_scene_render_views with frames -> tr_host_alloc memory, phong + darboux at 3840x2160, 64-angle
turntable, two host threads), for quick A/B of the sparse delivery kernel, e.g.
    TINYR_STORE_CTAS=64 python tools/e2e_probe.py       TINYR_NO_SPARSE_D2H=1 python tools/e2e_probe.py
Run under gpurun."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import bench  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

model, _ = bench.bench_model()
scenes = {p: Scene(bench.W4K, bench.H4K, model, p, device=0, ring=bench.RING) for p in bench.PIPES}
views = bench.view_batch()
for s in scenes.values():
    s.render_views(views[:16], want_checksums=False)
r = bench.run_e2e(torch, None, 0, 1, 0, scenes, views, int(os.environ.get("E2E_STEPS", "10")))
print(json.dumps({"store_ctas": os.environ.get("TINYR_STORE_CTAS", "32 (default)"), "sparse": os.environ.get("TINYR_NO_SPARSE_D2H") is None,
                  "gpx_per_s": r["value"] / 1e3, "MB_per_frame": r["d2h_bytes_per_frame"] / 1e6, "link_GBps": r["roofline"]["achieved"],
                  "plain_copy_GBps": r["roofline"]["peak"], "frac": r["roofline"]["frac"], "exact": r["frames_exact"]}))

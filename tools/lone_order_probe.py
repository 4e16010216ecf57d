"""Diagnostic (TR_TIMELINE build): what would a cost-sorted launch order buy a lone 4K frame?  Pass 1 records the per-tile
times of a lone frame in the product's centre-out order, pass 2 re-runs the same view with the tiles launched longest
first (TINYR_ORDER_PERM), pass 3 with 4-us cost buckets (stable inside a bucket).  Run by tools/lone_order_probe.sh."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")


def run(tag, env):
    e = dict(os.environ, TIMELINE_DUMP=OUT, **env)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "lone_timeline.py"), PIPE], env=e, capture_output=True, text=True)
    line = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if not line:
        print(tag, "failed", r.stderr[-400:])
        return None
    d = json.loads(line[-1])
    print(json.dumps({"order": tag, "pipeline": PIPE, "frame_us_events_median": d["frame_us_events_median"], "kernel_span_us": d["kernel_span_us"],
                      "after_release_us": d["after_release_us"], "touched_in_flight": d["touched_in_flight"]}))
    return np.load(os.path.join(OUT, "timeline_%s.npy" % PIPE)).astype(np.int64)


for PIPE in sys.argv[1:] or ["phong", "darboux"]:
    t = run("centre-out", {"TINYR_LONE_SORT": "0"})
    dur = (t[:, 3] - np.maximum(t[:, 1], t[:, 0])) / 1e3
    for tag, key in (("longest first", -dur), ("4 us buckets", -np.floor(dur / 4.0)), ("tiles with geometry first, centre-out inside", -(t[:, 5] > 0).astype(np.int64))):
        perm = np.argsort(key, kind="stable").astype(np.uint32)
        f = os.path.join(OUT, "perm.bin")
        perm.tofile(f)
        run(tag, {"TINYR_LONE_SORT": "0", "TINYR_ORDER_PERM": f})

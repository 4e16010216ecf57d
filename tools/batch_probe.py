"""Device-timed view-batch throughput at 3840x2160 for the library found under the CURRENT DIRECTORY (so the same script can
be run from an A/B worktree): 64-angle turntable, phong alone, darboux alone, and both side by side on a stream each as in
bench.py.  Run under gpurun:  python tools/batch_probe.py;  (cd _ab/r2a && python ../../tools/batch_probe.py)"""
import json
import math
import os
import sys

ROOT = os.getcwd()
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import assets  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

W, H, RING, REPS = 3840, 2160, 8, int(os.environ.get("PROBE_REPS", "8"))
model = assets.get_model("diablo")
views = np.zeros((64, 12), np.float32)
for k in range(64):
    lf, la, up, light = assets.view_vectors(2 * math.pi * k / 64, 0.7)
    views[k, 0:3], views[k, 3:6], views[k, 6:9], views[k, 9:12] = lf, la, up, light
tstream = torch.cuda.Stream()
torch.cuda.set_stream(tstream)
out = {"root": ROOT}
for pipes in (("phong",), ("darboux",), ("phong", "darboux"), ("phong", "darboux")):
    scenes = {p: Scene(W, H, model, p, ring=RING) for p in pipes}
    ps = {p: torch.cuda.Stream() for p in pipes}
    for p in pipes:
        scenes[p].set_stream(ps[p].cuda_stream)

    def step():
        for _ in range(REPS):
            for p in pipes:
                scenes[p].render_views(views, want_checksums=False)
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for s in ps.values():
        s.wait_stream(tstream)
    step()
    for s in ps.values():
        tstream.wait_stream(s)
    e0.record()
    for s in ps.values():
        s.wait_stream(tstream)
    n = 6
    for _ in range(n):
        step()
    for s in ps.values():
        tstream.wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    key = "+".join(pipes)
    out.setdefault(key, []).append(round(n * REPS * 64 * len(pipes) * W * H / (ms * 1e-3) / 1e9, 1))
    for s in scenes.values():
        s.close()
print(json.dumps(out))

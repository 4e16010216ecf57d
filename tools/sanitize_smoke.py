"""Small run of every kernel path for compute-sanitizer (memcheck / racecheck): all seven pipelines at a ragged size,
a view batch over 3 slots, a many-light AO batch, the getters.  Run under gpurun:
    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import assets  # noqa: E402
import tiny_renderer_b200 as tr  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

m = assets.get_model("african_head")
w, h = 203, 131
for pipe in tr.PIPELINES:
    s = Scene(w, h, m, pipe, ring=3)
    views = np.zeros((5, 12), np.float32)
    for k in range(5):
        lf, la, up, l = assets.view_vectors(0.4 * k, 0.7)
        views[k, 0:3], views[k, 3:6], views[k, 6:9], views[k, 9:12] = lf, la, up, l
    frames = np.zeros((5, h, w, 3), np.uint8)
    cs = s.render_views(views, want_checksums=True, frames_out=frames)
    s.set_camera(views[1, 0:3], views[1, 3:6], views[1, 6:9]); s.set_light_direction(views[1, 9:12])
    s.render()                     # render without clear on top of the last frame
    s.raw_visibility(); s.raw_z(); s.raw_shadow(); s.get_z_buffer(); s.get_shadow_buffer(); s.raw_frame()
    if pipe == "occlusion":
        lights, _ = tr.generate_ao_lights(m.positions, 5)
        s.render_ao(lights)
    print(pipe, int(cs[0]) & 0xffff, flush=True)
    s.close()
# whole model inside one tile (multi-round staging path)
s = Scene(40, 30, assets.get_model("diablo"), "phong")
s.clear(); s.render(); s.get_frame_buffer(); s.close()
print("ok")

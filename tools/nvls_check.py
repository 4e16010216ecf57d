"""AO exchange of the native multi-GPU driver (tr_multi_*), one process over G devices: result against a single scene, which
exchange runs ("nvls" = multimem reductions inside the NVSwitch, or the peer-memory kernel) and how long one exchange of
the 33 MB of 4K sums takes (tr_multi_time_exchange: 20 back to back, slowest device, CUDA events).
    TINYR_NVLS=1 python tools/nvls_check.py [G];   python tools/nvls_check.py [G]        (run under gpurun --gpus G)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import assets  # noqa: E402
import tiny_renderer_b200 as tr  # noqa: E402
from tiny_renderer_b200.scene import MultiScene, Scene  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else tr.device_count()
model = assets.get_model("diablo")
lights, _ = tr.generate_ao_lights(model.positions, 4 * G)
lf, la, up, _ = assets.view_vectors(0.0, 0.0)
one = Scene(3840, 2160, model, "occlusion", ring=2)
one.set_camera(lf, la, up)
want = one.render_ao(lights)
one.close()
m = MultiScene(3840, 2160, model, "occlusion", list(range(G)), ring=2)
m.set_camera(lf, la, up)
kind = m.exchange_kind()
ok = all(bool(np.array_equal(m.render_ao(lights), want)) for _ in range(3))
every = all(bool(np.array_equal(m.scene(i).read_ao(), want)) for i in range(G))
us = [m.time_exchange(20) for _ in range(3)]
print(json.dumps({"devices": G, "exchange": kind, "equals_single_scene": ok, "every_device_holds_the_total": every,
                  "us_per_exchange": [round(x, 1) for x in us], "TINYR_NVLS": os.environ.get("TINYR_NVLS")}))
m.close()

"""Small many-light AO run (BASELINE config 4 shape, 16 lights, one frame slot so that kernels run one after the
other) for ncu captures and per-kernel CUDA-event times.  Run under gpurun."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import assets  # noqa: E402
import tiny_renderer_b200 as tr  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

model = assets.get_model("diablo")
lights, _ = tr.generate_ao_lights(model.positions, 16)
# AO_PROFILE_RING=2: the light passes overlap on two slots, i.e. the tile kernel runs in its batch shape (128 threads) as in
# the bench's AO batch — used for the ncu capture (which serialises kernels anyway); the CUDA-event times need ring 1.
s = Scene(3840, 2160, model, "occlusion", ring=int(os.environ.get("AO_PROFILE_RING", "1")))
lf, la, up, _ = assets.view_vectors(0.0, 0.0)
s.set_camera(lf, la, up)
s.render_ao(lights[:4], read=False)
s.set_profiling(True)
s.kernel_times()
s.render_ao(lights, read=False)
kt = s.kernel_times()
print({k: round(t / n * 1e3, 1) for k, (t, n) in kt.items() if n and t > 0})
s.close()

"""Multi-GPU parity of the AO exchange: needs two or more GPUs on the box, skipped otherwise.
Each rank is a torchrun process running tools/ao_peer_check.py: the sharded many-light AO batch summed by the
library's own kernel over NVLink peer memory must equal the NCCL all-reduce of the same accumulators and the unsharded
batch, bit for bit, on every rank."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_ao_peer_reduction_matches_nccl_and_single_gpu(world):
    if _gpu_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "ao_peer_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    scratch = os.path.join(ROOT, "gpurun_out")             # kept when run under gpurun (timings of the two exchanges)
    if os.path.isdir(scratch):
        with open(os.path.join(scratch, "ao_peer_check_n%d.json" % world), "w") as f:
            f.write(line + "\n")
    assert out["ok"] and out["world"] == world
    assert all(c["identical_on_all_ranks"] for c in out["cases"])


@pytest.mark.gpu
def test_peer_reduction_with_one_rank_is_the_identity(tinyr):
    import numpy as np
    from assets import get_model
    from tiny_renderer_b200 import multi
    from tiny_renderer_b200.scene import Scene
    m = get_model("african_head")
    g = Scene(161, 119, m, "occlusion")                   # 161 * 119 is not a multiple of 4: the padded tail is exercised
    lights, _ = tinyr.generate_ao_lights(m.positions, 3)
    ref = g.render_ao(lights)
    with pytest.raises(tinyr.TinyrError):
        g.ao_reduce_peers()                               # not connected yet
    multi.connect_ao_peers(g, 0, 1)
    g.ao_reduce_peers()
    assert np.array_equal(g.read_ao(), ref)
    assert len(g.ao_ipc_handle()) == 64
    with pytest.raises(tinyr.TinyrError):
        g.ao_open_peers([b"\0" * 64] * 17, 0)             # more ranks than the kernel's flag area holds
    g.ao_close_peers()
    g.close()


# ---- the native driver (tr_multi_*): one process, one scene per device -------------------------------------------
# On a single-GPU box the device list names GPU 0 several times: the same code path (one host thread and one scene per
# entry, accumulators summed by k_ao_reduce_peers) runs with every "device" on one GPU.

def _device_lists():
    n = _gpu_count()
    lists = [[0], [0, 0]]           # (two scenes per GPU is the limit: the exchange kernels of all entries must be co-resident)
    if n >= 2:
        lists += [[0, 1]]
    if n >= 4:
        lists += [[0, 1, 2, 3]]
    if n >= 8:
        lists += [list(range(8))]
    return lists


@pytest.mark.gpu
def test_native_multi_ao_equals_single_scene_and_oracle():
    """tr_multi_render_ao: lights in contiguous blocks per device entry, accumulators summed over peer memory by the
    library's kernel — equal, bit for bit, to one scene rendering all lights, for every device list, and to the oracle."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import assets
    import oracle_lib
    import tiny_renderer_b200 as tr
    from tiny_renderer_b200.scene import MultiScene, Scene
    m = assets.get_model("diablo")
    w, h = 640, 360
    lights, _ = tr.generate_ao_lights(m.positions, 13)
    lf, la, up, _ = assets.view_vectors(0.0, 0.0)
    one = Scene(w, h, m, "occlusion", ring=2)
    one.set_camera(lf, la, up)
    want = one.render_ao(lights)
    o = oracle_lib.OracleScene(w, h, m, "occlusion")
    o.set_camera(lf, la, up)
    assert np.array_equal(want, o.render_ao(lights))
    kinds = set()
    for devices in _device_lists():
        # default = the peer-memory kernel (the faster one); TINYR_NVLS=1 sums inside the NVSwitch where the box can
        for nvls in ((None, "1") if len(set(devices)) == len(devices) > 1 else (None,)):
            os.environ.pop("TINYR_NVLS", None)
            if nvls:
                os.environ["TINYR_NVLS"] = nvls
            g = MultiScene(w, h, m, "occlusion", devices, ring=2)
            assert g.device_count() == len(devices)
            g.set_camera(lf, la, up)
            kind = g.exchange_kind()
            assert kind == "nvls" or kind.startswith("peer: "), kind
            assert not (nvls is None and kind == "nvls")
            kinds.add(kind.split(":")[0])
            for rep in range(2):            # the barrier flags of the exchange kernel carry over from call to call
                got = g.render_ao(lights)
                assert np.array_equal(got, want), (devices, kind, rep, int((got != want).sum()))
            for i in range(len(devices)):   # every device's accumulator holds the total
                assert np.array_equal(g.scene(i).read_ao(), want), (devices, kind, i)
            g.close()
    os.environ.pop("TINYR_NVLS", None)
    assert "peer" in kinds
    one.close()


@pytest.mark.gpu
def test_native_multi_views_equal_single_scene():
    """tr_multi_render_views: contiguous blocks of views per device entry; checksums and frames land in view order."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import math
    import numpy as np
    import assets
    import tiny_renderer_b200 as tr
    from tiny_renderer_b200.scene import MultiScene, Scene
    m = assets.get_model("african_head")
    w, h, n = 512, 384, 11
    views = np.zeros((n, 12), np.float32)
    for k in range(n):
        lf, la, up, l = assets.view_vectors(2 * math.pi * k / n, 0.7)
        views[k, 0:3], views[k, 3:6], views[k, 6:9], views[k, 9:12] = lf, la, up, l
    one = Scene(w, h, m, "specular", ring=3)
    want_frames = np.zeros((n, h, w, 3), np.uint8)
    want = one.render_views(views, want_checksums=True, frames_out=want_frames)
    for devices in _device_lists():
        g = MultiScene(w, h, m, "specular", devices, ring=3)
        frames = tr.pinned_empty((n, h, w, 3), np.uint8)            # sparse delivery from several scenes into one block
        got = g.render_views(views, want_checksums=True, frames_out=frames)
        assert np.array_equal(got, want), devices
        assert np.array_equal(frames, want_frames), devices
        g.close()
    one.close()

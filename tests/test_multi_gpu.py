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

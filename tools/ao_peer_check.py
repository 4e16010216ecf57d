"""One rank of the multi-GPU AO exchange check (launched by torchrun, one process per GPU; tests/test_multi_gpu.py
and tools/capture_profiles.sh run it).  For a small and a 4K frame it renders a sharded many-light AO batch twice —
summed by the library's own kernel over NVLink peer memory (k_peer.cuh) and by an NCCL all-reduce — and compares both
with the unsharded batch rendered on this rank; then it times the two exchanges on their own (CUDA events on the
stream each runs on, max over ranks).  Rank 0 prints one JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    import torch
    import torch.distributed as dist
    import tiny_renderer_b200 as tr
    from tiny_renderer_b200 import multi
    from tiny_renderer_b200.scene import Scene
    import assets

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    model = assets.get_model("diablo")
    out = {"world": world, "cases": []}
    ok = True
    for (w, h, n_lights, reps) in ((322, 241, 2 * world + 1, 1), (3840, 2160, 4 * world, 20)):
        lights, _ = tr.generate_ao_lights(model.positions, n_lights)
        s = Scene(w, h, model, "occlusion", device=local, ring=4)
        lf, la, up, _ = assets.view_vectors(0.2, 0.1)
        s.set_camera(lf, la, up)
        single = s.render_ao(lights)                                   # unsharded, on this rank
        multi.connect_ao_peers(s, rank, world)
        peer = multi.render_ao_sharded(s, lights, rank, world, exchange="peer")
        again = multi.render_ao_sharded(s, lights, rank, world, exchange="peer")      # flags carry over between calls
        nccl = multi.render_ao_sharded(s, lights, rank, world, exchange="allreduce")
        same = bool(np.array_equal(peer, single) and np.array_equal(again, single) and np.array_equal(nccl, single))
        # the exchanges alone: the accumulators hold the totals now; summing them again multiplies by `world`, mod 2^32
        acc = multi.ao_accumulator_tensor(s)
        scratch = acc.clone()
        t_peer = t_nccl = None
        if reps > 1:
            dist.barrier()
            torch.cuda.synchronize()
            for _ in range(3):
                s.ao_reduce_peers()
            s.synchronize()
            dist.barrier()
            s.timer_begin()
            for _ in range(reps):
                s.ao_reduce_peers()
            t_peer = s.timer_end() / reps
            for _ in range(3):
                dist.all_reduce(scratch)
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                dist.all_reduce(scratch)
            e1.record()
            torch.cuda.synchronize()
            t_nccl = e0.elapsed_time(e1) / reps
            t = torch.tensor([t_peer, t_nccl], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_peer, t_nccl = float(t[0]), float(t[1])
            # (3 + reps) more sums of equal accumulators: total * world^(3 + reps) mod 2^32
            expect = (single.astype(np.uint64) * pow(world, 3 + reps, 1 << 32)) & 0xFFFFFFFF
            same = same and bool(np.array_equal(s.read_ao().astype(np.uint64), expect))
        flag = torch.tensor([1 if same else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = ok and bool(flag.item())
        out["cases"].append({"size": [w, h], "lights": n_lights, "identical_on_all_ranks": bool(flag.item()),
                             "checksum": int(single.astype(np.uint64).sum()), "bytes": 4 * w * h,
                             "peer_kernel_ms": t_peer, "nccl_allreduce_ms": t_nccl})
        del acc, scratch
        multi.disconnect_ao_peers(s, world)
        s.close()
    out["ok"] = ok
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

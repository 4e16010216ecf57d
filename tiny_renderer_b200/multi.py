"""Multi-GPU sharding of the two batch modes that shard naturally; one process per GPU,
`torch.distributed` for the plumbing (NCCL over NVLink on the GPU box, gloo in the CPU tests).

* view batches (BASELINE config 5): independent units, contiguous blocks of views per rank, NO data-path
  collective; only the 8-byte per-view checksums are gathered (control plane).
* many-light ambient occlusion (BASELINE config 4): lights sharded, every rank accumulates its slice into its own
  u32 buffer (camera visibility is recomputed per rank: cheaper than broadcasting 66 MB of keys), then ONE sum of
  the W*H u32 values over the ranks.  Integer sums make the result independent of the rank count and of the
  exchange's algorithm, hence bit-identical for 1/2/4/8 GPUs.  Two exchanges: the library's own kernel over NVLink
  peer memory (connect_ao_peers + Scene.ao_reduce_peers: no host synchronisation, stays on the scene stream; the
  default on GPUs) and a torch.distributed all-reduce (allreduce_ao: NCCL on GPUs, gloo in the CPU tests).

Single frames (configs 1-3) do not shard: replicas only.
"""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous block [lo, hi) of n units owned by `rank`: lo = rank*n // world."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank %r / world %r" % (rank, world))
    return (rank * n) // world, ((rank + 1) * n) // world


class _DevicePtr:
    """Exposes a raw device allocation to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def ao_accumulator_tensor(scene):
    """The scene's device AO accumulator (u32, W*H) as an int32 torch tensor sharing its memory.  Two's-complement
    addition of int32 equals addition of u32 mod 2^32, and int32 is the dtype every backend reduces."""
    import torch
    ptr, n = scene.ao_device_buffer()
    return torch.as_tensor(_DevicePtr(ptr, n, "<i4"), device="cuda")


def allreduce_ao(accum, group=None):
    """The one collective of the product: in-place sum of the per-rank AO accumulators (a torch int32 tensor)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM, group=group)
    return accum


def ao_sharded(render_slice, light_dirs, rank, world, group=None, reduce=allreduce_ao):
    """Backend-independent core of config 4: `render_slice(lights[lo:hi])` returns this rank's accumulator and
    `reduce(accumulator, group)` sums it over the ranks (default: an int32 torch tensor on the process group's
    device, all-reduced in place)."""
    lights = np.ascontiguousarray(light_dirs, np.float32).reshape(-1, 3)
    lo, hi = shard_range(len(lights), rank, world)
    return reduce(render_slice(lights[lo:hi]), group)


def exchange_handles(mine, group=None):
    """Every rank's 64-byte IPC handle, in rank order (control plane: one all_gather_object)."""
    import torch.distributed as dist
    handles = [None] * dist.get_world_size(group)
    dist.all_gather_object(handles, bytes(mine), group=group)
    if any(h is None or len(h) != len(mine) for h in handles):
        raise RuntimeError("a rank sent no handle")
    return handles


def connect_ao_peers(scene, rank, world, group=None):
    """Maps the other ranks' AO accumulators into `scene` (once per scene; collective).  Afterwards
    scene.ao_reduce_peers() sums the accumulators with the library's own kernel over NVLink."""
    if world == 1:
        scene.ao_open_peers([b"\0" * 64], 0)
        return
    import torch.distributed as dist
    scene.ao_open_peers(exchange_handles(scene.ao_ipc_handle(), group), rank)
    dist.barrier(group)                 # nobody starts a reduction before every rank has mapped every accumulator


def disconnect_ao_peers(scene, world, group=None):
    """Unmaps the peers' accumulators; collective, and to be called before any rank closes its scene (an exporter
    must not free memory that other processes still have mapped)."""
    scene.ao_close_peers()
    if world > 1:
        import torch.distributed as dist
        dist.barrier(group)


def render_ao_sharded(scene, light_dirs, rank, world, group=None, read=True, exchange="peer"):
    """Config 4: this rank renders lights [lo, hi) into the scene's accumulator, then the accumulators are summed
    over the ranks.  Returns the (H, W) u32 total on every rank (or None when read=False; the total then stays in
    the scene's device accumulator).  exchange="peer": the library's kernel over NVLink peer memory (needs
    connect_ao_peers first); "allreduce": torch.distributed."""
    if exchange == "peer":
        def render_slice(lights):
            scene.render_ao(lights, reset=True, read=False)
            return scene

        ao_sharded(render_slice, light_dirs, rank, world, group, reduce=lambda s, _g: s.ao_reduce_peers() if world > 1 else None)
    elif exchange == "allreduce":
        import torch

        def render_slice(lights):
            scene.render_ao(lights, reset=True, read=False)
            scene.synchronize()         # the scene renders on its own stream; the collective runs on torch's
            return ao_accumulator_tensor(scene)

        ao_sharded(render_slice, light_dirs, rank, world, group)
        torch.cuda.current_stream().synchronize()
    else:
        raise ValueError("exchange is 'peer' or 'allreduce'")
    return scene.read_ao() if read else None


def render_views_sharded(scene, views, rank, world, group=None, gather=True):
    """Config 5: this rank renders views [lo, hi); returns the u64 checksums of ALL views (gathered over the ranks,
    8 bytes per view — no frame data crosses GPUs) or only this rank's when gather=False."""
    v = np.ascontiguousarray(views, np.float32).reshape(-1, 12)
    lo, hi = shard_range(len(v), rank, world)
    mine = scene.render_views(v[lo:hi], want_checksums=True)
    if not gather or world == 1:
        return mine
    return gather_checksums(mine, len(v), rank, world, group)


def gather_checksums(mine, n_total, rank, world, group=None, device=None):
    """All ranks' per-view checksums in view order.  u64 travels as int64 (bit pattern preserved)."""
    import torch
    import torch.distributed as dist
    if device is None:
        device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    sizes = [shard_range(n_total, r, world)[1] - shard_range(n_total, r, world)[0] for r in range(world)]
    pad = max(sizes) if sizes else 0
    buf = torch.zeros(pad, dtype=torch.int64, device=device)
    buf[:len(mine)] = torch.from_numpy(np.ascontiguousarray(mine, np.uint64).view(np.int64)).to(device)
    parts = [torch.zeros(pad, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    out = np.concatenate([p[:sizes[r]].cpu().numpy() for r, p in enumerate(parts)]) if world else np.zeros(0, np.int64)
    return out.view(np.uint64)

#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 rasterization backend (contract: see the task statement).

    python bench.py --gpus 1 --steps K --warmup W            # own arm
    python bench.py --impl reference --gpus 1 --steps K ...  # the reference's CPU path (oracle port) on host cores
    torchrun ... bench.py --gpus N ...                       # one rank per GPU, weak scaling over view batches

Metric (BASELINE.json): Mpixels/s shaded, device-timed, at 3840x2160.  Workload at every N = BASELINE config 2:
diablo, phong + darboux, 3840x2160, 8 camera angles 2*pi*k/8, light angle 0.7 rad -> one step = 16 frames per
GPU, each `clear; set_light_direction; set_camera; render` (app.rs:170,208-210).  N > 1 shards independent view
batches (no data-path collective) and additionally reports the many-light AO mode (config 4: lights sharded,
one NCCL all-reduce of the u32 accumulators) under "ao".
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W4K, H4K = 3840, 2160
METRIC = "Mpixels/s shaded (device-timed) at 3840x2160"
PIPES = ("phong", "darboux")
N_ANGLES = 8
LIGHT_ANGLE = 0.7
PREROLL_STEPS = 2   # untimed steps queued directly in front of the timed ones (see run_gpu)
RING = int(os.environ.get("TINYR_RING", "8"))   # frame slots per scene: 8 x (8 B keys + 3 B rgb) x 8.29 Mpx = 730 MB of render targets > 126 MB L2


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def bench_model():
    import assets
    m = assets.get_model("diablo")
    data = "reference asset diablo (staged copy of /root/reference/assets), synthetic views" if m.name == "diablo" \
        else "synthetic procedural model (reference assets not staged), synthetic views"
    return m, data


def view_batch(rank=0):
    """8 camera angles (rank offsets the phase so ranks do not render identical frames), app.rs:200-207."""
    import assets
    v = np.zeros((N_ANGLES, 12), np.float32)
    for k in range(N_ANGLES):
        lf, la, up, light = assets.view_vectors(2 * math.pi * k / N_ANGLES + 0.01 * rank, LIGHT_ANGLE)
        v[k, 0:3], v[k, 3:6], v[k, 6:9], v[k, 9:12] = lf, la, up, light
    return v


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region, sampled through NVML from a thread every ~2 ms
    (the timed region of a default run is well under a second, too short for `nvidia-smi -lms`)."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        self.index, self.rows, self.thread, self.stop_flag, self.max_mhz = index, [], None, False, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        except Exception:
            return

        def pump():
            while not self.stop_flag:
                try:
                    self.rows.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
                                      pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)))
                except Exception:
                    pass
                time.sleep(0.002)

        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=2)
        sm = [float(r[0]) for r in self.rows]
        mask = 0
        for r in self.rows:
            mask |= int(r[1])
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": [n for n, bit in self.REASONS if mask & bit], "samples": len(sm)}


# ---- CPU side: the reference's own CPU path, as restated by the oracle (Rust cannot be compiled here) --------

def cpu_render_frames(model, jobs, threads):
    """jobs: list of (pipeline, view_row).  Each worker owns its OracleScenes (ctypes releases the GIL)."""
    import oracle_lib
    lock, pos = threading.Lock(), [0]

    def worker():
        scenes = {}
        while True:
            with lock:
                i = pos[0]
                pos[0] += 1
            if i >= len(jobs):
                return
            pipe, v = jobs[i]
            s = scenes.get(pipe)
            if s is None:
                s = scenes[pipe] = oracle_lib.OracleScene(W4K, H4K, model, pipe)
            s.clear(); s.set_light_direction(v[9:12]); s.set_camera(v[0:3], v[3:6], v[6:9]); s.render()
            s.get_frame_buffer()

    # scene construction (buffer allocation) happens inside the timed region for the first job of each worker;
    # it is ~10 ms against >= 300 ms per 4K frame.
    t0 = time.perf_counter()
    ts = [threading.Thread(target=worker) for _ in range(threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    return time.perf_counter() - t0


def cpu_baseline(model, n_frames_per_pipe=2):
    """Faithful single-thread baseline (the reference's render loop is single-threaded, scene.rs:151) on a
    bounded sample of the same workload: n views of each pipeline at 3840x2160."""
    views = view_batch()
    jobs = [(p, views[k]) for p in PIPES for k in range(n_frames_per_pipe)]
    secs = cpu_render_frames(model, jobs, 1)
    return {"value": len(jobs) * W4K * H4K / secs / 1e6, "unit": "Mpixels/s", "cores": 1, "kind": "port",
            "sample": "%d frames (%d views x phong+darboux) at 3840x2160, oracle C++ port of scene.rs/shader.rs, 1 thread, %.2f s"
                      % (len(jobs), n_frames_per_pipe, secs)}


def cpu_ao_baseline(model, n_lights, threads):
    """BASELINE config 4 on the host: per light the reference's sequence set_light_direction / clear / render with the
    occlusion pipeline (both passes, shader.rs:806-963) at 3840x2160, camera angle 0 — what tr_scene_render_ao sums.  A
    bounded sample of the 1,024 lights (the first n of the same seeded list), one light per host thread."""
    import assets
    import tiny_renderer_b200 as tr
    lights, _ = tr.generate_ao_lights(model.positions, n_lights)
    lf, la, up, _ = assets.view_vectors(0.0, 0.0)
    jobs = []
    for l in lights:
        v = np.zeros(12, np.float32)
        v[0:3], v[3:6], v[6:9], v[9:12] = lf, la, up, l
        jobs.append(("occlusion", v))
    secs = cpu_render_frames(model, jobs, threads)
    return {"value": n_lights / secs, "unit": "passes/s", "cores": threads, "kind": "port",
            "sample": "%d of the %d light directions at 3840x2160, oracle C++ port of the occlusion pipeline, %d thread%s, %.2f s"
                      % (n_lights, AO_LIGHTS, threads, "" if threads == 1 else "s", secs)}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (the oracle port — the Rust crate cannot
    be built on this box) on all host threads it can use, same config/metric/unit."""
    if rank != 0:
        return
    model, data = bench_model()
    cores = os.cpu_count() or 1
    views = view_batch()
    # One reference step = a bounded sample of the GPU arm's step (16 frames): 2..8 of the 8 views of each pipeline, so
    # that any --steps K the driver passes ends within minutes (a 4K frame costs ~0.15-0.5 s of one host core).
    sample_views = max(2, min(N_ANGLES, cores // len(PIPES)))      # enough frames to occupy every host thread
    jobs = [(p, views[k]) for p in PIPES for k in range(sample_views)]
    threads = min(cores, len(jobs))
    times = []
    for i in range(args.warmup + args.steps):
        t = cpu_render_frames(model, jobs, threads)
        if i >= args.warmup:
            times.append(t)
    total = sum(times)
    value = args.steps * len(jobs) * W4K * H4K / total / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mpixels/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": data,
            "config": workload_config(),
            "cpu_baseline": {"value": value, "unit": "Mpixels/s", "cores": threads, "kind": "port",
                             "sample": "%d frames per step (%d views x phong+darboux) at 3840x2160, oracle C++ port of "
                                       "scene.rs/shader.rs, one frame per host thread, %d threads" % (len(jobs), sample_views, threads)},
            "e2e": {"value": value, "unit": "Mpixels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    # the second half of BASELINE's metric, once per run (not per step): one light per host thread
    ao = cpu_ao_baseline(model, threads, threads)
    line["ao"] = {"metric": "AO depth passes/s", "value": ao["value"], "unit": "passes/s", "cpu_baseline": ao}
    print(json.dumps(line), flush=True)


def workload_config():
    return {"workload": "BASELINE config 2: diablo, phong + darboux, 3840x2160, 8 camera angles 2*pi*k/8, light 0.7 rad "
                        "(16 frames per step per GPU)", "width": W4K, "height": H4K, "frames_per_step_per_gpu": len(PIPES) * N_ANGLES,
            "triangles": 5022, "l2": "render targets rotate through a ring of %d buffer sets per scene (%d MB of keys + RGB > 126 MB L2)" % (RING, RING * W4K * H4K * 11 // 1000000),
            "parallelism": "independent view batches per GPU, no data-path collective; on a GPU the phong scene and the "
                           "darboux scene run on a stream each, frames of a scene overlap on its %d frame slots" % RING,
            "preroll_steps": PREROLL_STEPS}


# ---- GPU arm -----------------------------------------------------------------------------------------------

def algorithmic_bytes(n_px, covered, n_tri, pipe):
    """SURVEY.md §8(d) / BASELINE.md §3: B_frame = 19*P + n_maps*min(3*C, 3 MiB) + geometry, split per kernel.
    K1 vertex_setup: static geometry (96 B) read, setup record (64 B) + bbox (16 B) + varyings written per polygon.
    K3+K5 (one fused tile kernel since r1b): the visibility + shading part of the formula, 19*P (key write 8*P,
    key read 8*P, RGB write 3*P — the fused kernel keeps the key read on chip and skips the keys of empty
    tiles, so its DRAM traffic is BELOW this figure; the algorithmic figure is kept as the common yardstick)
    + compulsory texels + setup records/varyings of the polygons read once."""
    n_maps = {"phong": 1, "darboux": 2}[pipe]
    tex = n_maps * min(3 * covered, 3 << 20)
    k1 = n_tri * (96 + 64 + 16 + (96 if pipe == "darboux" else 16))
    k35 = 19 * n_px + tex + n_tri * (64 + 16 + 96)
    return {"K1_vertex_setup": k1, "K3_raster_visibility": k35, "frame": 19 * n_px + tex + k1}


def ncu_kernel_summary(want):
    """Entry of the kernel whose name contains `want` in the committed newest `ncu --set full` summary
    (profiles/r*_traffic.json, written by tools/ncu_summary.py: DRAM bytes, duration and warp instructions per launch);
    None when absent."""
    import glob
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")))
    if not found:
        return None
    # "k_raster<4,128>" matches ncu's spellings "k_raster<4, 128>" and "k_raster<(int)4, (int)128>"
    def norm(name):
        return name.replace("(int)", "").replace(" ", "")
    with open(found[-1]) as f:
        for k, v in json.load(f).items():
            if want.replace(" ", "") in norm(k):
                return dict(v, source=os.path.basename(found[-1]))
    return None


def ncu_traffic(pipe):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the fused tile kernel (None when not captured)."""
    e = ncu_kernel_summary("k_raster<%d,256>" % {"phong": 1, "darboux": 4}[pipe])
    return e["dram_bytes_per_launch"] if e else None


def issue_roofline(kernels, us, n_sm, sm_mhz):
    """The bound these kernels actually run against: one warp instruction per cycle per SM sub-partition, which for
    FP32 work is also the FP32 pipe's rate (tools/probe/ffma2_probe.cu: scalar FMUL+FADD reaches 95 % of it, packed
    FFMA2 is half rate).  Warp instructions per launch come from the committed ncu capture of the named kernels
    (they do not depend on timing), `us` is measured live.  None when a kernel has no capture."""
    entries = [ncu_kernel_summary(k) for k in kernels]
    if any(e is None for e in entries) or not sm_mhz:
        return None
    instr = sum(e["warp_instructions"] for e in entries)
    peak = n_sm * 4 * sm_mhz * 1e6
    ach = instr / (us * 1e-6)
    return {"bound": "issue (1 warp instruction / cycle / SM sub-partition = the FP32 pipe rate)", "kernels": kernels,
            "warp_instructions": instr, "us": us, "achieved": ach / 1e9, "peak": peak / 1e9, "unit": "G warp instr/s",
            "frac": ach / peak, "source": entries[0]["source"]}


def batch_roofline(us_per_frame, n_px, covered, n_tri, peak):
    """Average algorithmic frame bytes of the step's pipelines over the measured time per frame of the timed region."""
    nbytes = sum(algorithmic_bytes(n_px, covered[p], n_tri, p)["frame"] for p in PIPES) / len(PIPES)
    ach = nbytes / (us_per_frame * 1e-6) / 1e9
    return {"algorithmic_bytes_per_frame": nbytes, "us_per_frame": us_per_frame, "achieved": ach, "frac": ach / peak}


AO_LIGHTS = 1024


def run_ao(args, rank, world, dev, model, dist, stream, sm_mhz=None):
    """BASELINE config 4: diablo, occlusion pipeline, 1024 light-direction depth passes at 3840x2160, lights sharded
    over the ranks (strong scaling: the light count is fixed), ONE sum of the u32 accumulators over the ranks at the
    end (the library's kernel over NVLink peer memory).  Device-timed (CUDA events on the launching stream), max over ranks."""
    import torch
    import tiny_renderer_b200 as tr
    from tiny_renderer_b200 import multi
    from tiny_renderer_b200.scene import Scene
    lights, drawn = tr.generate_ao_lights(model.positions, AO_LIGHTS)
    s = Scene(W4K, H4K, model, "occlusion", device=dev, ring=RING)
    s.set_stream(stream)
    lf, la, up, _ = __import__("assets").view_vectors(0.0, 0.0)
    s.set_camera(lf, la, up)
    lo, hi = multi.shard_range(len(lights), rank, world)
    # The sum over the ranks is the library's own kernel over NVLink peer memory (k_peer.cuh), on the scene stream
    # behind the light passes; the NCCL all-reduce of the same buffer is timed beside it on a copy, for comparison.
    multi.connect_ao_peers(s, rank, world)
    s.render_ao(lights[lo:min(hi, lo + 8)], reset=True, read=False)          # warm-up
    acc = multi.ao_accumulator_tensor(s)
    nccl_ms = None
    if dist is not None:
        s.ao_reduce_peers()
        s.synchronize()
        scratch = acc.clone()
        for _ in range(3):
            dist.all_reduce(scratch)
        torch.cuda.synchronize()
        dist.barrier()
        n0, n1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0.record()
        for _ in range(5):
            dist.all_reduce(scratch)
        n1.record()
        torch.cuda.synchronize()
        nccl_ms = n0.elapsed_time(n1) / 5
        del scratch
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    s.render_ao(lights[lo:hi], reset=True, read=False)
    e1.record()
    if world > 1:
        s.ao_reduce_peers()
    e2.record()
    torch.cuda.synchronize()
    ms, ar_ms = e0.elapsed_time(e2), e1.elapsed_time(e2)
    if dist is not None:
        t = torch.tensor([ms, ar_ms, nccl_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ar_ms, nccl_ms = float(t[0].item()), float(t[1].item()), float(t[2].item())
    total = s.read_ao()
    covered = int((s.raw_visibility() != 0xFFFFFFFF).sum())
    diag = s.diagnostics()
    del acc
    multi.disconnect_ao_peers(s, world)
    s.close()
    n_px = W4K * H4K
    per_pass_bytes = 8 * n_px + 8 * covered              # shadow write + one read + u32 accumulate RMW
    peak, _ = load_peaks()
    us_per_pass = ms * 1e3 / max(1, hi - lo)
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    issue = issue_roofline(["k_raster<-2,128>", "k_ao_accumulate"], us_per_pass, n_sm, sm_mhz) if rank == 0 else None
    return {"metric": "AO depth passes/s", "value": AO_LIGHTS / (ms * 1e-3), "unit": "passes/s", "n_gpus": world,
            "scaling": "strong", "lights": AO_LIGHTS, "lights_per_gpu": hi - lo, "ms": ms, "exchange": "k_ao_reduce_peers (own kernel, NVLink peer memory)" if world > 1 else "none",
            "exchange_ms": ar_ms, "nccl_allreduce_ms": nccl_ms, "exchange_bytes": 4 * n_px, "us_per_pass": us_per_pass, "checksum": int(total.astype(np.uint64).sum()),
            "shadow_index_oob": diag["shadow_index_oob"], "shadow_x_wrap": diag["shadow_x_wrap"],
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_pass": per_pass_bytes,
                         "achieved": per_pass_bytes / (us_per_pass * 1e-6) / 1e9, "peak": peak,
                         "frac": per_pass_bytes / (us_per_pass * 1e-6) / 1e9 / peak, "issue": issue},
            "config": "BASELINE config 4: diablo, occlusion, %d light directions (SplitMix64 seed 0x5EED0A0, %d drawn), "
                      "3840x2160, camera angle 0" % (AO_LIGHTS, drawn)}


def bind_to_gpu_numa_node(dev):
    """Pin this rank's host threads (and with them the pinned frame buffers it allocates next) to the NUMA node the GPU
    hangs off, so that N ranks copying frames to the host do not all cross one socket.  Best effort: returns the
    node or None when the box does not expose the topology."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(dev).pci_bus_id
        dom = torch.cuda.get_device_properties(dev).pci_domain_id
        devid = torch.cuda.get_device_properties(dev).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, devid)
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def run_gpu(args, rank, world, local_rank):
    import torch
    import tiny_renderer_b200 as tr
    from tiny_renderer_b200.scene import Scene
    numa_node = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank
    torch.cuda.set_device(dev)
    # All device work of the bench (scenes, events, the NCCL all-reduce) is ordered on ONE explicit stream: a
    # scene given the NULL stream would fall back to its own stream and escape the CUDA events below.
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    model, data = bench_model()
    n_px = W4K * H4K
    n_tri = len(model.tri_idx)
    views = view_batch(rank)
    scenes = {p: Scene(W4K, H4K, model, p, device=dev, ring=RING) for p in PIPES}
    # The two scenes of the step are independent objects (one pipeline per Scene, as in the reference), so each is
    # driven on a stream of its own: their view batches run side by side instead of draining one before the other
    # starts.  Both streams are fenced by the timing events on the main stream (fork() / join() below).
    pstreams = {p: torch.cuda.Stream(device=dev) for p in PIPES}
    for p in PIPES:
        scenes[p].set_stream(pstreams[p].cuda_stream)

    def fork():
        for ps in pstreams.values():
            ps.wait_stream(tstream)

    def join():
        for ps in pstreams.values():
            tstream.wait_stream(ps)

    def step():
        for p in PIPES:
            scenes[p].render_views(views, want_checksums=False)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    for s in scenes.values():
        s.kernel_times()                      # reset launch counters
    sampler = ClockSampler(dev)          # every rank samples its own GPU: `value` is set by the slowest one
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # Two more untimed steps behind the barrier, joined before e0: the device has drained them when the timed steps
    # start, but the host has long since queued the timed steps behind them.  Without this the first timed launches
    # are fed by a host thread that has just left the barrier; with 8 ranks on one host a single late wake-up showed
    # up as 10 % of a 20-step (9 ms) region on one rank — and `value` is the slowest rank's.
    fork()
    for _ in range(PREROLL_STEPS):
        step()
    join()
    e0.record()
    fork()
    for _ in range(args.steps):
        step()
    join()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    launches = sum(n for s in scenes.values() for (_, n) in s.kernel_times().values())
    if dist is not None:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, {"rank": rank, "ms_per_step": ms / args.steps, "sm_mhz": clocks["sm_mhz"], "reasons": clocks["reasons"]})
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        # the line's clocks describe the whole job: lowest median clock and every throttle reason any rank saw
        mhz = [r["sm_mhz"] for r in per_rank if r["sm_mhz"]]
        clocks = dict(clocks, sm_mhz=min(mhz) if mhz else None, reasons=sorted({x for r in per_rank for x in r["reasons"]}), per_rank=per_rank)
    frames_per_step = len(PIPES) * N_ANGLES
    value = world * args.steps * frames_per_step * n_px / (ms * 1e-3) / 1e6

    # ---- e2e: the reference-facing calls with a HOST destination, D2H inside the timed region ----
    # One pinned host buffer per frame of the step: tr_scene_render_views copies every finished frame to the host
    # (asynchronously, on the frame's own stream, so copies overlap the next frames' rendering) and returns when
    # all of them have landed.
    host = tr.pinned_empty((len(PIPES), N_ANGLES, H4K, W4K, 3), np.uint8)
    def e2e_step():
        for i, p in enumerate(PIPES):
            scenes[p].render_views(views, want_checksums=False, frames_out=host[i])
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 3))
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = world * e2e_steps * frames_per_step * n_px / e2e_s / 1e6
    # What bounds e2e: the device->host link.  A plain cudaMemcpyAsync of one step's frame bytes into pinned memory,
    # timed the same way, is the yardstick (rank 0, single-GPU runs only: ranks of one box share the host's link).
    pcie = None
    if rank == 0 and world == 1:
        nbytes = frames_per_step * n_px * 3
        src = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        dst = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        copy_s = (time.perf_counter() - t0) / 3
        del src, dst
        pcie = {"bound": "pcie d2h", "achieved": nbytes * e2e_steps / e2e_s / 1e9, "peak": nbytes / copy_s / 1e9, "unit": "GB/s",
                "frac": (nbytes * e2e_steps / e2e_s) / (nbytes / copy_s),
                "peak_source": "plain pinned device->host copy of one step's frame bytes, measured in this run"}

    line = {"metric": METRIC, "value": value, "unit": "Mpixels/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": data, "config": workload_config(),
            "e2e": {"value": e2e_value, "unit": "Mpixels/s", "h2d_bytes_per_step": frames_per_step * 48,
                    "d2h_bytes_per_step": frames_per_step * n_px * 3,
                    "note": "tr_scene_render_views(views, frames -> pinned host): per frame clear/set_light_direction/set_camera/"
                            "render + device->host copy of the RGB8 frame, all inside the timed region; inputs are the 12 "
                            "floats of camera+light per frame (kernel arguments); PCIe-bound: 24.9 MB per frame",
                    "host_numa_node": numa_node, "roofline": pcie},
            "gpu_launches": int(launches), "clocks": clocks}

    if rank == 0:
        # ---- roofline of the dominant kernel: per-kernel CUDA events (profiling mode), separate from `value` ----
        peak, peak_src = load_peaks()
        per = {}
        covered = {}
        for p in PIPES:
            s = scenes[p]
            # untimed first launch: the lone-frame CTA shape of the tile kernel has not run yet in this process, and
            # lazy module loading would otherwise land in the first timed launch
            v = views[0]
            s.clear(); s.set_light_direction(v[9:12]); s.set_camera(v[0:3], v[3:6], v[6:9]); s.render(); s.synchronize()
            s.set_profiling(True)
            s.kernel_times()
            # one frame at a time on one stream (no overlap between frames): the kernel's own duration
            for _ in range(3):
                for v in views:
                    s.clear(); s.set_light_direction(v[9:12]); s.set_camera(v[0:3], v[3:6], v[6:9]); s.render()
            kt = s.kernel_times()
            s.set_profiling(False)
            covered[p] = int((s.raw_visibility() != 0xFFFFFFFF).sum())
            per[p] = {k: (t / n * 1e3 if n and t > 0 else 0.0) for k, (t, n) in kt.items() if n and t > 0}     # us per launch
        line["kernel_us"] = per
        best = None
        for p in PIPES:
            ab = algorithmic_bytes(n_px, covered[p], n_tri, p)
            frame_us = sum(per[p].get(k, 0.0) for k in ("K1_vertex_setup", "K3_raster_visibility"))
            for k in ("K3_raster_visibility",):
                us = per[p].get(k, 0.0)
                if best is None or us > best[0]:
                    best = (us, p, k, ab[k], ab["frame"], frame_us)
        us, p, k, nbytes, frame_bytes, frame_us = best
        achieved = nbytes / (us * 1e-6) / 1e9
        line["roofline"] = {"bound": "hbm", "kernel": "k_raster<%s> = K3 raster_visibility + K5 shade fused" % p, "achieved": achieved, "peak": peak, "unit": "GB/s",
                            "frac": achieved / peak, "traffic": ncu_traffic(p), "peak_source": peak_src,
                            "algorithmic_bytes_per_launch": nbytes, "us_per_launch": us,
                            "note": "kernel timed one frame at a time (CUDA events on the launching stream, profiling mode); "
                                    "`batch` is the same byte count over the overlapped view batch of the timed region",
                            "frame_serial": {"algorithmic_bytes": frame_bytes, "us": frame_us,
                                             "achieved": frame_bytes / (frame_us * 1e-6) / 1e9,
                                             "frac": frame_bytes / (frame_us * 1e-6) / 1e9 / peak},
                            "batch": batch_roofline(ms / args.steps / frames_per_step * 1e3, n_px, covered, n_tri, peak)}
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        # the same kernel and the whole overlapped batch against the instruction-issue / FP32 rate (what actually binds them)
        line["roofline"]["issue"] = issue_roofline(["k_raster<%d,256>" % {"phong": 1, "darboux": 4}[p]], us, n_sm, clocks.get("sm_mhz"))
        line["roofline"]["batch"]["issue"] = issue_roofline(["k_raster<1,128>", "k_raster<4,128>", "k_vertex_setup<0>", "k_vertex_setup<0>"],
                                                            2 * ms / args.steps / frames_per_step * 1e3, n_sm, clocks.get("sm_mhz"))
        if world == 1:
            # what a rasteriser resolving visibility with global 64-bit atomicMin could reach on this box
            # (distinct addresses, every atomic updates, one atomic per fragment); k_raster resolves in registers instead
            rate = tr.bench_atomic_min(W4K, H4K, rounds=4, repeats=3, device=dev)
            line["atomic_min_peak"] = {"atomics_per_s": rate, "equivalent_GB_per_s": rate * 8 / 1e9,
                                       "us_per_4k_depth_layer": n_px / rate * 1e6,
                                       "note": "micro-benchmark only (tr_bench_atomic_min); not on the render path"}
            line["cpu_baseline"] = cpu_baseline(model)
            ao_cpu = cpu_ao_baseline(model, 3, 1)
    for s in scenes.values():
        s.close()
    line["ao"] = run_ao(args, rank, world, dev, model, dist, stream, sm_mhz=(clocks or {}).get("sm_mhz") if rank == 0 else None)
    if rank == 0 and world == 1:
        line["ao"]["cpu_baseline"] = ao_cpu
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="tinyr", choices=["tinyr", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

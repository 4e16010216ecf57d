#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 rasterization backend (contract: see the task statement).

    python bench.py --gpus 1 --steps K --warmup W            # own arm
    python bench.py --impl reference --gpus 1 --steps K ...  # the reference's CPU path (oracle port) on host cores
    torchrun ... bench.py --gpus N ...                       # one rank per GPU, weak scaling over view batches

Metric (BASELINE.json): Mpixels/s shaded, device-timed, at 3840x2160.  Workload at every N = BASELINE config 2 as a
turntable: diablo, phong + darboux, 3840x2160, 64 camera angles 2*pi*k/64, light angle 0.7 rad; one step = REPS
turns of both pipelines = 2 * 64 * REPS frames per GPU, each `clear; set_light_direction; set_camera; render`
(app.rs:170,208-210).  With 64 angles on 8 frame slots every slot renders a different view each time it comes round,
so tile-state flips, zero-fills and the read-back's damage tracking all run inside the timed region (r1 used 8 angles
on 8 slots: every slot re-rendered the same frame; that figure is reported once as `aliased_8_angles`).  N > 1 shards
independent view batches (no data-path collective).  Beside the headline the line carries BASELINE's other configs:
`shadow4k` (config 3), `ao` (config 4: 1024 lights sharded, one sum of the u32 accumulators) and `views8k` (config 5:
4096-view specular turntable at 7680x4320, views sharded, strong scaling).
"""
import argparse
import json
import math
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W4K, H4K = 3840, 2160
W8K, H8K = 7680, 4320
METRIC = "Mpixels/s shaded (device-timed) at 3840x2160"
PIPES = ("phong", "darboux")
PIPE_ID = {"default": 0, "phong": 1, "normal_map": 2, "specular": 3, "darboux": 4, "shadow": 5, "occlusion": 6}
N_ANGLES = 64
LIGHT_ANGLE = 0.7
QUICK = os.environ.get("TINYR_BENCH_QUICK") == "1"          # smoke-test sizes (never a bench value)
REPS = int(os.environ.get("TINYR_BENCH_REPS", "1" if QUICK else "8"))   # turns of the 64-angle turntable per step
PREROLL_STEPS = 1   # untimed steps queued directly in front of the timed ones (see timed_batches)
RING = int(os.environ.get("TINYR_RING", "8"))   # frame slots per scene: 8 x (8 B keys + 3 B rgb) x 8.29 Mpx = 730 MB of render targets > 126 MB L2
E2E_HOST_FRAMES = 16    # pinned host frames per scene that the e2e leg delivers into, round robin
VIEWS_8K = 64 if QUICK else 4096
AO_LIGHTS = 32 if QUICK else 1024


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def bench_model(name="diablo"):
    import assets
    m = assets.get_model(name)
    data = ("reference asset %s (staged copy of /root/reference/assets), synthetic views" % name) if m.name == name \
        else "synthetic procedural model (reference assets not staged), synthetic views"
    return m, data


def view_batch(n_angles=N_ANGLES, rank=0, light_angle=LIGHT_ANGLE):
    """Turntable camera angles 2*pi*k/n (rank offsets the phase so ranks do not render identical frames), app.rs:200-207."""
    import assets
    v = np.zeros((n_angles, 12), np.float32)
    for k in range(n_angles):
        lf, la, up, light = assets.view_vectors(2 * math.pi * k / n_angles + 0.01 * rank, light_angle)
        v[k, 0:3], v[k, 3:6], v[k, 6:9], v[k, 9:12] = lf, la, up, light
    return v


def views_8k(lo, hi, n_per_model):
    """Config 5: view j of a model uses camera angle 2*pi*j/n_per_model, light angle 0.7 rad."""
    import assets
    v = np.zeros((hi - lo, 12), np.float32)
    for i, j in enumerate(range(lo, hi)):
        lf, la, up, light = assets.view_vectors(2 * math.pi * j / n_per_model, LIGHT_ANGLE)
        v[i, 0:3], v[i, 3:6], v[i, 6:9], v[i, 9:12] = lf, la, up, light
    return v


class ClockSampler:
    """SM clock, power and throttle reasons DURING the timed region, sampled through NVML from a thread every ~2 ms."""
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
                                      pynvml.nvmlDeviceGetCurrentClocksEventReasons(h),
                                      pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0))
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
                "reasons": [n for n, bit in self.REASONS if mask & bit], "samples": len(sm),
                "power_w_max": max((r[2] for r in self.rows), default=None)}


# ---- CPU side: the reference's own CPU path, as restated by the oracle (Rust cannot be compiled here) --------

def use_timing_oracle():
    """The CPU legs time oracle/liboracle_timing.so: the oracle's loops without its extras (visibility store, per-fragment
    counters), i.e. the work the reference's scene.rs:240-263 does.  Falls back to the checker build when it is absent."""
    p = os.path.join(ROOT, "oracle", "liboracle_timing.so")
    if os.path.exists(p):
        os.environ.setdefault("TINYR_ORACLE_LIB", p)
        return "liboracle_timing.so"
    return "liboracle.so"


class CpuPool:
    """Persistent host worker threads, each owning its OracleScenes (built BEFORE any timed region; ctypes releases the
    GIL during a render).  run(jobs) renders every job — (scene key, view row) — and returns the wall time."""

    def __init__(self, model_of, scene_keys, threads):
        self.oracle_build = use_timing_oracle()
        import oracle_lib
        self.threads = threads
        self.jobs, self.pos, self.lock = [], [0], threading.Lock()
        self.start, self.done = threading.Barrier(threads + 1), threading.Barrier(threads + 1)
        self.quit = False
        ready = threading.Barrier(threads + 1)

        def worker():
            scenes = {key: oracle_lib.OracleScene(key[2], key[3], model_of[key[0]], key[1]) for key in scene_keys}
            for s in scenes.values():       # touch every page of the buffers once
                s.clear()
            ready.wait()
            while True:
                self.start.wait()
                if self.quit:
                    return
                while True:
                    with self.lock:
                        i = self.pos[0]
                        self.pos[0] += 1
                    if i >= len(self.jobs):
                        break
                    key, v = self.jobs[i]
                    s = scenes[key]
                    s.clear(); s.set_light_direction(v[9:12]); s.set_camera(v[0:3], v[3:6], v[6:9]); s.render()
                    s.get_frame_buffer()
                self.done.wait()

        self.workers = [threading.Thread(target=worker, daemon=True) for _ in range(threads)]
        [t.start() for t in self.workers]
        ready.wait()

    def run(self, jobs):
        self.jobs, self.pos[0] = jobs, 0
        t0 = time.perf_counter()
        self.start.wait()
        self.done.wait()
        return time.perf_counter() - t0

    def close(self):
        self.quit = True
        self.start.wait()
        [t.join() for t in self.workers]


def ao_lights_host(model, n):
    """The seeded light list of config 4 WITHOUT the product library: the committed fixture for diablo, else the Python
    twin of tr_ao_generate_lights (tests/ao_lights_py.py)."""
    fixture = os.path.join(ROOT, "tests", "golden", "ao_lights_diablo_1024.npy")
    if model.name == "diablo" and os.path.exists(fixture):
        return np.load(fixture)[:n]
    import ao_lights_py
    return ao_lights_py.generate(model.positions, n)[0]


def ao_jobs(model, n_lights, key):
    import assets
    lf, la, up, _ = assets.view_vectors(0.0, 0.0)
    jobs = []
    for l in ao_lights_host(model, n_lights):
        v = np.zeros(12, np.float32)
        v[0:3], v[3:6], v[6:9], v[9:12] = lf, la, up, l
        jobs.append((key, v))
    return jobs


def cpu_baselines(model, head):
    """Faithful single-thread baselines (the reference's render loop is single-threaded, scene.rs:151; its ThreadPool is
    never used, scene.rs:72-74) on bounded samples of the same workloads, scenes built before the timed region."""
    views = view_batch()
    k4 = [("diablo", p, W4K, H4K) for p in PIPES] + [("diablo", "shadow", W4K, H4K), ("diablo", "occlusion", W4K, H4K)]
    k8 = [("head", "specular", W8K, H8K), ("diablo", "specular", W8K, H8K)]
    pool = CpuPool({"diablo": model, "head": head}, k4 + k8, 1)
    out = {}
    jobs = [(k4[i], views[8 * k]) for i in range(2) for k in range(2)]
    secs = pool.run(jobs)
    out["main"] = {"value": len(jobs) * W4K * H4K / secs / 1e6, "unit": "Mpixels/s", "cores": 1, "kind": "port",
                   "sample": "%d frames (2 views x phong+darboux) at 3840x2160, oracle C++ port of scene.rs/shader.rs, 1 thread, %.2f s"
                             % (len(jobs), secs)}
    jobs = [(k4[2], view_batch(light_angle=2.1)[8]), (k4[2], views[24])]
    secs = pool.run(jobs)
    out["shadow4k"] = {"value": len(jobs) * W4K * H4K / secs / 1e6, "unit": "Mpixels/s", "cores": 1, "kind": "port",
                       "sample": "%d shadow frames at 3840x2160 (both passes), oracle port, 1 thread, %.2f s" % (len(jobs), secs)}
    n = 3
    secs = pool.run(ao_jobs(model, n, k4[3]))
    out["ao"] = {"value": n / secs, "unit": "passes/s", "cores": 1, "kind": "port",
                 "sample": "%d of the %d light directions at 3840x2160, oracle C++ port of the occlusion pipeline "
                           "(set_light_direction / clear / render per light, shader.rs:806-963), 1 thread, %.2f s" % (n, AO_LIGHTS, secs)}
    v8 = views_8k(0, 2048, 2048)
    jobs = [(k8[0], v8[0]), (k8[0], v8[700]), (k8[1], v8[0]), (k8[1], v8[700])]
    secs = pool.run(jobs)
    out["views8k"] = {"value": len(jobs) * W8K * H8K / secs / 1e6, "unit": "Mpixels/s", "cores": 1, "kind": "port",
                      "sample": "4 of the 4096 views (2 african_head + 2 diablo, specular) at 7680x4320, oracle port, 1 thread, %.2f s" % secs}
    for v in out.values():
        v["oracle_build"] = pool.oracle_build       # liboracle_timing.so: the loops without the checker's extras
    pool.close()
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (the oracle port — the Rust crate cannot
    be built on this box) on all host threads it can use, same config/metric/unit.  Worker threads and their scenes
    are created before the timed loop; nothing of the product library is loaded."""
    if rank != 0:
        return
    model, data = bench_model()
    cores = os.cpu_count() or 1
    views = view_batch()
    # One reference step = a bounded sample of the GPU arm's step: one frame per host thread, spread evenly over the 64
    # angles and alternating between the two pipelines, so that any --steps K ends within minutes (a 4K frame costs
    # ~0.15-0.5 s of one host core).
    n_jobs = max(4, cores)
    keys = [("diablo", p, W4K, H4K) for p in PIPES]
    jobs = [(keys[i % 2], views[((i // 2) * N_ANGLES // max(1, n_jobs // 2)) % N_ANGLES]) for i in range(n_jobs)]
    threads = min(cores, len(jobs))
    occ = ("diablo", "occlusion", W4K, H4K)
    pool = CpuPool({"diablo": model}, keys + [occ], threads)
    times = []
    for i in range(args.warmup + args.steps):
        t = pool.run(jobs)
        if i >= args.warmup:
            times.append(t)
    total = sum(times)
    value = args.steps * len(jobs) * W4K * H4K / total / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mpixels/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": data,
            "config": workload_config(),
            "cpu_baseline": {"value": value, "unit": "Mpixels/s", "cores": threads, "kind": "port",
                             "sample": "%d frames per step (phong + darboux alternating, angles spread over the turntable) at "
                                       "3840x2160, oracle C++ port of scene.rs/shader.rs, one frame per host thread, %d persistent "
                                       "threads, scenes built before the timed loop" % (len(jobs), threads)},
            "e2e": {"value": value, "unit": "Mpixels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    # the second half of BASELINE's metric, once per run (not per step): one light per host thread
    secs = pool.run(ao_jobs(model, threads, occ))
    pool.close()
    ao = {"value": threads / secs, "unit": "passes/s", "cores": threads, "kind": "port",
          "sample": "%d of the %d light directions at 3840x2160, oracle port of the occlusion pipeline, one light per host thread, %.2f s"
                    % (threads, AO_LIGHTS, secs)}
    line["ao"] = {"metric": "AO depth passes/s", "value": ao["value"], "unit": "passes/s", "cpu_baseline": ao}
    line["cpu_baseline"]["oracle_build"] = ao["oracle_build"] = pool.oracle_build
    print(json.dumps(line), flush=True)


def workload_config():
    return {"workload": "BASELINE config 2 as a turntable: diablo, phong + darboux, 3840x2160, %d camera angles 2*pi*k/%d, light 0.7 rad, "
                        "%d turns per step (%d frames per step per GPU)" % (N_ANGLES, N_ANGLES, REPS, len(PIPES) * N_ANGLES * REPS),
            "width": W4K, "height": H4K, "frames_per_step_per_gpu": len(PIPES) * N_ANGLES * REPS,
            "triangles": 5022, "l2": "render targets rotate through a ring of %d buffer sets per scene (%d MB of keys + RGB > 126 MB L2); "
                                     "view k lands on slot k mod %d, so a slot never renders the same view twice in a row" % (RING, RING * W4K * H4K * 11 // 1000000, RING),
            "parallelism": "independent view batches per GPU, no data-path collective; on a GPU the phong scene and the "
                           "darboux scene run on a stream each, frames of a scene overlap on its %d frame slots" % RING,
            "preroll_steps": PREROLL_STEPS}


# ---- GPU arm -----------------------------------------------------------------------------------------------

N_MAPS = {"default": 1, "phong": 1, "shadow": 1, "normal_map": 2, "darboux": 2, "specular": 3, "occlusion": 0}


def algorithmic_bytes(n_px, covered, n_tri, pipe):
    """SURVEY.md §8(d) / BASELINE.md §3: B_frame = 19*P + n_maps*min(3*C, 3 MiB) + geometry (+ 8*P for the shadow
    pipeline's light-depth buffer: one write, one read), split per kernel.
    K1 vertex_setup: static geometry (96 B) read, setup record (64 B) + bbox (16 B) + varyings written per polygon.
    K3+K5 (one fused tile kernel): the visibility + shading part of the formula, 19*P (key write 8*P, key read 8*P,
    RGB write 3*P — the fused kernel keeps the key read on chip and skips the keys of empty tiles, so its DRAM
    traffic is BELOW this figure; the algorithmic figure is kept as the common yardstick) + compulsory texels + setup
    records/varyings of the polygons read once."""
    tex = N_MAPS[pipe] * min(3 * covered, 3 << 20)
    k1 = n_tri * (96 + 64 + 16 + (96 if pipe == "darboux" else 16))
    k35 = 19 * n_px + tex + n_tri * (64 + 16 + 96)
    extra = 8 * n_px + n_tri * (96 + 64 + 16) if pipe == "shadow" else 0
    return {"K1_vertex_setup": k1, "K3_raster_visibility": k35, "frame": 19 * n_px + tex + k1 + extra}


def ncu_summaries():
    """The committed newest `ncu` summary (profiles/r*_traffic.json, written by tools/ncu_summary.py)."""
    import glob
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")))
    if not found:
        return {}, None
    with open(found[-1]) as f:
        return json.load(f), os.path.basename(found[-1])


PIPE_STRUCT = {"default": "PipelineDefault", "phong": "PipelinePhong", "normal_map": "PipelineNormalMap", "specular": "PipelineSpecular",
               "darboux": "PipelineDarboux", "shadow": "PipelineShadow", "occlusion": "PipelineOcclusion",
               "light_depth": "PassLightDepth", "visibility": "PassVisibility"}
LEGACY_ID = dict(PIPE_ID, light_depth=-2, visibility=-1)


def raster_kernel(pipe, threads):
    """Name of the tile kernel instantiated with `pipe`'s fragment role (pipeline.cuh) in a CTA shape."""
    return "k_raster<%s,%s>" % (PIPE_STRUCT[pipe], threads)


def setup_kernel(pipe):
    """The vertex set-up kernel: one instantiation for the camera pass of every pipeline, one for the light pass."""
    return "k_vertex_setup<true,false>" if pipe == "light_depth" else "k_vertex_setup<false,false>"     # <LIGHT_PASS, SORT> (k_setup.cuh)


def ncu_kernel_summary(want):
    """Entry of the kernel whose name contains `want` (DRAM bytes, duration and warp instructions per launch).  Also finds
    captures taken before the kernels were templated on pipeline structs (r1: k_raster<4,128>, k_vertex_setup<0>)."""
    def norm(name):     # ncu spells "k_raster<tr::PipelineDarboux, 128>" / "k_raster<(int)4, (int)128>"
        return name.replace("(int)", "").replace("(bool)", "").replace("tr::", "").replace(" ", "")
    wants = [want.replace(" ", "")]
    for pipe, st in PIPE_STRUCT.items():
        if st in want:
            wants.append(wants[0].replace("k_raster<" + st, "k_raster<%d" % LEGACY_ID[pipe]))
    if wants[0] == "k_vertex_setup<false,false>":     # earlier spellings of the same kernel in committed captures
        wants += ["k_vertex_setup<0,0>", "k_vertex_setup<false>", "k_vertex_setup<0>", "k_vertex_setup<PipelineDarboux>", "k_vertex_setup<PipelinePhong>"]
    if wants[0] == "k_vertex_setup<true,false>":
        wants += ["k_vertex_setup<1,0>", "k_vertex_setup<true>", "k_vertex_setup<1>", "k_vertex_setup<PassLightDepth>"]
    table, src = ncu_summaries()
    for w in wants:
        for k, v in table.items():
            if isinstance(v, dict) and w in norm(k):
                return dict(v, source=src)
    return None


def issue_roofline(kernels, us, n_sm, sm_mhz, optional=()):
    """The bound these kernels actually run against: one warp instruction per cycle per SM sub-partition, which for
    FP32 work is also the FP32 pipe's rate (tools/probe/ffma2_probe.cu).  Warp instructions per launch come from the
    committed ncu capture of the named kernels (they do not depend on timing), `us` is measured live."""
    entries = [ncu_kernel_summary(k) for k in kernels]
    if any(e is None or "warp_instructions" not in e for e in entries) or not sm_mhz:
        return None
    extra = [(k, ncu_kernel_summary(k)) for k in optional]          # counted when the committed capture has them
    kernels = list(kernels) + [k for k, e in extra if e]
    entries += [e for _, e in extra if e]
    instr = sum(e["warp_instructions"] for e in entries)
    peak = n_sm * 4 * sm_mhz * 1e6
    ach = instr / (us * 1e-6)
    return {"bound": "issue (1 warp instruction / cycle / SM sub-partition = the FP32 pipe rate)", "kernels": kernels,
            "warp_instructions": instr, "us": us, "achieved": ach / 1e9, "peak": peak / 1e9, "unit": "G warp instr/s",
            "frac": ach / peak, "source": entries[0]["source"]}


def lone_frame_us(scene, views, reps=2):
    """Per-kernel device time of lone frames (profiling mode: CUDA events around every launch on the scene stream,
    one frame at a time, no overlap between frames)."""
    v = views[0]
    scene.clear(); scene.set_light_direction(v[9:12]); scene.set_camera(v[0:3], v[3:6], v[6:9]); scene.render(); scene.synchronize()
    scene.set_profiling(True)
    scene.kernel_times()
    for _ in range(reps):
        for v in views:
            scene.clear(); scene.set_light_direction(v[9:12]); scene.set_camera(v[0:3], v[3:6], v[6:9]); scene.render()
    kt = scene.kernel_times()
    scene.set_profiling(False)
    return {k: (t / n * 1e3) for k, (t, n) in kt.items() if n and t > 0}


def settled_frame_us(scene, views, warm=4, frames=2):
    """As lone_frame_us, for a view that stands still: every view is rendered `warm` times unprofiled, then `frames` times
    profiled.  This is the reference's app loop with no key down (app.rs:166-247 re-renders continuously; with a key down
    the camera turns by 3 rad/s x frame time, i.e. < 0.1 degree per frame).  From the fifth frame of a view on a lone
    frame launches its tiles in the order of their cost in the previous frames (k_setup.cuh sort_tiles, built twice) and
    re-uses the set-up kernel's output of the identical view instead of launching it again (tinyr.cu run_setup)."""
    scene.kernel_times()
    for v in views:
        for i in range(warm + frames):
            if i == warm:
                scene.synchronize()
                scene.set_profiling(True)
            scene.clear(); scene.set_light_direction(v[9:12]); scene.set_camera(v[0:3], v[3:6], v[6:9]); scene.render()
        scene.synchronize()
        scene.set_profiling(False)
    kt = scene.kernel_times()       # (the launch counts include the unprofiled frames: divide by the profiled ones)
    return {k: (t / (len(views) * frames) * 1e3) for k, (t, n) in kt.items() if n and t > 0}


def back_to_back_us(scene, v, frames=100, reps=5):
    """SURVEY.md 8(d) config 2 as it is defined: `frames` back-to-back frames of ONE view (clear / set_* / render per frame,
    nothing else between them), CUDA events on the scene stream around all of them, median over `reps`: us per frame."""
    def drive(n):
        for _ in range(n):
            scene.clear(); scene.set_light_direction(v[9:12]); scene.set_camera(v[0:3], v[3:6], v[6:9]); scene.render()
    drive(8)
    scene.synchronize()
    out = []
    for _ in range(reps):
        scene.timer_begin()
        drive(frames)
        out.append(scene.timer_end() / frames * 1e3)
    return statistics.median(out)


def timed_batches(torch, dist, tstream, pstreams, call, steps, preroll=1):
    """`steps` x call() between CUDA events on tstream, the scenes' streams forked from / joined into it; a barrier and a
    device synchronize on both sides; returns ms (this rank)."""
    def fork():
        for ps in pstreams:
            ps.wait_stream(tstream)

    def join():
        for ps in pstreams:
            tstream.wait_stream(ps)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # Untimed steps behind the barrier, joined before e0: the device has drained them when the timed steps start, but
    # the host has long since queued the timed steps behind them (r1: without this the first timed launches are fed by
    # a host thread that has just left the barrier).
    fork()
    for _ in range(preroll):
        call()
    join()
    e0.record()
    fork()
    for _ in range(steps):
        call()
    join()
    e1.record()
    barrier()
    return e0.elapsed_time(e1)


def max_over_ranks(torch, dist, *vals):
    if dist is not None:
        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        vals = tuple(float(x) for x in t.tolist())
    return vals if len(vals) > 1 else vals[0]


def sum_over_ranks(torch, dist, val):
    if dist is None:
        return val
    t = torch.tensor([val], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def weighted_checksum(a):
    """Position-weighted 64-bit hash of a u32 image (the tr_checksum_rgb8 scheme): sum of w_i * (2 i + 1) mod 2^64 — a
    permutation of the values changes it, unlike a plain sum."""
    w = np.ascontiguousarray(a, np.uint32).reshape(-1).astype(np.uint64)
    idx = np.arange(w.size, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return int((w * (idx * np.uint64(2) + np.uint64(1))).sum(dtype=np.uint64))


def run_ao(args, rank, world, dev, model, dist, tstream, sm_mhz=None):
    """BASELINE config 4: diablo, occlusion pipeline, 1024 light-direction depth passes at 3840x2160, lights sharded
    over the ranks (strong scaling: the light count is fixed), ONE sum of the u32 accumulators over the ranks at the
    end.  Device-timed (CUDA events on the launching stream), max over ranks.  The exchange is timed both ways, apples to
    apples (same buffer, from a barrier, 5 back-to-back operations, max over ranks): the library's kernel over NVLink
    peer memory and NCCL's all-reduce; the timed batch uses whichever was faster."""
    import torch
    import tiny_renderer_b200 as tr
    from tiny_renderer_b200 import multi
    from tiny_renderer_b200.scene import Scene
    import assets
    stream = tstream.cuda_stream
    lights, drawn = tr.generate_ao_lights(model.positions, AO_LIGHTS)
    s = Scene(W4K, H4K, model, "occlusion", device=dev, ring=RING)
    s.set_stream(stream)
    lf, la, up, _ = assets.view_vectors(0.0, 0.0)
    s.set_camera(lf, la, up)
    lo, hi = multi.shard_range(len(lights), rank, world)
    multi.connect_ao_peers(s, rank, world)
    s.render_ao(lights[lo:min(hi, lo + 8)], reset=True, read=False)          # warm-up
    acc = multi.ao_accumulator_tensor(s)
    peer_ms = nccl_ms = None
    use_peer = True
    if dist is not None:
        def time_exchange(op):
            for _ in range(2):
                op()
            torch.cuda.synchronize()
            dist.barrier()
            n0, n1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n0.record()
            for _ in range(5):
                op()
            n1.record()
            torch.cuda.synchronize()
            return max_over_ranks(torch, dist, n0.elapsed_time(n1) / 5)
        peer_ms = time_exchange(s.ao_reduce_peers)
        nccl_ms = time_exchange(lambda: dist.all_reduce(acc))
        use_peer = peer_ms <= nccl_ms
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    s.render_ao(lights[lo:hi], reset=True, read=False)
    e1.record()
    if world > 1:
        if use_peer:
            s.ao_reduce_peers()
        else:
            dist.all_reduce(acc)            # on tstream, behind the light passes
    e2.record()
    torch.cuda.synchronize()
    ms, ar_ms = max_over_ranks(torch, dist, e0.elapsed_time(e2), e1.elapsed_time(e2))
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
    issue = issue_roofline([raster_kernel("light_depth", 128), "k_ao_accumulate"], us_per_pass, n_sm, sm_mhz, optional=[setup_kernel("light_depth")]) if rank == 0 else None
    return {"metric": "AO depth passes/s", "value": AO_LIGHTS / (ms * 1e-3), "unit": "passes/s", "n_gpus": world,
            "scaling": "strong", "lights": AO_LIGHTS, "lights_per_gpu": hi - lo, "ms": ms,
            "exchange": ("none" if world == 1 else "k_ao_reduce_peers (own kernel, NVLink peer memory)" if use_peer else "NCCL all-reduce (faster than the own kernel in this run)"),
            "exchange_ms_in_batch": ar_ms, "exchange_ms_alone": {"k_ao_reduce_peers": peer_ms, "nccl_allreduce": nccl_ms,
                                                                 "method": "5 back-to-back operations on the accumulator from a barrier, CUDA events, max over ranks"},
            "exchange_bytes": 4 * n_px, "us_per_pass": us_per_pass,
            "checksum": weighted_checksum(total), "checksum_kind": "sum of accum[i] * (2 i + 1) mod 2^64 (position-weighted)",
            "shadow_index_oob": diag["shadow_index_oob"], "shadow_x_wrap": diag["shadow_x_wrap"],
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_pass": per_pass_bytes,
                         "achieved": per_pass_bytes / (us_per_pass * 1e-6) / 1e9, "peak": peak,
                         "frac": per_pass_bytes / (us_per_pass * 1e-6) / 1e9 / peak, "issue": issue},
            "config": "BASELINE config 4: diablo, occlusion, %d light directions (SplitMix64 seed 0x5EED0A0, %d drawn), "
                      "3840x2160, camera angle 0" % (AO_LIGHTS, drawn)}


def run_views8k(rank, world, dev, model, head, dist, tstream):
    """BASELINE config 5 exactly as SURVEY.md §8(d) defines it: specular at 7680x4320, 4096-view turntable, views 0-2047
    african_head and 2048-4095 diablo (view j of a model: camera angle 2*pi*j/2048, light 0.7 rad), per-view 64-bit
    checksums computed on the device.  Strong scaling: every rank renders a contiguous block of each model's views
    (balanced: the head covers 1.7x the pixels of diablo), no data-path collective; only checksums are gathered."""
    import torch
    from tiny_renderer_b200 import multi
    from tiny_renderer_b200.scene import Scene
    n_per_model = VIEWS_8K // 2
    n_px = W8K * H8K
    lo, hi = multi.shard_range(n_per_model, rank, world)
    scenes = {"african_head": Scene(W8K, H8K, head, "specular", device=dev, ring=RING),
              "diablo": Scene(W8K, H8K, model, "specular", device=dev, ring=RING)}
    pstreams = [torch.cuda.Stream(device=dev) for _ in scenes]
    for s, ps in zip(scenes.values(), pstreams):
        s.set_stream(ps.cuda_stream)
    v = views_8k(lo, hi, n_per_model)
    for s in scenes.values():
        s.render_views(v[:min(len(v), 2 * RING)], want_checksums=True)      # warm-up
    sums = {}

    def call():
        # the two scenes are driven from a host thread each (a scene is single-threaded, distinct scenes are not), so
        # their batches run side by side on the device and the checksum read-back of one does not drain the other
        def job(name):
            sums[name] = scenes[name].render_views(v, want_checksums=True)
        ts = [threading.Thread(target=job, args=(name,)) for name in scenes]
        [t.start() for t in ts]
        [t.join() for t in ts]

    ms = max_over_ranks(torch, dist, timed_batches(torch, dist, tstream, pstreams, call, 1, preroll=0))
    mine = 0
    for name in scenes:
        with np.errstate(over="ignore"):
            mine = (mine + int(sums[name].sum(dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
    if dist is not None:
        parts = [None] * world
        dist.all_gather_object(parts, mine)
        mine = sum(parts) & 0xFFFFFFFFFFFFFFFF
    out = {"metric": "Mpixels/s shaded (device-timed) at 7680x4320", "value": VIEWS_8K * n_px / (ms * 1e-3) / 1e6, "unit": "Mpixels/s",
           "n_gpus": world, "scaling": "strong", "views": VIEWS_8K, "views_per_gpu": 2 * (hi - lo), "ms": ms,
           "us_per_view": ms * 1e3 / max(1, 2 * (hi - lo)), "checksum_of_checksums": mine,
           "config": "BASELINE config 5: african_head (views 0-%d) + diablo (views %d-%d), specular, 7680x4320, camera angle 2*pi*j/%d, "
                     "light 0.7 rad, per-view checksums on; rank r renders views [r n/N, (r+1) n/N) of each model"
                     % (n_per_model - 1, n_per_model, VIEWS_8K - 1, n_per_model)}
    if rank == 0:
        peak, _ = load_peaks()
        cov = {}
        for name, s in scenes.items():
            v0 = views_8k(0, 1, n_per_model)[0]
            s.clear(); s.set_light_direction(v0[9:12]); s.set_camera(v0[0:3], v0[3:6], v0[6:9]); s.render()
            cov[name] = int((s.raw_visibility() != 0xFFFFFFFF).sum())
        tri = {"african_head": len(head.tri_idx), "diablo": len(model.tri_idx)}
        nbytes = sum(algorithmic_bytes(n_px, cov[k], tri[k], "specular")["frame"] for k in scenes) / 2
        us = out["us_per_view"]                  # device time per view on one GPU (each rank renders its own block)
        out["roofline"] = {"bound": "hbm", "algorithmic_bytes_per_view": nbytes, "us_per_view_per_gpu": us,
                           "achieved": nbytes / (us * 1e-6) / 1e9, "peak": peak,
                           "unit": "GB/s", "frac": nbytes / (us * 1e-6) / 1e9 / peak, "covered_px_view0": cov,
                           "note": "overlapped view batch (k_raster<3,128>); bytes = 19 P + 3 maps x min(3 C, 3 MiB) + geometry, C of view 0"}
        out["covered_mpix_per_s"] = out["value"] * (sum(cov.values()) / 2) / n_px
    for s in scenes.values():
        s.close()
    return out


def run_shadow4k(rank, world, dev, model, dist, tstream, steps):
    """BASELINE config 3: diablo, shadow pipeline (light depth pass + phong shade), 3840x2160, the 64 turntable views under
    light angles 0.7 and 2.1 rad: one step = 128 frames through tr_scene_render_views."""
    import torch
    from tiny_renderer_b200.scene import Scene
    s = Scene(W4K, H4K, model, "shadow", device=dev, ring=RING)
    ps = torch.cuda.Stream(device=dev)
    s.set_stream(ps.cuda_stream)
    views = np.concatenate([view_batch(rank=rank, light_angle=0.7), view_batch(rank=rank, light_angle=2.1)])
    s.render_views(views[:16], want_checksums=False)
    ms = timed_batches(torch, dist, tstream, [ps], lambda: s.render_views(views, want_checksums=False), steps)
    ms = max_over_ranks(torch, dist, ms)
    n_px = W4K * H4K
    out = {"metric": METRIC, "value": world * steps * len(views) * n_px / (ms * 1e-3) / 1e6, "unit": "Mpixels/s", "n_gpus": world,
           "scaling": "weak", "frames_per_step_per_gpu": len(views), "steps": steps, "us_per_frame": ms * 1e3 / (steps * len(views)),
           "config": "BASELINE config 3: diablo, shadow, 3840x2160, 64 camera angles x light angles {0.7, 2.1} rad"}
    if rank == 0:
        peak, _ = load_peaks()
        lone = lone_frame_us(s, views[[8, 72]], reps=4)
        diag = s.diagnostics()
        covered = int((s.raw_visibility() != 0xFFFFFFFF).sum())
        nbytes = algorithmic_bytes(n_px, covered, len(model.tri_idx), "shadow")["frame"]
        frame_us = sum(lone.get(k, 0.0) for k in ("K1_vertex_setup", "K3_raster_visibility", "K4_shadow_depth"))
        if "K1_vertex_setup" in lone:
            frame_us += lone["K1_vertex_setup"]          # two set-up launches per frame (light, camera); the figure is per launch
        out["lone_frame"] = {"kernel_us": lone, "us": frame_us, "algorithmic_bytes": nbytes,
                             "frac": nbytes / (frame_us * 1e-6) / 1e9 / peak if frame_us else None,
                             "note": "K1(light) + K4 + K1(camera) + K3/K5, 256-thread CTA shape, one frame at a time"}
        out["roofline"] = {"bound": "hbm", "algorithmic_bytes_per_frame": nbytes, "achieved": nbytes / (out["us_per_frame"] * 1e-6) / 1e9,
                           "peak": peak, "unit": "GB/s", "frac": nbytes / (out["us_per_frame"] * 1e-6) / 1e9 / peak,
                           "note": "overlapped batch (k_raster<-2,128> + k_raster<5,128>); bytes = 27 P + min(3 C, 3 MiB) + geometry of both passes"}
        b2b = back_to_back_us(s, views[8])
        s.set_debug_flags(4)
        b2b_all = back_to_back_us(s, views[8])
        s.set_debug_flags(0)
        out["single_view"] = {"us_per_frame": b2b, "us_per_frame_every_pass_run": b2b_all, "frames": 100, "reps": 5,
                              "frac": nbytes / (b2b * 1e-6) / 1e9 / peak, "frac_every_pass_run": nbytes / (b2b_all * 1e-6) / 1e9 / peak,
                              "note": "100 back-to-back lone frames of one view under one light, events around all of them, median of 5: frames 2.. "
                                      "re-use the shadow buffer (no light set-up, no depth raster), the camera pass's set-up output and a "
                                      "cost-sorted tile order; `*_every_pass_run` (debug bit 2): K1 + K4 + K1 + K3/K5 every frame, static order"}
        out["shadow_index_oob"] = diag["shadow_index_oob"]
    s.close()
    return out


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


def run_e2e(torch, dist, rank, world, dev, scenes, views, steps):
    """The reference-facing call with HOST destinations: tr_scene_render_views(views, frames -> host), per frame
    clear / set_light_direction / set_camera / render + delivery of the top-down RGB8 frame, all inside the timed region
    (wall clock around the calls, barrier + synchronize on both sides).  Each scene delivers into E2E_HOST_FRAMES pinned
    host frames (tr_host_alloc), 16 views per call, from a host thread of its own; a host frame receives a different view
    every time it comes round, so the damage tracking of the sparse delivery (k_host.cuh) is exercised, not bypassed."""
    import tiny_renderer_b200 as tr
    n_px = W4K * H4K
    host = {p: tr.pinned_empty((E2E_HOST_FRAMES, H4K, W4K, 3), np.uint8) for p in PIPES}
    chunks = [views[i:i + E2E_HOST_FRAMES] for i in range(0, len(views), E2E_HOST_FRAMES)]

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def e2e_step():
        def job(p):
            for c in chunks:
                scenes[p].render_views(c, want_checksums=False, frames_out=host[p][:len(c)])
        ts = [threading.Thread(target=job, args=(p,)) for p in PIPES]
        [t.start() for t in ts]
        [t.join() for t in ts]

    e2e_step()
    for s in scenes.values():
        s.transfer_stats(reset=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    barrier()
    secs = max_over_ranks(torch, dist, time.perf_counter() - t0)
    st = [s.transfer_stats() for s in scenes.values()]
    frames = steps * len(PIPES) * len(views)
    d2h = sum(x["sparse_bytes"] + x["copied_bytes"] for x in st)
    # exactness spot check, outside the timed region: the last delivered host frames against a plain full copy
    exact = True
    for p in PIPES:
        last = chunks[-1]
        for k in (0, len(last) - 1):
            v = last[k]
            s = scenes[p]
            s.clear(); s.set_light_direction(v[9:12]); s.set_camera(v[0:3], v[3:6], v[6:9]); s.render()
            exact = exact and bool(np.array_equal(host[p][k], s.get_frame_buffer()))
    # What bounds e2e: the device->host link.  A plain cudaMemcpyAsync into pinned memory, all ranks at once, is the yardstick.
    nbytes = 16 * n_px * 3
    src = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dst = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    dst.copy_(src, non_blocking=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    copy_s = max_over_ranks(torch, dist, (time.perf_counter() - t0) / 3)
    del src, dst
    d2h_all = sum_over_ranks(torch, dist, float(d2h))
    link = world * nbytes / copy_s / 1e9
    ach = d2h_all / secs / 1e9
    return {"value": world * frames * n_px / secs / 1e6, "unit": "Mpixels/s", "h2d_bytes_per_step": len(PIPES) * len(views) * 48,
            "d2h_bytes_per_step": d2h / steps, "d2h_bytes_per_frame": d2h / frames, "full_frame_bytes": n_px * 3,
            "frames_sparse": sum(x["frames_sparse"] for x in st), "frames_copied": sum(x["frames_copied"] for x in st),
            "frames_exact": exact, "steps": steps, "frames_per_step_per_gpu": len(PIPES) * len(views), "seconds": secs,
            "note": "tr_scene_render_views(views, frames -> tr_host_alloc memory): the device stores only tile rows that are non-zero "
                    "now or were non-zero after the previous delivery to the same host frame (k_frame_to_host); the caller reads "
                    "complete, exact top-down RGB8 frames (frames_exact: spot check against tr_scene_get_frame_buffer into ordinary "
                    "memory, outside the timed region); inputs are the 12 floats of camera+light per frame (kernel arguments)",
            "roofline": {"bound": "pcie d2h", "achieved": ach, "peak": link, "unit": "GB/s", "frac": ach / link,
                         "peak_source": "plain pinned device->host copy of 16 frames, all ranks at once, measured in this run",
                         "equivalent_full_frame_GBps": world * frames * n_px * 3 / secs / 1e9}}


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
    assert tstream.cuda_stream != 0
    model, data = bench_model()
    head, _ = bench_model("african_head")
    n_px = W4K * H4K
    n_tri = len(model.tri_idx)
    views = view_batch(rank=rank)
    scenes = {p: Scene(W4K, H4K, model, p, device=dev, ring=RING) for p in PIPES}
    # The two scenes of the step are independent objects (one pipeline per Scene, as in the reference), so each is
    # driven on a stream of its own: their view batches run side by side instead of draining one before the other
    # starts.  Both streams are fenced by the timing events on the main stream (timed_batches).
    pstreams = {p: torch.cuda.Stream(device=dev) for p in PIPES}
    for p in PIPES:
        scenes[p].set_stream(pstreams[p].cuda_stream)

    def make_step(v, reps):
        def step():
            for _ in range(reps):
                for p in PIPES:
                    scenes[p].render_views(v, want_checksums=False)
        return step

    step = make_step(views, REPS)
    for _ in range(max(args.warmup, 3)):
        step()
    for s in scenes.values():
        s.kernel_times()                      # reset launch counters
    sampler = ClockSampler(dev)          # every rank samples its own GPU: `value` is set by the slowest one
    sampler.start()
    ms = timed_batches(torch, dist, tstream, list(pstreams.values()), step, args.steps, preroll=PREROLL_STEPS)
    clocks = sampler.stop()
    launches = sum(n for s in scenes.values() for (_, n) in s.kernel_times().values())
    launches = launches * args.steps // (args.steps + PREROLL_STEPS)         # the pre-roll step is outside the events
    if dist is not None:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, {"rank": rank, "ms_per_step": ms / args.steps, "sm_mhz": clocks["sm_mhz"], "reasons": clocks["reasons"]})
        ms = max_over_ranks(torch, dist, ms)
        # the line's clocks describe the whole job: lowest median clock and every throttle reason any rank saw
        mhz = [r["sm_mhz"] for r in per_rank if r["sm_mhz"]]
        clocks = dict(clocks, sm_mhz=min(mhz) if mhz else None, reasons=sorted({x for r in per_rank for x in r["reasons"]}), per_rank=per_rank)
    frames_per_step = len(PIPES) * N_ANGLES * REPS
    value = world * args.steps * frames_per_step * n_px / (ms * 1e-3) / 1e6

    # the r1 workload once, for comparison: 8 angles on 8 slots, every slot re-renders the same frame (no tile-state flips)
    v8 = view_batch(8, rank)
    step8 = make_step(v8, 8 * REPS)
    step8()
    ms8 = max_over_ranks(torch, dist, timed_batches(torch, dist, tstream, list(pstreams.values()), step8, 3))
    aliased = world * 3 * len(PIPES) * 8 * 8 * REPS * n_px / (ms8 * 1e-3) / 1e6

    # each pipeline's batch on its own (the same overlapped 128-thread kernels, one scene at a time): the sustained time per
    # launch of each tile kernel, which `roofline` quotes for the dominant one
    alone_us = {}
    for p in PIPES:
        def one(p=p):
            for _ in range(REPS):
                scenes[p].render_views(views, want_checksums=False)
        t = max_over_ranks(torch, dist, timed_batches(torch, dist, tstream, [pstreams[p]], one, 3))
        alone_us[p] = t * 1e3 / (3 * REPS * N_ANGLES)

    e2e = run_e2e(torch, dist, rank, world, dev, scenes, views, max(2, min(args.steps, 10)))
    e2e["host_numa_node"] = numa_node

    line = {"metric": METRIC, "value": value, "unit": "Mpixels/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": data, "config": workload_config(),
            "timed_region_ms": ms, "aliased_8_angles": {"value": aliased, "unit": "Mpixels/s",
                                                        "note": "r1's workload: 8 angles on 8 frame slots, each slot re-renders the identical frame"},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}

    if rank == 0:
        # ---- coverage of the 64 views (light-independent; identical for both pipelines) ----
        s = scenes["phong"]
        cov = []
        for v in views:
            s.clear(); s.set_light_direction(v[9:12]); s.set_camera(v[0:3], v[3:6], v[6:9]); s.render()
            cov.append(int((s.raw_visibility() != 0xFFFFFFFF).sum()))
        mean_cov = float(np.mean(cov))
        line["covered_mpix_per_s"] = value * mean_cov / n_px
        line["covered_fraction"] = mean_cov / n_px
        # ---- roofline of the dominant kernel: per-kernel CUDA events (profiling mode), separate from `value` ----
        peak, peak_src = load_peaks()
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        sm_mhz = clocks.get("sm_mhz")
        sample = views[::8]                                  # 8 of the 64 angles
        per = {}
        for p in PIPES:
            s = scenes[p]
            per[p] = {"256": lone_frame_us(s, sample)}
            s.set_debug_flags(2)                             # lone frames in the batch CTA shape (128 threads)
            per[p]["128"] = lone_frame_us(s, sample)
            s.set_debug_flags(0)
            per[p]["256_settled"] = settled_frame_us(s, sample)
        line["kernel_us"] = per
        b2b = {p: back_to_back_us(scenes[p], views[8]) for p in PIPES}
        b2b_all = {}
        for p in PIPES:                                      # the same with every pass run every frame (debug bit 2)
            scenes[p].set_debug_flags(4)
            b2b_all[p] = back_to_back_us(scenes[p], views[8])
            scenes[p].set_debug_flags(0)
        cov8 = float(np.mean(cov[::8]))
        ab = {p: algorithmic_bytes(n_px, cov8, n_tri, p) for p in PIPES}
        dom = "darboux"                                      # the slower pipeline: its tile kernel dominates the step
        ab_v8 = {p: algorithmic_bytes(n_px, float(cov[8]), n_tri, p)["frame"] for p in PIPES}
        line["config2_single_view"] = {"us_per_frame": b2b, "mpix_per_s": {p: n_px / b2b[p] for p in PIPES}, "frames": 100, "reps": 5,
                                       "us_per_frame_every_pass_run": b2b_all,
                                       "frac_every_pass_run": {p: ab_v8[p] / (b2b_all[p] * 1e-6) / 1e9 / peak for p in PIPES},
                                       "algorithmic_bytes": ab_v8,
                                       "achieved_GBps": {p: ab_v8[p] / (b2b[p] * 1e-6) / 1e9 for p in PIPES},
                                       "frac": {p: ab_v8[p] / (b2b[p] * 1e-6) / 1e9 / peak for p in PIPES},
                                       "note": "SURVEY.md 8(d) config 2 as defined: 100 back-to-back lone frames of one view (camera 2*pi/8, light 0.7) per "
                                               "pipeline, events around all of them, median of 5; no overlap between frames; frames 2.. of an unchanged view "
                                               "re-use the vertex set-up output and a cost-sorted tile order (tinyr.cu run_setup), "
                                               "`*_every_pass_run` is the same with K1 + K3/K5 run every frame in the static order; "
                                               "frac = byte model of that view / time against the measured HBM peak"}

        def kernel_fig(p, shape):
            us = per[p][shape].get("K3_raster_visibility", 0.0)
            nb = ab[p]["K3_raster_visibility"]
            e = ncu_kernel_summary(raster_kernel(p, shape))
            return {"kernel": "k_raster<%s,%s>" % (p, shape), "cta_threads": int(shape), "us_per_launch": us, "algorithmic_bytes_per_launch": nb,
                    "achieved": nb / (us * 1e-6) / 1e9 if us else None, "frac": nb / (us * 1e-6) / 1e9 / peak if us else None,
                    "traffic": e["dram_bytes_per_launch"] if e else None,
                    "issue": issue_roofline([raster_kernel(p, shape)], us, n_sm, sm_mhz) if us else None}

        def frame_fig(p, shape):
            us = sum(per[p][shape].get(k, 0.0) for k in ("K1_vertex_setup", "K3_raster_visibility"))
            return {"cta_threads": int(shape.split("_")[0]), "algorithmic_bytes": ab[p]["frame"], "us": us,
                    "achieved": ab[p]["frame"] / (us * 1e-6) / 1e9, "frac": ab[p]["frame"] / (us * 1e-6) / 1e9 / peak}

        k = kernel_fig(dom, "128")
        sustained = {p: {"kernel": raster_kernel(p, 128), "us_per_frame": alone_us[p], "mpix_per_s": n_px / alone_us[p],
                         "algorithmic_bytes_per_launch": ab[p]["K3_raster_visibility"],
                         "achieved": ab[p]["K3_raster_visibility"] / (alone_us[p] * 1e-6) / 1e9,
                         "frac": ab[p]["K3_raster_visibility"] / (alone_us[p] * 1e-6) / 1e9 / peak} for p in PIPES}
        batch_bytes = sum(ab[p]["frame"] for p in PIPES) / len(PIPES)
        us_frame = ms / args.steps / frames_per_step * 1e3
        table, tsrc = ncu_summaries()
        steady = table.get("batch_steady_state") if isinstance(table, dict) else None
        line["roofline"] = {
            "bound": "hbm", "kernel": k["kernel"] + " = K3 raster_visibility + K5 shade fused, the kernel and CTA shape that dominate the timed region",
            "achieved": sustained[dom]["achieved"], "peak": peak, "unit": "GB/s", "frac": sustained[dom]["frac"], "peak_source": peak_src,
            "how": "sustained: the darboux scene's view batch on its own (64-angle turntable x %d turns x 3 steps, frames overlapped on the %d "
                   "slot streams exactly as in the timed region), CUDA events on the launching stream, time / launches of this kernel "
                   "(the vertex set-up launches in between are inside that time); `alone` is the same kernel timed one launch at a time"
                   % (REPS, RING),
            "us_per_launch_sustained": alone_us[dom], "sustained": sustained,
            "note": "time-to-solution against SURVEY.md 8(d)'s byte model (19 P + texels + geometry), not a DRAM utilisation: the tile "
                    "kernel never writes the keys of untouched tiles nor rewrites zero tiles, so it moves far fewer bytes than the model "
                    "(`traffic`) and a light pipeline can exceed 1.0 (phong); what binds it is instruction issue (`issue`)",
            "alone": {"achieved": k["achieved"], "frac": k["frac"], "us_per_launch": k["us_per_launch"],
                      "note": "one frame at a time on one stream (profiling mode): a lone launch leaves SMs idle behind its heaviest tiles"},
            "traffic": (steady or {}).get("dram_bytes_per_frame", k["traffic"]),
            "traffic_source": ("%s: dram__bytes summed over %s consecutive tile-kernel launches of the de-aliased overlapped batch / frames"
                               % (tsrc, (steady or {}).get("launches"))) if steady else "per-launch ncu capture of this kernel (serialised)",
            "algorithmic_bytes_per_launch": k["algorithmic_bytes_per_launch"], "us_per_launch": k["us_per_launch"], "issue": k["issue"],
            "kernels": {p + "," + sh: kernel_fig(p, sh) for p in PIPES for sh in ("128", "256")},
            "frame_serial": dict(frame_fig(dom, "256_settled"), cta_threads=256,
                                 note="lone darboux frame K1 + K3/K5, 256-thread shape (what tr_scene_render launches), view settled as in the "
                                      "reference's app loop (< 0.1 degree per frame): tiles launched in the order of their cost in the previous frame"),
            "frame_serial_phong": dict(frame_fig("phong", "256_settled"), cta_threads=256),
            "frame_serial_new_view": dict(frame_fig(dom, "256"), note="the same with every frame 45 degrees from the last: static centre-out launch order"),
            "frame_serial_new_view_phong": frame_fig("phong", "256"),
            "batch": {"algorithmic_bytes_per_frame": batch_bytes, "us_per_frame": us_frame, "achieved": batch_bytes / (us_frame * 1e-6) / 1e9,
                      "frac": batch_bytes / (us_frame * 1e-6) / 1e9 / peak,
                      "note": "the timed region: both pipelines' frames overlapped, 128-thread shape",
                      "issue": issue_roofline([raster_kernel("phong", 128), raster_kernel("darboux", 128), setup_kernel("phong"), setup_kernel("darboux")],
                                              2 * us_frame, n_sm, sm_mhz)}}
        if world == 1 and not QUICK:
            # what a rasteriser resolving visibility with global 64-bit atomicMin could reach on this box
            rate = tr.bench_atomic_min(W4K, H4K, rounds=4, repeats=3, device=dev)
            line["atomic_min_peak"] = {"atomics_per_s": rate, "equivalent_GB_per_s": rate * 8 / 1e9,
                                       "us_per_4k_depth_layer": n_px / rate * 1e6,
                                       "note": "micro-benchmark only (tr_bench_atomic_min); not on the render path"}
    for s in scenes.values():
        s.close()
    sm_mhz0 = (clocks or {}).get("sm_mhz") if rank == 0 else None
    line["shadow4k"] = run_shadow4k(rank, world, dev, model, dist, tstream, max(2, min(args.steps, 10)))
    line["ao"] = run_ao(args, rank, world, dev, model, dist, tstream, sm_mhz=sm_mhz0)
    line["views8k"] = run_views8k(rank, world, dev, model, head, dist, tstream)
    if rank == 0 and world == 1 and not QUICK:
        cpu = cpu_baselines(model, head)
        line["cpu_baseline"] = cpu["main"]
        line["shadow4k"]["cpu_baseline"] = cpu["shadow4k"]
        line["ao"]["cpu_baseline"] = cpu["ao"]
        line["views8k"]["cpu_baseline"] = cpu["views8k"]
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
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

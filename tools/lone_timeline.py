"""Per-CTA timeline of the tile kernel of a LONE 4K frame (diagnostic build: `make -C tiny_renderer_b200/csrc
EXTRA=-DTR_TIMELINE`; tools/lone_timeline.sh builds it on the box, runs this, and restores the product build).
Shows where a lone frame's time goes: when CTAs pass the wait for the vertex set-up kernel, how long the polygon filter
and the raster + shade phases of the touched tiles take, how many heavy tiles each SM ends up with and how many CTAs are
still busy near the end (the tail).  Run under gpurun."""
import ctypes
import json
import math
import os
import statistics
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import assets  # noqa: E402
import parity  # noqa: E402
import tiny_renderer_b200 as tr  # noqa: E402
from tiny_renderer_b200.scene import Scene  # noqa: E402

W, H = 3840, 2160
WORDS = 8
L = tr.lib()
L.tr_debug_timeline.argtypes = [ctypes.c_void_p, ctypes.c_uint32]
L.tr_debug_timeline.restype = ctypes.c_int


def pct(a, q):
    return float(np.percentile(a, q)) if len(a) else None


for pipe in sys.argv[1:] or ["phong", "darboux"]:
    model = assets.get_model("diablo")
    lf, la, up, l = assets.view_vectors(2 * math.pi / 8, 0.7)
    g = Scene(W, H, model, pipe)
    if os.environ.get("TINYR_DEBUG_FLAGS"):
        g.set_debug_flags(int(os.environ["TINYR_DEBUG_FLAGS"]))
    n_tiles = ((W + 63) // 64) * ((H + 31) // 32)
    times = []
    for _ in range(8):
        g.timer_begin()
        parity.drive(g, lf, la, up, l)
        times.append(g.timer_end())
    g.synchronize()
    tl = np.zeros((n_tiles, WORDS), np.uint64)
    assert L.tr_debug_timeline(tl.ctypes.data, n_tiles) == 0
    ts = np.zeros(512, np.uint64)
    if hasattr(L, "tr_debug_timeline_setup"):
        L.tr_debug_timeline_setup.argtypes = [ctypes.c_void_p]
        L.tr_debug_timeline_setup(ts.ctypes.data)
    ph = np.zeros(1024, np.uint64)
    if hasattr(L, "tr_debug_timeline_phase"):
        L.tr_debug_timeline_phase.argtypes = [ctypes.c_void_p]
        L.tr_debug_timeline_phase(ph.ctypes.data)
    g.close()
    if os.environ.get("TIMELINE_DUMP"):
        np.save(os.path.join(os.environ["TIMELINE_DUMP"], "timeline_%s.npy" % pipe), tl)
    t = tl.astype(np.int64)
    t0 = t[:, 0].min()
    entry, waited, filt, end, sm, n_list = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3, (t[:, 2] - t0) / 1e3, (t[:, 3] - t0) / 1e3, t[:, 4], t[:, 5]
    touched = n_list > 0
    dur = end - waited                    # a tile's own time, from the moment the set-up kernel's data is there
    out = {"pipeline": pipe, "frame_us_events_median": round(statistics.median(times) * 1e3, 1), "ctas": int(n_tiles),
           "touched_tiles": int(touched.sum()), "kernel_span_us": round(float(end.max()), 1),
           "wait_released_us": round(float(waited.min()), 1),
           "ctas_resident_before_release": int((entry < waited.min()).sum()),
           "after_release_us": round(float(end.max() - waited.min()), 1)}
    k1 = ts.astype(np.int64).reshape(256, 2)
    used = k1[:255, 0] > 0
    if used.any():      # the vertex set-up kernel in front, on the tile kernel's clock (us, 0 = first tile CTA's entry)
        out["setup_kernel_us"] = {"blocks": int(used.sum()), "first_entry": round(float((k1[:255, 0][used].min() - t0) / 1e3), 1),
                                  "last_vertex_block_exit": round(float((k1[:255, 1][used].max() - t0) / 1e3), 1),
                                  "sort_block_entry_exit": [round(float((k1[255, 0] - t0) / 1e3), 1), round(float((k1[255, 1] - t0) / 1e3), 1)] if k1[255, 0] > 0 else None}
    if used.any() and ph.any():     # thread 0 of every vertex block: us after the block's entry
        p4 = ph.astype(np.int64).reshape(256, 4)[:255][used]
        rel = (p4 - k1[:255, 0][used][:, None]) / 1e3
        ent = (k1[:255, 0][used] - k1[:255, 0][used].min()) / 1e3
        life = (k1[:255, 1][used] - k1[:255, 0][used]) / 1e3
        out["setup_blocks_entry_us_min_med_max"] = [round(float(ent.min()), 2), round(float(np.median(ent)), 2), round(float(ent.max()), 2)]
        out["setup_blocks_life_us_min_med_max"] = [round(float(life.min()), 2), round(float(np.median(life)), 2), round(float(life.max()), 2)]
        slow = np.argsort(-life)[:4]
        out["setup_slowest_blocks"] = [{"block": int(b), "entry": round(float(ent[b]), 2), "life": round(float(life[b]), 2),
                                        "phases": [round(float(x), 2) for x in rel[b]]} for b in slow]
        if k1[255, 0] > 0:
            sp = (ph.astype(np.int64).reshape(256, 4)[255] - k1[255, 0]) / 1e3
            out["sort_block_phases_us"] = {"zeroed_and_loads_issued": round(float(sp[0]), 2), "counted": round(float(sp[1]), 2), "prefix": round(float(sp[2]), 2),
                                           "scattered": round(float(sp[3]), 2), "exit": round(float((k1[255, 1] - k1[255, 0]) / 1e3), 2)}
        out["setup_block_phases_us_median"] = {"geometry_loaded": round(float(np.median(rel[:, 0])), 2), "transformed": round(float(np.median(rel[:, 1])), 2),
                                               "binned": round(float(np.median(rel[:, 2])), 2), "records_issued": round(float(np.median(rel[:, 3])), 2),
                                               "exit": round(float(np.median((k1[:255, 1][used] - k1[:255, 0][used]) / 1e3)), 2)}
    d = dur[touched]
    out["touched_tile_us"] = {"sum": round(float(d.sum()), 0), "p10": round(pct(d, 10), 1), "p50": round(pct(d, 50), 1), "p90": round(pct(d, 90), 1),
                              "max": round(float(d.max()), 1)}
    f = (filt - waited)[touched]
    out["filter_phase_us"] = {"p50": round(pct(f, 50), 1), "p90": round(pct(f, 90), 1), "max": round(float(f.max()), 1)}
    e = dur[~touched]
    out["untouched_cta_us"] = {"n": int((~touched).sum()), "p50": round(pct(e, 50), 2), "p90": round(pct(e, 90), 2), "sum": round(float(e.sum()), 0)}
    # per SM: touched tiles, their summed time, the time its last touched tile ends
    sms = np.unique(sm)
    per = [(int((touched & (sm == s)).sum()), float(dur[touched & (sm == s)].sum()), float(end[touched & (sm == s)].max()) if (touched & (sm == s)).any() else 0.0) for s in sms]
    cnt, work, last = np.array([p[0] for p in per]), np.array([p[1] for p in per]), np.array([p[2] for p in per])
    out["per_sm"] = {"sms": int(len(sms)), "touched_tiles_min_med_max": [int(cnt.min()), float(np.median(cnt)), int(cnt.max())],
                     "tile_time_sum_us_min_med_max": [round(float(work.min()), 0), round(float(np.median(work)), 0), round(float(work.max()), 0)],
                     "last_touched_end_us_p10_p50_p90_max": [round(pct(last, 10), 1), round(pct(last, 50), 1), round(pct(last, 90), 1), round(float(last.max()), 1)]}
    # touched CTAs in flight over time (4 us bins from the release of the wait)
    r = float(waited.min())
    bins = np.arange(r, float(end.max()) + 4.0, 4.0)
    out["touched_in_flight"] = [[round(float(b - r), 0), int(((waited <= b) & (end > b) & touched).sum()), int(len(np.unique(sm[(waited <= b) & (end > b) & touched])))] for b in bins]
    late = np.argsort(-end)[:8]
    out["latest_tiles"] = [{"pos": int(i), "start_us": round(float(waited[i]), 1), "end_us": round(float(end[i]), 1), "polys": int(n_list[i]), "sm": int(sm[i])} for i in late]
    print(json.dumps(out))

"""Turns `ncu --set full` reports into the summaries committed under profiles/.

  python tools/ncu_summary.py <tag> gpurun_out/<tag>_full.ncu-rep gpurun_out/<tag>_full_ao.ncu-rep

writes profiles/<tag>_ncu_full.csv (one row per captured launch, the raw page filtered to the roofline / occupancy /
issue / stall metrics) and profiles/<tag>_traffic.json (per kernel: launches, DRAM bytes per launch =
dram__bytes_read.sum + dram__bytes_write.sum, duration under ncu, warp instructions).  bench.py reads
`roofline.traffic` from the newest *_traffic.json.  Also prints a launch-list share table when
gpurun_out/<tag>_launches.csv exists, and — when gpurun_out/<tag>_steady.csv exists (dram__bytes per launch over a long
run of consecutive tile-kernel launches of the de-aliased, ring-rotating view batch, `--cache-control none`, so L2
contents carry over from launch to launch as they do in the timed region) — adds `batch_steady_state`: DRAM bytes per
frame in steady state.
"""
import collections
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KEEP = re.compile(
    r"^(dram__bytes_(read|write)\.sum|gpu__dram_throughput|gpu__time_duration\.sum|l1tex__throughput|lts__throughput|"
    r"lts__t_sector_hit_rate|launch__(block_size|grid_size|occupancy|registers_per_thread|waves_per_multiprocessor|shared_mem_per_block)|"
    r"sm__throughput|sm__warps_active|sm__inst_executed\.sum$|sm__inst_executed\.avg\.per_cycle|smsp__inst_executed\.sum$|"
    r"smsp__issue_active|smsp__inst_issued|sm__inst_issued|smsp__warp_issue_stalled.*_per_warp_active|"
    r"smsp__average_warp.*_per_issue_active|smsp__thread_inst_executed_per_inst_executed|sm__cycles_active\.avg$|"
    r"sm__cycles_elapsed\.max$|smsp__cycles_active\.avg$|sm__pipe_.*pct_of_peak_sustained_active|sm__inst_executed_pipe_.*pct)")


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    body = [r for r in rows[2:] if len(r) == len(hdr)]
    return hdr, dict(zip(hdr, rows[1])), body


TIME_US = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    tag, reps = sys.argv[1], sys.argv[2:]
    cols, table = None, []
    per = collections.OrderedDict()
    units = {}
    for rep in reps:
        hdr, un, body = raw_rows(rep)
        units = units or un
        rd, wr, tm = (un.get(c, "") for c in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
        idx = {h: i for i, h in enumerate(hdr)}
        keep = [h for h in hdr if KEEP.match(h)]
        if cols is None:
            cols = keep
        for r in body:
            name = r[idx["Kernel Name"]]
            table.append([name] + [(r[idx[c]] + ("" if un.get(c, "") == units.get(c, "") else " " + un.get(c, ""))) if c in idx else "" for c in cols])
            d = per.setdefault(name, {"launches": 0, "dram": 0.0, "us": 0.0, "inst": 0.0})
            d["launches"] += 1
            d["dram"] += (num(r[idx["dram__bytes_read.sum"]]) or 0) * BYTES[rd] + (num(r[idx["dram__bytes_write.sum"]]) or 0) * BYTES[wr]
            d["us"] += (num(r[idx["gpu__time_duration.sum"]]) or 0) * TIME_US[tm]
            key = "smsp__inst_executed.sum" if "smsp__inst_executed.sum" in idx else "sm__inst_executed.sum"
            d["inst"] += (num(r[idx[key]]) or 0) if key in idx else 0
    with open(os.path.join(ROOT, "profiles", f"{tag}_ncu_full.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["Kernel Name"] + cols)
        w.writerow([""] + [units.get(c, "") for c in cols])
        w.writerows(table)
    traffic = {}
    for name, d in per.items():
        n = d["launches"]
        traffic[name] = {"launches": n, "dram_bytes_per_launch": d["dram"] / n,
                         "us_per_launch_under_ncu": d["us"] / n, "warp_instructions": d["inst"] / n}
    steady = os.path.join(ROOT, "gpurun_out", f"{tag}_steady.csv")
    if os.path.exists(steady):
        lines = [ln for ln in open(steady, errors="replace") if ln.startswith('"')]
        per_id = collections.OrderedDict()
        for r in csv.DictReader(io.StringIO("".join(lines))):
            m = r.get("Metric Name", "")
            v = num(r.get("Metric Value", "")) or 0.0
            e = per_id.setdefault(r["ID"], {"name": r["Kernel Name"], "dram": 0.0, "us": 0.0})
            if m.startswith("dram__bytes"):
                e["dram"] += v * BYTES.get(r.get("Metric Unit", "byte"), 1.0)
            elif m == "gpu__time_duration.sum":
                e["us"] += v * TIME_US.get(r.get("Metric Unit", "us"), 1.0)
        rows = list(per_id.values())
        if rows:
            names = collections.Counter(e["name"].split("(")[0] for e in rows)
            traffic["batch_steady_state"] = {
                "launches": len(rows), "frames": len(rows), "kernels": dict(names),
                "dram_bytes_per_frame": sum(e["dram"] for e in rows) / len(rows),
                "dram_bytes_min_max": [min(e["dram"] for e in rows), max(e["dram"] for e in rows)],
                "us_per_launch_under_ncu": sum(e["us"] for e in rows) / len(rows),
                "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --cache-control none "
                       "--clock-control none -k regex:k_raster over consecutive tile-kernel launches of the 64-angle view batch "
                       "(8 frame slots per scene; one launch = one frame)"}
    with open(os.path.join(ROOT, "profiles", f"{tag}_traffic.json"), "w") as f:
        json.dump(traffic, f, indent=1)
    print(json.dumps(traffic, indent=1))

    launches = os.path.join(ROOT, "gpurun_out", f"{tag}_launches.csv")
    if os.path.exists(launches):
        agg = collections.OrderedDict()
        lines = [ln for ln in open(launches, errors="replace") if ln.startswith('"')]
        rd = csv.DictReader(io.StringIO("".join(lines)))
        for r in rd:
            if r.get("Metric Name") != "gpu__time_duration.sum":
                continue
            v = num(r["Metric Value"]) or 0.0
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r.get("Metric Unit", "us"), 1.0)
            a = agg.setdefault(r["Kernel Name"], [0, 0.0])
            a[0] += 1
            a[1] += v
        total = sum(a[1] for a in agg.values()) or 1.0
        print("| kernel | launches | avg us (under ncu) | share |\n|---|---|---|---|")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print(f"| `{k[:70]}` | {n} | {t / n:.1f} | {100 * t / total:.1f} % |")


if __name__ == "__main__":
    main()

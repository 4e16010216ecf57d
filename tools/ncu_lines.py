#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares of one kernel from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_lines.py report.ncu-rep <kernel-id> [top]   (ncu -i ... --page source --print-source cuda,sass)"""
import csv, subprocess, sys
rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-id", ":::" + kid],
                     capture_output=True, text=True).stdout
cur, agg, srcs = None, {}, {}
ie = iss = None
for r in csv.reader(out.splitlines()):
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": print(r[1][:100]); continue
    if r[0] == "Line No": ie = r.index("Instructions Executed"); iss = r.index("Warp Stall Sampling (All Samples)"); continue
    try: ln, ex, st = int(r[0]), int(r[ie]), int(r[iss])
    except (ValueError, TypeError, IndexError): continue
    a = agg.setdefault((cur, ln), [0, 0, 0]); a[0] += ex; a[1] += st; a[2] += 1
    srcs[(cur, ln)] = r[1]
tot = sum(a[0] for a in agg.values()) or 1; tots = sum(a[1] for a in agg.values()) or 1
print("warp instructions %d, stall samples %d" % (tot, tots))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-14s %4d %5.1f%% instr %5.1f%% samples sass=%3d | %s" % (k[0], k[1], a[0] / tot * 100, a[1] / tots * 100, a[2], srcs[k].strip()[:100]))

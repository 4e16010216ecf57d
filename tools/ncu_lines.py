#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares of one kernel from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_lines.py report.ncu-rep <kernel name regex>[:nth] [top]   (ncu -i ... --page source --print-source cuda,sass)"""
import csv
import subprocess
import sys

rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
kid = kid if kid.rsplit(":", 1)[-1].isdigit() else kid + ":1"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-id", "::regex:" + kid],
                     capture_output=True, text=True).stdout
cur, agg, srcs = None, {}, {}
ie = iss = None
line = None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        print(r[1][:110])
        continue
    if r[0] == "Line No":
        ie = r.index("Instructions Executed")
        iss = r.index("Warp Stall Sampling (All Samples)")
        continue
    if ie is None or len(r) <= ie:
        continue
    if r[0].isdigit():                  # a source line (its own totals are repeated in the SASS rows below it)
        line = (cur, int(r[0]))
        srcs[line] = r[1]
        continue
    if r[0] == "" and line is not None and r[ie].isdigit():     # a SASS row belonging to `line`
        a = agg.setdefault(line, [0, 0, 0])
        a[0] += int(r[ie])
        a[1] += int(r[iss]) if r[iss].isdigit() else 0
        a[2] += 1
tot = sum(a[0] for a in agg.values()) or 1
tots = sum(a[1] for a in agg.values()) or 1
print("warp instructions %d, stall samples %d" % (tot, tots))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-14s %4d %5.1f%% instr %5.1f%% samples sass=%3d | %s" % (k[0], k[1], a[0] / tot * 100, a[1] / tots * 100, a[2], srcs[k].strip()[:100]))

"""Registers / spills of every kernel from the build's `ptxas -v` log (tiny_renderer_b200/csrc/ptxas_tinyr.log, git-ignored)
and the number of SASS instructions per kernel from cuobjdump, as one committed text table:

    python tools/ptxas_summary.py > profiles/<tag>_ptxas.txt
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LOG = os.path.join(ROOT, "tiny_renderer_b200", "csrc", "ptxas_tinyr.log")
LIB = os.path.join(ROOT, "tiny_renderer_b200", "libtinyr.so")


def main():
    rows, cur = [], None
    for ln in open(LOG):
        m = re.search(r"Compiling entry function '(\S+)'", ln)
        if m:
            cur = {"sym": m.group(1), "spill": None, "regs": None}
            rows.append(cur)
        elif cur and "bytes stack frame" in ln and cur["spill"] is None:
            cur["spill"] = ln.strip()
        elif cur and "Used" in ln and cur["regs"] is None:
            cur["regs"] = ln.split(":", 1)[1].strip() if ":" in ln else ln.strip()
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, sym = {}, None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            sym = m.group(1)
            counts[sym] = 0
        elif sym and re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            counts[sym] += 1
    names = subprocess.run(["c++filt"] + [r["sym"] for r in rows], capture_output=True, text=True).stdout.splitlines()
    print("# kernel | SASS instructions | ptxas: registers, barriers, shared | stack / spills")
    for n, r in zip(names, rows):
        n = re.sub(r"\(.*", "", n).replace("void ", "")
        print("%-50s | %6s | %s | %s" % (n, counts.get(r["sym"], "?"), r["regs"], r["spill"]))


if __name__ == "__main__":
    sys.exit(main())

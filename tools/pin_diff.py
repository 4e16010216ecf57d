"""Compares the output of rust/pin_oracle (the REFERENCE's Rust Scene rendering the golden cases) with the oracle's own
hashes in tests/golden/oracle_golden.json.

    python tools/pin_diff.py pinned.json

Exit status 0 = every frame / z view / shadow view of every case is byte-identical: the C++ oracle is pinned to the Rust
reference on these cases.  Otherwise the differing case and buffer are named (start with oracle/na32.h)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    if len(sys.argv) < 2:
        sys.exit(__doc__)
    pinned = json.load(open(sys.argv[1]))
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "oracle_golden.json")))
    by_case = {tuple(g["case"][:4]): g for g in golden}
    bad = 0
    for p in pinned:
        key = tuple(p["case"][:4])
        g = by_case.get(key)
        if g is None:
            print("?? case not in the golden file:", key)
            bad += 1
            continue
        for buf in ("frame_topdown", "z_gray", "shadow_gray"):
            same = p[buf] == g[buf]
            bad += 0 if same else 1
            print("%-3s %-60s %s" % ("ok" if same else "DIFF", " ".join(map(str, key)), buf))
    missing = set(by_case) - {tuple(p["case"][:4]) for p in pinned}
    for key in sorted(missing):
        print("?? case not rendered by pin_oracle:", key)
    bad += len(missing)
    print("pinned: oracle == Rust reference on %d cases" % len(pinned) if bad == 0 else "%d differences" % bad)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()

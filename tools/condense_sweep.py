#!/usr/bin/env python
"""Condense a raw tools/sweep.py CSV (thousands of rows) into a digest that is worth keeping in the tree:
per (kernel, dtype, log2n, mode) the 12 fastest shapes and the slowest one.

    python tools/condense_sweep.py profiles/r01/sweep_f64.csv [more.csv ...]   ->  <name>_top.csv next to each input
"""
import csv
import sys
from collections import defaultdict


def condense(path: str, keep: int = 12) -> str:
    rows = list(csv.DictReader(open(path)))
    groups = defaultdict(list)
    for r in rows:
        groups[(r.get("kernel"), r.get("dtype"), r.get("log2n"), r.get("mode", ""))].append(r)
    out = path[:-4] + "_top.csv"
    with open(out, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=["rank", "of"] + list(rows[0].keys()))
        w.writeheader()
        for key in sorted(groups, key=lambda k: tuple(str(x) for x in k)):
            g = sorted(groups[key], key=lambda r: -float(r["gbs"]))
            for i, r in enumerate(g[:keep]):
                w.writerow({"rank": i + 1, "of": len(g), **r})
            if len(g) > keep:
                w.writerow({"rank": len(g), "of": len(g), **g[-1]})
    return out


if __name__ == "__main__":
    for p in sys.argv[1:]:
        print(condense(p))

#!/usr/bin/env python
"""Result tooling: size-sweep CSV (tools/size_sweep.py) -> markdown report.

    python tools/report.py profiles/r01/size_sweep_2p20_2p30_f64_f32_b200_vs_ref_cuda.csv > profiles/r01/REPORT_size_sweep.md

For every (dtype, size): GB/s per kernel for b200-stream and for the reference's cuda-stream (both
driver-timed by the unmodified src/main.cpp), their ratio, and b200-stream's triad as a fraction of
each peak denominator.  Marks where the three arrays stop fitting in L2.
Also prints a block in the layout of the reference's sample results (results/v3.3/titanxp/cuda.txt).
"""
import csv
import json
import os
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KERNELS = ["copy", "mul", "add", "triad", "dot", "nstream"]
# FORMULA: what cuda-stream itself prints on the box, "3996 MHz x 7680-bit = 7672.32 GB/s PEAK"
# (src/cuda/CUDAStream.cu:107-108; this B200 SKU reports a 7680-bit bus, not 8192)
NOMINAL, FORMULA = 8000.0, 7672.3


def measured_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6551.0


def main():
    path = sys.argv[1]
    rows = list(csv.DictReader(open(path)))
    data = defaultdict(dict)
    for r in rows:
        data[(r["dtype"], int(r["log2n"]), r["model"])][r["kernel"]] = (float(r["gbs"]), float(r["min_us"]), int(r["l2_resident"]))
    peak = measured_peak()
    print(f"# Array-size sweep (BASELINE config 3), one B200, driver-timed\n")
    print(f"Source: `{os.path.relpath(path, ROOT)}`. GB/s = weight x sizeof(T) x N / min time, exactly as `src/main.cpp:197-199,319-327` "
          f"prints it. `ref` = the reference's `cuda` model (`-arch=sm_100`, its defaults) on the same GPU, runs interleaved. "
          f"Peaks: nominal {NOMINAL:.0f}, reference formula {FORMULA:.1f}, measured copy {peak:.0f} GB/s.\n")
    for dtype in ("f64", "f32"):
        sizes = sorted({k[1] for k in data if k[0] == dtype})
        if not sizes:
            continue
        print(f"## {dtype}\n")
        print("| log2 N | where | " + " | ".join(f"{k} b200 / ref (x)" for k in KERNELS) + " | triad % nominal / formula / measured |")
        print("|---|---|" + "---|" * (len(KERNELS) + 1))
        prev_res = None
        for l in sizes:
            ours, ref = data.get((dtype, l, "b200"), {}), data.get((dtype, l, "ref_cuda"), {})
            if not ours:
                continue
            res = next(iter(ours.values()))[2]
            where = "L2" if res else ("**knee -> HBM**" if prev_res else "HBM")
            prev_res = res
            cells = []
            for k in KERNELS:
                o = ours.get(k, (0, 0, 0))[0]
                r = ref.get(k, (0, 0, 0))[0]
                cells.append(f"{o:.0f} / {r:.0f} ({o / r:.2f})" if r else f"{o:.0f} / -")
            t = ours.get("triad", (0, 0, 0))[0]
            print(f"| {l} | {where} | " + " | ".join(cells) + f" | {100 * t / NOMINAL:.1f} / {100 * t / FORMULA:.1f} / {100 * t / peak:.1f} |")
        print()
    # reference-style block for the largest fp64 size
    big = max((k[1] for k in data if k[0] == "f64" and k[2] == "b200"), default=None)
    if big:
        print(f"## Sample output block, fp64, N = 2^{big} (layout of results/v3.3/titanxp/cuda.txt; MB/s)\n")
        print("```")
        print("BabelStream\nVersion: 5.0\nImplementation: B200 (sm_100a CUDA, variant of BabelStream)")
        print(f"Running kernels 30 times\nPrecision: double\nArray size: {8 * 2 ** big / 1e6:.1f} MB (={8 * 2 ** big / 1e9:.1f} GB)")
        print(f"Total size: {24 * 2 ** big / 1e6:.1f} MB (={24 * 2 ** big / 1e9:.1f} GB)")
        print("Function    MBytes/sec  Min (sec)")
        for k in ("copy", "mul", "add", "triad", "dot"):
            g, us, _ = data[("f64", big, "b200")][k]
            print(f"{k.capitalize():<12}{g * 1000:<12.3f}{us * 1e-6:<.5f}")
        print("```")


if __name__ == "__main__":
    main()

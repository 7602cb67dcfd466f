#!/usr/bin/env python
"""BASELINE config 3: array-size sweep 2^20 .. 2^30, fp32 and fp64, one B200, driver-timed.

Runs build/b200-stream (reference main.cpp + our model) and oracle/_ref/cuda-stream (the reference's
own CUDA model, -arch=sm_100) back to back at every size with identical flags and parses the
driver's --csv output.  Classic kernels plus `--only Nstream` (fp32 `--only All` overflows dot to inf
in the reference itself at 100 iterations).  Marks the L2 -> HBM knee.

    python tools/size_sweep.py --out gpurun_out/size_sweep.csv [--lo 20 --hi 30] [--ngpus 1]
"""
import argparse
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import parse_stream_csv  # noqa: E402


def run(exe, n, num_times, extra, env=None):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run([exe, "--csv", "--gigabytes", "-s", str(n), "-n", str(num_times), *extra],
                       capture_output=True, text=True, timeout=1200, env=e)
    if p.returncode != 0:
        return None, p.stderr[-300:]
    return parse_stream_csv(p.stdout), ""


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/size_sweep.csv")
    ap.add_argument("--lo", type=int, default=20)
    ap.add_argument("--hi", type=int, default=30)
    ap.add_argument("--ngpus", type=int, default=1)
    ap.add_argument("--l2-mib", type=float, default=126.0)
    args = ap.parse_args()
    exes = {"b200": os.path.join(ROOT, "build", "b200-stream"), "ref_cuda": os.path.join(ROOT, "oracle", "_ref", "cuda-stream")}
    rows = []
    for dtype, eb, flags in (("f64", 8, []), ("f32", 4, ["--float"])):
        for log2n in range(args.lo, args.hi + 1):
            n = 1 << log2n
            # the driver reports the MINIMUM over iterations: L2-resident calls take ~10 us with +-1 us of
            # launch/wait jitter, so they get more samples (for both binaries alike)
            num_times = 500 if log2n <= 22 else 100 if log2n <= 27 else 30
            resident = 3 * n * eb <= args.l2_mib * (1 << 20)
            for name, exe in exes.items():
                if not os.path.exists(exe):
                    continue
                if name == "ref_cuda" and args.ngpus > 1:
                    continue
                env = {"BABELSTREAM_B200_NGPUS": str(args.ngpus)} if name == "b200" else None
                for sel in ([], ["--only", "Nstream"]):
                    res, err = run(exe, n, num_times, flags + sel, env)
                    if res is None:
                        print("FAILED", name, dtype, log2n, sel, err)
                        continue
                    for k, v in res.items():
                        rows.append(dict(model=name, dtype=dtype, log2n=log2n, kernel=k, gbs=round(v["gbs"], 1),
                                         min_us=round(v["min_s"] * 1e6, 2), avg_us=round(v["avg_s"] * 1e6, 2),
                                         l2_resident=int(resident), ngpus=args.ngpus if name == "b200" else 1))
            last = {(r["model"], r["kernel"]): r["gbs"] for r in rows if r["dtype"] == dtype and r["log2n"] == log2n}
            print(dtype, f"2^{log2n}", "L2" if resident else "HBM",
                  " ".join(f"{k}:{last.get(('b200', k), 0):.0f}/{last.get(('ref_cuda', k), 0):.0f}" for k in
                           ("copy", "mul", "add", "triad", "dot", "nstream")), flush=True)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader()
        w.writerows(rows)
    print("wrote", args.out)


if __name__ == "__main__":
    main()

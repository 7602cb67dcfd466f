#!/usr/bin/env python
"""Launch-shape sweep on a B200: times every compiled variant of a kernel through LEVEL 1 of the C ABI.

    python tools/sweep.py --kernels triad,copy --log2n 29 --dtype f64 --out gpurun_out/sweep.csv [--quick]

Timing: CUDA events on the launching stream (torch's current stream, whose handle is passed to the
launcher), 3 warm-up launches, then `reps` launches back to back between two events.  Arrays are
larger than L2 for every size of interest (3 x 2^25 x 8 B = 805 MB vs 126 MB), which is stated in
the CSV as `l2_note`.  GB/s uses BabelStream's algorithmic bytes (src/benchmark.h:33-38).
"""
from __future__ import annotations

import argparse
import csv
import itertools
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import babelstream_b200 as bs  # noqa: E402


def time_launch(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps  # ms


def time_blocking(fn, reps, warm=3, before=None):
    """The driver's way (src/main.cpp:109-116,154-165): host clock around launch + wait, best of reps.
    `before` (untimed, synchronised) runs ahead of every timed call to reproduce the cache state the
    kernel meets in the driver's sequence, e.g. triad's dirty lines still in L2 when dot starts."""
    stream = torch.cuda.current_stream()
    for _ in range(warm):
        fn()
    stream.synchronize()
    best = 1e30
    for _ in range(reps):
        if before is not None:
            before()
            stream.synchronize()
        t0 = time.perf_counter()
        fn()
        stream.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3


def configs(kernel, quick, hints_override=None):
    if kernel == "copy_async":
        for variant, stages, chunk, grid in itertools.product(("bulk", "tma"), (2, 3, 4, 6, 8), (8192, 16384, 32768, 49152, 65536),
                                                              (-1, -2, -3, -4)):
            if stages * chunk > 220 * 1024 // max(1, -grid) or (variant == "tma" and chunk % 2048):
                continue
            for hint in ((1, 3) if variant == "bulk" else (1,)):
                yield dict(variant=variant, stages=stages, chunk_bytes=chunk, grid=grid, cache_hint=hint)
        return
    vec_unroll = [(8, 1), (8, 2), (8, 4), (16, 1), (16, 2), (16, 4), (16, 8), (32, 1), (32, 2), (32, 4)]
    hints = (1, 2, 3, 4, 5)
    if kernel == "dot":
        blocks = (128, 256, 512)
        grids = (-1, -2, -4, -8, -16)
        variants = ("ticket", "two_pass")
    else:
        blocks = (128, 256, 512, 1024)
        grids = (2 ** 31 - 1, -4) if not quick else (2 ** 31 - 1,)
        variants = ("ldst",)
    if quick:
        hints = (1, 3)
        blocks = (256, 512, 1024) if kernel != "dot" else (256, 512)
    if hints_override:
        hints = hints_override
    for (v, u), b, g, h, var in itertools.product(vec_unroll, blocks, grids, hints, variants):
        if kernel == "dot" and b == 1024 and v * u >= 128:
            continue
        if kernel != "dot" and (b * (-g if g < 0 else 1) > 2048 and g < 0):
            continue  # more threads per SM than can be resident: pointless for a persistent grid
        if kernel != "dot" and b == 1024 and v * u > 32:
            continue
        yield dict(vec_bytes=v, unroll=u, block=b, grid=g, cache_hint=h, variant=var)


def pad_test(log2n, out_path):
    """Do the three streams collide in DRAM banks/channels when a, b, c are laid out at identical offsets modulo
    large powers of two?  Time triad with b and c displaced by a pad inside one big allocation."""
    n, eb = 1 << log2n, 8
    st = torch.cuda.current_stream().cuda_stream
    rows = []
    big = torch.zeros(3 * n + 3 * (1 << 22), dtype=torch.float64, device="cuda")
    for pad in (0, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 1 << 14, 1 << 16, 1 << 17, 1 << 18, 1 << 20, 3 << 18):
        a = big[0:n]
        b = big[n + pad: 2 * n + pad]
        c = big[2 * n + 2 * pad: 3 * n + 2 * pad]
        for name, kw in (("v32u2", dict(vec_bytes=32, unroll=2)), ("v16u8", dict(vec_bytes=16, unroll=8))):
            cfg = bs.LaunchCfg.make(block=256, cache_hint=2, **kw)
            ms = time_launch(lambda: bs.launch("triad", a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, cfg=cfg, stream=st), 20)
            rows.append(dict(pad_elems=pad, pad_bytes=pad * 8, cfg=name, ms=round(ms, 6), gbs=round(3 * n * eb / ms / 1e6, 1)))
            print(rows[-1])
    with open(out_path, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader(); w.writerows(rows)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pad-test", type=int, default=0, help="log2 n for the array-displacement experiment")
    ap.add_argument("--kernels", default="triad,copy,add,mul,nstream,dot,copy_async")
    ap.add_argument("--log2n", default="25,29")
    ap.add_argument("--dtype", default="f64")
    ap.add_argument("--out", default="gpurun_out/sweep.csv")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--blocking", action="store_true", help="time launch+wait per call with the host clock (min), like the driver")
    ap.add_argument("--budget-s", type=float, default=1e9)
    ap.add_argument("--after", default="", help="with --blocking: run this kernel (default config) before every timed call")
    ap.add_argument("--alternate", action="store_true", help="consecutive launches sweep in opposite directions (B200_LAUNCH_REVERSE)")
    ap.add_argument("--hints", default="", help="comma-separated cache hints to try (default: all)")
    ap.add_argument("--top", type=int, default=0, help="write only the best N rows per kernel and size")
    args = ap.parse_args()

    if args.pad_test:
        pad_test(args.pad_test, args.out)
        return
    tdt = torch.float64 if args.dtype == "f64" else torch.float32
    eb = 8 if args.dtype == "f64" else 4
    st = torch.cuda.current_stream().cuda_stream
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    t_start = time.time()
    rows = []
    for log2n in [int(x) for x in args.log2n.split(",")]:
        n = 1 << log2n
        a = torch.full((n,), 0.1, dtype=tdt, device="cuda")
        b = torch.full((n,), 0.2, dtype=tdt, device="cuda")
        c = torch.zeros(n, dtype=tdt, device="cuda")
        ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
        out = torch.zeros(1, dtype=tdt, device="cuda")
        reps = max(5, min(50, int(2e10 / (3 * n * eb))))
        for kernel in args.kernels.split(","):
            base = "copy" if kernel == "copy_async" else kernel
            nbytes = bs.WEIGHTS[base] * n * eb
            best = []
            first_row = len(rows)
            for kw in configs(kernel, args.quick, tuple(int(h) for h in args.hints.split(",")) if args.hints else None):
                if time.time() - t_start > args.budget_s:
                    break
                pair = [bs.LaunchCfg.make(**kw), bs.LaunchCfg.make(flags=bs.LAUNCH_REVERSE if args.alternate else 0, **kw)]
                turn = [0]

                def next_cfg(pair=pair, turn=turn):
                    turn[0] ^= 1
                    return pair[turn[0]]
                if base == "dot":
                    fn = lambda: bs.launch_dot(a.data_ptr(), b.data_ptr(), n, eb, ws.data_ptr(), out.data_ptr(), cfg=next_cfg(), stream=st)
                else:
                    fn = lambda: bs.launch(base, a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, cfg=next_cfg(), stream=st)
                try:
                    before = (lambda: bs.launch(args.after, a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, stream=st)) if args.after else None
                    ms = time_blocking(fn, max(reps, 20), before=before) if args.blocking else time_launch(fn, reps)
                except bs.B200Error as e:
                    print("skip", kernel, kw, e)
                    continue
                gbs = nbytes / ms / 1e6
                row = dict(kernel=kernel, dtype=args.dtype, log2n=log2n, mode=("blocking" if args.blocking else "back_to_back") + ("_alternate" if args.alternate else ""), ms=round(ms, 6), gbs=round(gbs, 1), reps=reps,
                           l2_note="arrays>L2" if 3 * n * eb > 2 * (126 << 20) else "L2-resident", **{k: kw.get(k, "") for k in
                           ("vec_bytes", "unroll", "block", "grid", "cache_hint", "variant", "stages", "chunk_bytes")})
                rows.append(row)
                best.append((gbs, kw))
                if base in ("triad", "nstream", "add", "mul"):
                    # keep the values bounded over thousands of launches
                    a.fill_(0.1); b.fill_(0.2); c.zero_()
            if args.top:
                rows[first_row:] = sorted(rows[first_row:], key=lambda r: -r["gbs"])[:args.top]
            best.sort(key=lambda t: -t[0])
            print(f"== {kernel} {args.dtype} 2^{log2n}: {len(best)} configs, top 8:")
            for gbs, kw in best[:8]:
                print(f"   {gbs:8.1f} GB/s  {kw}")
            if best:
                print(f"   worst {best[-1][0]:8.1f} GB/s  {best[-1][1]}")
            sys.stdout.flush()
        del a, b, c
        torch.cuda.empty_cache()
    with open(args.out, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader()
        w.writerows(rows)
    print("wrote", args.out, len(rows), "rows in", round(time.time() - t_start, 1), "s")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""First-call cost of every blocking kernel call on a fresh context, with and without b200_warmup (ms)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import babelstream_b200 as bs

n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 22)
warm = len(sys.argv) > 2 and sys.argv[2] == "warm"
t0 = time.perf_counter()
s = bs.B200Stream(n, dtype=np.float64)
t1 = time.perf_counter()
if warm:
    s.warmup()
t2 = time.perf_counter()
s.init_arrays()
print(f"n=2^{n.bit_length()-1} warm={warm}: create {1e3*(t1-t0):.1f} ms, warmup {1e3*(t2-t1):.2f} ms")
for k in ("copy", "mul", "add", "triad", "dot", "nstream"):
    ts = []
    for _ in range(5):
        a = time.perf_counter(); s.run(k); ts.append(1e3 * (time.perf_counter() - a))
    print(f"  {k:8s} " + " ".join(f"{t:8.3f}" for t in ts))
s.close()

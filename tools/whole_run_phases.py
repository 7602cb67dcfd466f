#!/usr/bin/env python
"""Where the wall time of a whole `b200-stream -s 2^25 -n 100 --only All` run goes: times the phases of the same call
sequence through the C ABI (ctypes), then the whole programs (ours and the reference's models) a few times each."""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def phases(n, mem):
    import numpy as np
    import babelstream_b200 as bs
    t = [time.perf_counter()]
    s = bs.B200Stream(n, dtype=np.float64, mem=mem)
    t.append(time.perf_counter())
    s.init_arrays()
    t.append(time.perf_counter())
    for _ in range(100):
        for k in ("copy", "mul", "add", "triad", "dot", "nstream"):
            s.run(k)
    t.append(time.perf_counter())
    s.get_arrays()
    t.append(time.perf_counter())
    s.get_arrays()
    t.append(time.perf_counter())
    s.close()
    t.append(time.perf_counter())
    names = ["create (CUDA context, module load, cudaMalloc)", "init_arrays", "600 blocking kernel calls", "get_arrays #1 (host mirrors + D2H)",
             "get_arrays #2 (D2H only)", "destroy"]
    for nm, a, b in zip(names, t, t[1:]):
        print(f"  {nm:55s} {1e3 * (b - a):9.1f} ms")


def whole(cmd, env=None, reps=3):
    out = []
    for _ in range(reps):
        e = dict(os.environ); e.update(env or {})
        t0 = time.perf_counter()
        subprocess.run(cmd, capture_output=True, env=e, check=False)
        out.append(time.perf_counter() - t0)
    return " ".join(f"{x:.2f}" for x in out)


if __name__ == "__main__":
    n = 1 << 25
    if len(sys.argv) > 1 and sys.argv[1] == "phases":
        phases(n, sys.argv[2] if len(sys.argv) > 2 else "device")
        sys.exit(0)
    for mem in ("device", "managed"):
        print(f"phases in a fresh process, mem={mem}:")
        sys.stdout.flush()
        subprocess.run([sys.executable, __file__, "phases", mem])
    args = ["-s", str(n), "-n", "100", "--only", "All"]
    print("whole program wall, s (3 runs each):")
    print("  b200-stream            ", whole([os.path.join(ROOT, "build", "b200-stream")] + args))
    print("  b200-stream mem=managed", whole([os.path.join(ROOT, "build", "b200-stream")] + args, {"BABELSTREAM_B200_MEM": "managed"}))
    print("  ref cuda-stream        ", whole([os.path.join(ROOT, "oracle", "_ref", "cuda-stream")] + args))
    print("  ref thrust-stream      ", whole([os.path.join(ROOT, "oracle", "_ref", "thrust-stream")] + args))
    print("  b200-stream -n 1       ", whole([os.path.join(ROOT, "build", "b200-stream"), "-s", str(n), "-n", "2", "--only", "All"]))
    print("  cuda-stream --list     ", whole([os.path.join(ROOT, "oracle", "_ref", "cuda-stream"), "--list"]))
    print("  b200-stream --list     ", whole([os.path.join(ROOT, "build", "b200-stream"), "--list"]))

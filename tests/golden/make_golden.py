#!/usr/bin/env python
"""Generate the golden fixtures from the UNMODIFIED reference (run in the CPU container only).

    python tests/golden/make_golden.py

Needs oracle/_ref/libref_serial_{fast,off}.so, which oracle/Makefile builds from
/root/reference/src/serial/SerialStream.cpp behind oracle/ref_shim.cpp.  /root/reference does
not exist on the GPU box, so the outputs are committed:

  known_answers.json   bit patterns of a, b, c and the serial dot after driving the reference
                       serial model through the loop of src/main.cpp:150-180 with the driver's
                       constants (0.1, 0.2, 0.0, scalar 0.4), N = 1024, for every
                       (dtype, selection, order, num_times, contraction mode).
  random_vectors.npz   seeded random a, b, c (N = 1031, deliberately ragged) and what each single
                       reference kernel makes of them, per dtype and contraction mode.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
N_KNOWN = 1024
N_RANDOM = 1031


def bits(x) -> str:
    x = np.asarray(x)
    if x.dtype == np.float32:
        return "0x%08x" % int(x.view(np.uint32))
    return "0x%016x" % int(x.view(np.uint64))


def known_answers(orc: O.Oracle):
    rows = []
    for dtype in ("float64", "float32"):
        for mode in ("fast", "off"):
            for sel_name in ("Classic", "All", "Copy", "Mul", "Add", "Triad", "Nstream", "Dot"):
                for order_name, order in (("Classic", O.ORDER_CLASSIC), ("Isolated", O.ORDER_ISOLATED)):
                    for num_times in (2, 10, 100):
                        ref = O.Reference(f"serial_{mode}", dtype, N_KNOWN)
                        ref.init()  # the driver's second init_arrays, src/main.cpp:280
                        total = ref.run(O.SELECTIONS[sel_name], order, num_times, orc)
                        a, b, c = ref.get()
                        for arr in (a, b, c):  # the reference keeps every element identical
                            assert (arr == arr[0]).all()
                        rows.append({
                            "dtype": dtype, "contract": mode, "selection": sel_name, "order": order_name,
                            "num_times": num_times, "n": N_KNOWN,
                            "a": bits(a[0]), "b": bits(b[0]), "c": bits(c[0]), "dot": bits(total),
                        })
                        ref.close()
    return rows


def random_vectors():
    out = {}
    for dtype in (np.float64, np.float32):
        rng = np.random.default_rng(20261104)
        # mixed magnitudes and signs so that fused and unfused roundings really differ
        def draw():
            return (rng.standard_normal(N_RANDOM) * np.exp(rng.uniform(-6, 6, N_RANDOM))).astype(dtype)
        a0, b0, c0 = draw(), draw(), draw()
        tag = np.dtype(dtype).name
        out[f"{tag}/in/a"], out[f"{tag}/in/b"], out[f"{tag}/in/c"] = a0, b0, c0
        for mode in ("fast", "off"):
            for kern in ("copy", "mul", "add", "triad", "nstream", "dot"):
                ref = O.Reference(f"serial_{mode}", dtype, N_RANDOM)
                ref.set(a0, b0, c0)
                r = getattr(ref, kern)()
                a, b, c = ref.get()
                if kern == "dot":
                    out[f"{tag}/{mode}/dot"] = np.array([r], dtype=dtype)
                else:
                    out[f"{tag}/{mode}/{kern}/a"] = a
                    out[f"{tag}/{mode}/{kern}/b"] = b
                    out[f"{tag}/{mode}/{kern}/c"] = c
                ref.close()
    return out


def main():
    if not O.Reference.available("serial_fast"):
        sys.exit("oracle/_ref is not built: run `make -C oracle ref` where /root/reference exists")
    orc = O.Oracle()
    rows = known_answers(orc)
    with open(os.path.join(HERE, "known_answers.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py",
                   "source": "unmodified /root/reference/src/serial/SerialStream.cpp via oracle/ref_shim.cpp",
                   "init": [O.START_A, O.START_B, O.START_C], "scalar": O.START_SCALAR,
                   "rows": rows}, f, indent=0)
    np.savez_compressed(os.path.join(HERE, "random_vectors.npz"), **random_vectors())
    print(f"wrote {len(rows)} known-answer rows and random_vectors.npz")


if __name__ == "__main__":
    main()

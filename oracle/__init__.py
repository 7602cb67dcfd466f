"""oracle -- TEST INFRASTRUCTURE (the checker), never the product path.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import this package.  ``babelstream_b200`` never does.

Two things live here:

* ``Oracle``   -- ctypes bindings for ``babelstream_oracle.c``, a plain-C restatement of the
  reference's serial / OpenMP kernels and of ``check_solution`` (citations in the C file).
* ``Reference`` -- ctypes bindings for ``oracle/_ref/libref_*.so``: the UNMODIFIED reference
  sources compiled in place by ``oracle/Makefile`` behind ``ref_shim.cpp``.  Present only when
  the artefacts were built (in the CPU container, then shipped to the GPU box).

Parity status: pinned (see ``tests/test_oracle.py`` and ``tests/golden/``).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "libbabelstream_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")

# BenchId values, /root/reference/src/benchmark.h:18
COPY, MUL, ADD, TRIAD, NSTREAM, DOT, CLASSIC, ALL = range(8)
SELECTIONS = {"Copy": COPY, "Mul": MUL, "Add": ADD, "Triad": TRIAD, "Nstream": NSTREAM,
              "Dot": DOT, "Classic": CLASSIC, "All": ALL}
ORDER_CLASSIC, ORDER_ISOLATED = 0, 1

# Stream.h:24-27
START_A, START_B, START_C, START_SCALAR = 0.1, 0.2, 0.0, 0.4


def build(verbose: bool = False) -> None:
    """Compile the C restatement (and oracle/_ref when the reference tree is present)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=out)
    subprocess.check_call(["make", "-C", HERE, "ref"], stdout=out)


def _np_dtype(dtype):
    dt = np.dtype(dtype)
    if dt not in (np.dtype(np.float32), np.dtype(np.float64)):
        raise TypeError(f"BabelStream is float/double only, got {dt}")
    return dt


def _suf(dt) -> str:
    return "f32" if dt == np.dtype(np.float32) else "f64"


def _ct(dt):
    return C.c_float if dt == np.dtype(np.float32) else C.c_double


def _ptr(x: np.ndarray):
    assert x.flags["C_CONTIGUOUS"]
    return x.ctypes.data_as(C.c_void_p)


class Oracle:
    """The C restatement.  All array arguments are 1-D contiguous numpy arrays, modified in place."""

    def __init__(self, path: str = ORACLE_SO):
        if not os.path.exists(path):
            build()
        self.lib = C.CDLL(path)
        L = self.lib
        ll, vp, ci = C.c_longlong, C.c_void_p, C.c_int
        for suf, T in (("f32", C.c_float), ("f64", C.c_double)):
            getattr(L, f"oracle_init_{suf}").argtypes = [vp, vp, vp, ll, T, T, T]
            getattr(L, f"oracle_copy_{suf}").argtypes = [vp, vp, ll]
            getattr(L, f"oracle_mul_{suf}").argtypes = [vp, vp, ll, T]
            getattr(L, f"oracle_add_{suf}").argtypes = [vp, vp, vp, ll]
            getattr(L, f"oracle_triad_{suf}").argtypes = [vp, vp, vp, ll, T, ci]
            getattr(L, f"oracle_nstream_{suf}").argtypes = [vp, vp, vp, ll, T, ci]
            for name in ("dot", "dot_omp"):
                f = getattr(L, f"oracle_{name}_{suf}")
                f.argtypes = [vp, vp, ll, ci]
                f.restype = T
            f = getattr(L, f"oracle_gold_{suf}")
            f.argtypes = [ci, ci, ll, ll, ci, T, T, T, T, vp]
            f = getattr(L, f"oracle_check_solution_{suf}")
            f.argtypes = [ci, ci, ll, ll, ci, vp, vp, vp, T, T, T, T, T, vp]
            f.restype = ll
            f = getattr(L, f"oracle_run_{suf}")
            f.argtypes = [ci, ci, ll, vp, vp, vp, ll, T, ci, ci]
            f.restype = T
        L.oracle_run_benchmark.argtypes = [ci, ci]
        L.oracle_num_threads.restype = ci

    # -- kernels ---------------------------------------------------------------------------
    def init(self, a, b, c, ia=START_A, ib=START_B, ic=START_C):
        dt = _np_dtype(a.dtype)
        getattr(self.lib, f"oracle_init_{_suf(dt)}")(_ptr(a), _ptr(b), _ptr(c), a.size, ia, ib, ic)

    def copy(self, a, c):
        dt = _np_dtype(a.dtype)
        getattr(self.lib, f"oracle_copy_{_suf(dt)}")(_ptr(a), _ptr(c), a.size)

    def mul(self, b, c, scalar=START_SCALAR):
        dt = _np_dtype(b.dtype)
        getattr(self.lib, f"oracle_mul_{_suf(dt)}")(_ptr(b), _ptr(c), b.size, dt.type(scalar))

    def add(self, a, b, c):
        dt = _np_dtype(a.dtype)
        getattr(self.lib, f"oracle_add_{_suf(dt)}")(_ptr(a), _ptr(b), _ptr(c), a.size)

    def triad(self, a, b, c, scalar=START_SCALAR, contract=True):
        dt = _np_dtype(a.dtype)
        getattr(self.lib, f"oracle_triad_{_suf(dt)}")(_ptr(a), _ptr(b), _ptr(c), a.size,
                                                      dt.type(scalar), int(contract))

    def nstream(self, a, b, c, scalar=START_SCALAR, contract=True):
        dt = _np_dtype(a.dtype)
        getattr(self.lib, f"oracle_nstream_{_suf(dt)}")(_ptr(a), _ptr(b), _ptr(c), a.size,
                                                        dt.type(scalar), int(contract))

    def dot(self, a, b, contract=True, omp=False):
        dt = _np_dtype(a.dtype)
        name = "dot_omp" if omp else "dot"
        return dt.type(getattr(self.lib, f"oracle_{name}_{_suf(dt)}")(_ptr(a), _ptr(b), a.size, int(contract)))

    # -- driver logic ------------------------------------------------------------------------
    def run_benchmark(self, selection: int, pos: int) -> bool:
        return bool(self.lib.oracle_run_benchmark(selection, pos))

    def gold(self, dtype, selection, order, num_times, array_size, contract=True,
             ia=START_A, ib=START_B, ic=START_C, scalar=START_SCALAR):
        dt = _np_dtype(dtype)
        g = np.zeros(4, dtype=dt)
        getattr(self.lib, f"oracle_gold_{_suf(dt)}")(selection, order, num_times, array_size, int(contract),
                                                     dt.type(ia), dt.type(ib), dt.type(ic), dt.type(scalar), _ptr(g))
        return g

    def check_solution(self, selection, order, num_times, a, b, c, total, contract=True,
                       ia=START_A, ib=START_B, ic=START_C, scalar=START_SCALAR, array_size=None):
        """Number of failed comparisons, exactly as src/main.cpp:331-421 counts them."""
        dt = _np_dtype(a.dtype)
        n = a.size if array_size is None else array_size
        return int(getattr(self.lib, f"oracle_check_solution_{_suf(dt)}")(
            selection, order, num_times, n, int(contract), _ptr(a), _ptr(b), _ptr(c), dt.type(total),
            dt.type(ia), dt.type(ib), dt.type(ic), dt.type(scalar), None))

    def run(self, selection, order, num_times, a, b, c, scalar=START_SCALAR, contract=True, omp_dot=False):
        dt = _np_dtype(a.dtype)
        return dt.type(getattr(self.lib, f"oracle_run_{_suf(dt)}")(
            selection, order, num_times, _ptr(a), _ptr(b), _ptr(c), a.size, dt.type(scalar),
            int(contract), int(omp_dot)))

    def num_threads(self) -> int:
        return int(self.lib.oracle_num_threads())


class Reference:
    """The unmodified reference model behind oracle/ref_shim.cpp.  kind: serial_fast | serial_off | omp."""

    def __init__(self, kind: str, dtype, n: int):
        path = os.path.join(REF_DIR, f"libref_{kind}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = L = C.CDLL(path)
        self.dt = _np_dtype(dtype)
        self.is_float = int(self.dt == np.dtype(np.float32))
        self.n = n
        vp, ci, ll, d = C.c_void_p, C.c_int, C.c_longlong, C.c_double
        L.ref_impl_string.restype = C.c_char_p
        L.ref_create.argtypes = [ci, ll]
        L.ref_create.restype = vp
        L.ref_destroy.argtypes = [vp, ci]
        L.ref_init.argtypes = [vp, ci, d, d, d]
        L.ref_set.argtypes = [vp, ci, vp, vp, vp]
        L.ref_get.argtypes = [vp, ci, vp, vp, vp]
        for k in ("copy", "mul", "add", "triad", "nstream"):
            getattr(L, f"ref_{k}").argtypes = [vp, ci]
        L.ref_dot.argtypes = [vp, ci]
        L.ref_dot.restype = d
        self.h = L.ref_create(self.is_float, n)

    @staticmethod
    def available(kind: str = "serial_fast") -> bool:
        return os.path.exists(os.path.join(REF_DIR, f"libref_{kind}.so"))

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h, self.is_float)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def impl(self) -> str:
        return self.lib.ref_impl_string().decode()

    def init(self, ia=START_A, ib=START_B, ic=START_C):
        self.lib.ref_init(self.h, self.is_float, ia, ib, ic)

    def set(self, a, b, c):
        for x in (a, b, c):
            assert x.dtype == self.dt and x.size == self.n
        self.lib.ref_set(self.h, self.is_float, _ptr(a), _ptr(b), _ptr(c))

    def get(self):
        a, b, c = (np.empty(self.n, dtype=self.dt) for _ in range(3))
        self.lib.ref_get(self.h, self.is_float, _ptr(a), _ptr(b), _ptr(c))
        return a, b, c

    def copy(self): self.lib.ref_copy(self.h, self.is_float)
    def mul(self): self.lib.ref_mul(self.h, self.is_float)
    def add(self): self.lib.ref_add(self.h, self.is_float)
    def triad(self): self.lib.ref_triad(self.h, self.is_float)
    def nstream(self): self.lib.ref_nstream(self.h, self.is_float)

    def dot(self):
        return self.dt.type(self.lib.ref_dot(self.h, self.is_float))

    def run(self, selection: int, order: int, num_times: int, oracle: Oracle):
        """Drive the reference model through the loop of src/main.cpp:150-180; returns the last dot."""
        total = self.dt.type(0)
        calls = {COPY: self.copy, MUL: self.mul, ADD: self.add, TRIAD: self.triad, NSTREAM: self.nstream}
        order_ids = [COPY, MUL, ADD, TRIAD, DOT, NSTREAM]  # benchmark.h:32-39
        outer, inner = (num_times, 6) if order == ORDER_CLASSIC else (6, num_times)
        for o in range(outer):
            for i in range(inner):
                pos = i if order == ORDER_CLASSIC else o
                if not oracle.run_benchmark(selection, pos):
                    continue
                bid = order_ids[pos]
                if bid == DOT:
                    total = self.dot()
                else:
                    calls[bid]()
        return total

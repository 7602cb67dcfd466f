"""CPU tests that PIN the oracle: the C restatement must reproduce, bit for bit,

  (1) the known-answer table generated from the unmodified reference serial model
      (tests/golden/known_answers.json; the rows of SURVEY.md section 8c are a subset),
  (2) the per-kernel outputs of the reference on seeded random inputs
      (tests/golden/random_vectors.npz), and
  (3) oracle/_ref/libref_serial_*.so / libref_omp.so called live, when those artefacts exist.
"""
import numpy as np
import pytest

import oracle as O

DTYPES = [np.float64, np.float32]
KERNELS = ["copy", "mul", "add", "triad", "nstream"]


def bits(x):
    x = np.asarray(x)
    return ("0x%08x" % int(x.view(np.uint32))) if x.dtype == np.float32 else ("0x%016x" % int(x.view(np.uint64)))


def same_bits(x, y):
    u = np.uint32 if x.dtype == np.float32 else np.uint64
    return np.array_equal(np.ascontiguousarray(x).view(u), np.ascontiguousarray(y).view(u))


def apply_kernel(orc, kern, a, b, c, contract):
    if kern == "copy": orc.copy(a, c)
    elif kern == "mul": orc.mul(b, c)
    elif kern == "add": orc.add(a, b, c)
    elif kern == "triad": orc.triad(a, b, c, contract=contract)
    elif kern == "nstream": orc.nstream(a, b, c, contract=contract)
    else: raise KeyError(kern)


def test_survey_table_rows_present(golden_rows):
    """The rows quoted in SURVEY.md 8(c) must be in the generated fixture unchanged."""
    want = [
        ("float64", "Classic", 100, "fast", "0x3f5ba3ecbf7bd6d2", "0x3f47089a9f91ddaf", "0x3f6427874b9fa1f9", "0x3f53e540b388b94e"),
        ("float64", "Classic", 100, "off", "0x3f5ba3ecbf7bd6d0", "0x3f47089a9f91ddad", "0x3f6427874b9fa1f7", "0x3f53e540b388b94e"),
        ("float64", "Classic", 2, "fast", "0x3fb797cc39ffd610", "0x3fa3a92a30553263", "0x3fc13404ea4a8c16", "0x400cfdb417c18ae7"),
        ("float64", "All", 100, "fast", "0x459ba3ecbf7bd6d2", "0x4577089a9f91ddaf", "0x459427874b9fa1f9", "0x4bb3e540b388b94e"),
        ("float64", "All", 100, "off", "0x459ba3ecbf7bd6d0", "0x4577089a9f91ddad", "0x459427874b9fa1f7", "0x4bb3e540b388b94e"),
        ("float64", "All", 2, "off", "0x3fd797cc39ffd610", "0x3fb3a92a30553263", "0x3fd13404ea4a8c16", "0x402cfdb417c18ae7"),
        ("float32", "Classic", 100, "fast", "0x3add1f85", "0x3a3844ef", "0x3b213c51", "0x3a9f2a6a"),
        ("float32", "Classic", 100, "off", "0x3add1f8c", "0x3a3844f5", "0x3b213c56", "0x3a9f2a6b"),
        ("float32", "All", 100, "fast", "0x6cdd1f85", "0x6bb844ef", "0x6ca13c51", "0x7f800000"),
        ("float32", "All", 100, "off", "0x6cdd1f8c", "0x6bb844f5", "0x6ca13c56", "0x7f800000"),
    ]
    index = {(r["dtype"], r["selection"], r["num_times"], r["contract"], r["order"]): r for r in golden_rows}
    for dt, sel, nt, mode, a, b, c, d in want:
        r = index[(dt, sel, nt, mode, "Classic")]
        assert (r["a"], r["b"], r["c"], r["dot"]) == (a, b, c, d)


def test_oracle_reproduces_every_known_answer(orc, golden_rows):
    assert len(golden_rows) == 192
    for r in golden_rows:
        dt = np.dtype(r["dtype"])
        n = r["n"]
        a, b, c = (np.empty(n, dtype=dt) for _ in range(3))
        orc.init(a, b, c)
        order = O.ORDER_CLASSIC if r["order"] == "Classic" else O.ORDER_ISOLATED
        with np.errstate(over="ignore"):
            total = orc.run(O.SELECTIONS[r["selection"]], order, r["num_times"], a, b, c,
                            contract=(r["contract"] == "fast"))
        got = (bits(a[0]), bits(b[0]), bits(c[0]), bits(total))
        assert got == (r["a"], r["b"], r["c"], r["dot"]), r
        assert (a == a[0]).all() and (b == b[0]).all() and (c == c[0]).all()


def test_oracle_gold_recurrence_matches_arrays(orc, golden_rows):
    """check_solution's scalar recurrence (main.cpp:342-379) lands on the same bits as the arrays."""
    for r in golden_rows:
        dt = np.dtype(r["dtype"])
        order = O.ORDER_CLASSIC if r["order"] == "Classic" else O.ORDER_ISOLATED
        with np.errstate(over="ignore"):
            g = orc.gold(dt, O.SELECTIONS[r["selection"]], order, r["num_times"], r["n"],
                         contract=(r["contract"] == "fast"))
        assert (bits(g[0]), bits(g[1]), bits(g[2])) == (r["a"], r["b"], r["c"]), r


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("mode", ["fast", "off"])
def test_oracle_matches_reference_random_fixture(orc, golden_random, dtype, mode):
    tag = np.dtype(dtype).name
    a0, b0, c0 = (golden_random[f"{tag}/in/{k}"] for k in "abc")
    assert a0.size == 1031
    for kern in KERNELS:
        a, b, c = a0.copy(), b0.copy(), c0.copy()
        apply_kernel(orc, kern, a, b, c, contract=(mode == "fast"))
        for name, arr in (("a", a), ("b", b), ("c", c)):
            assert same_bits(arr, golden_random[f"{tag}/{mode}/{kern}/{name}"]), (kern, name)
    d = orc.dot(a0, b0, contract=(mode == "fast"))
    assert bits(d) == bits(golden_random[f"{tag}/{mode}/dot"][0])


def test_fixture_distinguishes_contraction(golden_random):
    """The random fixture must actually tell fused from unfused rounding apart."""
    for tag in ("float64", "float32"):
        for kern in ("triad", "nstream"):
            assert not same_bits(golden_random[f"{tag}/fast/{kern}/a"], golden_random[f"{tag}/off/{kern}/a"])
        # copy/mul/add have a single rounding: both modes agree
        assert same_bits(golden_random[f"{tag}/fast/mul/b"], golden_random[f"{tag}/off/mul/b"])
        assert same_bits(golden_random[f"{tag}/fast/add/c"], golden_random[f"{tag}/off/add/c"])


@pytest.mark.skipif(not O.Reference.available("serial_fast"), reason="oracle/_ref not built")
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("mode", ["fast", "off"])
@pytest.mark.parametrize("n", [1, 7, 1000, 65537])
def test_oracle_matches_live_reference(orc, dtype, mode, n):
    rng = np.random.default_rng(n)
    a0, b0, c0 = (rng.standard_normal(n).astype(dtype) for _ in range(3))
    for kern in KERNELS:
        ref = O.Reference(f"serial_{mode}", dtype, n)
        ref.set(a0, b0, c0)
        getattr(ref, kern)()
        ra, rb, rc = ref.get()
        ref.close()
        a, b, c = a0.copy(), b0.copy(), c0.copy()
        apply_kernel(orc, kern, a, b, c, contract=(mode == "fast"))
        assert same_bits(a, ra) and same_bits(b, rb) and same_bits(c, rc), kern
    ref = O.Reference(f"serial_{mode}", dtype, n)
    ref.set(a0, b0, c0)
    assert bits(ref.dot()) == bits(orc.dot(a0, b0, contract=(mode == "fast")))
    ref.close()


@pytest.mark.skipif(not O.Reference.available("omp"), reason="oracle/_ref not built")
@pytest.mark.parametrize("dtype", DTYPES)
def test_reference_omp_agrees_with_serial(orc, dtype):
    """omp and serial models are the same arithmetic; only dot's order differs (tolerance)."""
    n = 100003
    rng = np.random.default_rng(5)
    a0, b0, c0 = (rng.standard_normal(n).astype(dtype) for _ in range(3))
    for kern in KERNELS:
        ref = O.Reference("omp", dtype, n)
        assert ref.impl() == "OpenMP"
        ref.set(a0, b0, c0)
        getattr(ref, kern)()
        ra, rb, rc = ref.get()
        ref.close()
        a, b, c = a0.copy(), b0.copy(), c0.copy()
        apply_kernel(orc, kern, a, b, c, contract=True)
        assert same_bits(a, ra) and same_bits(b, rb) and same_bits(c, rc), kern
    ref = O.Reference("omp", dtype, n)
    ref.set(np.abs(a0), np.abs(b0), c0)
    d_omp = float(ref.dot())
    ref.close()
    d_ser = float(orc.dot(np.abs(a0), np.abs(b0)))
    eps = np.finfo(dtype).eps
    assert abs(d_omp - d_ser) <= 1e7 * eps * max(abs(d_omp), abs(d_ser))  # main.cpp:384


@pytest.mark.parametrize("dtype", DTYPES)
def test_numpy_restatement_agrees_unfused(orc, dtype):
    """numpy never fuses: it is an independent second statement of the contract=off arithmetic."""
    rng = np.random.default_rng(11)
    n = 4099
    a0, b0, c0 = (rng.standard_normal(n).astype(dtype) for _ in range(3))
    s = dtype(O.START_SCALAR)
    a, b, c = a0.copy(), b0.copy(), c0.copy()
    orc.mul(b, c); assert same_bits(b, s * c0)
    a, b, c = a0.copy(), b0.copy(), c0.copy()
    orc.add(a, b, c); assert same_bits(c, a0 + b0)
    a, b, c = a0.copy(), b0.copy(), c0.copy()
    orc.triad(a, b, c, contract=False); assert same_bits(a, b0 + s * c0)
    a, b, c = a0.copy(), b0.copy(), c0.copy()
    orc.nstream(a, b, c, contract=False); assert same_bits(a, a0 + (b0 + s * c0))
    a, b, c = a0.copy(), b0.copy(), c0.copy()
    orc.copy(a, c); assert same_bits(c, a0)


@pytest.mark.parametrize("dtype", DTYPES)
def test_check_solution_accepts_and_rejects(orc, dtype):
    """Restated check_solution (main.cpp:331-421): passes a clean run, counts corrupted elements, fails NaN."""
    n, nt = 2048, 10
    a, b, c = (np.empty(n, dtype=dtype) for _ in range(3))
    for sel in ("Classic", "All", "Triad", "Nstream", "Dot"):
        orc.init(a, b, c)
        total = orc.run(O.SELECTIONS[sel], O.ORDER_CLASSIC, nt, a, b, c)
        assert orc.check_solution(O.SELECTIONS[sel], O.ORDER_CLASSIC, nt, a, b, c, total) == 0
    a2 = a.copy(); a2[17] *= dtype(1.001); a2[18] = np.nan
    assert orc.check_solution(O.SELECTIONS["Dot"], O.ORDER_CLASSIC, nt, a2, b, c, total) == 2
    # a wrong sum is caught only beyond 1e7*eps relative (main.cpp:384) ...
    bad = total * dtype(1 + 3e7 * np.finfo(dtype).eps)
    ok = total * dtype(1 + 1e6 * np.finfo(dtype).eps)
    n_bad = orc.check_solution(O.SELECTIONS["Dot"], O.ORDER_CLASSIC, nt, a, b, c, bad)
    if dtype is np.float64:
        assert n_bad == 1
    else:
        # ... which for float is 1e7 * 1.19e-7 = 1.19 > 1: the reference's float dot check can
        # only ever trip on NaN or a sign flip.  Restated faithfully.
        assert n_bad == 0
        assert orc.check_solution(O.SELECTIONS["Dot"], O.ORDER_CLASSIC, nt, a, b, c, dtype(np.nan)) == 1
        assert orc.check_solution(O.SELECTIONS["Dot"], O.ORDER_CLASSIC, nt, a, b, c, -total) == 1
    assert orc.check_solution(O.SELECTIONS["Dot"], O.ORDER_CLASSIC, nt, a, b, c, ok) == 0


def test_run_benchmark_table(orc):
    """benchmark.h:32-39,62-66: order Copy Mul Add Triad Dot Nstream; Classic excludes Nstream."""
    L = orc.lib
    assert [L.oracle_bench_id(i) for i in range(6)] == [O.COPY, O.MUL, O.ADD, O.TRIAD, O.DOT, O.NSTREAM]
    assert [L.oracle_bench_weight(i) for i in range(6)] == [2, 2, 3, 3, 2, 4]
    assert [orc.run_benchmark(O.CLASSIC, i) for i in range(6)] == [True] * 5 + [False]
    assert all(orc.run_benchmark(O.ALL, i) for i in range(6))
    assert [orc.run_benchmark(O.TRIAD, i) for i in range(6)] == [False, False, False, True, False, False]

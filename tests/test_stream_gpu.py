"""GPU parity tests, LEVEL 2 of the C ABI: the Stream<T> object, driven the way src/main.cpp drives a model,
against (1) the committed known answers of the unmodified reference serial model, (2) the oracle's
check_solution, and (3) the b200-stream binary = reference main.cpp + our model, when it was built."""
import os
import subprocess

import numpy as np
import pytest

import babelstream_b200 as bs
import oracle as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "build", "b200-stream")
ORDER = ["copy", "mul", "add", "triad", "dot", "nstream"]  # src/benchmark.h:32-39


def bits(x):
    x = np.asarray(x)
    return ("0x%08x" % int(x.view(np.uint32))) if x.dtype == np.float32 else ("0x%016x" % int(x.view(np.uint64)))


def drive(stream, orc, selection, order, num_times):
    """The loop of src/main.cpp:150-180 over a B200Stream."""
    total = stream.dtype.type(0)
    outer, inner = (num_times, 6) if order == O.ORDER_CLASSIC else (6, num_times)
    for o in range(outer):
        for i in range(inner):
            pos = i if order == O.ORDER_CLASSIC else o
            if not orc.run_benchmark(selection, pos):
                continue
            r = stream.run(ORDER[pos])
            if ORDER[pos] == "dot":
                total = r
    return total


def test_known_answers_every_row(orc, golden_rows):
    """192 rows: (dtype, selection, order, num_times, contraction) -> a, b, c bit patterns."""
    streams = {}
    for r in golden_rows:
        dt = np.dtype(r["dtype"])
        key = (r["dtype"], r["contract"])
        if key not in streams:
            streams[key] = bs.B200Stream(r["n"], dtype=dt, contract=r["contract"])
        s = streams[key]
        s.init_arrays()  # the driver's timed init, src/main.cpp:280
        order = O.ORDER_CLASSIC if r["order"] == "Classic" else O.ORDER_ISOLATED
        total = drive(s, orc, O.SELECTIONS[r["selection"]], order, r["num_times"])
        a, b, c = s.get_arrays()
        assert (bits(a[0]), bits(b[0]), bits(c[0])) == (r["a"], r["b"], r["c"]), r
        assert (a == a[0]).all() and (b == b[0]).all() and (c == c[0]).all()
        want = np.array(int(r["dot"], 16), dtype=np.uint64 if dt == np.float64 else np.uint32).view(dt)
        if np.isfinite(want) and orc.run_benchmark(O.SELECTIONS[r["selection"]], 4):
            # The reference's serial chain of n equal terms carries up to n*eps/2 of rounding error, the
            # GPU's tree far less: compare with the serial answer within the serial error bound ...
            eps = np.finfo(dt).eps
            assert abs(float(total) - float(want)) <= r["n"] * eps * abs(float(want)), r
            # ... and with check_solution's exact gold a*b*n (src/main.cpp:349) tightly.
            g = orc.gold(dt, O.SELECTIONS[r["selection"]], order, r["num_times"], r["n"], contract=(r["contract"] == "fast"))
            assert abs(float(total) - float(g[3])) <= 16 * eps * abs(float(g[3])), r
    for s in streams.values():
        s.close()


def test_warmup_pays_first_use_costs_and_leaves_results_alone(orc):
    """b200_warmup (called by the model's constructor): afterwards the FIRST call of every kernel costs what the later
    ones do -- --order Isolated (src/main.cpp:168-175) averages over all of them -- and a run gives the usual bits."""
    import time
    n = (1 << 20) + 3
    with bs.B200Stream(n, dtype=np.float64) as s:
        s.warmup()
        s.init_arrays()
        first = {}
        for k in ORDER:
            t0 = time.perf_counter()
            s.run(k)
            first[k] = time.perf_counter() - t0
        # unwarmed: 1.7-3.9 ms per elementwise kernel, 38 ms for dot (profiles/r02/first_call_costs_session_n.txt)
        assert max(first.values()) < 1e-3, first
        s.init_arrays()
        total = drive(s, orc, O.ALL, O.ORDER_ISOLATED, 6)
        a, b, c = s.get_arrays()
        g = orc.gold(np.float64, O.ALL, O.ORDER_ISOLATED, 6, n)
        assert bits(a[0]) == bits(g[0]) and bits(b[-1]) == bits(g[1]) and bits(c[n // 2]) == bits(g[2])
        assert orc.check_solution(O.ALL, O.ORDER_ISOLATED, 6, a, b, c, total) == 0
        s.set_option("sweep", "forward")      # plans are rebuilt on their next use; a second warm-up is legal any time
        s.warmup(0.4)
        s.init_arrays()
        total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 3)
        a, b, c = s.get_arrays()
        assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 3, a, b, c, total) == 0


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("n", [1, 100, 33554432 // 32 + 5])
def test_check_solution_passes(orc, dtype, n):
    for sel in ("Classic", "All", "Triad", "Nstream", "Dot", "Copy"):
        for order in (O.ORDER_CLASSIC, O.ORDER_ISOLATED):
            nt = 4 if dtype is np.float32 and sel == "All" else 10
            with bs.B200Stream(n, dtype=dtype) as s:
                s.init_arrays()
                total = drive(s, orc, O.SELECTIONS[sel], order, nt)
                a, b, c = s.get_arrays()
                assert orc.check_solution(O.SELECTIONS[sel], order, nt, a, b, c, total) == 0, (sel, order)
                # stronger than the reference's 100*eps: bit-exact against the oracle's gold recurrence
                g = orc.gold(dtype, O.SELECTIONS[sel], order, nt, n)
                assert bits(a[0]) == bits(g[0]) and bits(b[-1]) == bits(g[1]) and bits(c[n // 2]) == bits(g[2])


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_full_size_all_kernels_bit_exact(orc, dtype):
    """BASELINE config 2: 2^25 elements, all six kernels; every element compared with the gold bits."""
    n, nt = 1 << 25, 5
    with bs.B200Stream(n, dtype=dtype) as s:
        s.init_arrays()
        total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, nt)
        a, b, c = s.get_arrays()
        g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, nt, n)
        u = np.uint64 if dtype is np.float64 else np.uint32
        for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
            assert (arr.view(u) == np.asarray(gold).view(u)).all()
        assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, nt, a, b, c, total) == 0
        assert abs(float(total) - float(g[3])) <= 64 * np.finfo(dtype).eps * abs(float(g[3]))


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_write_read_roundtrip_and_random_parity(orc, dtype):
    n = 100003
    rng = np.random.default_rng(3)
    a0, b0, c0 = (rng.standard_normal(n).astype(dtype) for _ in range(3))
    for contract in ("fast", "off"):
        with bs.B200Stream(n, dtype=dtype, contract=contract) as s:
            s.write_arrays(a0, b0, c0)
            ra, rb, rc = s.read_arrays()
            assert np.array_equal(ra, a0) and np.array_equal(rb, b0) and np.array_equal(rc, c0)
            a, b, c = a0.copy(), b0.copy(), c0.copy()
            for k in ("copy", "mul", "add", "triad", "nstream", "triad", "add"):
                s.run(k)
                {"copy": lambda: orc.copy(a, c), "mul": lambda: orc.mul(b, c), "add": lambda: orc.add(a, b, c),
                 "triad": lambda: orc.triad(a, b, c, contract=(contract == "fast")),
                 "nstream": lambda: orc.nstream(a, b, c, contract=(contract == "fast"))}[k]()
            ga, gb, gc = s.get_arrays()
            u = np.uint64 if dtype is np.float64 else np.uint32
            assert np.array_equal(ga.view(u), a.view(u)) and np.array_equal(gb.view(u), b.view(u)) and np.array_equal(gc.view(u), c.view(u))
            d = float(s.dot())
            want = float(np.dot(a.astype(np.float64), b.astype(np.float64)))
            scale = float(np.dot(np.abs(a).astype(np.float64), np.abs(b).astype(np.float64)))
            assert abs(d - want) <= 64 * np.finfo(dtype).eps * scale


def test_run_sequence_is_the_same_blocking_calls(orc):
    """b200_run_sequence: the driver's loop over bench[] in one call -- same bits as six separate calls."""
    n = (1 << 21) + 7
    with bs.B200Stream(n, dtype=np.float64) as s, bs.B200Stream(n, dtype=np.float64) as t:
        for _ in range(4):
            v = s.run_sequence(ORDER)
            for k in ORDER:
                w = t.run(k)
                if k == "dot":
                    assert np.asarray(v).tobytes() == np.asarray(w).tobytes()
        for x, y in zip(s.get_arrays(), t.get_arrays()):
            assert x.tobytes() == y.tobytes()
        assert s.run_sequence(["copy", "add"]) is None
        with pytest.raises(bs.B200Error):
            bs.stream._check(bs.lib().b200_run_sequence(s._h, None, 3, 0.4, None))


def test_options_and_variants(orc):
    n = 1 << 20
    ref = None
    for opts in ({}, {"copy_variant": "bulk"}, {"copy_variant": "tma"}, {"dot_variant": "two_pass"},
                 {"sync": "spin"}, {"sync": "stream"}, {"sync": "doorbell"}, {"sync": "doorbell", "dot_variant": "two_pass"},
                 {"launch": "direct"}, {"launch": "direct", "sync": "doorbell"}, {"launch": "graph", "dot_variant": "two_pass"},
                 {"sweep": "forward"}, {"sweep": "forward", "launch": "direct"}, {"tune": "all:keep=-1"}, {"tune": "all:hint=3,keep=2;dot:hint=3,keep=1"},
                 {"timing": "1"}, {"timing": "1", "sync": "doorbell"}, {"tune": "all:vec=16,unroll=4,hint=3,grid=-4;dot:vec=32,unroll=2,block=256"}):
        with bs.B200Stream(n, dtype=np.float64, **opts) as s:
            total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 3)
            a, b, c = s.get_arrays()
            cur = (bits(a[0]), bits(b[0]), bits(c[0]), bits(a[-1]))
            ref = ref or cur
            assert cur == ref, opts
            assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 3, a, b, c, total) == 0
            if opts.get("timing"):
                assert s.last_kernel_ms() > 0
            else:
                assert s.last_kernel_ms() < 0
            assert "triad" in s.describe()


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_wait_modes_bit_exact_at_every_size_class(orc, dtype):
    """sync=auto rings a doorbell for dot and uses cudaStreamSynchronize for the elementwise kernels;
    forcing either mode for every kernel at any size must give the same bits (and the same dot)."""
    u = np.uint64 if dtype is np.float64 else np.uint32
    for n in (1, 1000, (1 << 20) + 3, (1 << 24) + 9):
        ref = None
        for sync in ("auto", "stream", "doorbell", "spin"):
            with bs.B200Stream(n, dtype=dtype, sync=sync) as s:
                assert "wait:" in s.describe()
                total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 5)
                a, b, c = s.get_arrays()
                g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, 5, n)
                for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
                    assert (arr.view(u) == np.asarray(gold).view(u)).all(), (n, sync)
                cur = total.tobytes()
                ref = ref or cur
                assert cur == ref, (n, sync)



def test_plans_replay_and_follow_the_arguments(orc):
    """Every LEVEL 2 call replays a captured one-kernel graph (launch=graph, the default).  The plan is keyed on the
    scalar and on the options: changing either must rebuild it, and replays are counted as launches."""
    n = (1 << 16) + 3
    with bs.B200Stream(n, dtype=np.float64) as s, bs.B200Stream(n, dtype=np.float64, launch="direct") as d:
        assert "CUDA graph" in s.describe() and "direct" in d.describe()
        c0 = bs.launch_count()
        for _ in range(5):
            s.triad()
        assert bs.launch_count() - c0 == 5, "one kernel per call, captured or replayed"
        for scalar in (0.4, 0.25, -3.0, 0.4):
            for x in (s, d):
                x.init_arrays()
                x.copy(); x.mul(scalar); x.add(); x.triad(scalar); x.nstream(scalar)
            for ga, gd in zip(s.get_arrays(), d.get_arrays()):
                assert np.array_equal(ga.view(np.uint64), gd.view(np.uint64)), scalar
            assert s.dot().tobytes() == d.dot().tobytes()
        # an option change between calls (new launch shape) is picked up by the next call
        s.set_option("tune", "triad:vec=8,unroll=4,block=128,hint=1")
        d.set_option("tune", "triad:vec=8,unroll=4,block=128,hint=1")
        s.triad(); d.triad()
        assert np.array_equal(s.get_arrays()[0].view(np.uint64), d.get_arrays()[0].view(np.uint64))
        # many dots in a row: the doorbell value lives on the device and must stay in step with the host
        want = d.dot().tobytes()
        for _ in range(300):
            assert s.dot().tobytes() == want


def test_gpu_dot_matches_reference_omp_model_on_random_data(orc):
    """north_star: "match the reference's serial and OpenMP models".  dot of the UNMODIFIED OMPStream.cpp
    (oracle/_ref/libref_omp.so) on seeded random data, within the driver's own bound (src/main.cpp:384)."""
    if not O.Reference.available("omp"):
        pytest.skip("oracle/_ref/libref_omp.so not built")
    for dtype in (np.float64, np.float32):
        n = (1 << 20) + 11
        rng = np.random.default_rng(2025)
        a0 = rng.uniform(0.5, 1.5, n).astype(dtype)
        b0 = rng.uniform(0.5, 1.5, n).astype(dtype)
        c0 = rng.uniform(0.5, 1.5, n).astype(dtype)
        ref = O.Reference("omp", dtype, n)
        ref.set(a0, b0, c0)
        want = float(ref.dot())
        ref.close()
        with bs.B200Stream(n, dtype=dtype) as s:
            s.write_arrays(a0, b0, c0)
            got = float(s.dot())
        eps = np.finfo(dtype).eps
        assert abs(got - want) <= 1e7 * eps * max(abs(got), abs(want)), (dtype, got, want)
        # and far tighter than that against the exact sum: both are sums of n positive terms
        exact = float(np.dot(a0.astype(np.float64), b0.astype(np.float64)))
        assert abs(got - exact) <= 64 * eps * np.log2(n) * abs(exact), (dtype, got, exact)


def test_dot_allreduce_timeout_is_an_error_not_a_nan(orc):
    """A rank that never sends its partial: the fused all-reduce gives up after the configured bound,
    and b200_dot reports B200_ETIMEOUT instead of handing back a NaN that looks like data.  One GPU is enough:
    rank 1's mailbox is a second allocation on the same device and nobody ever writes to it."""
    import ctypes as C
    n = 4096
    bs.set_p2p_timeout_ms(50.0)
    try:
        mine, silent = bs.Mailbox(), bs.Mailbox()
        with bs.B200Stream(n, dtype=np.float64) as s:
            alone = s.dot()
            s.join_group(2, 0, [mine.ptr, silent.ptr])
            with pytest.raises(bs.B200Error) as e:
                s.dot()
            assert e.value.code == bs.ETIMEOUT and "timed out" in str(e.value)
            with pytest.raises(bs.B200Error):   # and again: the status word was cleared, the next call times out afresh
                s.dot()
            s.join_group(1, 0, [])
            assert s.dot().tobytes() == alone.tobytes(), "leaving the group restores the single-GPU dot"
        mine.close(); silent.close()
    finally:
        bs.set_p2p_timeout_ms(2000.0)

@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_managed_memory_mode(orc, dtype):
    """mem=managed (the reference's MEM=MANAGED build, src/cuda/model.cmake:6-10): same bits; get_arrays
    returns the managed ranges themselves (CUDAStream.cu:217-221) and the kernels keep working afterwards."""
    u = np.uint64 if dtype is np.float64 else np.uint32
    for n in (1000, (1 << 22) + 17):
        with bs.B200Stream(n, dtype=dtype, mem="managed") as s:
            assert "MANAGED" in s.describe()
            total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 3)
            a, b, c = s.get_arrays()
            g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, 3, n)
            for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
                assert (arr.view(u) == np.asarray(gold).view(u)).all(), n
            assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 3, a, b, c, total) == 0
            a2, _, _ = s.get_arrays()
            assert a2.ctypes.data == a.ctypes.data          # the same range, no mirror
            # pages are on the CPU now: the next kernels bring them home and continue the recurrence
            drive(s, orc, O.ALL, O.ORDER_CLASSIC, 2)
            a, b, c = s.get_arrays()
            g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, 5, n)
            for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
                assert (arr.view(u) == np.asarray(gold).view(u)).all(), n
            ra, rb, rc = s.read_arrays()
            assert np.array_equal(ra.view(u), a.view(u)) and np.array_equal(rc.view(u), c.view(u))
            with pytest.raises(bs.B200Error):
                s.set_option("mem", "device")   # fixed at allocation time
    with pytest.raises(bs.B200Error):
        bs.B200Stream(16, mem="bogus")
    with pytest.raises(bs.B200Error):
        bs.B200Stream(16, nonsense="1")


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_pagefault_memory_mode(orc, dtype):
    """mem=pagefault (the reference's MEM=PAGEFAULT build, src/cuda/CUDAStream.cu:39-40,217-221): plain system
    allocations reached by the GPU through page faults.  Same bits as mem=device; get_arrays returns the ranges.
    Where the device cannot touch pageable memory the constructor must refuse (no silent fallback)."""
    u = np.uint64 if dtype is np.float64 else np.uint32
    try:
        probe = bs.B200Stream(1000, dtype=dtype, mem="pagefault")
    except bs.B200Error as e:
        assert e.code == -6 and "pageable" in str(e)     # B200_EUNSUPPORTED
        pytest.skip("device has no pageable memory access (no HMM/ATS on this box)")
    probe.close()
    for n in (1000, (1 << 21) + 17):
        with bs.B200Stream(n, dtype=dtype, mem="pagefault") as s:
            assert "PAGEFAULT" in s.describe()
            total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 3)
            a, b, c = s.get_arrays()
            g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, 3, n)
            for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
                assert (arr.view(u) == np.asarray(gold).view(u)).all(), n
            assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 3, a, b, c, total) == 0
            a2, _, _ = s.get_arrays()
            assert a2.ctypes.data == a.ctypes.data and a.ctypes.data % (2 << 20) == 0
            drive(s, orc, O.ALL, O.ORDER_CLASSIC, 2)      # the CPU touched the pages: the kernels fault them back
            a, b, c = s.get_arrays()
            g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, 5, n)
            for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
                assert (arr.view(u) == np.asarray(gold).view(u)).all(), n
            ra, _, rc = s.read_arrays()
            assert np.array_equal(ra.view(u), a.view(u)) and np.array_equal(rc.view(u), c.view(u))
            with pytest.raises(bs.B200Error):
                s.set_option("mem", "managed")


def test_constructor_errors():
    with pytest.raises(bs.B200Error) as e:
        bs.B200Stream(1024, device=99)
    assert "Invalid device index" in str(e.value)  # the reference's message, CUDAStream.cu:88
    with pytest.raises(bs.B200Error) as e:
        bs.B200Stream(1 << 40, dtype=np.float64)
    assert "enough memory" in str(e.value)         # CUDAStream.cu:139
    with pytest.raises(bs.B200Error):
        bs.B200Stream(0)
    with pytest.raises(TypeError):
        bs.B200Stream(16, dtype=np.int32)


def test_bad_option_values_are_refused_not_mapped_to_auto(orc):
    """A benchmark that silently runs another configuration than the one asked for prints wrong tables: every
    misspelt value is B200_EINVAL with the accepted list (the reference refuses unknown flags, src/main.cpp:423-603)."""
    n = 1 << 16
    with bs.B200Stream(n, dtype=np.float64) as s:
        before = s.describe()
        for key, bad, word in (("copy_variant", "tmaa", "ldst"), ("dot_variant", "tickets", "two_pass"),
                               ("dot_combine", "ncl", "nccl"), ("sync", "doorbel", "doorbell"), ("contract", "fats", "off"),
                               ("timing", "yes", "off"), ("sweep", "alt", "forward"), ("launch", "graphs", "direct"),
                               ("tune", "triad:vec=16,unrol=4", "unknown field"), ("tune", "triad:vec=1x6", "not an integer"),
                               ("tune", "triad:vec", "field=value"), ("tune", "tirad:vec=16", "unknown kernel"),
                               ("tune", "triad:vec=24", "vec_bytes"), ("tune", "dot:block=1024", "block"),
                               ("tune", "triad", "<kernel>:"), ("nosuchkey", "1", "unknown option")):
            with pytest.raises(bs.B200Error) as e:
                s.set_option(key, bad)
            assert word in str(e.value), (key, bad, str(e.value))
        assert s.describe() == before, "a refused option must leave the configuration as it was"
        total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 2)
        a, b, c = s.get_arrays()
        assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 2, a, b, c, total) == 0
    with pytest.raises(bs.B200Error, match="copy_variant must be"):
        bs.B200Stream(n, dtype=np.float64, copy_variant="tmaa")


def test_dot_survives_a_refused_tune(orc):
    """a dot shape the launcher would refuse (block 1024, a grid beyond the workspace) used to be accepted by
    set_option and to fail at the next dot, after which the host waited for a doorbell value the device never reached.
    Such shapes now fail at set time and leave the context as it was; the host's doorbell count also follows the
    device's only when a launch went through."""
    n = 1 << 18
    with bs.B200Stream(n, dtype=np.float64) as s:
        s.init_arrays()
        good = s.dot()
        for bad in ("dot:block=1024", "dot:hint=9", "dot:vec=32,unroll=8"):
            with pytest.raises(bs.B200Error):
                s.set_option("tune", bad)
            assert bits(np.float64(s.dot())) == bits(np.float64(good)), bad
        s.set_option("tune", "dot:vec=16,unroll=2,block=256")  # a legal one still goes through, and dot keeps working
        for _ in range(3):
            got = s.dot()
            assert abs(float(got) - float(good)) <= 64 * np.finfo(np.float64).eps * abs(float(good))


needs_bin = pytest.mark.skipif(not os.path.exists(BIN), reason="build/b200-stream not built (needs the reference tree)")


CMAKE_BIN = os.path.join(ROOT, "build", "cmake", "b200-stream")


def run_bin(*args, env=None, exe=BIN):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([exe, *args], capture_output=True, text=True, timeout=600, env=e)


@pytest.mark.parametrize("exe", [BIN, CMAKE_BIN], ids=["makefile", "cmake-overlay"])
def test_driver_ci_sanity_case(exe):
    """The one run-time check of the reference's CI (src/ci-test-compile.sh:143-147), on our model -- for the
    Makefile binary and for the one registered through the reference's cmake/register_models.cmake."""
    if not os.path.exists(exe):
        pytest.skip(f"{os.path.relpath(exe, ROOT)} not built (needs the reference tree" + (" and cmake)" if exe == CMAKE_BIN else ")"))
    p = run_bin("-s", "1048576", "-n", "10", exe=exe)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "Implementation: B200" in p.stdout and "Triad" in p.stdout and "FAILED" not in p.stderr
    # the banner lines of the reference's model (src/cuda/CUDAStream.cu:103-121)
    assert "ECC is O" in p.stdout and "Memory Allocation: DEFAULT" in p.stdout, p.stdout
    q = run_bin("-s", "1048576", "-n", "10", "--only", "All", "--float", exe=exe)
    assert q.returncode == 0, q.stdout + q.stderr


@needs_bin
@pytest.mark.parametrize("args", [
    ["--float"], ["--only", "Triad"], ["--only", "Nstream"], ["--only", "All"], ["--only", "Dot"],
    ["--order", "Isolated"], ["--only", "All", "--order", "Isolated", "--float", "-n", "4"], ["-s", "1000003"],
    ["-s", "1", "-n", "2"],
])
def test_driver_cli_matrix(args):
    """CLI flags of src/main.cpp:423-603 reach the model unchanged; check_solution decides the exit code."""
    base = [] if "-s" in args else ["-s", "2097152"]
    base += [] if "-n" in args else ["-n", "10"]
    p = run_bin(*base, *args)
    assert p.returncode == 0, p.stdout + p.stderr


@needs_bin
def test_driver_managed_memory():
    p = run_bin("-s", str((1 << 22) + 3), "-n", "10", "--only", "All", env={"BABELSTREAM_B200_MEM": "managed"})
    assert p.returncode == 0, p.stdout + p.stderr
    assert "Memory Allocation: MANAGED" in p.stdout


@needs_bin
def test_driver_pagefault_memory_or_loud_refusal():
    """BABELSTREAM_B200_MEM=pagefault: validates like any other run where the GPU can touch pageable memory,
    otherwise the constructor throws (the reference's convention, CUDAStream.cu:87-88) -- never a fallback."""
    p = run_bin("-s", str((1 << 21) + 3), "-n", "10", "--only", "All", env={"BABELSTREAM_B200_MEM": "pagefault"})
    if p.returncode == 0:
        assert "PAGEFAULT" in p.stdout
    else:
        assert "mem=pagefault needs GPU access to pageable system memory" in p.stderr, p.stdout + p.stderr


@needs_bin
def test_driver_csv_and_contract_off():
    p = run_bin("-s", "4194304", "-n", "20", "--csv", "--only", "All", env={"BABELSTREAM_B200_CONTRACT": "off"})
    assert p.returncode == 0, p.stdout + p.stderr
    rows = [l.split(",") for l in p.stdout.splitlines() if l and l.split(",")[0] in ("Copy", "Mul", "Add", "Triad", "Dot", "Nstream")]
    assert [r[0] for r in rows] == ["Copy", "Mul", "Add", "Triad", "Dot", "Nstream"]
    assert all(int(r[2]) == 4194304 and int(r[3]) == 8 and float(r[4]) > 0 for r in rows)
    header = [l for l in p.stdout.splitlines() if l.startswith("function,")]
    assert header and header[0].split(",")[4] == "max_MB_per_sec"


@needs_bin
@pytest.mark.parametrize("args", [
    ["-s", str(1 << 29), "-n", "3", "--only", "All"],    # the north-star size: 3 x 4.29 GB
    ["-s", str(1 << 30), "-n", "3", "--only", "Triad"],  # BASELINE config 4 on one GPU: 3 x 8.59 GB
])
def test_driver_full_sizes_pass_check_solution(args):
    """BASELINE.json's full sizes through the size-independent property the domain offers: the reference's
    own check_solution (src/main.cpp:331-421) walks EVERY element of a, b, c against the gold recurrence
    and checks dot against goldA*goldB*n; the exit code is its verdict."""
    p = run_bin(*args)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "FAILED validation" not in p.stderr and "Triad" in p.stdout


@needs_bin
def test_driver_list_and_bad_device():
    p = run_bin("--list")
    assert p.returncode == 0 and "0:" in p.stdout
    p = run_bin("--device", "99", "-s", "1024", "-n", "2")
    assert p.returncode != 0 and "Invalid device index" in p.stderr


@needs_bin
def test_driver_refuses_a_misspelt_option():
    p = run_bin("-s", "65536", "-n", "2", env={"BABELSTREAM_B200_COPY_VARIANT": "tmaa"})
    assert p.returncode != 0 and "copy_variant must be" in (p.stderr + p.stdout), p.stdout + p.stderr


multi = pytest.mark.skipif(bs.lib() is None or bs.device_count() < 2, reason="needs >= 2 GPUs")


@multi
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("combine", ["nccl", "host", "p2p"])
def test_multi_gpu_sharded_bit_exact(orc, dtype, combine):
    os.environ["BABELSTREAM_B200_DOT_COMBINE"] = combine
    try:
        top = min(bs.device_count(), 8)
        for ngpus in sorted({2, min(3, top), top}):           # smallest, a non-power-of-two, all of them
            for n in (ngpus - 1, 1000, (1 << 22) + 17):      # fewer elements than GPUs, ragged, ragged large
                nt = 4
                with bs.B200Stream(n, dtype=dtype, ngpus=ngpus) as s:
                    assert s.ngpus == ngpus
                    total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, nt)
                    a, b, c = s.get_arrays()
                    g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, nt, n)
                    u = np.uint64 if dtype is np.float64 else np.uint32
                    for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
                        assert (arr.view(u) == np.asarray(gold).view(u)).all(), (ngpus, n)
                    assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, nt, a, b, c, total) == 0
                    assert abs(float(total) - float(g[3])) <= 64 * np.finfo(dtype).eps * abs(float(g[3]))
                    t2 = s.dot()
                    t3 = s.dot()
                    assert t2.tobytes() == t3.tobytes()
    finally:
        os.environ.pop("BABELSTREAM_B200_DOT_COMBINE", None)


@multi
def test_multi_gpu_threading_options(orc):
    """In-process layout: the caller drives GPU 0, pinned workers drive the others; pin/launch/sweep change how the
    calls are issued, never the bits.  The default dot combine inside one process is the host sum."""
    ngpus = min(bs.device_count(), 8)
    n = (1 << 21) + 5
    ref = None
    for opts in ({}, {"pin": "off"}, {"launch": "direct"}, {"sweep": "forward", "pin": "off"}, {"dot_combine": "p2p"}):
        with bs.B200Stream(n, dtype=np.float64, ngpus=ngpus, **opts) as s:
            text = s.describe()
            assert "the caller drives GPU 0" in text
            assert ("unpinned" in text) == (opts.get("pin") == "off") or "unpinned" in text   # pinning may be refused by the box
            if "dot_combine" not in opts:
                assert "host sum in rank order" in text
            total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 5)
            a, b, c = s.get_arrays()
            cur = (a.tobytes(), b.tobytes(), c.tobytes(), np.asarray(total).tobytes())
            ref = ref or cur
            assert cur == ref, opts
            assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 5, a, b, c, total) == 0
            with pytest.raises(bs.B200Error):
                s.set_option("pin", "off")        # fixed when the worker threads start
    with pytest.raises(bs.B200Error):
        bs.B200Stream(n, dtype=np.float64, ngpus=ngpus, pin="sometimes")


@multi
def test_two_contexts_do_not_share_worker_cpus():
    """the worker threads of two multi-GPU contexts living side by side in one process are pinned to different
    CPUs (a process-wide registry, released by destroy), instead of both contexts choosing the same cores."""
    import re

    def cpus(s):
        return [int(c) for c in re.findall(r"\d+->cpu(\d+)", s.describe())]

    n = 1 << 20
    with bs.B200Stream(n, dtype=np.float64, ngpus=2) as s:
        first = cpus(s)
        with bs.B200Stream(n, dtype=np.float64, ngpus=2) as d:
            second = cpus(d)
            if first and second:   # pinning may be unavailable (fewer CPUs than threads, restricted affinity)
                assert not set(first) & set(second), (first, second)
        with bs.B200Stream(n, dtype=np.float64, ngpus=2) as e:   # d's CPUs were released
            assert cpus(e) == second or not second


@multi
@pytest.mark.parametrize("combine", ["p2p", "host"])
def test_multi_gpu_managed_memory(orc, combine):
    """mem=managed over several GPUs: one managed range per array, shards on 2 MiB boundaries."""
    top = min(bs.device_count(), 8)
    for ngpus in sorted({2, top}):
        for n in (1000, (1 << 23) + 17):
            with bs.B200Stream(n, dtype=np.float64, ngpus=ngpus, mem="managed", dot_combine=combine) as s:
                total = drive(s, orc, O.ALL, O.ORDER_CLASSIC, 3)
                a, b, c = s.get_arrays()
                g = orc.gold(np.float64, O.ALL, O.ORDER_CLASSIC, 3, n)
                for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
                    assert (arr.view(np.uint64) == np.asarray(gold).view(np.uint64)).all(), (ngpus, n)
                assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 3, a, b, c, total) == 0
                if n > (1 << 23):
                    info = [s.shard_info(g_) for g_ in range(ngpus)]
                    assert all(i["begin"] * 8 % (2 << 20) == 0 for i in info if i["count"])  # trailing shards may be empty
                    assert sum(i["count"] for i in info) == n


@multi
def test_dot_combine_modes_agree_and_switch_at_run_time(orc):
    """nccl, p2p (fused peer-memory all-reduce inside the dot kernel) and host combine the same G partials:
    p2p and host add them in rank order and must agree bit for bit; nccl within a few ulp."""
    n, ngpus = (1 << 23) + 5, min(bs.device_count(), 8)
    rng = np.random.default_rng(1)
    a0, b0 = np.abs(rng.standard_normal(n)), np.abs(rng.standard_normal(n))
    with bs.B200Stream(n, dtype=np.float64, ngpus=ngpus, dot_combine="host") as s:
        s.write_arrays(a0, b0, None)
        host = s.dot()
        s.set_option("dot_combine", "p2p")
        p2p = [s.dot() for _ in range(5)]
        s.set_option("dot_combine", "nccl")
        nccl = s.dot()
        s.set_option("dot_combine", "p2p")
        again = s.dot()
        assert "peer-memory" in s.describe()
    assert all(x.tobytes() == host.tobytes() for x in p2p + [again])
    exact = float(np.dot(a0, b0))
    assert abs(float(host) - exact) <= 64 * np.finfo(np.float64).eps * exact
    assert abs(float(nccl) - float(host)) <= 8 * np.finfo(np.float64).eps * exact


@multi
@needs_bin
def test_driver_multi_gpu():
    for combine in ("nccl", "p2p", "host"):
        p = run_bin("-s", str((1 << 24) + 3), "-n", "10", "--only", "All",
                    env={"BABELSTREAM_B200_NGPUS": "2", "BABELSTREAM_B200_DOT_COMBINE": combine})
        assert p.returncode == 0, p.stdout + p.stderr
        assert "GPUs: 2" in p.stdout


GROUP_WORKER = os.path.join(ROOT, "tests", "group_worker.py")


@multi
@pytest.mark.parametrize("world", [2, 8])
def test_process_group_fused_dot_allreduce(world):
    """One process per GPU (the bench.py / torchrun layout): mailboxes exchanged as CUDA IPC handles,
    b200_join_group, then every rank's dot() is the all-rank sum from the fused peer-memory all-reduce:
    identical bits on all ranks, equal to the rank-order sum of the per-rank partials."""
    import sys
    world = min(world, bs.device_count())
    port = 29600 + world
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), GROUP_WORKER],
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert f"group ok world={world}" in p.stdout, p.stdout[-3000:]

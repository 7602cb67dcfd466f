"""CPU tests (no GPU, no compute calls): the C ABI loads and exports every declared symbol, the host
logic behind it (shard partition, launch-shape resolution, error behaviour), the build glue, and the
N > 1 plumbing with world_size-2 gloo."""
import os
import re
import socket
import sys

import numpy as np
import pytest

import babelstream_b200 as bs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_stream.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 35 and "b200_launch_triad" in names and "b200_create" in names
    L = bs.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert b"sm_100a" in L.b200_version()


def test_python_binding_covers_the_header():
    """stream.py binds (argtypes/restype) every function the header declares."""
    L = bs.lib()
    unbound = [n for n in declared_symbols() if getattr(L, n).argtypes is None]
    assert not unbound, unbound


@pytest.mark.skipif(bs.device_count() > 0, reason="only meaningful on a machine without a GPU")
def test_no_cpu_fallback():
    with pytest.raises(bs.B200Error) as e:
        bs.B200Stream(1024)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(bs.B200Error):
        bs.launch("triad", 64, 64, 64, 16, 8)


def test_missing_library_fails_loudly(monkeypatch):
    import babelstream_b200.stream as st
    monkeypatch.setattr(st, "_lib", None)
    monkeypatch.setenv("BABELSTREAM_B200_LIB", "/nonexistent/libb200stream.so")
    with pytest.raises(bs.B200Error) as e:
        st.lib()
    assert "no CPU fallback" in str(e.value)
    monkeypatch.delenv("BABELSTREAM_B200_LIB")
    monkeypatch.setattr(st, "_lib", None)
    assert st.lib() is not None


@pytest.mark.parametrize("elem_bytes", [4, 8])
def test_shard_range_partitions_exactly(elem_bytes):
    align = 128 // elem_bytes
    for n in [1, 2, 15, 16, 17, 1000, 4097, 1 << 20, (1 << 25) + 3, 1 << 31, (1 << 33) + 5]:
        for g in range(1, 9):
            pos = 0
            sizes = []
            for r in range(g):
                b, c = bs.shard_range(n, elem_bytes, g, r)
                assert b == pos, "shards must be contiguous and in rank order"
                assert b % align == 0 or c == 0, "every non-empty shard starts on a 128-byte boundary"
                pos += c
                sizes.append(c)
            assert pos == n
            assert max(sizes) - min(s for s in sizes if s or True) <= max(sizes)  # trailing shards may be short
            if n % (g * align) == 0:
                assert len(set(sizes)) == 1, "divisible sizes split evenly (all BASELINE configs are powers of two)"
    with pytest.raises(bs.B200Error):
        bs.shard_range(10, 8, 0, 0)
    with pytest.raises(bs.B200Error):
        bs.shard_range(10, 8, 2, 2)
    with pytest.raises(bs.B200Error):
        bs.shard_range(10, 3, 2, 0)


def test_resolve_cfg_defaults_and_overrides():
    big, grid = bs.resolve_cfg("triad", 1 << 29, 8)
    assert big.variant == bs.VARIANTS["ldst"] and big.contract == 1 and big.cache_hint != bs.HINTS["default"]
    tile = big.block * big.unroll * (big.vec_bytes // 8)
    assert grid == ((1 << 29) + tile - 1) // tile or grid % 148 == 0
    small, _ = bs.resolve_cfg("triad", 1 << 20, 8)   # 3 x 8 MB lives in the 126 MB L2
    assert small.cache_hint == bs.HINTS["default"], "no evict-first / no-allocate hints for an L2-resident working set"
    dot, grid = bs.resolve_cfg("dot", 1 << 25, 8)
    assert dot.variant in (1, 2) and 7 * 148 < grid <= 8 * 148, "8 blocks per SM asked for, rounded to whole chunks of tiles"
    o, grid = bs.resolve_cfg("copy", 1 << 25, 4, cfg=bs.LaunchCfg.make(vec_bytes=16, unroll=8, block=128, grid=-3,
                                                                      cache_hint="stream", variant="bulk", contract="off"))
    assert (o.vec_bytes, o.unroll, o.block, o.grid, o.cache_hint, o.variant, o.contract) == (16, 8, 128, -3, 4, 2, 2)
    assert grid == 3 * 148
    _, grid = bs.resolve_cfg("add", 1000, 8, num_sms=148, cfg=bs.LaunchCfg.make(grid=-8, vec_bytes=16, unroll=1, block=256))
    assert grid == 2, "never more blocks than tiles (500 16-byte vectors / 256 threads = 2 tiles)"
    scalar, _ = bs.resolve_cfg("mul", 1000, 4, cfg=bs.LaunchCfg.make(vec_bytes=4))
    assert scalar.unroll == 4 and scalar.cache_hint == 1, "the float scalar path has one compiled shape"
    wide, _ = bs.resolve_cfg("triad", 1 << 29, 8, cfg=bs.LaunchCfg.make(block=1024))
    assert wide.block == 1024, "1024-thread blocks are allowed while vec_bytes*unroll <= 32"


@pytest.mark.parametrize("bad", [dict(vec_bytes=24), dict(unroll=3), dict(unroll=8, vec_bytes=32), dict(block=100),
                                 dict(block=1024, unroll=4), dict(block=2048), dict(cache_hint=9), dict(contract=5)])
def test_resolve_cfg_rejects(bad):
    with pytest.raises(bs.B200Error):
        bs.resolve_cfg("triad", 1 << 20, 8, cfg=bs.LaunchCfg.make(**bad))


def test_resolve_cfg_size_classes():
    """The per-size table of csrc/tuning.cu: four classes by the footprint of the three arrays against L2."""
    l2 = 126 << 20
    shapes = {}
    for log2n in range(18, 31):
        t, _ = bs.resolve_cfg("triad", 1 << log2n, 8, num_sms=148, l2_bytes=l2)
        shapes[log2n] = (t.vec_bytes, t.unroll, t.block, t.cache_hint)
    assert shapes[20] == shapes[22] and shapes[20][3] == bs.HINTS["default"], "L2-resident: default cache policy"
    assert shapes[23][3] == bs.HINTS["default"] and shapes[24][3] == bs.HINTS["default"], "knee: keep the lines in L2"
    assert shapes[25][:3] == shapes[26][:3] and shapes[27] == shapes[29], "the tile grows between 2^26 and 2^27 (fp64)"
    assert shapes[25][3] == bs.HINTS["evict_first"], "a few times L2: evict-first body, the tail of the sweep kept"
    assert shapes[26][3] == bs.HINTS["default"] and shapes[29][3] == bs.HINTS["noalloc"], "larger: one L2 priority for all lines"
    keep = {k: bs.resolve_cfg("add", 1 << k, 8, num_sms=148, l2_bytes=l2)[0].keep_mb for k in (22, 24, 25, 26, 29)}
    assert keep == {22: 0, 24: 0, 25: 128, 26: 0, 29: 0}, "only the evict-first class pins the end of a sweep"
    tile = lambda s: s[0] * s[1] * s[2]
    assert tile(shapes[25]) < tile(shapes[29]), "short HBM-bound calls run smaller tiles (ramp-up and tail) than 2^29"
    # the class depends on BYTES, not elements: fp32 at 2n is the same class as fp64 at n
    for log2n in range(19, 30):
        d, _ = bs.resolve_cfg("add", 1 << log2n, 8, num_sms=148, l2_bytes=l2)
        f, _ = bs.resolve_cfg("add", 1 << (log2n + 1), 4, num_sms=148, l2_bytes=l2)
        assert (d.vec_bytes, d.unroll, d.block, d.cache_hint) == (f.vec_bytes, f.unroll, f.block, f.cache_hint)
    # flags travel through the resolver untouched
    o, _ = bs.resolve_cfg("triad", 1 << 25, 8, cfg=bs.LaunchCfg.make(flags=bs.LAUNCH_PDL))
    assert o.flags == bs.LAUNCH_PDL


def test_resolve_cfg_kernel_specific_rules():
    with pytest.raises(bs.B200Error):
        bs.resolve_cfg("triad", 1 << 20, 8, cfg=bs.LaunchCfg.make(variant="tma"))   # only copy has async variants
    bs.resolve_cfg("copy", 1 << 20, 8, cfg=bs.LaunchCfg.make(variant="tma"))
    bs.resolve_cfg("dot", 1 << 20, 8, cfg=bs.LaunchCfg.make(block=512, vec_bytes=16, unroll=2))
    with pytest.raises(bs.B200Error):   # dot blocks stop at 512 threads (three of them per SM, register-capped)
        bs.resolve_cfg("dot", 1 << 20, 8, cfg=bs.LaunchCfg.make(block=1024, vec_bytes=16, unroll=2))
    with pytest.raises(bs.B200Error):
        bs.resolve_cfg("dot", 1 << 20, 2)


def test_gen_stream_models_inserts_exactly_the_b200_branch():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_stream_models as g
    fake = ('#pragma once\n#include <memory>\n\n#if defined(CUDA)\n#include "CUDAStream.h"\n#elif defined(OMP)\n#include "OMPStream.h"\n#endif\n\n'
            'template <typename T, typename...Args>\nstd::unique_ptr<Stream<T>> make_stream(Args... args) {\n#if defined(CUDA)\n'
            '  // Use the CUDA implementation\n  return std::make_unique<CUDAStream<T>>(args...);\n#else\n  #error unknown benchmark\n#endif\n}\n')
    out = g.transform(fake)
    assert out.count("B200Stream") == 2 and out.count("defined(B200)") == 2
    # removing the inserted lines gives the input back: nothing else changed
    stripped = out.replace('#if defined(B200)\n#include "B200Stream.h"\n#elif defined(CUDA)', "#if defined(CUDA)")
    stripped = stripped.replace("#if defined(B200)\n  // Use the B200 (sm_100a) implementation\n  return std::make_unique<B200Stream<T>>(args...);\n\n#elif defined(CUDA)", "#if defined(CUDA)")
    assert stripped == fake
    with pytest.raises(SystemExit):
        g.transform("nothing to anchor on")
    ref = "/root/reference/src/StreamModels.h"
    if os.path.exists(ref):
        real = g.transform(open(ref).read())
        assert real.count("B200Stream") == 2 and "CUDAStream" in real and "SerialStream" in real


def test_bench_csv_parser_and_step_metric():
    sys.path.insert(0, ROOT)
    import bench
    text = ("BabelStream\nfunction,num_times,n_elements,sizeof,max_GB_per_sec,min_runtime,max_runtime,avg_runtime\n"
            "Copy,100,1024,8,10.0,0.001,0.002,0.0015\nMul,100,1024,8,10.0,0.001,0.002,0.0015\nAdd,100,1024,8,10.0,0.001,0.002,0.0015\n"
            "Triad,100,1024,8,12.5,0.001,0.002,0.0015\nDot,100,1024,8,10.0,0.001,0.002,0.0015\nNstream,100,1024,8,10.0,0.001,0.002,0.0015\n")
    rows = bench.parse_stream_csv(text)
    assert list(rows) == ["copy", "mul", "add", "triad", "dot", "nstream"] and rows["triad"]["gbs"] == 12.5
    gbs, t = bench.step_gbs_from_csv(rows, 1024)
    assert abs(t - 0.009) < 1e-12 and abs(gbs - 16 * 8 * 1024 / 0.009 / 1e9) < 1e-12
    assert bench.STEP_WEIGHT == 16 and bench.WEIGHT == {k: bs.WEIGHTS[k] for k in bench.KERNEL_ORDER}
    assert bench.KERNEL_ORDER == bs.KERNELS


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "omp-stream")), reason="oracle/_ref not built")
def test_bench_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference`: the reference's own omp-stream on the host cores, one JSON line with the
    same metric / unit / config keys as the b200 arm (the driver computes the ratio itself)."""
    import json
    import subprocess
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                        "--no-extras"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    sys.path.insert(0, ROOT)
    import bench
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == bench.UNIT and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["dtype"] == "f64" and d["vs_baseline"] is None
    assert d["config"]["elements_per_gpu"] == 1 << bench.LOG2N and d["config"]["kernels"] == bench.KERNEL_ORDER
    assert d["config"]["bytes_per_step_per_gpu"] == 16 * 8 * (1 << bench.LOG2N)
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] > 0 and "OMPStream.cpp" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": bench.UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # both arms print the SAME config (key for key) and the same warm-up count, so that the driver's
    # same_config / same_steps checks hold
    assert d["config"] == bench.make_config(bench.LOG2N, 1) and d["warmup"] == 3
    # value = algorithmic bytes of a step / time of a step
    assert abs(d["value"] - 16 * 8 * (1 << bench.LOG2N) / (d["ms_per_step"] * 1e-3) / 1e9) < 0.05 * d["value"]


def test_bench_gold_recurrence_matches_the_oracle_bit_for_bit():
    """bench.py's parity key computes check_solution's gold values itself (exact rational fma); it must agree with
    the oracle's restatement of src/main.cpp:331-370 in both contraction modes."""
    sys.path.insert(0, ROOT)
    import bench
    import oracle as O
    orc = O.Oracle()
    for steps in (1, 2, 3, 10, 100):
        for contract in (True, False):
            g = orc.gold(np.float64, O.ALL, O.ORDER_CLASSIC, steps, 1 << 20, contract=contract)
            a, b, c, a_dot, b_dot = bench.gold_recurrence(steps, contract)
            assert [float(x).hex() for x in g[:3]] == [a.hex(), b.hex(), c.hex()], (steps, contract)
            assert (a_dot * b_dot * float(1 << 20)).hex() == float(g[3]).hex()


def test_report_tool_renders_sweep_csv(tmp_path):
    """tools/report.py: size-sweep CSV -> markdown with the L2 knee marked and a sample block in
    the layout of the reference's results/ files."""
    import subprocess
    rows = ["model,dtype,log2n,kernel,gbs,min_us,l2_resident"]
    for model, scale in (("b200", 1.0), ("ref_cuda", 0.8)):
        for l, res in ((22, 1), (23, 0), (24, 0)):
            for k, g in (("copy", 6000), ("mul", 6000), ("add", 6500), ("triad", 6600), ("dot", 6400), ("nstream", 7000)):
                rows.append(f"{model},f64,{l},{k},{g * scale},{100.0},{res}")
    f = tmp_path / "sweep.csv"
    f.write_text("\n".join(rows) + "\n")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "report.py"), str(f)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    out = p.stdout
    assert "| 22 | L2 |" in out and "| 23 | **knee -> HBM** |" in out and "| 24 | HBM |" in out
    assert "6600 / 5280 (1.25)" in out                       # triad b200 / ref (ratio)
    assert f"{100 * 6600 / 8000:.1f} / {100 * 6600 / 7672.3:.1f}" in out
    assert "Function    MBytes/sec  Min (sec)" in out and "Triad       6600000.000 0.00010" in out


# ---- N > 1 on CPU: world_size 2, gloo -----------------------------------------------------------------
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, n, q):
    """What one rank of the sharded path does, with the oracle standing in for the GPU kernels:
    its contiguous shard, no data-path collective for the elementwise kernels, ONE all-reduce of one
    element for dot, and the max-over-ranks reduction bench.py uses for timing."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import oracle as O
    import babelstream_b200 as bsx
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = O.Oracle()
    begin, count = bsx.shard_range(n, 8, world, rank)
    a, b, c = (np.empty(count, dtype=np.float64) for _ in range(3))
    orc.init(a, b, c)
    total = None
    for _ in range(3):
        orc.copy(a, c); orc.mul(b, c); orc.add(a, b, c); orc.triad(a, b, c)
        part = torch.tensor([float(orc.dot(a, b))], dtype=torch.float64)
        dist.all_reduce(part)                      # the path's only exchange step
        total = part.item()
        orc.nstream(a, b, c)
    ms = torch.tensor([10.0 + rank], dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)      # timing: max over ranks
    gathered = [None] * world
    dist.all_gather_object(gathered, (begin, count, a.tobytes(), b.tobytes(), c.tobytes()))
    if rank == 0:
        q.put((total, ms.item(), gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_path_matches_unsharded_oracle(orc):
    import torch.multiprocessing as mp
    import oracle as O
    n, world = 100003, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    total, ms, gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ms == 11.0
    a, b, c = (np.empty(n, dtype=np.float64) for _ in range(3))
    orc.init(a, b, c)
    want = None
    for _ in range(3):
        orc.copy(a, c); orc.mul(b, c); orc.add(a, b, c); orc.triad(a, b, c)
        want = float(orc.dot(a, b))
        orc.nstream(a, b, c)
    pos = 0
    for begin, count, ra, rb, rc in gathered:
        assert begin == pos
        for full, raw in ((a, ra), (b, rb), (c, rc)):
            assert np.array_equal(full[begin:begin + count].view(np.uint64), np.frombuffer(raw, dtype=np.uint64))
        pos += count
    assert pos == n
    assert abs(total - want) <= 1e7 * np.finfo(np.float64).eps * abs(want)   # src/main.cpp:384
    assert orc.check_solution(O.ALL, O.ORDER_CLASSIC, 3, a, b, c, np.float64(total)) == 0


def _mailbox_rank_main(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from babelstream_b200.group import connect_mailboxes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mb = connect_mailboxes(dist, rank, world)
    q.put((rank, mb is None, connect_mailboxes.last_error))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(bs.device_count() > 0, reason="only meaningful on a machine without a GPU")
def test_mailbox_exchange_fails_collectively_without_gpus():
    """bench.py's N > 1 bring-up: when no rank can create a mailbox (no device here), EVERY rank gets None
    (so all of them take the NCCL fallback together) and the reason names each rank."""
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_mailbox_rank_main, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [g[1] for g in got] == [True, True]
    assert all("rank 0" in g[2] and "rank 1" in g[2] for g in got)


def test_mailbox_bank_protocol_model():
    """A sequential model of the two-bank mailbox protocol of csrc/kernels.cuh, explored over random
    interleavings: per GPU the program is  send(1) collect(1) send(2) collect(2) ...  (collect k before
    send k+1, the documented rule); a collect may only run when all its slots carry its sequence number.
    Whatever the schedule, no slot is overwritten before every peer has read it, and no rank deadlocks."""
    import random
    for world in (2, 3, 8):
        for trial in range(200):
            rng = random.Random(world * 1000 + trial)
            steps = 6
            slots = [[[0] * world for _ in range(2)] for _ in range(world)]   # [owner][bank][sender] = seq
            read = [[[0] * world for _ in range(2)] for _ in range(world)]    # last seq the owner consumed
            pc = [0] * world                                                   # 2*(k-1) = send k, 2*(k-1)+1 = collect k
            while any(p < 2 * steps for p in pc):
                ready = []
                for g in range(world):
                    if pc[g] >= 2 * steps:
                        continue
                    k = pc[g] // 2 + 1
                    if pc[g] % 2 == 0 or all(slots[g][k & 1][s] == k for s in range(world)):
                        ready.append(g)
                assert ready, ("deadlock", world, trial, pc)
                g = rng.choice(ready)
                k = pc[g] // 2 + 1
                if pc[g] % 2 == 0:       # send k: overwrite my slot in everyone's bank k&1
                    for owner in range(world):
                        old = slots[owner][k & 1][g]
                        assert old == 0 or read[owner][k & 1][g] == old, ("overwrote an unread partial", world, trial)
                        slots[owner][k & 1][g] = k
                else:                    # collect k
                    for s in range(world):
                        assert slots[g][k & 1][s] == k
                        read[g][k & 1][s] = k
                pc[g] += 1


# ---------------------------------------------------------------------------------------------------
# The walk of a launch over its tiles (csrc/tuning.cu: plan_sweep), the host half of the alternating sweep
# ---------------------------------------------------------------------------------------------------
from hypothesis import given, settings, strategies as st  # noqa: E402

_shapes = st.sampled_from([(8, 1), (8, 2), (8, 4), (16, 1), (16, 2), (16, 4), (16, 8), (32, 1), (32, 2), (32, 4)])


@settings(max_examples=300, deadline=None)
@given(kernel=st.sampled_from(["copy", "mul", "add", "triad", "nstream", "init"]), log2n=st.integers(0, 33), jitter=st.integers(0, 4099),
       shape=_shapes, block=st.sampled_from([32, 128, 256, 512]), grid=st.sampled_from([0, -1, -4, 7, 1000]),
       reverse=st.booleans(), keep=st.sampled_from([0, -1, 1, 96, 4096]), hint=st.sampled_from([0, 1, 2, 3]), eb=st.sampled_from([4, 8]))
def test_elementwise_sweep_covers_every_tile_once(kernel, log2n, jitter, shape, block, grid, reverse, keep, hint, eb):
    n = (1 << log2n) + jitter
    cfg = bs.LaunchCfg.make(vec_bytes=shape[0], unroll=shape[1], block=block, grid=grid, cache_hint=hint, keep_mb=keep,
                            flags=bs.LAUNCH_REVERSE if reverse else 0)
    try:
        sp = bs.plan_sweep(kernel, n, eb, cfg=cfg)
    except bs.B200Error:
        return   # a shape the resolver refuses (covered by test_resolve_cfg_rejects)
    r, _ = bs.resolve_cfg(kernel, n, eb, cfg=cfg)
    tv = sp.tile_vecs
    assert tv == r.block * sp.unroll and sp.ntiles == -(-sp.nvec // tv) and sp.nvec == n * eb // sp.vec_bytes
    assert 1 <= sp.grid <= max(1, sp.ntiles) and sp.reverse == int(reverse)
    if sp.ntiles:
        # dispatch index p -> first vector origin + stride*p: p = 0 and p = ntiles-1 are the two ends, the map is affine,
        # so every tile start t*tv, t in [0, ntiles), is hit exactly once
        first, last = sp.origin, sp.origin + sp.stride * (sp.ntiles - 1)
        assert abs(sp.stride) == tv
        assert (first, last) == (((sp.ntiles - 1) * tv, 0) if reverse else (0, (sp.ntiles - 1) * tv))
    # the keep region: only with the evict-first hint, the LAST tiles in dispatch order, keep_mb MB of traffic (or everything)
    arrays = {"copy": 2, "mul": 2, "nstream": 4}.get(kernel, 3)
    none = (1 << 64) - 1
    if sp.cache_hint != bs.HINTS["evict_first"] or r.keep_mb <= 0 or sp.ntiles == 0:
        assert sp.keep_from == none
    else:
        kept = sp.ntiles - sp.keep_from
        want = r.keep_mb * 2 ** 20 / (tv * sp.vec_bytes * arrays)
        assert 0 <= sp.keep_from <= sp.ntiles and (kept == sp.ntiles or want - 1 < kept <= want)


@settings(max_examples=300, deadline=None)
@given(log2n=st.integers(0, 33), jitter=st.integers(0, 4099), shape=_shapes, block=st.sampled_from([32, 128, 256, 512]),
       grid=st.sampled_from([0, -1, -3, -8, -27, 5, 16384, 100000]), reverse=st.booleans(), stride_mode=st.booleans(),
       eb=st.sampled_from([4, 8]), sms=st.sampled_from([1, 132, 148]))
def test_dot_chunks_partition_the_tiles(log2n, jitter, shape, block, grid, reverse, stride_mode, eb, sms):
    n = (1 << log2n) + jitter
    flags = (bs.LAUNCH_REVERSE if reverse else 0) | (bs.LAUNCH_DOT_GRIDSTRIDE if stride_mode else 0)
    cfg = bs.LaunchCfg.make(vec_bytes=shape[0], unroll=shape[1], block=block, grid=grid, flags=flags)
    try:
        sp = bs.plan_sweep("dot", n, eb, num_sms=sms, cfg=cfg)
    except bs.B200Error as e:
        assert "block must be" in str(e) or "unroll must be" in str(e) or "exceeds the workspace capacity" in str(e), str(e)
        return
    _, grid_out = bs.resolve_cfg("dot", n, eb, num_sms=sms, cfg=cfg)
    assert grid_out == sp.grid, "b200_resolve_cfg reports the grid that is launched"
    assert 1 <= sp.grid <= 16384, "the partials must fit the workspace (kDotMaxBlocks)"
    if stride_mode or sp.ntiles == 0:
        assert sp.tiles_per_block == 0
        return
    k = sp.tiles_per_block
    assert k >= 1 and sp.grid * k >= sp.ntiles > (sp.grid - 1) * k, "every tile has exactly one owner and no block is empty"

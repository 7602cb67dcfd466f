"""GPU parity tests, LEVEL 1 of the C ABI: each CUDA kernel on raw device pointers against the oracle.

Bar: BIT-EXACT for init, copy, mul, add, triad and nstream in both rounding modes
(contract = fused: the reference's default -ffp-contract=fast build; unfused: -ffp-contract=off);
dot within BabelStream's own relative check (1e7*eps, src/main.cpp:384) -- in fact far tighter --
and bit-reproducible run to run.  torch is used for device memory and streams only.
"""
import math

import numpy as np
import pytest

import babelstream_b200 as bs
import oracle as O

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

DTYPES = [np.float64, np.float32]
EW = ["copy", "mul", "add", "triad", "nstream"]
SIZES = [1, 2, 3, 5, 31, 32, 33, 255, 1000, 4097, 65536 + 7, (1 << 20) + 3]


def tdtype(dt):
    return torch.float64 if np.dtype(dt) == np.float64 else torch.float32


def same_bits(x, y):
    u = np.uint32 if x.dtype == np.float32 else np.uint64
    return np.array_equal(np.ascontiguousarray(x).view(u), np.ascontiguousarray(y).view(u))


def rand(n, dtype, seed):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(n) * np.exp(rng.uniform(-4, 4, n))).astype(dtype)


def oracle_apply(orc, kern, a, b, c, contract):
    if kern == "copy": orc.copy(a, c)
    elif kern == "mul": orc.mul(b, c)
    elif kern == "add": orc.add(a, b, c)
    elif kern == "triad": orc.triad(a, b, c, contract=contract)
    elif kern == "nstream": orc.nstream(a, b, c, contract=contract)
    elif kern == "init": orc.init(a, b, c)


def gpu_apply(kern, a0, b0, c0, cfg=None, offset=(0, 0, 0), pad=16):
    """Run one kernel on device copies placed `offset` elements into padded buffers; returns host results
    plus the untouched padding (to catch out-of-bounds writes)."""
    n, dt = a0.size, a0.dtype
    dev = []
    for x, off in zip((a0, b0, c0), offset):
        buf = torch.full((n + 2 * pad + 8,), 7.0, dtype=tdtype(dt), device="cuda")
        view = buf[pad + off: pad + off + n]
        view.copy_(torch.from_numpy(x))
        dev.append((buf, view, pad + off))
    st = torch.cuda.current_stream().cuda_stream
    bs.launch(kern, dev[0][1].data_ptr(), dev[1][1].data_ptr(), dev[2][1].data_ptr(), n, dt.itemsize, cfg=cfg, stream=st)
    torch.cuda.synchronize()
    outs = []
    for buf, view, start in dev:
        full = buf.cpu().numpy()
        assert (full[:start] == 7.0).all() and (full[start + n:] == 7.0).all(), "kernel wrote outside its array"
        outs.append(full[start:start + n].copy())
    return outs


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kern", EW + ["init"])
def test_elementwise_auto_cfg_bit_exact(orc, kern, dtype):
    for n in SIZES:
        for contract in (True, False):
            a0, b0, c0 = rand(n, dtype, n), rand(n, dtype, n + 1), rand(n, dtype, n + 2)
            a, b, c = a0.copy(), b0.copy(), c0.copy()
            oracle_apply(orc, kern, a, b, c, contract)
            ga, gb, gc = gpu_apply(kern, a0, b0, c0, cfg=bs.LaunchCfg.make(contract=contract))
            assert same_bits(ga, a) and same_bits(gb, b) and same_bits(gc, c), (kern, n, contract)


CFG_GRID = [dict(vec_bytes=v, unroll=u, cache_hint=h, block=blk, grid=g)
            for v, us in ((8, (1, 2, 4)), (16, (1, 2, 4, 8)), (32, (1, 2, 4))) for u in us
            for h in (1, 2, 3, 4, 5) for blk, g in ((128, 0), (256, -2), (1024 if v * u <= 32 else 512, 3))]


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kern", EW + ["init"])
def test_elementwise_every_compiled_variant(orc, kern, dtype):
    """All (vector width, unroll, cache hint) instantiations x three grid shapes, ragged n."""
    n = 50021
    a0, b0, c0 = rand(n, dtype, 1), rand(n, dtype, 2), rand(n, dtype, 3)
    want = {}
    for contract in (True, False):
        a, b, c = a0.copy(), b0.copy(), c0.copy()
        oracle_apply(orc, kern, a, b, c, contract)
        want[contract] = (a, b, c)
    for kw in CFG_GRID:
        for contract in ((True, False) if kern in ("triad", "nstream") else (True,)):
            ga, gb, gc = gpu_apply(kern, a0, b0, c0, cfg=bs.LaunchCfg.make(contract=contract, **kw))
            a, b, c = want[contract]
            assert same_bits(ga, a) and same_bits(gb, b) and same_bits(gc, c), (kern, kw, contract)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kern", EW)
def test_elementwise_misaligned_pointers(orc, kern, dtype):
    """Congruent offsets take the vector path with a scalar head; incongruent ones the scalar fallback."""
    n = 10007
    a0, b0, c0 = rand(n, dtype, 4), rand(n, dtype, 5), rand(n, dtype, 6)
    a, b, c = a0.copy(), b0.copy(), c0.copy()
    oracle_apply(orc, kern, a, b, c, True)
    for offset in [(1, 1, 1), (3, 3, 3), (1, 2, 3), (0, 1, 0), (2, 0, 5)]:
        for vec in (16, 32):
            ga, gb, gc = gpu_apply(kern, a0, b0, c0, cfg=bs.LaunchCfg.make(vec_bytes=vec), offset=offset)
            assert same_bits(ga, a) and same_bits(gb, b) and same_bits(gc, c), (kern, offset, vec)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("variant", ["bulk", "tma"])
def test_copy_async_variants_bit_exact(variant, dtype):
    """cp.async.bulk and TMA-tensor copy: sizes around the chunk / row boundaries, several ring shapes."""
    eb = np.dtype(dtype).itemsize
    for stages, chunk in ((2, 2048), (4, 32768), (3, 16384), (8, 8192)):
        for nbytes in (8, 2048, 2048 * 3 + 16, chunk - 8, chunk, chunk + 8, 5 * chunk + 2048 + 24,
                       37 * chunk + 1024, (1 << 22) + 4096 + 8):
            n = max(1, nbytes // eb)
            a0 = rand(n, dtype, nbytes % 1000)
            junk = rand(n, dtype, 99)
            _, _, gc = gpu_apply("copy", a0, junk, junk, cfg=bs.LaunchCfg.make(variant=variant, stages=stages, chunk_bytes=chunk, grid=-1))
            assert same_bits(gc, a0), (variant, stages, chunk, n)
    # a misaligned pointer silently takes the register path and is still exact
    a0 = rand(5001, dtype, 3)
    _, _, gc = gpu_apply("copy", a0, a0, a0, cfg=bs.LaunchCfg.make(variant=variant), offset=(1, 0, 1))
    assert same_bits(gc, a0)


@pytest.mark.parametrize("dtype", DTYPES)
def test_ieee_edge_cases_match_cpu(orc, dtype):
    """Denormals (no flush-to-zero), signed zeros, infinities, overflow and NaN: the GPU result must
    carry the same bits as the CPU reference wherever that is a number, and be NaN wherever that is NaN
    (NaN payloads are the one thing x86 and NVIDIA hardware legitimately differ in)."""
    fi = np.finfo(dtype)
    tiny, big = fi.tiny, fi.max
    specials = np.array([0.0, -0.0, tiny, -tiny, tiny / 4, -tiny / 8, fi.smallest_subnormal, big, -big, big / 2,
                         np.inf, -np.inf, np.nan, 1.0, -1.0, 0.4, 2.5, 1e-30 if dtype is np.float32 else 1e-300],
                        dtype=dtype)
    grid = np.array(np.meshgrid(specials, specials, specials)).reshape(3, -1)
    a0, b0, c0 = (np.ascontiguousarray(grid[i]).astype(dtype) for i in range(3))
    u = np.uint32 if dtype is np.float32 else np.uint64
    for kern in EW:
        for contract in (True, False):
            a, b, c = a0.copy(), b0.copy(), c0.copy()
            with np.errstate(all="ignore"):
                oracle_apply(orc, kern, a, b, c, contract)
            ga, gb, gc = gpu_apply(kern, a0, b0, c0, cfg=bs.LaunchCfg.make(contract=contract))
            for want, got in ((a, ga), (b, gb), (c, gc)):
                nan = np.isnan(want)
                assert np.array_equal(np.isnan(got), nan), (kern, contract)
                assert np.array_equal(got.view(u)[~nan], want.view(u)[~nan]), (kern, contract)


def gpu_dot(a0, b0, cfg=None, repeats=1):
    n, dt = a0.size, a0.dtype
    a = torch.from_numpy(a0).cuda()
    b = torch.from_numpy(b0).cuda()
    ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
    out = torch.full((1,), float("nan"), dtype=tdtype(dt), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    res = []
    for _ in range(repeats):
        out.fill_(float("nan"))
        bs.launch_dot(a.data_ptr(), b.data_ptr(), n, dt.itemsize, ws.data_ptr(), out.data_ptr(), cfg=cfg, stream=st)
        torch.cuda.synchronize()
        res.append(out.cpu().numpy()[0])
    assert int(ws.cpu().numpy().view(np.uint32)[-64:].sum()) == 0, "dot must leave its ticket zeroed"
    return res


@pytest.mark.parametrize("dtype", DTYPES)
def test_dot_tolerance_and_determinism(orc, dtype):
    eps = np.finfo(dtype).eps
    for n in SIZES + [(1 << 22) + 11]:
        a0, b0 = np.abs(rand(n, dtype, n)), np.abs(rand(n, dtype, n + 9))
        exact = math.fsum(float(x) * float(y) for x, y in zip(a0[:200000], b0[:200000])) if n <= 200000 else \
            float(np.dot(a0.astype(np.float64), b0.astype(np.float64)))
        serial = float(orc.dot(a0, b0, contract=True))
        for contract in (True, False):
            r = gpu_dot(a0, b0, cfg=bs.LaunchCfg.make(contract=contract), repeats=3)
            assert r[0].tobytes() == r[1].tobytes() == r[2].tobytes(), "dot is not run-to-run deterministic"
            got = float(r[0])
            # the reference's own bar, src/main.cpp:384 ...
            assert abs(got - serial) <= 1e7 * eps * max(abs(got), abs(serial))
            # ... and the honest one: a tree reduction of positive terms is good to a few eps*log2(n)
            assert abs(got - exact) <= 4 * eps * max(4.0, math.log2(n)) * abs(exact), (n, got, exact)


@pytest.mark.parametrize("dtype", DTYPES)
def test_dot_variants_agree_bitwise(dtype):
    """two_pass and ticket run the same second pass: identical bits; every compiled shape is sane."""
    n = 300007
    a0, b0 = rand(n, dtype, 1), rand(n, dtype, 2)
    exact = float(np.dot(a0.astype(np.float64), b0.astype(np.float64)))
    scale = float(np.dot(np.abs(a0).astype(np.float64), np.abs(b0).astype(np.float64)))
    eps = np.finfo(dtype).eps
    for v, us in ((8, (1, 2, 4)), (16, (1, 2, 4, 8)), (32, (1, 2, 4))):
        for u in us:
            for h in (1, 2, 3, 4, 5):
                for blk, g, fl in ((256, -2, 0), (512, 7, bs.LAUNCH_DOT_GRIDSTRIDE), (128, -1, 0), (512, -40, 0)):
                    kw = dict(vec_bytes=v, unroll=u, cache_hint=h, block=blk, grid=g)
                    r1 = gpu_dot(a0, b0, cfg=bs.LaunchCfg.make(variant="two_pass", flags=fl, **kw))[0]
                    r2 = gpu_dot(a0, b0, cfg=bs.LaunchCfg.make(variant="ticket", flags=fl, **kw))[0]
                    assert r1.tobytes() == r2.tobytes(), kw
                    assert abs(float(r1) - exact) <= 8 * eps * 20 * scale, kw
                    # the sweep direction (and the L2 keep region) changes WHEN a chunk is read, never the bits
                    r3 = gpu_dot(a0, b0, cfg=bs.LaunchCfg.make(variant="ticket", flags=fl | bs.LAUNCH_REVERSE, keep_mb=1, **kw))[0]
                    assert r3.tobytes() == r2.tobytes(), kw


def test_dot_misaligned_and_empty(orc):
    for dtype in DTYPES:
        n = 7001
        a0, b0 = np.abs(rand(n, dtype, 1)), np.abs(rand(n, dtype, 2))
        at = torch.from_numpy(np.concatenate([np.zeros(3, dtype), a0])).cuda()
        bt = torch.from_numpy(np.concatenate([np.zeros(3, dtype), b0])).cuda()
        ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
        out = torch.zeros(1, dtype=tdtype(dtype), device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        for off_a, off_b in ((3, 3), (1, 1), (1, 2), (0, 3)):
            bs.launch_dot(at[off_a:].data_ptr(), bt[off_b:].data_ptr(), n - 3, a0.itemsize, ws.data_ptr(), out.data_ptr(), stream=st)
            torch.cuda.synchronize()
            want = float(np.dot(at[off_a:off_a + n - 3].cpu().numpy().astype(np.float64), bt[off_b:off_b + n - 3].cpu().numpy().astype(np.float64)))
            assert abs(float(out.item()) - want) <= 64 * np.finfo(dtype).eps * abs(want) + 1e-30
        bs.launch_dot(at.data_ptr(), bt.data_ptr(), 0, a0.itemsize, ws.data_ptr(), out.data_ptr(), stream=st)
        torch.cuda.synchronize()
        assert float(out.item()) == 0.0


def test_elementwise_n_zero_is_a_no_op():
    t = torch.ones(8, dtype=torch.float64, device="cuda")
    for k in EW + ["init"]:
        bs.launch(k, t.data_ptr(), t.data_ptr(), t.data_ptr(), 0, 8)
    torch.cuda.synchronize()
    assert (t == 1).all()


def test_bad_arguments_raise():
    t = torch.ones(64, dtype=torch.float64, device="cuda")
    with pytest.raises(bs.B200Error):
        bs.launch("triad", t.data_ptr(), t.data_ptr(), t.data_ptr(), 64, 2)  # elem_bytes
    with pytest.raises(bs.B200Error):
        bs.launch("triad", t.data_ptr(), t.data_ptr(), t.data_ptr(), 64, 8, cfg=bs.LaunchCfg.make(vec_bytes=64))
    with pytest.raises(bs.B200Error):
        bs.launch("triad", t.data_ptr(), t.data_ptr(), t.data_ptr(), 64, 8, cfg=bs.LaunchCfg.make(unroll=3))
    with pytest.raises(bs.B200Error):
        bs.launch("triad", t.data_ptr(), t.data_ptr(), t.data_ptr(), 64, 8, cfg=bs.LaunchCfg.make(variant="tma"))
    with pytest.raises(bs.B200Error):
        bs.launch("triad", 0, t.data_ptr(), t.data_ptr(), 64, 8)


@pytest.mark.parametrize("dtype", DTYPES)
def test_golden_random_fixture_on_gpu(golden_random, dtype):
    """The committed reference outputs (tests/golden/random_vectors.npz) straight against the CUDA path."""
    tag = np.dtype(dtype).name
    a0, b0, c0 = (golden_random[f"{tag}/in/{k}"] for k in "abc")
    for mode, contract in (("fast", True), ("off", False)):
        for kern in EW:
            ga, gb, gc = gpu_apply(kern, a0, b0, c0, cfg=bs.LaunchCfg.make(contract=contract))
            for name, got in (("a", ga), ("b", gb), ("c", gc)):
                assert same_bits(got, golden_random[f"{tag}/{mode}/{kern}/{name}"]), (kern, mode, name)


@pytest.mark.parametrize("dtype", DTYPES)
def test_full_size_properties(dtype):
    """BASELINE config size (2^25): size-independent properties instead of a CPU replay.
    copy is the identity; mul by 1 is the identity; add is commutative; triad(s=0) copies b;
    nstream(s) == a + triad(s); dot(a, 1) equals the tree sum in float64 within tolerance."""
    n = 1 << 25
    eb = np.dtype(dtype).itemsize
    g = torch.Generator(device="cuda").manual_seed(7)
    a = torch.randn(n, dtype=tdtype(dtype), device="cuda", generator=g)
    b = torch.randn(n, dtype=tdtype(dtype), device="cuda", generator=g)
    c = torch.empty_like(a)
    st = torch.cuda.current_stream().cuda_stream
    bs.launch("copy", a.data_ptr(), 0, c.data_ptr(), n, eb, stream=st)
    assert torch.equal(c, a)
    b2 = torch.empty_like(a)
    bs.launch("mul", 0, b2.data_ptr(), a.data_ptr(), n, eb, scalar=1.0, stream=st)
    assert torch.equal(b2, a)
    c2 = torch.empty_like(a)
    bs.launch("add", a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, stream=st)
    bs.launch("add", b.data_ptr(), a.data_ptr(), c2.data_ptr(), n, eb, stream=st)
    assert torch.equal(c, c2) and torch.equal(c, a + b)
    t = torch.empty_like(a)
    bs.launch("triad", t.data_ptr(), b.data_ptr(), a.data_ptr(), n, eb, scalar=0.0, stream=st)
    assert torch.equal(t, b)
    s = 0.4
    bs.launch("triad", t.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, scalar=s, stream=st)   # t = b + s*c
    a2 = a.clone()
    bs.launch("nstream", a2.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, scalar=s, stream=st)  # a2 = a + (b + s*c)
    assert torch.equal(a2, a + t)
    # fused triad against torch's own fma-free arithmetic: at most 1 ulp apart, and exactly addcmul in fp64
    ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
    out = torch.zeros(1, dtype=tdtype(dtype), device="cuda")
    ones = torch.ones_like(a)
    a_abs = a.abs()
    bs.launch_dot(a_abs.data_ptr(), ones.data_ptr(), n, eb, ws.data_ptr(), out.data_ptr(), stream=st)
    want = a_abs.double().sum().item()
    assert abs(out.item() - want) <= 64 * np.finfo(dtype).eps * want


def _spin_until(flag_np, seq, seconds=5.0):
    import time
    t0 = time.perf_counter()
    while int(flag_np[0]) != seq:
        if time.perf_counter() - t0 > seconds:
            return False
    return True


@pytest.mark.parametrize("dtype", DTYPES)
def test_doorbell_rings_after_the_results_are_complete(orc, dtype):
    """LEVEL 1 doorbell (b200_doorbell): the last block to finish stores seq to mapped pinned host memory.
    The host waits on that word ONLY (no synchronize before reading the flag); results stay bit-exact,
    the ticket returns to zero, and consecutive launches with increasing seq ring in order."""
    flag = torch.zeros(1, dtype=torch.int64).pin_memory()   # cudaHostAlloc: device-visible under UVA
    flag_np = flag.numpy()
    ticket = torch.zeros(64, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    seq = 0
    for n in (1, 33, 4097, (1 << 20) + 3, (1 << 23) + 5):
        a0, b0, c0 = rand(n, dtype, 1), rand(n, dtype, 2), rand(n, dtype, 3)
        for kern in EW + ["init"]:
            a, b, c = a0.copy(), b0.copy(), c0.copy()
            oracle_apply(orc, kern, a, b, c, True)
            da, db, dc = (torch.from_numpy(x).cuda() for x in (a0, b0, c0))
            torch.cuda.synchronize()
            seq += 1
            bell = bs.Doorbell(ticket.data_ptr(), flag.data_ptr(), seq)
            bs.launch_doorbell(kern, da.data_ptr(), db.data_ptr(), dc.data_ptr(), n, a0.itemsize, bell, stream=st)
            assert _spin_until(flag_np, seq), (kern, n, "doorbell never rang")
            torch.cuda.synchronize()
            for got, want in ((da, a), (db, b), (dc, c)):
                assert same_bits(got.cpu().numpy(), want), (kern, n)
            assert int(ticket[0].item()) == 0
    # dot: the ring follows the result into host memory
    res = torch.zeros(2, dtype=tdtype(dtype)).pin_memory()
    ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
    n = (1 << 21) + 7
    a0, b0 = np.abs(rand(n, dtype, 5)), np.abs(rand(n, dtype, 6))
    da, db = torch.from_numpy(a0).cuda(), torch.from_numpy(b0).cuda()
    out = torch.zeros(1, dtype=tdtype(dtype), device="cuda")
    torch.cuda.synchronize()
    for variant in ("ticket", "two_pass"):
        seq += 1
        res.zero_()
        bell = bs.Doorbell(None, flag.data_ptr(), seq)
        bs.launch_dot_doorbell(da.data_ptr(), db.data_ptr(), n, a0.itemsize, ws.data_ptr(), out.data_ptr(), bell,
                               result_host_mapped=res.data_ptr(), cfg=bs.LaunchCfg.make(variant=variant), stream=st)
        assert _spin_until(flag_np, seq), "dot doorbell never rang"
        got = float(res.numpy()[0])        # read BEFORE any synchronize: release/acquire through the flag
        torch.cuda.synchronize()
        assert got == float(out.item())
        want = float(np.dot(a0.astype(np.float64), b0.astype(np.float64)))
        assert abs(got - want) <= 64 * np.finfo(dtype).eps * want
    # argument checks
    with pytest.raises(bs.B200Error):
        bs.launch_doorbell("copy", da.data_ptr(), 0, db.data_ptr(), n, a0.itemsize, bs.Doorbell(None, flag.data_ptr(), 1), stream=st)
    with pytest.raises(bs.B200Error):
        bs.launch_doorbell("copy", da.data_ptr(), 0, db.data_ptr(), n, a0.itemsize, bs.Doorbell(ticket.data_ptr(), flag.data_ptr(), 0), stream=st)
    with pytest.raises(bs.B200Error):
        bs.launch_doorbell("copy", da.data_ptr(), 0, db.data_ptr(), 0, a0.itemsize, bs.Doorbell(ticket.data_ptr(), flag.data_ptr(), 9), stream=st)


def test_in_place_calls_are_rejected():
    """The written array must not overlap a read array (they are loaded through ld.global.nc): B200_EINVAL."""
    n, eb = 4096, 8
    t = torch.zeros(3 * n, dtype=torch.float64, device="cuda")
    a, b, c = t[:n], t[n:2 * n], t[2 * n:]
    st = torch.cuda.current_stream().cuda_stream
    with pytest.raises(bs.B200Error):
        bs.launch("triad", a.data_ptr(), a.data_ptr(), c.data_ptr(), n, eb, stream=st)      # a = a + s*c
    with pytest.raises(bs.B200Error):
        bs.launch("add", a.data_ptr(), b.data_ptr(), a.data_ptr(), n, eb, stream=st)        # a = a + b
    with pytest.raises(bs.B200Error):
        bs.launch("copy", a.data_ptr(), 0, a.data_ptr() + 64, n - 8, eb, stream=st)         # overlapping ranges
    bs.launch("triad", a.data_ptr(), c.data_ptr(), c.data_ptr(), n, eb, stream=st)          # two READ arrays may alias
    bs.launch("nstream", a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, stream=st)        # a is read+written by design
    torch.cuda.synchronize()


@pytest.mark.parametrize("dtype", DTYPES)
def test_programmatic_dependent_launch_sequence_bit_exact(orc, dtype):
    """cfg.flags = B200_LAUNCH_PDL: consecutive kernels of a stream overlap their launch with the predecessor's tail
    and wait in griddepcontrol.wait before the first access.  Each kernel of the driver's sequence reads what the
    previous one wrote, so any early access would show up as wrong bits."""
    tt = torch.float64 if dtype is np.float64 else torch.float32
    eb = np.dtype(dtype).itemsize
    u = np.uint64 if dtype is np.float64 else np.uint32
    st = torch.cuda.current_stream().cuda_stream
    for n in (1000, (1 << 22) + 5):
        a = torch.empty(n, dtype=tt, device="cuda"); b = torch.empty_like(a); c = torch.empty_like(a)
        ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
        out = torch.zeros(1, dtype=tt, device="cuda")
        cfg = bs.LaunchCfg.make(flags=bs.LAUNCH_PDL)
        bs.launch("init", a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, cfg=cfg, stream=st)
        steps = 6
        for _ in range(steps):
            for k in ("copy", "mul", "add", "triad"):
                bs.launch(k, a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, cfg=cfg, stream=st)
            bs.launch_dot(a.data_ptr(), b.data_ptr(), n, eb, ws.data_ptr(), out.data_ptr(), cfg=cfg, stream=st)
            bs.launch("nstream", a.data_ptr(), b.data_ptr(), c.data_ptr(), n, eb, cfg=cfg, stream=st)
        torch.cuda.synchronize()
        g = orc.gold(dtype, O.ALL, O.ORDER_CLASSIC, steps, n)
        for arr, gold in ((a, g[0]), (b, g[1]), (c, g[2])):
            assert (arr.cpu().numpy().view(u) == np.asarray(gold).view(u)).all(), n
        eps = np.finfo(dtype).eps
        assert abs(float(out.item()) - float(g[3])) <= 64 * eps * abs(float(g[3]))


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kern", EW + ["init"])
def test_sweep_direction_and_keep_region_bit_exact(orc, kern, dtype):
    """B200_LAUNCH_REVERSE walks the tiles from the top down, keep_mb switches the cache policy of the last tiles of the
    sweep: neither may change a bit, for any size (ragged edges are handled by logical tile, not by block index)."""
    for n in (1, 33, 4097, 65536 + 7, (1 << 21) + 3):
        a0, b0, c0 = rand(n, dtype, 11), rand(n, dtype, 12), rand(n, dtype, 13)
        want = [a0.copy(), b0.copy(), c0.copy()]
        oracle_apply(orc, kern, *want, True)
        for flags in (0, bs.LAUNCH_REVERSE):
            for keep in (-1, 1, 4096):
                for shape in (dict(vec_bytes=16, unroll=1, block=256), dict(vec_bytes=32, unroll=2, block=128),
                              dict(vec_bytes=8, unroll=4, block=512, grid=-2)):
                    cfg = bs.LaunchCfg.make(cache_hint="evict_first", flags=flags, keep_mb=keep, **shape)
                    got = gpu_apply(kern, a0, b0, c0, cfg=cfg)
                    for g, w in zip(got, want):
                        assert same_bits(g, w), (kern, n, flags, keep, shape)

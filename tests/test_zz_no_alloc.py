"""b200_launch_dot_static: dot for callers that cannot allocate a workspace.

Who needs it: b200_launch_dot asks the caller for a zeroed cudaMalloc'ed workspace.  A language binding that is
only handed device addresses and carries no CUDA runtime of its own cannot make one; a service that bans cudaMalloc /
cudaFree from its steady state (both synchronise the whole device and stall its other streams) will not; a process
whose device memory was fully handed out at start-up would only see the allocation fail.  The static variant keeps the
partials in an array inside the library image, so the caller passes the two vectors and the result and nothing else.

What is checked here, through the C ABI: the static entry point returns the SAME BITS as b200_launch_dot (same kernel,
same summation order -- the bar the other dot tests use for variants of one algorithm is bit equality), on two calls in
a row with different data (the array is shared and keeps the previous call's contents), for both element types, for
aligned and misaligned vectors, n = 0 included; that a launch shape beyond the array's capacity is refused with a
message rather than run; and that torch's allocator was not asked for anything in between.

This file is named so that pytest collects it last: it must pass alone in a fresh process (where its first call is
the one that looks the array's address up) as well as after every other GPU test has used the library.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
import babelstream_b200 as bs  # noqa: E402

pytestmark = pytest.mark.gpu


def _tdtype(dt):
    return torch.float64 if np.dtype(dt) == np.float64 else torch.float32


def _both(a, b, n, eb, dt, cfg=None):
    """(bits of b200_launch_dot, bits of b200_launch_dot_static) for device tensors a, b."""
    st = torch.cuda.current_stream().cuda_stream
    ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
    want = torch.full((1,), float("nan"), dtype=_tdtype(dt), device="cuda")
    got = torch.full((1,), float("nan"), dtype=_tdtype(dt), device="cuda")
    bs.launch_dot(a.data_ptr(), b.data_ptr(), n, eb, ws.data_ptr(), want.data_ptr(), cfg=cfg, stream=st)
    torch.cuda.synchronize()
    # from here to the synchronize nothing may be allocated: outputs exist already
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    bs.launch_dot_static(a.data_ptr(), b.data_ptr(), n, eb, got.data_ptr(), cfg=cfg, stream=st)
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
    return want.cpu().numpy().tobytes(), got.cpu().numpy().tobytes()


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_static_dot_matches_workspace_dot_on_consecutive_calls(dtype):
    dt = np.dtype(dtype)
    rng = np.random.default_rng(20261106)
    # first call: a large ragged vector; second call, straight after it on the same shared array: a small one with
    # other data (a stale partial or ticket of call 1 would show); then misaligned, tiny and empty inputs
    for n, off in (((1 << 22) + 11, 0), (300007, 0), (300007, 3), (1, 0), (0, 0), ((1 << 22) + 11, 1)):
        a0 = rng.uniform(0.5, 1.5, n + 8).astype(dt)
        b0 = rng.uniform(-1.0, 1.0, n + 8).astype(dt)
        a = torch.from_numpy(a0).cuda()[off:]
        b = torch.from_numpy(b0).cuda()[off:]
        for contract in (True, False):
            want, got = _both(a, b, n, dt.itemsize, dt, cfg=bs.LaunchCfg.make(contract=contract))
            assert got == want, (dt.name, n, off, contract)
        # and against the CPU, loosely, so that "both wrong in the same way" cannot pass: positive-ish tree sum
        exact = float(np.dot(a0[off:off + n].astype(np.float64), b0[off:off + n].astype(np.float64)))
        val = float(np.frombuffer(got, dtype=dt)[0])
        scale = float(np.dot(np.abs(a0[off:off + n]).astype(np.float64), np.abs(b0[off:off + n]).astype(np.float64)))
        assert abs(val - exact) <= 64 * np.finfo(dt).eps * max(scale, 1.0), (dt.name, n, val, exact)


def test_static_dot_back_to_back_without_host_sync():
    """Two calls queued on one stream with no synchronize between them: stream order alone keeps them apart."""
    dt = np.dtype(np.float64)
    rng = np.random.default_rng(7)
    n1, n2 = (1 << 21) + 5, 12345
    a1, b1 = (torch.from_numpy(rng.standard_normal(n1)).cuda() for _ in range(2))
    a2, b2 = (torch.from_numpy(rng.standard_normal(n2)).cuda() for _ in range(2))
    st = torch.cuda.current_stream().cuda_stream
    ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
    want = torch.zeros(2, dtype=torch.float64, device="cuda")
    got = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
    bs.launch_dot(a1.data_ptr(), b1.data_ptr(), n1, 8, ws.data_ptr(), want[0:].data_ptr(), stream=st)
    bs.launch_dot(a2.data_ptr(), b2.data_ptr(), n2, 8, ws.data_ptr(), want[1:].data_ptr(), stream=st)
    bs.launch_dot_static(a1.data_ptr(), b1.data_ptr(), n1, 8, got[0:].data_ptr(), stream=st)
    bs.launch_dot_static(a2.data_ptr(), b2.data_ptr(), n2, 8, got[1:].data_ptr(), stream=st)
    torch.cuda.synchronize()
    assert got.cpu().numpy().tobytes() == want.cpu().numpy().tobytes()
    assert dt.itemsize == 8


def test_static_dot_refuses_what_does_not_fit():
    n = 1 << 22
    a = torch.ones(n, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    # small tiles, one block per tile: more block partials than the array holds
    small = dict(vec_bytes=8, unroll=1, block=32)
    ntiles = int(bs.plan_sweep("dot", n, 8, cfg=bs.LaunchCfg.make(**small)).ntiles)
    assert ntiles > bs.DOT_STATIC_MAX_BLOCKS
    with pytest.raises(bs.B200Error, match="scratch array holds 16384"):
        bs.launch_dot_static(a.data_ptr(), a.data_ptr(), n, 8, out.data_ptr(), cfg=bs.LaunchCfg.make(grid=ntiles, **small),
                             stream=st)
    with pytest.raises(bs.B200Error, match="nowhere to put the sum"):
        bs.launch_dot_static(a.data_ptr(), a.data_ptr(), n, 8, 0, stream=st)
    with pytest.raises(bs.B200Error):
        bs.launch_dot_static(a.data_ptr(), a.data_ptr(), n, 2, out.data_ptr(), stream=st)
    # the refusals left the array usable
    bs.launch_dot_static(a.data_ptr(), a.data_ptr(), n, 8, out.data_ptr(), stream=st)
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == float(n)

"""ctypes binding of include/b200_stream.h and the host-side mirror of the reference's Stream<T>.

``B200Stream`` has the reference interface (/root/reference/src/Stream.h:29-49): ``copy mul add
triad nstream dot init_arrays get_arrays``, blocking, constructed as
``B200Stream(array_size, device, dtype)`` the way ``make_stream`` constructs a model
(src/main.cpp:273-274).  Errors raise ``B200Error`` carrying the library's message; the C++ host
class turns the same codes into the reference's ``std::runtime_error`` / ``exit`` convention.

The module-level ``launch`` / ``launch_dot`` are LEVEL 1 of the ABI: one kernel on raw device
pointers (e.g. ``torch.Tensor.data_ptr()``) and a stream handle, asynchronous.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

# BenchId order of src/benchmark.h:18, plus init
KERNEL_IDS = {"copy": 0, "mul": 1, "add": 2, "triad": 3, "nstream": 4, "dot": 5, "init": 6}
KERNELS = ["copy", "mul", "add", "triad", "dot", "nstream"]  # run order, src/benchmark.h:32-39
# arrays moved per element, src/benchmark.h:33-38 (init writes three)
WEIGHTS = {"copy": 2, "mul": 2, "add": 3, "triad": 3, "dot": 2, "nstream": 4, "init": 3}
HINTS = {"auto": 0, "default": 1, "noalloc": 2, "evict_first": 3, "stream": 4, "prefetch": 5}
VARIANTS = {"auto": 0, "ldst": 1, "bulk": 2, "tma": 3, "two_pass": 1, "ticket": 2}
LAUNCH_PDL = 1   # B200_LAUNCH_PDL: programmatic dependent launch (cfg.flags)
LAUNCH_REVERSE = 2          # B200_LAUNCH_REVERSE: sweep the arrays from the top down
LAUNCH_DOT_GRIDSTRIDE = 4   # B200_LAUNCH_DOT_GRIDSTRIDE: dot with the round-1 grid-stride assignment
ETIMEOUT = -7    # B200_ETIMEOUT

START_A, START_B, START_C, START_SCALAR = 0.1, 0.2, 0.0, 0.4  # src/Stream.h:24-27


IPC_HANDLE_BYTES = 64  # B200_IPC_HANDLE_BYTES


class B200Error(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[b200 error {code}] {message}")
        self.code = code


class LaunchCfg(C.Structure):
    """b200_launch_cfg.  All zero = auto."""
    _fields_ = [("vec_bytes", C.c_int), ("unroll", C.c_int), ("block", C.c_int), ("grid", C.c_int),
                ("cache_hint", C.c_int), ("variant", C.c_int), ("contract", C.c_int), ("stages", C.c_int),
                ("chunk_bytes", C.c_int), ("flags", C.c_int), ("keep_mb", C.c_int),
                ("reserved", C.c_int * 1)]

    @classmethod
    def make(cls, **kw) -> "LaunchCfg":
        c = cls()
        for k, v in kw.items():
            if k == "cache_hint" and isinstance(v, str):
                v = HINTS[v]
            if k == "variant" and isinstance(v, str):
                v = VARIANTS[v]
            if k == "contract" and isinstance(v, (bool, str)):
                v = 1 if v in (True, "fast", "fused") else 2
            setattr(c, k, int(v))
        return c

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class SweepPlan(C.Structure):
    """b200_sweep_plan: the walk a launch makes over its tiles (host logic, no GPU needed)."""
    _fields_ = [("grid", C.c_uint), ("tiles_per_block", C.c_uint), ("reverse", C.c_int), ("vec_bytes", C.c_int),
                ("unroll", C.c_int), ("cache_hint", C.c_int), ("keep_from", C.c_ulonglong), ("origin", C.c_ulonglong), ("stride", C.c_longlong),
                ("ntiles", C.c_ulonglong), ("nvec", C.c_ulonglong), ("tile_vecs", C.c_ulonglong)]


class Doorbell(C.Structure):
    """b200_doorbell: the kernel's last block stores `seq` to a mapped pinned host word."""
    _fields_ = [("ticket", C.c_void_p), ("flag_host_mapped", C.c_void_p), ("seq", C.c_ulonglong)]


def lib_path() -> str:
    return os.environ.get("BABELSTREAM_B200_LIB", os.path.join(_HERE, "csrc", "libb200stream.so"))


_lib = None


def lib() -> C.CDLL:
    """The loaded libb200stream.so.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise B200Error(-100, f"{path} is missing: build it with `make lib` (or __graft_entry__.build()); "
                              "there is no CPU fallback")
    L = C.CDLL(path)
    vp, ci, sz, d, cp = C.c_void_p, C.c_int, C.c_size_t, C.c_double, C.c_char_p
    cfgp = C.POINTER(LaunchCfg)
    sig = {
        "b200_version": (cp, []),
        "b200_last_error": (cp, []),
        "b200_device_count": (ci, []),
        "b200_device_name": (ci, [ci, cp, ci]),
        "b200_driver_version": (ci, [ci]),
        "b200_device_info": (ci, [ci, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci), C.POINTER(sz), C.POINTER(ci),
                                  C.POINTER(ci), C.POINTER(ci)]),
        "b200_launch_count": (C.c_uint64, []),
        "b200_shard_range": (ci, [sz, ci, ci, ci, C.POINTER(sz), C.POINTER(sz)]),
        "b200_resolve_cfg": (ci, [ci, sz, ci, ci, ci, cfgp, cfgp, C.POINTER(C.c_uint)]),
        "b200_plan_sweep": (ci, [ci, sz, ci, ci, ci, cfgp, C.POINTER(SweepPlan)]),
        "b200_launch_init": (ci, [vp, vp, vp, sz, ci, d, d, d, cfgp, vp]),
        "b200_launch_copy": (ci, [vp, vp, sz, ci, cfgp, vp]),
        "b200_launch_mul": (ci, [vp, vp, sz, ci, d, cfgp, vp]),
        "b200_launch_add": (ci, [vp, vp, vp, sz, ci, cfgp, vp]),
        "b200_launch_triad": (ci, [vp, vp, vp, sz, ci, d, cfgp, vp]),
        "b200_launch_nstream": (ci, [vp, vp, vp, sz, ci, d, cfgp, vp]),
        "b200_dot_workspace_bytes": (sz, []),
        "b200_launch_dot": (ci, [vp, vp, sz, ci, vp, vp, vp, cfgp, vp]),
        "b200_launch_dot_static": (ci, [vp, vp, sz, ci, vp, vp, cfgp, vp]),
        "b200_dot_mailbox_bytes": (sz, []),
        "b200_launch_dot_allreduce_p2p": (ci, [vp, vp, sz, ci, vp, vp, vp, ci, ci, C.c_ulonglong, C.POINTER(vp), cfgp, vp]),
        "b200_launch_dot_send_p2p": (ci, [vp, vp, sz, ci, vp, vp, ci, ci, C.c_ulonglong, C.POINTER(vp), cfgp, vp]),
        "b200_launch_dot_collect_p2p": (ci, [ci, vp, vp, ci, ci, C.c_ulonglong, C.POINTER(vp), vp, ci, vp]),
        "b200_dot_status_offset": (sz, []),
        "b200_set_p2p_timeout_ms": (ci, [d]),
        "b200_device_ecc": (ci, [ci]),
        "b200_device_mem_clock_mhz": (ci, [ci]),
        "b200_launch_doorbell": (ci, [ci, vp, vp, vp, sz, ci, d, d, d, cfgp, C.POINTER(Doorbell), vp]),
        "b200_launch_dot_doorbell": (ci, [vp, vp, sz, ci, vp, vp, vp, cfgp, C.POINTER(Doorbell), vp]),
        "b200_dot_mailbox_create": (ci, [C.POINTER(vp)]),
        "b200_dot_mailbox_destroy": (ci, [vp]),
        "b200_ipc_export": (ci, [vp, vp]),
        "b200_ipc_open": (ci, [vp, C.POINTER(vp)]),
        "b200_ipc_close": (ci, [vp]),
        "b200_join_group": (ci, [vp, ci, ci, C.POINTER(vp)]),
        "b200_create": (vp, [ci, ci, sz, ci, C.POINTER(ci)]),
        "b200_create_opts": (vp, [ci, ci, sz, ci, C.POINTER(cp), C.POINTER(ci)]),
        "b200_destroy": (None, [vp]),
        "b200_set_option": (ci, [vp, cp, cp]),
        "b200_describe": (ci, [vp, cp, ci]),
        "b200_init_arrays": (ci, [vp, d, d, d]),
        "b200_warmup": (ci, [vp, d]),
        "b200_run_sequence": (ci, [vp, C.POINTER(ci), ci, d, C.POINTER(d)]),
        "b200_copy": (ci, [vp]),
        "b200_mul": (ci, [vp, d]),
        "b200_add": (ci, [vp]),
        "b200_triad": (ci, [vp, d]),
        "b200_nstream": (ci, [vp, d]),
        "b200_dot": (ci, [vp, C.POINTER(d)]),
        "b200_run": (ci, [vp, ci, d, C.POINTER(d)]),
        "b200_get_arrays": (ci, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
        "b200_read_arrays": (ci, [vp, vp, vp, vp]),
        "b200_write_arrays": (ci, [vp, vp, vp, vp]),
        "b200_num_gpus": (ci, [vp]),
        "b200_memory_mode": (cp, [vp]),
        "b200_array_size": (sz, [vp]),
        "b200_elem_bytes": (ci, [vp]),
        "b200_last_kernel_ms": (d, [vp]),
        "b200_shard_info": (ci, [vp, ci, C.POINTER(ci), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(sz),
                                 C.POINTER(sz), C.POINTER(vp)]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype, f.argtypes = res, args
    _lib = L
    return L


def _check(rc: int) -> int:
    if rc < 0:
        raise B200Error(rc, lib().b200_last_error().decode(errors="replace"))
    return rc


def _elem_bytes(dtype) -> int:
    dt = np.dtype(dtype)
    if dt == np.float32:
        return 4
    if dt == np.float64:
        return 8
    raise TypeError(f"BabelStream is float/double only (src/cuda/CUDAStream.cu:349-350), got {dt}")


def device_count() -> int:
    return _check(lib().b200_device_count())


def launch_count() -> int:
    return int(lib().b200_launch_count())


def dot_workspace_bytes() -> int:
    return int(lib().b200_dot_workspace_bytes())


def shard_range(n: int, elem_bytes: int, ngpus: int, g: int):
    b, c = C.c_size_t(), C.c_size_t()
    _check(lib().b200_shard_range(n, elem_bytes, ngpus, g, C.byref(b), C.byref(c)))
    return b.value, c.value


def resolve_cfg(kernel: str, n: int, elem_bytes: int, num_sms: int = 148, l2_bytes: int = 126 << 20,
                cfg: LaunchCfg | None = None):
    out, grid = LaunchCfg(), C.c_uint()
    _check(lib().b200_resolve_cfg(KERNEL_IDS[kernel], n, elem_bytes, num_sms, l2_bytes,
                                  C.byref(cfg) if cfg is not None else None, C.byref(out), C.byref(grid)))
    return out, grid.value


def plan_sweep(kernel: str, n: int, elem_bytes: int, num_sms: int = 148, l2_bytes: int = 126 << 20,
               cfg: LaunchCfg | None = None) -> SweepPlan:
    out = SweepPlan()
    _check(lib().b200_plan_sweep(KERNEL_IDS[kernel], n, elem_bytes, num_sms, l2_bytes,
                                 C.byref(cfg) if cfg is not None else None, C.byref(out)))
    return out


def launch(kernel: str, a: int, b: int, c: int, n: int, elem_bytes: int, *, scalar: float = START_SCALAR,
           init=(START_A, START_B, START_C), cfg: LaunchCfg | None = None, stream: int = 0) -> None:
    """LEVEL 1: enqueue one elementwise kernel on raw device pointers (ints).  Asynchronous."""
    L = lib()
    p = C.byref(cfg) if cfg is not None else None
    st = C.c_void_p(stream)
    if kernel == "init":
        rc = L.b200_launch_init(a, b, c, n, elem_bytes, init[0], init[1], init[2], p, st)
    elif kernel == "copy":
        rc = L.b200_launch_copy(a, c, n, elem_bytes, p, st)
    elif kernel == "mul":
        rc = L.b200_launch_mul(b, c, n, elem_bytes, scalar, p, st)
    elif kernel == "add":
        rc = L.b200_launch_add(a, b, c, n, elem_bytes, p, st)
    elif kernel == "triad":
        rc = L.b200_launch_triad(a, b, c, n, elem_bytes, scalar, p, st)
    elif kernel == "nstream":
        rc = L.b200_launch_nstream(a, b, c, n, elem_bytes, scalar, p, st)
    else:
        raise KeyError(kernel)
    _check(rc)


def launch_dot(a: int, b: int, n: int, elem_bytes: int, workspace: int, result_dev: int, *,
               result_host_mapped: int = 0, cfg: LaunchCfg | None = None, stream: int = 0) -> None:
    """LEVEL 1: enqueue dot; the sum (type T) lands in result_dev.  Asynchronous."""
    _check(lib().b200_launch_dot(a, b, n, elem_bytes, workspace, result_dev or None, result_host_mapped or None,
                                 C.byref(cfg) if cfg is not None else None, C.c_void_p(stream)))


DOT_STATIC_MAX_BLOCKS = 16384  # B200_DOT_STATIC_MAX_BLOCKS


def launch_dot_static(a: int, b: int, n: int, elem_bytes: int, result_dev: int, *, result_host_mapped: int = 0,
                      cfg: LaunchCfg | None = None, stream: int = 0) -> None:
    """LEVEL 1: dot for callers that have no workspace to give and may not allocate one.

    ``launch_dot`` needs a zeroed device buffer from the caller.  A binding that only forwards device addresses and
    has no CUDA runtime of its own, a service that keeps cudaMalloc (a device-wide synchronisation) out of its steady
    state, or a process whose device memory is fully handed out cannot produce one.  This call keeps the partials in
    an array inside the library image instead and allocates nothing; same bits as ``launch_dot``.  One array per GPU:
    do not let two of these calls overlap on the device, stay within ``DOT_STATIC_MAX_BLOCKS`` blocks, and make the
    first call on a GPU outside stream capture (it looks the array's address up).  Asynchronous on ``stream``.
    """
    _check(lib().b200_launch_dot_static(a, b, n, elem_bytes, result_dev or None, result_host_mapped or None,
                                        C.byref(cfg) if cfg is not None else None, C.c_void_p(stream)))


def launch_dot_allreduce_p2p(a: int, b: int, n: int, elem_bytes: int, workspace: int, result_dev: int, *, world: int,
                             rank: int, seq: int, mailboxes, result_host_mapped: int = 0, cfg: LaunchCfg | None = None,
                             stream: int = 0) -> None:
    """LEVEL 1: fused dot + all-reduce over peer memory, GPU `rank` of `world` (one call per GPU, each on its device)."""
    boxes = (C.c_void_p * len(mailboxes))(*mailboxes)
    _check(lib().b200_launch_dot_allreduce_p2p(a, b, n, elem_bytes, workspace, result_dev or None, result_host_mapped or None,
                                               world, rank, seq, boxes, C.byref(cfg) if cfg is not None else None,
                                               C.c_void_p(stream)))


def launch_dot_send_p2p(a: int, b: int, n: int, elem_bytes: int, workspace: int, *, world: int, rank: int, seq: int,
                        mailboxes, partial_dev: int = 0, cfg: LaunchCfg | None = None, stream: int = 0) -> None:
    """LEVEL 1: the sending half of the fused all-reduce (no wait); pair with launch_dot_collect_p2p."""
    boxes = (C.c_void_p * len(mailboxes))(*mailboxes)
    _check(lib().b200_launch_dot_send_p2p(a, b, n, elem_bytes, workspace, partial_dev or None, world, rank, seq, boxes,
                                          C.byref(cfg) if cfg is not None else None, C.c_void_p(stream)))


def launch_dot_collect_p2p(elem_bytes: int, result_dev: int, *, world: int, rank: int, seq: int, mailboxes,
                           result_host_mapped: int = 0, workspace: int = 0, flags: int = 0, stream: int = 0) -> None:
    """LEVEL 1: one-warp kernel that waits for the partials of call `seq` and adds them in rank order.
    `workspace` (optional): the dot workspace whose status word is set when a rank never arrives."""
    boxes = (C.c_void_p * len(mailboxes))(*mailboxes)
    _check(lib().b200_launch_dot_collect_p2p(elem_bytes, result_dev or None, result_host_mapped or None, world, rank, seq,
                                             boxes, workspace or None, flags, C.c_void_p(stream)))


def dot_status_offset() -> int:
    return int(lib().b200_dot_status_offset())


def set_p2p_timeout_ms(ms: float) -> None:
    _check(lib().b200_set_p2p_timeout_ms(float(ms)))


def launch_doorbell(kernel: str, a: int, b: int, c: int, n: int, elem_bytes: int, bell: Doorbell, *,
                    scalar: float = START_SCALAR, init=(START_A, START_B, START_C), cfg: LaunchCfg | None = None,
                    stream: int = 0) -> None:
    """LEVEL 1: an elementwise kernel whose last block rings `bell` (see b200_doorbell).  Asynchronous."""
    s0, s1, s2 = init if kernel == "init" else (scalar, 0.0, 0.0)
    _check(lib().b200_launch_doorbell(KERNEL_IDS[kernel], a or None, b or None, c or None, n, elem_bytes, s0, s1, s2,
                                      C.byref(cfg) if cfg is not None else None, C.byref(bell), C.c_void_p(stream)))


def launch_dot_doorbell(a: int, b: int, n: int, elem_bytes: int, workspace: int, result_dev: int, bell: Doorbell, *,
                        result_host_mapped: int = 0, cfg: LaunchCfg | None = None, stream: int = 0) -> None:
    _check(lib().b200_launch_dot_doorbell(a, b, n, elem_bytes, workspace, result_dev or None, result_host_mapped or None,
                                          C.byref(cfg) if cfg is not None else None, C.byref(bell), C.c_void_p(stream)))


class Mailbox:
    """This process's dot mailbox on the current device plus, once `connect`ed, its peers' mailboxes.

    One process per GPU (torchrun): ``mb = Mailbox(); handles = all_gather(mb.handle);
    ptrs = mb.connect(handles, rank)`` -- ``ptrs`` is what ``launch_dot_allreduce_p2p`` and
    ``B200Stream.join_group`` take.  The exchange of the 64-byte handles is the caller's business
    (``babelstream_b200.group.connect_mailboxes`` does it with torch.distributed).
    """

    def __init__(self):
        p = C.c_void_p()
        _check(lib().b200_dot_mailbox_create(C.byref(p)))
        self.ptr = p.value
        h = C.create_string_buffer(IPC_HANDLE_BYTES)
        _check(lib().b200_ipc_export(self.ptr, h))
        self.handle = h.raw
        self._opened = []
        self.ptrs = None

    def connect(self, handles, rank: int):
        ptrs = []
        for g, h in enumerate(handles):
            if g == rank:
                ptrs.append(self.ptr)
                continue
            p = C.c_void_p()
            _check(lib().b200_ipc_open(C.create_string_buffer(bytes(h), IPC_HANDLE_BYTES), C.byref(p)))
            self._opened.append(p.value)
            ptrs.append(p.value)
        self.ptrs = ptrs
        return ptrs

    def close(self):
        for p in self._opened:
            lib().b200_ipc_close(p)
        self._opened = []
        if self.ptr:
            lib().b200_dot_mailbox_destroy(self.ptr)
            self.ptr = None


class B200Stream:
    """Host-side mirror of ``Stream<T>`` (src/Stream.h:29-49) over LEVEL 2 of the C ABI.

    ``B200Stream(array_size, device=0, dtype=np.float64, ngpus=1)`` allocates the three arrays
    sharded over ``ngpus`` devices and, like every reference model's constructor
    (src/cuda/CUDAStream.cu:148), initialises them to (0.1, 0.2, 0.0).
    """

    def __init__(self, array_size: int, device: int = 0, dtype=np.float64, ngpus: int = 1,
                 init=(START_A, START_B, START_C), **options):
        self.dtype = np.dtype(dtype)
        self.elem_bytes = _elem_bytes(dtype)
        self.array_size = int(array_size)
        if self.array_size <= 0:
            raise B200Error(-1, "Invalid array size.")  # src/main.cpp:462-466
        st = C.c_int(0)
        # options are applied before the arrays are allocated (mem=managed needs that)
        opts = [f"{k}={v}".encode() for k, v in options.items()]
        arr = (C.c_char_p * (len(opts) + 1))(*opts, None)
        self._h = lib().b200_create_opts(device, ngpus, self.array_size, self.elem_bytes, arr, C.byref(st))
        if not self._h:
            raise B200Error(st.value, lib().b200_last_error().decode(errors="replace"))
        self.init_arrays(*init)

    # -- lifetime ----------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            lib().b200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- Stream<T> ---------------------------------------------------------------------------
    def init_arrays(self, init_a=START_A, init_b=START_B, init_c=START_C):
        _check(lib().b200_init_arrays(self._h, init_a, init_b, init_c))

    def copy(self):
        _check(lib().b200_copy(self._h))

    def mul(self, scalar=START_SCALAR):
        _check(lib().b200_mul(self._h, scalar))

    def add(self):
        _check(lib().b200_add(self._h))

    def triad(self, scalar=START_SCALAR):
        _check(lib().b200_triad(self._h, scalar))

    def nstream(self, scalar=START_SCALAR):
        _check(lib().b200_nstream(self._h, scalar))

    def dot(self):
        out = C.c_double()
        _check(lib().b200_dot(self._h, C.byref(out)))
        return self.dtype.type(out.value)  # exact: the library widened a T

    def run(self, kernel: str, scalar=START_SCALAR):
        if kernel == "dot":
            return self.dot()
        return getattr(self, kernel)(*(() if kernel in ("copy", "add") else (scalar,)))

    def run_sequence(self, kernels, scalar=START_SCALAR):
        """The given kernels as consecutive blocking calls issued from C (b200_run_sequence); returns the last dot or None."""
        key = tuple(kernels)
        cached = getattr(self, "_seq_cache", None)
        if cached is None or cached[0] != key:
            ids = (C.c_int * len(key))(*[KERNEL_IDS[k] for k in key])
            self._seq_cache = cached = (key, ids, C.c_double(), "dot" in key)
        _, ids, out, has_dot = cached
        _check(lib().b200_run_sequence(self._h, ids, len(key), scalar, C.byref(out)))
        return self.dtype.type(out.value) if has_dot else None

    def get_arrays(self):
        """Three numpy VIEWS of the library-owned host mirrors (valid until close), like get_arrays()."""
        pa, pb, pc = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _check(lib().b200_get_arrays(self._h, C.byref(pa), C.byref(pb), C.byref(pc)))
        ct = C.c_float if self.elem_bytes == 4 else C.c_double
        out = []
        for p in (pa, pb, pc):
            buf = (ct * self.array_size).from_address(p.value)
            out.append(np.frombuffer(buf, dtype=self.dtype, count=self.array_size))
        return tuple(out)

    # -- extras --------------------------------------------------------------------------------
    def read_arrays(self, out=None):
        """Gather the shards into host arrays: fresh ones, or the caller's `out = (a, b, c)` (e.g. views of
        pinned memory; C-contiguous, dtype and length of the model)."""
        if out is None:
            out = tuple(np.empty(self.array_size, dtype=self.dtype) for _ in range(3))
        for x in out:
            if x.dtype != self.dtype or x.size != self.array_size or not x.flags.c_contiguous or not x.flags.writeable:
                raise ValueError("read_arrays: out arrays must be writable, C-contiguous, of the model's dtype and length")
        a, b, c = out
        _check(lib().b200_read_arrays(self._h, a.ctypes.data, b.ctypes.data, c.ctypes.data))
        return a, b, c

    def write_arrays(self, a=None, b=None, c=None):
        ptrs = []
        keep = []
        for x in (a, b, c):
            if x is None:
                ptrs.append(None)
                continue
            x = np.ascontiguousarray(x, dtype=self.dtype)
            if x.size != self.array_size:
                raise ValueError("array length mismatch")
            keep.append(x)
            ptrs.append(x.ctypes.data)
        _check(lib().b200_write_arrays(self._h, *ptrs))

    def set_option(self, key: str, value) -> None:
        _check(lib().b200_set_option(self._h, key.encode(), str(value).encode()))

    def join_group(self, world: int, rank: int, mailboxes) -> None:
        """One process per GPU: dot() returns the all-rank sum from here on (fused peer-memory all-reduce)."""
        boxes = (C.c_void_p * max(1, len(mailboxes)))(*mailboxes)
        _check(lib().b200_join_group(self._h, world, rank, boxes))

    def warmup(self, scalar: float = START_SCALAR) -> None:
        """Pays every kernel's first-use cost now (code loading, launch plans); array contents become undefined."""
        _check(lib().b200_warmup(self._h, C.c_double(scalar)))

    def describe(self) -> str:
        buf = C.create_string_buffer(4096)
        _check(lib().b200_describe(self._h, buf, len(buf)))
        return buf.value.decode()

    @property
    def ngpus(self) -> int:
        return _check(lib().b200_num_gpus(self._h))

    def last_kernel_ms(self) -> float:
        return float(lib().b200_last_kernel_ms(self._h))

    def shard_info(self, g: int = 0):
        dev = C.c_int()
        a, b, c, st = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        beg, cnt = C.c_size_t(), C.c_size_t()
        _check(lib().b200_shard_info(self._h, g, C.byref(dev), C.byref(a), C.byref(b), C.byref(c), C.byref(beg),
                                     C.byref(cnt), C.byref(st)))
        return {"device": dev.value, "a": a.value, "b": b.value, "c": c.value, "begin": beg.value,
                "count": cnt.value, "stream": st.value or 0}

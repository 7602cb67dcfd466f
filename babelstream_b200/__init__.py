"""babelstream_b200 -- a B200-native (sm_100a) BabelStream model.

The product is ``csrc/libb200stream.so`` (hand-written CUDA kernels behind the C ABI of
``include/b200_stream.h``) and ``src/b200/B200Stream.{h,cpp}``, the C++ ``Stream<T>`` model that
the reference's unmodified ``src/main.cpp`` drives.  This Python package is the same boundary
seen from ctypes: it is what the tests, ``bench.py`` and ``__graft_entry__`` call.

There is no CPU fallback and nothing here imports ``oracle``.
"""
from .stream import (  # noqa: F401
    B200Error,
    B200Stream,
    Doorbell,
    LaunchCfg,
    Mailbox,
    HINTS,
    KERNELS,
    KERNEL_IDS,
    VARIANTS,
    WEIGHTS,
    lib,
    lib_path,
    launch,
    launch_dot,
    launch_dot_static,
    DOT_STATIC_MAX_BLOCKS,
    launch_dot_allreduce_p2p,
    launch_dot_send_p2p,
    launch_dot_collect_p2p,
    launch_doorbell,
    launch_dot_doorbell,
    dot_workspace_bytes,
    dot_status_offset,
    set_p2p_timeout_ms,
    LAUNCH_PDL,
    LAUNCH_REVERSE,
    LAUNCH_DOT_GRIDSTRIDE,
    ETIMEOUT,
    device_count,
    launch_count,
    resolve_cfg,
    plan_sweep,
    SweepPlan,
    shard_range,
)

__all__ = ["B200Error", "B200Stream", "LaunchCfg", "Doorbell", "Mailbox", "lib", "lib_path", "launch", "launch_dot", "launch_dot_static", "DOT_STATIC_MAX_BLOCKS",
           "launch_dot_allreduce_p2p", "launch_dot_send_p2p", "launch_dot_collect_p2p", "launch_doorbell", "launch_dot_doorbell",
           "dot_workspace_bytes", "dot_status_offset", "set_p2p_timeout_ms", "LAUNCH_PDL", "LAUNCH_REVERSE", "LAUNCH_DOT_GRIDSTRIDE", "ETIMEOUT", "device_count", "launch_count", "resolve_cfg", "plan_sweep", "SweepPlan", "shard_range",
           "KERNELS", "KERNEL_IDS", "WEIGHTS", "HINTS", "VARIANTS"]

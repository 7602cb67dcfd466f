"""One process per GPU: exchange the dot mailboxes of all ranks over torch.distributed.

torch.distributed is plumbing here (it carries 64-byte IPC handles once, at start-up); the data path
of the multi-GPU dot is the fused kernel of ``csrc/kernels.cuh`` writing into peer memory over NVLink.
"""
from __future__ import annotations

from .stream import B200Error, Mailbox


def connect_mailboxes(dist, rank: int, world: int):
    """Returns a connected ``Mailbox`` (``.ptrs`` = every rank's mailbox as mapped here), or ``None``
    when any rank could not create, export or map one (no peer access, IPC closed in this container):
    all ranks agree on the outcome, so the caller can fall back to an NCCL all-reduce collectively."""
    mb, err = None, ""
    try:
        mb = Mailbox()
    except B200Error as e:  # noqa: PERF203
        err = str(e)
    handles = [None] * world
    dist.all_gather_object(handles, mb.handle if mb is not None else None)
    ok = mb is not None and all(h is not None for h in handles)
    if ok:
        try:
            mb.connect(handles, rank)
        except B200Error as e:
            ok, err = False, str(e)
    flags = [None] * world
    dist.all_gather_object(flags, (bool(ok), err))
    if all(f[0] for f in flags):
        return mb
    if mb is not None:
        mb.close()
    connect_mailboxes.last_error = "; ".join(f"rank {g}: {f[1]}" for g, f in enumerate(flags) if not f[0])
    return None


connect_mailboxes.last_error = ""


def join_group(stream, dist, world: int, rank: int, mailbox) -> None:
    """Make the single-GPU ``B200Stream`` rank ``rank`` of ``world`` and wait until every rank has done so.
    ``b200_join_group`` wipes this rank's own mailbox and restarts its all-reduce sequence number; the barrier
    guarantees that no peer's first partial can land before that wipe."""
    stream.join_group(world, rank, mailbox.ptrs)
    dist.barrier()

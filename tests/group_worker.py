"""Worker of test_process_group_fused_dot_allreduce: run under torch.distributed.run, one rank per GPU.
(Not a test module itself: the name does not start with test_.)"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import babelstream_b200 as bs  # noqa: E402
from babelstream_b200.group import connect_mailboxes, join_group  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")   # the handles travel over gloo: no NCCL anywhere in this test
    mb = connect_mailboxes(dist, rank, world)
    assert mb is not None, "mailbox exchange failed: " + connect_mailboxes.last_error
    for dtype in (np.float64, np.float32):
        n_total = (1 << 22) + 37
        begin, count = bs.shard_range(n_total, np.dtype(dtype).itemsize, world, rank)
        rng = np.random.default_rng(7)
        a0 = np.abs(rng.standard_normal(n_total)).astype(dtype)
        b0 = np.abs(rng.standard_normal(n_total)).astype(dtype)
        with bs.B200Stream(max(count, 1), device=local, dtype=dtype) as s:
            s.write_arrays(a0[begin:begin + max(count, 1)], b0[begin:begin + max(count, 1)], None)
            partial = s.dot()                       # before joining: this rank's shard only
            parts = [None] * world
            dist.all_gather_object(parts, partial.tobytes())
            want = dtype(0)
            for p in parts:                          # rank order, in type T: what the kernel does
                want = dtype(want + np.frombuffer(p, dtype=dtype)[0])
            join_group(s, dist, world, rank, mb)     # join + barrier: the join wipes this rank's own mailbox
            assert "IPC mailboxes" in s.describe()
            got = [s.dot() for _ in range(7)]        # odd count: both mailbox banks, several rounds
            for g in got:
                assert g.tobytes() == want.tobytes(), (rank, g, want)
            # ranks drifting apart between calls must not matter
            if rank == 0:
                import time
                time.sleep(0.3)
            assert s.dot().tobytes() == want.tobytes()
            # interleaved with other kernels, through every wait mode
            for sync in ("stream", "doorbell", "auto"):
                s.set_option("sync", sync)
                s.copy(); s.mul(); s.add(); s.triad()
                v1 = s.dot()
                allv = [None] * world
                dist.all_gather_object(allv, v1.tobytes())
                assert all(x == allv[0] for x in allv), "ranks disagree on the all-reduced dot"
            dist.barrier()
            s.join_group(1, 0, [])
            # LEVEL 1, the split all-reduce: send in the dot kernel, collect later -- same bits as the fused call
            mb2 = connect_mailboxes(dist, rank, world)
            assert mb2 is not None
            s.write_arrays(a0[begin:begin + max(count, 1)], b0[begin:begin + max(count, 1)], None)
            info = s.shard_info(0)
            tt = torch.float64 if dtype is np.float64 else torch.float32
            ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
            out = torch.zeros(1, dtype=tt, device="cuda")
            part = torch.zeros(1, dtype=tt, device="cuda")
            st = torch.cuda.current_stream().cuda_stream
            eb = np.dtype(dtype).itemsize
            torch.cuda.synchronize()
            for seq in range(1, 6):
                bs.launch_dot_send_p2p(info["a"], info["b"], info["count"], eb, ws.data_ptr(), world=world, rank=rank,
                                       seq=seq, mailboxes=mb2.ptrs, partial_dev=part.data_ptr(), stream=st)
                if rank == world - 1 and seq == 3:
                    torch.cuda.synchronize()
                    import time
                    time.sleep(0.2)          # a late collector must still find its bank intact
                bs.launch_dot_collect_p2p(eb, out.data_ptr(), world=world, rank=rank, seq=seq, mailboxes=mb2.ptrs, stream=st)
                torch.cuda.synchronize()
                assert out.cpu().numpy()[0].tobytes() == want.tobytes(), (rank, seq)
                assert part.cpu().numpy()[0].tobytes() == partial.tobytes(), (rank, seq)
            dist.barrier()
            mb2.close()
        dist.barrier()
    mb.close()
    dist.barrier()
    if rank == 0:
        print(f"group ok world={world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

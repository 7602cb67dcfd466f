#!/usr/bin/env python
"""bench.py -- the BabelStream hot path on N B200s, one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--no-extras]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): fp64, 2^25 elements PER GPU, all six kernels.  One STEP is one
pass of the driver's kernel sequence  copy, mul, add, triad, dot, nstream  (src/benchmark.h:32-39)
over the rank's contiguous shard; with N > 1 the step's dot partials are combined INSIDE the dot
kernel by the fused peer-memory all-reduce (mailboxes in every GPU's HBM, mapped into the other ranks
through CUDA IPC, written over NVLink; no NCCL on the data path -- NCCL is the fallback when the
mailboxes cannot be mapped).  Scaling is weak (per-GPU work fixed, no data-path collective for the
elementwise kernels).  The arrays (3 x 268 MB per GPU) are 6.4x the 126 MB L2, so consecutive
kernels cannot be served from cache; that is the `l2` statement in `config`.

value   = algorithmic bytes of all ranks per step (16 * 8 * n * N, weights of src/benchmark.h:33-38)
          / device time per step: CUDA events on the launching stream around the K timed steps,
          arrays resident in HBM, kernels enqueued back to back through LEVEL 1 of the C ABI,
          max over ranks.
e2e     = the same bytes / host wall time per step through LEVEL 2 (the blocking Stream<T> calls
          that the reference's src/main.cpp times with its chrono clock): launch + wait per
          kernel, dot's result read back to the host every step.  This path has no per-step host
          input buffers by the reference's design (the arrays live on the device for the whole
          run, src/Stream.h:29-49), so h2d_bytes_per_step is 0 and d2h is the dot scalar.
kernels = per-kernel GB/s from events recorded at every kernel boundary INSIDE the timed region.
roofline= the triad kernel (BabelStream's headline) against the measured HBM copy peak.
cpu_baseline / --impl reference = the reference's own omp-stream (oracle/_ref, built from the
          unmodified sources) on this box's host cores, same element count and kernel list.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG2N = 25
KERNEL_ORDER = ["copy", "mul", "add", "triad", "dot", "nstream"]
WEIGHT = {"copy": 2, "mul": 2, "add": 3, "triad": 3, "dot": 2, "nstream": 4}
STEP_WEIGHT = sum(WEIGHT.values())  # 16 arrays moved per step
METRIC = "babelstream_all_kernels_gbytes_per_sec"
UNIT = "GB/s"


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.rows, self.proc, self.idx = [], None, device_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm, smax, reasons = [], [], set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "reasons": sorted(reasons), "samples": len(sm)}


def parse_stream_csv(text: str):
    """rows of the driver's --csv output (src/main.cpp:200-222) -> {kernel: {...}}"""
    out = {}
    for line in text.splitlines():
        f = line.split(",")
        if len(f) == 8 and f[0] in ("Copy", "Mul", "Add", "Triad", "Dot", "Nstream"):
            out[f[0].lower()] = {"gbs": float(f[4]), "min_s": float(f[5]), "max_s": float(f[6]), "avg_s": float(f[7]),
                                 "n": int(f[2]), "sizeof": int(f[3]), "num_times": int(f[1])}
    return out


def run_stream_binary(path, n, num_times, only="All", env=None, timeout=900):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run([path, "--csv", "--gigabytes", "-s", str(n), "-n", str(num_times), "--only", only],
                       capture_output=True, text=True, timeout=timeout, env=e)
    if p.returncode != 0:
        raise RuntimeError(f"{os.path.basename(path)} exited {p.returncode}: {p.stderr[-400:]}")
    return parse_stream_csv(p.stdout)


def step_gbs_from_csv(rows, n, key="avg_s"):
    t = sum(rows[k][key] for k in KERNEL_ORDER)
    return STEP_WEIGHT * 8 * n / t / 1e9, t


def cpu_reference(n, num_times):
    """The reference's CPU implementation of the path on this box's host cores.
    kind 'reference' = oracle/_ref/omp-stream (unmodified sources); else the oracle port."""
    cores = os.cpu_count() or 1
    exe = os.path.join(ROOT, "oracle", "_ref", "omp-stream")
    if os.path.exists(exe):
        env = {"OMP_NUM_THREADS": str(cores), "OMP_PROC_BIND": "spread", "OMP_PLACES": "threads"}
        rows = run_stream_binary(exe, n, num_times, env=env)
        gbs, t = step_gbs_from_csv(rows, n)
        return {"value": round(gbs, 2), "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": f"oracle/_ref/omp-stream --only All -s {n} -n {num_times} (fp64; first iteration discarded by the "
                          f"driver; step = sum of the six kernels' average runtimes = {t * 1e3:.2f} ms)",
                "kernels": {k: round(rows[k]["gbs"], 2) for k in KERNEL_ORDER}}, t
    # port: the C restatement with OpenMP (oracle/ may be executed here, and only here, by bench.py)
    import numpy as np
    import oracle as O
    orc = O.Oracle()
    a, b, c = (np.empty(n, dtype=np.float64) for _ in range(3))
    orc.init(a, b, c)
    orc.run(O.ALL, O.ORDER_CLASSIC, 1, a, b, c, omp_dot=True)
    t0 = time.perf_counter()
    orc.run(O.ALL, O.ORDER_CLASSIC, num_times, a, b, c, omp_dot=True)
    t = (time.perf_counter() - t0) / num_times
    return {"value": round(STEP_WEIGHT * 8 * n / t / 1e9, 2), "unit": UNIT, "cores": orc.num_threads(), "kind": "port",
            "sample": f"oracle C restatement (OpenMP), {num_times} steps of all six kernels at n={n} fp64"}, t


def main_reference(args, rank, world):
    if rank != 0:
        return 0
    n = 1 << LOG2N
    num_times = max(2, args.steps + 1)  # the driver discards its first iteration = our warm-up
    base, t = cpu_reference(n, num_times)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": 1, "ms_per_step": round(t * 1e3, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"BabelStream fp64, 2^{LOG2N} elements per GPU, all six kernels (copy mul add triad dot nstream)",
                       "elements_per_gpu": n, "elements_total": n * args.gpus, "kernels": KERNEL_ORDER,
                       "bytes_per_step_per_gpu": STEP_WEIGHT * 8 * n,
                       "run_as": f"reference omp-stream (unmodified sources) on the host CPU cores; each step is a bounded sample "
                                 f"of the workload: one pass of the six kernels over 2^{LOG2N} fp64 elements"},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the driver-binary and 2^29 north-star extras")
    ap.add_argument("--log2n", type=int, default=LOG2N, help="elements per GPU (log2); default = BASELINE config")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return main_reference(args, rank, world)

    import numpy as np
    import torch
    import babelstream_b200 as bs

    if not torch.cuda.is_available() or bs.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = max(3, args.warmup)
    K = args.steps
    n = 1 << args.log2n
    eb = 8
    step_bytes = STEP_WEIGHT * eb * n          # per rank
    st = torch.cuda.current_stream().cuda_stream

    a = torch.empty(n, dtype=torch.float64, device="cuda")
    b = torch.empty_like(a)
    c = torch.empty_like(a)
    ws = torch.zeros(bs.dot_workspace_bytes(), dtype=torch.uint8, device="cuda")
    dot_out = torch.zeros(1, dtype=torch.float64, device="cuda")
    pa, pb, pc = a.data_ptr(), b.data_ptr(), c.data_ptr()

    pending = [None]
    # multi-GPU dot: fused peer-memory all-reduce inside the dot kernel; NCCL only as the fallback
    boxes = boxes_e2e = None
    if dist is not None and os.environ.get("BABELSTREAM_B200_BENCH_COMBINE", "p2p") == "p2p":
        from babelstream_b200.group import connect_mailboxes
        boxes = connect_mailboxes(dist, rank, world)
        boxes_e2e = connect_mailboxes(dist, rank, world) if boxes is not None else None
        if boxes_e2e is None and boxes is not None:
            boxes.close()
            boxes = None
        if boxes is None and rank == 0:
            print("bench.py: peer mailboxes unavailable (" + connect_mailboxes.last_error + "): dot falls back to NCCL",
                  file=sys.stderr)
    combine = "none" if dist is None else ("p2p" if boxes is not None else "nccl")
    dot_seq = [0]

    side = torch.cuda.Stream() if combine == "p2p" else None   # where the collecting half runs
    ev_sent = torch.cuda.Event() if combine == "p2p" else None
    ev_collected = torch.cuda.Event() if combine == "p2p" else None

    def collect_async():
        # the second half of the split all-reduce: a one-warp kernel on a SIDE stream (as NCCL's kernel
        # would be), after the send; the main stream only waits for it before the next send
        ev_sent.record()
        side.wait_event(ev_sent)
        bs.launch_dot_collect_p2p(eb, dot_out.data_ptr(), world=world, rank=rank, seq=dot_seq[0], mailboxes=boxes.ptrs,
                                  stream=side.cuda_stream)
        ev_collected.record(side)

    def collect():
        if combine == "p2p" and pending[0] == "sent":
            torch.cuda.current_stream().wait_event(ev_collected)
            pending[0] = None

    def enqueue(kernel):
        if kernel == "dot" and combine == "p2p":
            # Nothing later in the step depends on the sum, so the all-reduce is split: the dot kernel SENDS
            # its partial to every peer's mailbox over NVLink and returns; a one-warp kernel on a side stream
            # collects the partials, and the main stream joins it only before the next dot (and at the end
            # of the timed region), so the wait for the slowest rank hides under nstream/copy/mul/add/triad.
            # (The blocking e2e path below uses the fused call.)
            collect()
            dot_seq[0] += 1   # same count on every rank
            bs.launch_dot_send_p2p(pa, pb, n, eb, ws.data_ptr(), world=world, rank=rank, seq=dot_seq[0],
                                   mailboxes=boxes.ptrs, stream=st)
            collect_async()
            pending[0] = "sent"
        elif kernel == "dot":
            if pending[0] is not None:
                pending[0].wait()      # the previous step's all-reduce still owns dot_out
            bs.launch_dot(pa, pb, n, eb, ws.data_ptr(), dot_out.data_ptr(), stream=st)
            if dist is not None:
                # The one collective of the path: one element over NVLink.  Nothing later in the step
                # depends on the sum, so it runs on NCCL's stream underneath nstream/copy/... and is
                # only waited for before dot_out is overwritten (and before the host reads it).
                pending[0] = dist.all_reduce(dot_out, async_op=True)
        else:
            bs.launch(kernel, pa, pb, pc, n, eb, stream=st)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-timed region ------------------------------------------------------------------
    bs.launch("init", pa, pb, pc, n, eb, stream=st)
    for _ in range(W):
        for k in KERNEL_ORDER:
            enqueue(k)
    collect()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    # re-initialise so that K steps of a' = 1.92 a cannot overflow, exactly as a fresh driver run
    bs.launch("init", pa, pb, pc, n, eb, stream=st)
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6 * K + 1)]
    launches0 = bs.launch_count()
    ev[0].record()
    i = 1
    for _ in range(K):
        for k in KERNEL_ORDER:
            enqueue(k)
            ev[i].record()
            i += 1
    if pending[0] == "sent":
        collect()
        ev[-1].record()            # the timed region ends when the last sum has been collected
    elif pending[0] is not None:
        pending[0].wait()
        ev[-1].record()            # the timed region ends when the last all-reduce has landed
    torch.cuda.synchronize()
    launches = bs.launch_count() - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    per_kernel_ms = {k: [] for k in KERNEL_ORDER}
    for j in range(6 * K):
        per_kernel_ms[KERNEL_ORDER[j % 6]].append(ev[j].elapsed_time(ev[j + 1]))
    dot_value = float(dot_out.item())
    barrier()

    # ---- end to end: the blocking Stream<T> calls, host clock -----------------------------------
    del a, b, c
    torch.cuda.empty_cache()
    s = bs.B200Stream(n, device=local, dtype=np.float64, ngpus=1)
    e2e_dot = torch.zeros(1, dtype=torch.float64, device="cuda")
    if combine == "p2p":
        s.join_group(world, rank, boxes_e2e.ptrs)   # dot() now returns the all-rank sum, one launch, no NCCL

    def e2e_step():
        for k in KERNEL_ORDER:
            if k == "dot":
                v = s.dot()                       # device -> host: the sum, every step
                if combine == "nccl":
                    e2e_dot.fill_(float(v))
                    dist.all_reduce(e2e_dot)
                    v = e2e_dot.item()
            else:
                s.run(k)
        return v

    for _ in range(W):
        e2e_step()
    s.init_arrays()
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        e2e_v = e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    # ---- supplementary: the same step fed from and drained to HOST buffers -------------------------
    # Not the reference's metric (Stream<T> has no per-step host inputs, see `e2e.how`), reported so that
    # the cost of moving the arrays over PCIe every step can be read next to it: pinned host a, b, c ->
    # b200_write_arrays (H2D) -> the six blocking kernels -> b200_read_arrays (D2H), host clock.
    host_buffers = None
    if world == 1:
        try:
            hk = 3
            pinned = [torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(3)]
            hv = [p.numpy() for p in pinned]
            for x, v in zip(hv, (0.1, 0.2, 0.0)):
                x.fill(v)

            def host_step():
                s.write_arrays(*hv)
                for k in KERNEL_ORDER:
                    s.run(k)
                s.read_arrays(out=hv)

            host_step()
            t0 = time.perf_counter()
            for _ in range(hk):
                host_step()
            hs = (time.perf_counter() - t0) / hk
            host_buffers = {"value": round(step_bytes / hs / 1e9, 1), "unit": UNIT, "ms_per_step": round(hs * 1e3, 3),
                            "steps": hk, "h2d_bytes_per_step": 3 * n * eb, "d2h_bytes_per_step": 3 * n * eb + eb,
                            "pcie_gbs": round(2 * 3 * n * eb / hs / 1e9, 1),
                            "how": "supplementary, PCIe-bound: every step copies a, b, c from pinned host memory "
                                   "(b200_write_arrays), runs the six blocking kernels and copies a, b, c back "
                                   "(b200_read_arrays); the reference never does this per step -- its arrays are "
                                   "device-resident and get_arrays runs once after the timed loop (src/main.cpp:291-294)"}
            del pinned, hv
        except Exception as e:  # noqa: BLE001
            host_buffers = {"error": str(e)[:200]}
    ga, gb, gc = s.get_arrays()
    # cheap sanity on the e2e arrays: every element equal (the inputs are constants), finite
    sane = bool(np.isfinite(ga[0]) and (ga == ga[0]).all() and (gb == gb[0]).all() and (gc == gc[0]).all())
    s.close()
    if boxes is not None:
        barrier()
        boxes.close()
        boxes_e2e.close()

    # ---- max over ranks ---------------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_s * 1e3] + [statistics.fmean(per_kernel_ms[k]) for k in KERNEL_ORDER],
                     dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t = t.cpu().tolist()
    total_ms, e2e_ms = t[0], t[1]
    kernel_ms = dict(zip(KERNEL_ORDER, t[2:]))
    if dist is not None:
        dist.destroy_process_group()
    if rank != 0:
        return 0

    peak, peak_src = measured_peak()
    # what the reference's cuda model prints as PEAK: 2 x memory clock x bus width (CUDAStream.cu:107-108)
    device_peak = None
    try:
        import ctypes as C
        clk, bus = C.c_int(), C.c_int()
        if bs.lib().b200_device_info(local, None, None, None, None, None, C.byref(clk), C.byref(bus)) == 0:
            device_peak = round(2.0 * (clk.value / 1000.0) * (bus.value / 8) / 1000.0, 1)
    except Exception:
        pass
    ms_per_step = total_ms / K
    value = world * step_bytes / (ms_per_step * 1e-3) / 1e9
    kernels = {k: round(world * WEIGHT[k] * eb * n / (kernel_ms[k] * 1e-3) / 1e9, 1) for k in KERNEL_ORDER}
    triad_gbs_per_gpu = WEIGHT["triad"] * eb * n / (kernel_ms["triad"] * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"triad_f64_2^{args.log2n}")
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": round(ms_per_step, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"BabelStream fp64, 2^{args.log2n} elements per GPU, all six kernels (copy mul add triad dot nstream)",
                   "elements_per_gpu": n, "elements_total": n * world, "kernels": KERNEL_ORDER,
                   "bytes_per_step_per_gpu": step_bytes,
                   "l2": f"inputs larger than L2: 3 arrays x {n * eb / 1e6:.0f} MB per GPU vs 126 MB L2, no flush needed",
                   "parallelism": (f"{world} ranks (one process per GPU), contiguous shards, no collective for the elementwise kernels; "
                                   + ("dot: partials exchanged by the dot kernel itself over NVLink peer memory (CUDA IPC mailboxes), no NCCL; device-timed path: send in the dot kernel, one-warp collect on a side stream, joined before the next dot; blocking e2e path: fused send+collect in one launch"
                                      if combine == "p2p" else
                                      "dot: one NCCL all-reduce of 1 element per step, overlapped with the following kernels"))
                                  if world > 1 else "1 GPU",
                   "dot_combine": combine,
                   "timing": "CUDA events on the launching stream, max over ranks"},
        "kernels_gbs": kernels,
        "kernels_frac_of_peak": {k: round(v / (world * peak), 4) for k, v in kernels.items()},
        "peak_gbs_per_gpu": peak, "peak_source": peak_src,
        "roofline": {"bound": "hbm", "kernel": "triad (ew_kernel<double, OP_TRIAD, ...>)", "achieved": round(triad_gbs_per_gpu, 1),
                     "peak": peak, "unit": "GB/s", "frac": round(triad_gbs_per_gpu / peak, 4),
                     "frac_of_nominal_8000": round(triad_gbs_per_gpu / 8000.0, 4),
                     "frac_of_device_formula_peak": round(triad_gbs_per_gpu / device_peak, 4) if device_peak else None,
                     "device_formula_peak": device_peak, "traffic": traffic,
                     "algorithmic_bytes_per_launch": WEIGHT["triad"] * eb * n, "avg_launch_ms": round(kernel_ms["triad"], 6),
                     "peak_source": peak_src},
        "e2e": {"value": round(world * step_bytes / (e2e_ms / K * 1e-3) / 1e9, 1), "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": eb, "ms_per_step": round(e2e_ms / K, 5),
                "how": "blocking b200_* Stream<T> calls timed with the host clock, as src/main.cpp:109-116 does; "
                       "kernel arguments only on the way in, dot's sum read back every step", "sane": sane},
        "gpu_launches": int(launches), "clocks": clocks, "dot_last": dot_value,
    }
    if host_buffers is not None:
        line["e2e_host_buffers"] = host_buffers

    # ---- CPU baseline and extras (rank 0, N = 1 only) ---------------------------------------------
    if world == 1:
        try:
            line["cpu_baseline"], _ = cpu_reference(n, 30)
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"] = {"error": str(e)[:200]}
        if not args.no_extras:
            extras = {}
            b200_bin = os.path.join(ROOT, "build", "b200-stream")
            cuda_bin = os.path.join(ROOT, "oracle", "_ref", "cuda-stream")
            thrust_bin = os.path.join(ROOT, "oracle", "_ref", "thrust-stream")  # second library baseline
            try:
                if os.path.exists(b200_bin):
                    extras["b200_stream_driver"] = {k: round(v["gbs"], 1) for k, v in run_stream_binary(b200_bin, n, 100).items()}
                if os.path.exists(cuda_bin):
                    extras["reference_cuda_stream_driver"] = {k: round(v["gbs"], 1) for k, v in run_stream_binary(cuda_bin, n, 100).items()}
                if os.path.exists(thrust_bin):
                    try:
                        extras["reference_thrust_stream_driver"] = {k: round(v["gbs"], 1) for k, v in run_stream_binary(thrust_bin, n, 100).items()}
                    except Exception as e:  # noqa: BLE001  (a library baseline must not cost the other extras)
                        extras["reference_thrust_stream_driver"] = {"error": str(e)[:200]}
                big = 1 << 29
                if os.path.exists(b200_bin):
                    r = run_stream_binary(b200_bin, big, 30, only="Triad")
                    extras["north_star_triad_2^29"] = {"b200_stream_driver_gbs": round(r["triad"]["gbs"], 1),
                                                      "frac_of_measured_peak": round(r["triad"]["gbs"] / peak, 4),
                                                      "frac_of_nominal_8000": round(r["triad"]["gbs"] / 8000, 4),
                                                      "frac_of_device_formula_peak": round(r["triad"]["gbs"] / device_peak, 4) if device_peak else None}
                    if os.path.exists(cuda_bin):
                        r2 = run_stream_binary(cuda_bin, big, 30, only="Triad")
                        extras["north_star_triad_2^29"]["reference_cuda_stream_driver_gbs"] = round(r2["triad"]["gbs"], 1)
                    if os.path.exists(thrust_bin) and "error" not in extras.get("reference_thrust_stream_driver", {}):
                        r3 = run_stream_binary(thrust_bin, big, 30, only="Triad")
                        extras["north_star_triad_2^29"]["reference_thrust_stream_driver_gbs"] = round(r3["triad"]["gbs"], 1)
            except Exception as e:  # noqa: BLE001
                extras["error"] = str(e)[:300]
            line["driver"] = extras
    print(json.dumps(line))
    return 0


if __name__ == "__main__":
    sys.exit(main())

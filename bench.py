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
          / device time per step: two CUDA events on the launching stream around the K timed steps,
          arrays resident in HBM, kernels enqueued back to back through LEVEL 1 of the C ABI with
          programmatic dependent launch (cfg.flags = B200_LAUNCH_PDL) and alternating sweep direction
          (B200_LAUNCH_REVERSE on every other launch, what LEVEL 2 does by default), max over ranks.
kernels = per-kernel GB/s from a SECOND timed region of K steps with an event at every kernel boundary
          (an event between two kernels forbids their overlap, so this region runs without PDL and its
          step time is reported as `value_with_kernel_events`).  `roofline` is triad's average launch
          duration in that region against the measured HBM copy peak, plus the north-star point
          (triad fp64 2^29 through the unmodified driver) as `roofline.north_star`.
e2e     = the same bytes / host wall time per step through LEVEL 2 (the blocking Stream<T> calls
          that the reference's src/main.cpp times with its chrono clock): launch + wait per
          kernel, dot's result read back to the host every step.  This path has no per-step host
          input buffers by the reference's design (the arrays live on the device for the whole
          run, src/Stream.h:29-49), so h2d_bytes_per_step is 0 and d2h is the dot scalar;
          `e2e_host_buffers` (arrays in and out over PCIe every step) and `e2e_whole_run` (wall time of
          the whole driver program: constructor, init, timed loop, get_arrays, check_solution) sit beside it.
parity  = after the timed e2e steps every rank compares its shard of a, b, c bit for bit with the gold
          recurrence of src/main.cpp:331-347 and the all-rank dot with goldA*goldB*n*N (src/main.cpp:349,384);
          a mismatch fails the run.
strong_scaling = BASELINE configs 4 and 5: triad and nstream at 2^30 TOTAL elements sharded over the N GPUs
          (and --only All at 2^31 on 8), through the unmodified driver + the in-process sharded model
          (check_solution decides its exit code) and through one process per GPU.
cpu_baseline / --impl reference = the reference's own OpenMP model (oracle/_ref, built from the
          unmodified sources) on this box's host cores, same element count, kernel list, W and K.
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
from fractions import Fraction

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG2N = 25
KERNEL_ORDER = ["copy", "mul", "add", "triad", "dot", "nstream"]
WEIGHT = {"copy": 2, "mul": 2, "add": 3, "triad": 3, "dot": 2, "nstream": 4}
STEP_WEIGHT = sum(WEIGHT.values())  # 16 arrays moved per step
METRIC = "babelstream_all_kernels_gbytes_per_sec"
UNIT = "GB/s"
START_A, START_B, START_C, SCALAR = 0.1, 0.2, 0.0, 0.4   # src/Stream.h:24-27


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def make_config(log2n: int, world: int) -> dict:
    """`config` of the JSON line: identical, key for key, in both arms."""
    n = 1 << log2n
    return {"workload": f"BabelStream fp64, 2^{log2n} elements per GPU, all six kernels (copy mul add triad dot nstream)",
            "elements_per_gpu": n, "elements_total": n * world, "kernels": KERNEL_ORDER,
            "bytes_per_step_per_gpu": STEP_WEIGHT * 8 * n,
            "l2": f"inputs larger than L2: 3 arrays x {n * 8 / 1e6:.0f} MB per GPU vs 126 MB L2, no flush needed",
            "parallelism": "1 GPU" if world == 1 else
                           f"{world} ranks (one process per GPU), contiguous shards, no collective for the elementwise kernels, "
                           "dot partials combined by one exchange step",
            "dot_combine": "none (1 GPU)" if world == 1 else
                           "p2p: peer-memory all-reduce over NVLink inside the dot kernel (nccl: one ncclAllReduce, when peer "
                           "access is missing) -- the mode that ran is the line's top-level dot_combine",
            "timing": "CUDA events on the launching stream, max over ranks"}


def fma(x: float, y: float, z: float) -> float:
    """Correctly rounded x*y+z (Python 3.12 has no math.fma): exact rational arithmetic, one rounding."""
    return float(Fraction(x) * Fraction(y) + Fraction(z))


def gold_recurrence(steps: int, contract: bool = True):
    """check_solution's gold values (src/main.cpp:331-370) for `steps` passes of copy mul add triad dot nstream,
    with the contraction the reference's default build applies (-ffp-contract=fast: b + s*c is one fma).
    Returns (a, b, c, a_at_dot, b_at_dot) as Python floats (= IEEE doubles)."""
    a, b, c, s = START_A, START_B, START_C, SCALAR
    a_dot = b_dot = 0.0
    axpy = (lambda bb, cc: fma(s, cc, bb)) if contract else (lambda bb, cc: bb + s * cc)
    for _ in range(steps):
        c = a                 # copy
        b = s * c             # mul
        c = a + b             # add
        a = axpy(b, c)        # triad
        a_dot, b_dot = a, b   # dot reads a and b here
        a = a + axpy(b, c)    # nstream
    return a, b, c, a_dot, b_dot


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


# CPUs this program may use, taken before any OpenMP runtime is loaded: libgomp binds the thread that loads it
# (OMP_PROC_BIND) and a child process would inherit that one-CPU mask.
_START_AFFINITY = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None


class _start_affinity:
    """Runs a child with the CPU mask this program was started with (the mask is inherited at fork from the calling
    thread; set around the call rather than in a preexec_fn, which is unsafe in a process with other threads)."""

    def __enter__(self):
        self.saved = None
        if _START_AFFINITY:
            try:
                self.saved = os.sched_getaffinity(0)
                os.sched_setaffinity(0, _START_AFFINITY)
            except OSError:
                self.saved = None

    def __exit__(self, *exc):
        if self.saved:
            try:
                os.sched_setaffinity(0, self.saved)
            except OSError:
                pass
        return False


def run_stream_binary(path, n, num_times, only="All", env=None, timeout=900, want_wall=False, extra_args=()):
    """Runs a BabelStream binary (ours or a reference model); a non-zero exit (check_solution failed) raises."""
    e = dict(os.environ)
    e.update(env or {})
    t0 = time.perf_counter()
    with _start_affinity():
        p = subprocess.run([path, "--csv", "--gigabytes", "-s", str(n), "-n", str(num_times), "--only", only, *extra_args],
                           capture_output=True, text=True, timeout=timeout, env=e)
    wall = time.perf_counter() - t0
    if p.returncode != 0:
        raise RuntimeError(f"{os.path.basename(path)} exited {p.returncode}: {p.stderr[-400:]}")
    rows = parse_stream_csv(p.stdout)
    return (rows, wall) if want_wall else rows


def config3_size_sweep(binaries, l2_bytes=126 << 20):
    """BASELINE config 3: fp32 and fp64, 2^20 ... 2^30 elements, --only All through the unmodified driver (its own clock,
    min time per kernel as it prints it), for each binary in `binaries` ({label: path}).  Every row says whether the three
    arrays fit the L2 (the knee between the L2-resident and the HBM-bound regime)."""
    rows = []
    for log2n in (20, 22, 24, 26, 28, 30):
        n = 1 << log2n
        for dt, eb, flag in (("f64", 8, ()), ("f32", 4, ("--float",))):
            foot = 3 * n * eb
            row = {"log2n": log2n, "dtype": dt, "footprint_mb": round(foot / 1e6, 1),
                   "regime": "L2-resident" if foot <= l2_bytes else ("knee" if foot <= 4 * l2_bytes else "HBM")}
            num_times = 100 if log2n <= 26 else 30 if log2n <= 28 else 10
            if dt == "f32":
                num_times = min(num_times, 30)   # --only All -n 100 overflows fp32's dot in the reference's own recurrence
            for label, path in binaries.items():
                try:
                    r = run_stream_binary(path, n, num_times, extra_args=flag)   # non-zero exit = check_solution failed
                    t = sum(r[k]["min_s"] for k in KERNEL_ORDER)
                    row[label] = {"step_gbs": round(STEP_WEIGHT * eb * n / t / 1e9, 1), "triad_gbs": round(r["triad"]["gbs"], 1),
                                  "dot_gbs": round(r["dot"]["gbs"], 1), "copy_gbs": round(r["copy"]["gbs"], 1)}
                except Exception as e:  # noqa: BLE001
                    row[label] = {"error": str(e)[:160]}
            rows.append(row)
    return {"what": "BASELINE config 3 through the unmodified driver: --only All, min time per kernel (src/main.cpp clock), "
                    "check_solution passed in every run that reports a figure; step_gbs = bytes of the six kernels / sum of their "
                    "min times; -n 100 up to 2^26 (fp32: 30, the recurrence overflows fp32's dot sum at 100 in every model), 30 at 2^28, "
                    "10 at 2^30",
            "l2_mb": round(l2_bytes / 1e6, 1), "rows": rows}


def step_gbs_from_csv(rows, n, key="avg_s"):
    t = sum(rows[k][key] for k in KERNEL_ORDER)
    return STEP_WEIGHT * 8 * n / t / 1e9, t


# ---------------------------------------------------------------------------------------------------
# The reference's CPU implementation of the path (oracle/ is executed here, and only here, by bench.py)
# ---------------------------------------------------------------------------------------------------
def omp_env():
    """All host threads the box offers: one OpenMP thread per LOGICAL CPU, pinned (OMP_PLACES=threads)."""
    threads = os.cpu_count() or 1
    return {"OMP_NUM_THREADS": str(threads), "OMP_PROC_BIND": "spread", "OMP_PLACES": "threads"}, threads


def cpu_reference(n, warmup, steps):
    """W untimed + K timed steps of the six blocking Stream<T> calls of the reference's UNMODIFIED OpenMP model
    (src/omp/OMPStream.cpp behind oracle/ref_shim.cpp, kind 'reference'), host clock -- the same loop the b200
    arm's e2e runs.  Falls back to the oracle's C restatement (kind 'port') when oracle/_ref was not built."""
    env, threads = omp_env()
    os.environ.update(env)   # libgomp reads these when it is loaded, i.e. below
    import numpy as np
    import oracle as O
    if O.Reference.available("omp"):
        ref = O.Reference("omp", np.float64, n)
        calls = {"copy": ref.copy, "mul": ref.mul, "add": ref.add, "triad": ref.triad, "dot": ref.dot, "nstream": ref.nstream}

        def step():
            for k in KERNEL_ORDER:
                calls[k]()
        for _ in range(warmup):
            step()
        ref.init()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        t = (time.perf_counter() - t0) / steps
        ref.close()
        kind, what = "reference", "oracle/_ref/libref_omp.so = unmodified src/omp/OMPStream.cpp"
    else:
        orc = O.Oracle()
        a, b, c = (np.empty(n, dtype=np.float64) for _ in range(3))
        orc.init(a, b, c)
        orc.run(O.ALL, O.ORDER_CLASSIC, max(1, warmup), a, b, c, omp_dot=True)
        orc.init(a, b, c)
        t0 = time.perf_counter()
        orc.run(O.ALL, O.ORDER_CLASSIC, steps, a, b, c, omp_dot=True)
        t = (time.perf_counter() - t0) / steps
        kind, what = "port", "oracle C restatement (OpenMP)"
    return {"value": round(STEP_WEIGHT * 8 * n / t / 1e9, 2), "unit": UNIT, "cores": threads,
            "cores_are": "logical CPUs (os.cpu_count()); OMP_NUM_THREADS=all of them, OMP_PROC_BIND=spread, OMP_PLACES=threads",
            "kind": kind,
            "sample": f"{what}: {warmup} untimed + {steps} timed steps of the six blocking calls at n={n} fp64, host clock "
                      f"({t * 1e3:.2f} ms per step)"}, t


def whole_run(binary, n, num_times, env=None):
    """Wall time of a whole driver program as a user runs it: process start, constructor, init_arrays, the timed loop,
    get_arrays (device -> host where there is a device) and check_solution."""
    rows, wall = run_stream_binary(binary, n, num_times, only="All", env=env, want_wall=True)
    loop_s = sum(rows[k]["avg_s"] for k in KERNEL_ORDER) * num_times
    return {"value": round(STEP_WEIGHT * 8 * n * num_times / wall / 1e9, 1), "unit": UNIT, "wall_s": round(wall, 3),
            "timed_loop_s": round(loop_s, 4), "command": f"{os.path.relpath(binary, ROOT)} -s {n} -n {num_times} --only All"}


def main_reference(args, rank, world):
    if rank != 0:
        return 0
    n = 1 << args.log2n
    W, K = max(3, args.warmup), args.steps
    base, t = cpu_reference(n, W, K)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": K, "warmup": W, "ms_per_step": round(t * 1e3, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": make_config(args.log2n, args.gpus),
            "elements_run": n,
            "elements_run_note": f"this arm runs 2^{args.log2n} elements on the host whatever --gpus says; config.elements_total "
                                 f"({n * args.gpus}) is the GPU arm's job at N = {args.gpus}, repeated here so that the two lines "
                                 "carry the same config",
            "run_as": "the reference's OpenMP model on the host CPU cores (no GPU); each step is a bounded sample of the "
                      f"workload: one pass of the six kernels over 2^{args.log2n} fp64 elements, whatever --gpus says",
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    exe = os.path.join(ROOT, "oracle", "_ref", "omp-stream")
    if os.path.exists(exe) and not args.no_extras:
        try:
            env, _ = omp_env()
            w = whole_run(exe, n, 100, env=env)
            w.update({"h2d_bytes": 0, "d2h_bytes": 0})
            line["e2e_whole_run"] = w
        except Exception as e:  # noqa: BLE001
            line["e2e_whole_run"] = {"error": str(e)[:200]}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------
# BASELINE configs 4 and 5: a FIXED total size sharded over the N GPUs
# ---------------------------------------------------------------------------------------------------
def strong_scaling(world, rank, local, dist, cpu_group, no_extras):
    """triad and nstream at 2^30 fp64 elements in total (25.8 GB), 2^30/N per GPU; at N = 8 also --only All at 2^31.
    Two layouts: (a) one process per GPU (this program's ranks, no collective for these kernels), device-timed
    back to back and as blocking calls; (b) ONE process, the unmodified driver + the in-process sharded model
    (BABELSTREAM_B200_NGPUS=N), run by rank 0 while the other ranks wait on the CPU -- its exit code is
    check_solution's verdict.  Efficiency = aggregate GB/s / (N x the N = 1 figure of the same layout-(b) binary)."""
    import numpy as np
    import torch
    import babelstream_b200 as bs

    total = 1 << 30
    per = total // world
    out = {"elements_total": total, "elements_per_gpu": per, "dtype": "f64"}

    def cpu_barrier():
        if dist is not None:
            dist.barrier(group=cpu_group)

    # ---- (a) one process per GPU ---------------------------------------------------------------
    st = torch.cuda.current_stream().cuda_stream
    pdl = bs.LaunchCfg.make(flags=bs.LAUNCH_PDL)
    res = {}
    with bs.B200Stream(per, device=local, dtype=np.float64) as s:
        info = s.shard_info(0)
        pa, pb, pc = info["a"], info["b"], info["c"]
        for k in ("triad", "nstream"):
            s.init_arrays()
            for _ in range(3):
                bs.launch(k, pa, pb, pc, per, 8, cfg=pdl, stream=st)
            torch.cuda.synchronize()
            cpu_barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 20
            e0.record()
            for _ in range(reps):
                bs.launch(k, pa, pb, pc, per, 8, cfg=pdl, stream=st)
            e1.record()
            torch.cuda.synchronize()
            dev_ms = e0.elapsed_time(e1) / reps
            s.init_arrays()
            for _ in range(3):
                s.run(k)
            cpu_barrier()
            best = 1e30
            for _ in range(reps):
                t0 = time.perf_counter()
                s.run(k)
                best = min(best, time.perf_counter() - t0)
            res[k] = (dev_ms, best * 1e3)
    torch.cuda.empty_cache()
    t = torch.tensor([res["triad"][0], res["triad"][1], res["nstream"][0], res["nstream"][1]], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t = t.cpu().tolist()
    out["one_process_per_gpu"] = {
        "triad": {"device_timed_gbs": round(3 * 8 * total / t[0] / 1e6, 1), "blocking_call_gbs": round(3 * 8 * total / t[1] / 1e6, 1)},
        "nstream": {"device_timed_gbs": round(4 * 8 * total / t[2] / 1e6, 1), "blocking_call_gbs": round(4 * 8 * total / t[3] / 1e6, 1)},
        "how": "each rank times its own shard (CUDA events over 20 launches; best of 20 blocking calls, host clock), "
               "aggregate = total bytes / max over ranks"}

    # ---- (b) the unmodified driver, in-process sharded model (rank 0) ------------------------------
    b200_bin = os.path.join(ROOT, "build", "b200-stream")
    if rank == 0 and os.path.exists(b200_bin) and not no_extras:
        drv = {}
        try:
            for only, key, w in (("Triad", "triad", 3), ("Nstream", "nstream", 4)):
                rows = run_stream_binary(b200_bin, total, 20, only=only, env={"BABELSTREAM_B200_NGPUS": str(world)})
                drv[key] = {"gbs": round(rows[key]["gbs"], 1), "check_solution": "passed"}
                if world > 1:
                    one = run_stream_binary(b200_bin, total, 20, only=only, env={"BABELSTREAM_B200_NGPUS": "1"})
                    drv[key]["gbs_1gpu"] = round(one[key]["gbs"], 1)
                    drv[key]["efficiency"] = round(rows[key]["gbs"] / (world * one[key]["gbs"]), 4)
            if world == 8:
                rows = run_stream_binary(b200_bin, 1 << 31, 20, only="All", env={"BABELSTREAM_B200_NGPUS": "8"}, timeout=1200)
                drv["all_2^31"] = {k: round(v["gbs"], 1) for k, v in rows.items()}
                drv["all_2^31"]["check_solution"] = "passed"
            if world > 1:
                # BASELINE config 5's "one NCCL allreduce over NVLink", next to the two other ways the in-process model can
                # combine the G partial sums; --only Dot, the driver's clock, check_solution decides the exit code
                combos = {}
                for nd, label in ((world << 25, f"2^25 x {world}"),) + (((1 << 31, "2^31"),) if world == 8 else ()):
                    per_mode = {}
                    for mode in ("host", "p2p", "nccl"):
                        try:
                            rows = run_stream_binary(b200_bin, nd, 100, only="Dot", timeout=600,
                                                     env={"BABELSTREAM_B200_NGPUS": str(world), "BABELSTREAM_B200_DOT_COMBINE": mode})
                            per_mode[mode] = {"gbs": round(rows["dot"]["gbs"], 1), "check_solution": "passed"}
                        except Exception as e:  # noqa: BLE001
                            per_mode[mode] = {"error": str(e)[:200]}
                    combos[label] = per_mode
                combos["how"] = (f"BABELSTREAM_B200_NGPUS={world} BABELSTREAM_B200_DOT_COMBINE=host|p2p|nccl build/b200-stream --only Dot "
                                 "-n 100: host = rank-order sum of the G doorbell-delivered partials on the caller's thread, p2p = "
                                 "all-reduce over peer memory inside the dot kernel, nccl = dot kernel + one ncclAllReduce per call")
                drv["dot_combine"] = combos
        except Exception as e:  # noqa: BLE001
            drv["error"] = str(e)[:300]
        drv["how"] = (f"BABELSTREAM_B200_NGPUS={world} build/b200-stream -s 2^30 -n 20 --only Triad|Nstream (the reference's "
                      "unmodified main.cpp; GB/s from its own clock; a failed check_solution is a non-zero exit = error)")
        out["driver_in_process"] = drv
    cpu_barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the driver-binary runs (2^29 north star, configs 4/5, whole-run)")
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
    dist = cpu_group = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        cpu_group = dist.new_group(backend="gloo")   # waits that must not occupy the GPU (and the join barrier)
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
                                  workspace=ws.data_ptr(), stream=side.cuda_stream)
        ev_collected.record(side)

    def collect():
        if combine == "p2p" and pending[0] == "sent":
            torch.cuda.current_stream().wait_event(ev_collected)
            pending[0] = None

    # Consecutive kernels sweep the arrays in opposite directions (B200_LAUNCH_REVERSE), as LEVEL 2 does by default:
    # each kernel starts on the data its predecessor touched last, i.e. on what is still in L2.
    sweep = [0]
    cfgs = {(p, r): bs.LaunchCfg.make(flags=(bs.LAUNCH_PDL if p else 0) | (bs.LAUNCH_REVERSE if r else 0))
            for p in (0, 1) for r in (0, 1)}

    def enqueue(kernel, pdl_on):
        cfg = cfgs[(1 if pdl_on else 0, sweep[0])]
        sweep[0] ^= 1
        if kernel == "dot" and combine == "p2p":
            # Nothing later in the step depends on the sum, so the all-reduce is split: the dot kernel SENDS
            # its partial to every peer's mailbox over NVLink and returns; a one-warp kernel on a side stream
            # collects the partials, and the main stream joins it only before the next dot (and at the end
            # of the timed region), so the wait for the slowest rank hides under nstream/copy/mul/add/triad.
            # (The blocking e2e path below uses the fused call.)
            collect()
            dot_seq[0] += 1   # same count on every rank
            bs.launch_dot_send_p2p(pa, pb, n, eb, ws.data_ptr(), world=world, rank=rank, seq=dot_seq[0],
                                   mailboxes=boxes.ptrs, cfg=cfg, stream=st)
            collect_async()
            pending[0] = "sent"
        elif kernel == "dot":
            if pending[0] is not None:
                pending[0].wait()      # the previous step's all-reduce still owns dot_out
            bs.launch_dot(pa, pb, n, eb, ws.data_ptr(), dot_out.data_ptr(), cfg=cfg, stream=st)
            if dist is not None:
                # The one collective of the path: one element over NVLink.  Nothing later in the step
                # depends on the sum, so it runs on NCCL's stream underneath nstream/copy/... and is
                # only waited for before dot_out is overwritten (and before the host reads it).
                pending[0] = dist.all_reduce(dot_out, async_op=True)
        else:
            bs.launch(kernel, pa, pb, pc, n, eb, cfg=cfg, stream=st)

    def finish_pending(ev_end):
        if pending[0] == "sent":
            collect()
            ev_end.record()            # the timed region ends when the last sum has been collected
        elif pending[0] is not None:
            pending[0].wait()
            pending[0] = None
            ev_end.record()            # the timed region ends when the last all-reduce has landed

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-timed region A: `value` ---------------------------------------------------------
    bs.launch("init", pa, pb, pc, n, eb, stream=st)
    for _ in range(W):
        for k in KERNEL_ORDER:
            enqueue(k, True)
    collect()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    # re-initialise so that K steps of a' = 1.92 a cannot overflow, exactly as a fresh driver run
    bs.launch("init", pa, pb, pc, n, eb, stream=st)
    barrier()
    ev_a0, ev_a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = bs.launch_count()
    ev_a0.record()
    for _ in range(K):
        for k in KERNEL_ORDER:
            enqueue(k, True)
    ev_a1.record()
    finish_pending(ev_a1)
    torch.cuda.synchronize()
    launches = bs.launch_count() - launches0
    total_ms = ev_a0.elapsed_time(ev_a1)
    dot_value = float(dot_out.item())
    # the region behind `value` proves itself: its a, b, c after the K timed steps against the gold recurrence, bit for
    # bit, and its last all-rank dot -- outside the timed region, before region B re-initialises the arrays
    ra_gold = gold_recurrence(K)
    ra_bits = [int(np.asarray([g], dtype=np.float64).view(np.int64)[0]) for g in ra_gold[:3]]
    ra_wrong = sum(int((x.view(torch.int64) != gbits).sum().item()) for x, gbits in zip((a, b, c), ra_bits))
    ra_sum = ra_gold[3] * ra_gold[4] * float(n * world)
    ra_rel = abs(dot_value - ra_sum) / max(abs(dot_value), abs(ra_sum))
    ra_dot_ok = bool(np.isfinite(dot_value)) and ra_rel <= 1e7 * float(np.finfo(np.float64).eps)
    barrier()

    # ---- device-timed region B: an event at every kernel boundary (per-kernel GB/s, roofline) ---------
    bs.launch("init", pa, pb, pc, n, eb, stream=st)
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6 * K + 1)]
    ev[0].record()
    i = 1
    for _ in range(K):
        for k in KERNEL_ORDER:
            enqueue(k, False)
            ev[i].record()
            i += 1
    finish_pending(ev[-1])
    torch.cuda.synchronize()
    total_b_ms = ev[0].elapsed_time(ev[-1])
    per_kernel_ms = {k: [] for k in KERNEL_ORDER}
    for j in range(6 * K):
        per_kernel_ms[KERNEL_ORDER[j % 6]].append(ev[j].elapsed_time(ev[j + 1]))
    barrier()

    # ---- end to end: the blocking Stream<T> calls, host clock -----------------------------------
    del a, b, c
    torch.cuda.empty_cache()
    s = bs.B200Stream(n, device=local, dtype=np.float64, ngpus=1)
    e2e_dot = torch.zeros(1, dtype=torch.float64, device="cuda")
    if combine == "p2p":
        from babelstream_b200.group import join_group
        join_group(s, dist, world, rank, boxes_e2e)   # dot() now returns the all-rank sum, one launch, no NCCL

    def e2e_step():
        if combine != "nccl":
            # six blocking calls issued from C (b200_run_sequence), like the driver's loop over bench[]: each kernel is
            # launched and waited for before the next; only the interpreter's per-call cost (~1 us) is left out
            return s.run_sequence(KERNEL_ORDER)
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

    # ---- parity: this rank's shard against the gold recurrence, the all-rank dot against goldA*goldB*n*N --------
    ga, gb, gc = s.get_arrays()
    gold_a, gold_b, gold_c, a_dot, b_dot = gold_recurrence(K)
    u64 = lambda x: np.asarray([x], dtype=np.float64).view(np.uint64)[0]
    wrong = int((ga.view(np.uint64) != u64(gold_a)).sum() + (gb.view(np.uint64) != u64(gold_b)).sum()
                + (gc.view(np.uint64) != u64(gold_c)).sum())
    gold_sum = a_dot * b_dot * float(n * world)          # src/main.cpp:349
    dot_rel = abs(float(e2e_v) - gold_sum) / max(abs(float(e2e_v)), abs(gold_sum))
    dot_ok = dot_rel <= 1e7 * np.finfo(np.float64).eps and np.isfinite(float(e2e_v))   # src/main.cpp:384
    pt = torch.tensor([float(wrong), 0.0 if dot_ok else 1.0, dot_rel, float(ra_wrong), 0.0 if ra_dot_ok else 1.0, ra_rel],
                      dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(pt, op=dist.ReduceOp.MAX)
    pt = pt.cpu().tolist()
    parity = {"ok": pt[0] == 0 and pt[1] == 0 and pt[3] == 0 and pt[4] == 0,
              "value_region": {"ok": pt[3] == 0 and pt[4] == 0,
                               "array_elements_differing_from_gold_bits_max_over_ranks": int(pt[3]),
                               "dot": dot_value, "dot_gold": ra_sum, "dot_rel_err_max_over_ranks": pt[5],
                               "how": "the device-timed region A itself (asynchronous launches, dependent launch, split "
                                      "send/collect dot at N > 1): its a, b, c after the K timed steps bit for bit against the "
                                      "gold recurrence and its last all-rank dot within 1e7*eps, checked before the arrays are "
                                      "re-initialised"}, "array_elements_differing_from_gold_bits_max_over_ranks": int(pt[0]),
              "arrays_checked_per_rank": 3 * n, "steps": K,
              "gold_bits": {"a": "0x%016x" % u64(gold_a), "b": "0x%016x" % u64(gold_b), "c": "0x%016x" % u64(gold_c)},
              "dot": float(e2e_v), "dot_gold": gold_sum, "dot_rel_err_max_over_ranks": pt[2], "dot_bound": 1e7 * float(np.finfo(np.float64).eps),
              "how": "a, b, c of every rank after the K timed e2e steps, bit for bit against check_solution's gold recurrence "
                     "(src/main.cpp:331-347, fma-contracted as the reference's default build); the all-rank dot of the last step "
                     "against goldA*goldB*n*N within 1e7*eps (src/main.cpp:349,384)"}

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
    s.close()
    if boxes is not None:
        barrier()
        boxes.close()
        boxes_e2e.close()

    # ---- BASELINE configs 4 and 5 ------------------------------------------------------------------
    strong = None
    if args.log2n == LOG2N:
        try:
            strong = strong_scaling(world, rank, local, dist, cpu_group, args.no_extras)
        except Exception as e:  # noqa: BLE001
            strong = {"error": str(e)[:300]}
            if dist is not None:
                raise

    # ---- max over ranks ---------------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_s * 1e3, total_b_ms] + [statistics.fmean(per_kernel_ms[k]) for k in KERNEL_ORDER],
                     dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t = t.cpu().tolist()
    total_ms, e2e_ms, total_b_ms = t[0], t[1], t[2]
    kernel_ms = dict(zip(KERNEL_ORDER, t[3:]))
    if dist is not None:
        dist.destroy_process_group()
    if rank != 0:
        return 0 if parity["ok"] else 1

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
    parallel_detail = "1 GPU"
    if world > 1:
        parallel_detail = (
            "dot: partials exchanged by the dot kernel itself over NVLink peer memory (CUDA IPC mailboxes), no NCCL; device-timed "
            "path: send in the dot kernel, one-warp collect on a side stream, joined before the next dot; blocking e2e path: fused "
            "send+collect in one launch" if combine == "p2p" else
            "dot: one NCCL all-reduce of 1 element per step, overlapped with the following kernels")
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": round(ms_per_step, 5), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": make_config(args.log2n, world),
        "dot_combine": combine, "parallelism_detail": parallel_detail,
        "value_is": "device time: two CUDA events around K steps enqueued back to back through LEVEL 1 with programmatic "
                    "dependent launch; the like-for-like number against the reference's host-timed metric is `e2e`",
        "value_with_kernel_events": round(world * step_bytes / (total_b_ms / K * 1e-3) / 1e9, 1),
        "kernels_gbs": kernels,
        "kernels_frac_of_peak": {k: round(v / (world * peak), 4) for k, v in kernels.items()},
        "peak_gbs_per_gpu": peak, "peak_source": peak_src,
        "roofline": {"bound": "hbm", "kernel": "triad (ew_kernel<double, OP_TRIAD, ...>)", "achieved": round(triad_gbs_per_gpu, 1),
                     "peak": peak, "unit": "GB/s", "frac": round(triad_gbs_per_gpu / peak, 4),
                     "frac_of_nominal_8000": round(triad_gbs_per_gpu / 8000.0, 4),
                     "frac_of_device_formula_peak": round(triad_gbs_per_gpu / device_peak, 4) if device_peak else None,
                     "device_formula_peak": device_peak, "traffic": traffic,
                     "algorithmic_bytes_per_launch": WEIGHT["triad"] * eb * n, "avg_launch_ms": round(kernel_ms["triad"], 6),
                     "measured_in": "region B (an event at every kernel boundary, K launches averaged)",
                     "peak_source": peak_src},
        "e2e": {"value": round(world * step_bytes / (e2e_ms / K * 1e-3) / 1e9, 1), "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": eb, "ms_per_step": round(e2e_ms / K, 5),
                "how": "blocking b200_* Stream<T> calls (six per step, issued by b200_run_sequence) timed with the host clock, as src/main.cpp:109-116 does; "
                       "kernel arguments only on the way in, dot's sum read back every step; the arrays are device-resident "
                       "by the reference's design (src/Stream.h:29-49), see e2e_host_buffers and e2e_whole_run for the PCIe-inclusive views"},
        "parity": parity,
        "gpu_launches": int(launches), "clocks": clocks, "dot_last": dot_value,
    }
    if host_buffers is not None:
        line["e2e_host_buffers"] = host_buffers
    if strong is not None:
        line["strong_scaling"] = strong

    # ---- CPU baseline and extras (rank 0, N = 1 only) ---------------------------------------------
    if world == 1:
        try:
            line["cpu_baseline"], _ = cpu_reference(n, 2, 30)
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"] = {"error": str(e)[:200]}
        if not args.no_extras:
            extras = {}
            b200_bin = os.path.join(ROOT, "build", "b200-stream")
            cuda_bin = os.path.join(ROOT, "oracle", "_ref", "cuda-stream")
            thrust_bin = os.path.join(ROOT, "oracle", "_ref", "thrust-stream")  # second library baseline
            omp_bin = os.path.join(ROOT, "oracle", "_ref", "omp-stream")
            try:
                if os.path.exists(b200_bin):
                    w = whole_run(b200_bin, n, 100)
                    w.update({"h2d_bytes": 0, "d2h_bytes": 3 * n * eb + 100 * eb,
                              "what": "wall time of the whole program a user runs: CUDA context + constructor, init_arrays, 100 x 6 blocking "
                                      "kernels, get_arrays (805 MB device -> host), check_solution; value = kernel bytes of the timed loop / wall"})
                    if os.path.exists(omp_bin):
                        env, _ = omp_env()
                        w["reference_omp_stream"] = whole_run(omp_bin, n, 100, env=env)
                    line["e2e_whole_run"] = w
                    extras["b200_stream_driver"] = {k: round(v["gbs"], 1) for k, v in run_stream_binary(b200_bin, n, 100).items()}
                    extras["b200_stream_driver_sweep_forward"] = {
                        k: round(v["gbs"], 1)
                        for k, v in run_stream_binary(b200_bin, n, 100, env={"BABELSTREAM_B200_SWEEP": "forward"}).items()}
                if os.path.exists(cuda_bin):
                    extras["reference_cuda_stream_driver"] = {k: round(v["gbs"], 1) for k, v in run_stream_binary(cuda_bin, n, 100).items()}
                if os.path.exists(thrust_bin):
                    try:
                        extras["reference_thrust_stream_driver"] = {k: round(v["gbs"], 1) for k, v in run_stream_binary(thrust_bin, n, 100).items()}
                    except Exception as e:  # noqa: BLE001  (a library baseline must not cost the other extras)
                        extras["reference_thrust_stream_driver"] = {"error": str(e)[:200]}
                big = 1 << 29
                if os.path.exists(b200_bin):
                    r = max((run_stream_binary(b200_bin, big, 30, only="Triad") for _ in range(2)), key=lambda x: x["triad"]["gbs"])
                    ns = {"workload": "triad fp64 2^29 elements through the unmodified driver (src/main.cpp clock, best of 2 runs of 30)",
                          "b200_stream_driver_gbs": round(r["triad"]["gbs"], 1),
                          "frac_of_measured_peak": round(r["triad"]["gbs"] / peak, 4),
                          "frac_of_nominal_8000": round(r["triad"]["gbs"] / 8000, 4),
                          "frac_of_device_formula_peak": round(r["triad"]["gbs"] / device_peak, 4) if device_peak else None}
                    # the same with every sweep bottom-up: how much of the figure is the previous call's tail still in L2
                    rf = run_stream_binary(b200_bin, big, 30, only="Triad", env={"BABELSTREAM_B200_SWEEP": "forward"})
                    ns["b200_stream_driver_gbs_sweep_forward"] = round(rf["triad"]["gbs"], 1)
                    ns["sweep"] = ("default: consecutive calls sweep the arrays in opposite directions, so a call starts on the "
                                   "<= 126 MB its predecessor left in L2 (of 12.9 GB moved per call at this size); "
                                   "sweep_forward: every call bottom-up, as the reference's kernels run")
                    if os.path.exists(cuda_bin):
                        r2 = run_stream_binary(cuda_bin, big, 30, only="Triad")
                        ns["reference_cuda_stream_driver_gbs"] = round(r2["triad"]["gbs"], 1)
                    if os.path.exists(thrust_bin) and "error" not in extras.get("reference_thrust_stream_driver", {}):
                        r3 = run_stream_binary(thrust_bin, big, 30, only="Triad")
                        ns["reference_thrust_stream_driver_gbs"] = round(r3["triad"]["gbs"], 1)
                    line["roofline"]["north_star"] = ns
                    extras["north_star_triad_2^29"] = ns
                bins = {k: v for k, v in (("b200_stream", b200_bin), ("reference_cuda_stream", cuda_bin)) if os.path.exists(v)}
                if "b200_stream" in bins:
                    extras["config3_size_sweep"] = config3_size_sweep(bins)
            except Exception as e:  # noqa: BLE001
                extras["error"] = str(e)[:300]
            line["driver"] = extras
    print(json.dumps(line))
    if not parity["ok"]:
        print("bench.py: PARITY FAILED -- " + json.dumps(parity), file=sys.stderr)
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())

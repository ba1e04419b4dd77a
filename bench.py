#!/usr/bin/env python
"""bench.py -- CAF cells/s of the fused frequency-search fastXcorr on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c4]

A "step" is one pass of the hot path over one synthetic batch:
  N = 1 : BASELINE.json configs[1] (C2): 8192-sample cutout, 2^20 shifts, 8192-bin FFT per shift.
  N > 1 : configs[4] (C5) scaled weakly: every rank searches 2^23 shifts of a shard (block + 8191-sample
          halo) of one N*2^23-sample record, so N = 8 is exactly the 2^26-sample C5 record; the only
          exchange is the allgather of the per-rank peak tuples (NCCL).
`value` = cells (shifts x bins) of all ranks / device time with inputs resident in HBM (CUDA events,
max over ranks); `e2e` = the same through the public API with pinned HOST buffers (H2D + D2H inside).
The reference arm (`--impl reference`) times the CPU restatement of the reference's numpy/scipy path
(oracle/) on all host cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "caf_cells_per_sec"
UNIT = "cells/s"
N_CUT = 8192
K_C2 = 1 << 20
K_RANK_MULTI = 1 << 23
ALGO_BYTES_PER_SHIFT = 16.0        # one new rx sample (8 B) + one result (f32 qf2 + u32 bin)
# dram__bytes_read.sum + dram__bytes_write.sum of one C2 launch of fused_fft_staged_kernel<8192>, from the ncu --set full
# capture summarised in profiles/r1_c2_staged_ncu.md (12.89 MB read, 0 written: the 8 MB of results stay in L2)
NCU_DRAM_BYTES_C2 = 12.89e6
FLOPS_PER_CELL = 75.0              # 5*log2(8192) + 10


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured"
    return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------------
# synthetic workload (seeded; block-keyed so that shards of different ranks agree on the halo)
# ------------------------------------------------------------------------------------------------
def noise_block(seed, b, n):
    rng = np.random.default_rng([seed, b])
    w = rng.standard_normal(2 * n, dtype=np.float32)
    return (np.float32(np.sqrt(0.5)) * w).view(np.complex64)


def make_shard(seed, rank, world, k_rank, N, p0, k0):
    """rx[rank*k_rank : (rank+1)*k_rank + N - 1] of the global record + the QPSK cutout."""
    from pydsproutines_b200 import signals
    cutout = signals.qpsk(np.random.default_rng([seed, 10 ** 6]), N)
    first = rank * k_rank
    a = noise_block(seed, rank, k_rank)
    b = noise_block(seed, rank + 1, k_rank)[:N - 1]
    rx = np.concatenate((a, b))
    lo, hi = max(p0, first), min(p0 + N, first + rx.size)
    if hi > lo:
        planted = cutout * signals.tone(N, k0, N)
        rx[lo - first:hi - first] += planted[lo - p0:hi - p0]
    return cutout, rx, first


class ClockSampler:
    """nvidia-smi clock / throttle-reason samples during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (numpy restatement of xcorrRoutines.fastXcorr) on host cores
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    seed, lo, hi = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from oracle import xcorr_oracle as orc
    from pydsproutines_b200 import signals
    cutout, rx = signals.make_fastxcorr_case(seed, N_CUT, (hi - lo) + N_CUT - 1, (hi - lo) // 2, 1234)
    t0 = time.perf_counter()
    q, f = orc.fastXcorr(cutout, rx, freqsearch=True)
    return time.perf_counter() - t0, int(np.argmax(q)), int(f[np.argmax(q)])


def cpu_time_sample(n_shifts, procs):
    """Times oracle.fastXcorr(freqsearch=True) over n_shifts shifts split over `procs` processes;
    returns cells/s (wall clock of the slowest worker)."""
    per = max(1, n_shifts // procs)
    jobs = [(1002 + i, i * per, (i + 1) * per) for i in range(procs)]
    if procs == 1:
        res = [_cpu_worker(jobs[0])]
        wall = res[0][0]
    else:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(procs) as pool:
            pool.map(_cpu_worker, [(1, 0, 8)] * procs)       # import + page-in outside the timing
            t0 = time.perf_counter()
            res = pool.map(_cpu_worker, jobs)
            wall = time.perf_counter() - t0
    for _, k, b in res:
        assert k == per // 2 and b == 1234, "CPU oracle lost the planted peak"
    return per * procs * N_CUT / wall, per * procs


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    procs = max(1, min(cores, 64))
    sample = 256 * procs
    vals = []
    for i in range(args.warmup + args.steps):
        v, n = cpu_time_sample(sample, procs)
        if i >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals))
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * sample * N_CUT / value, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": "C2 fastXcorr freqsearch: 8192-sample cutout, 8192-bin FFT per shift "
                                   "(bounded sample of the 2^20 shifts)", "cutout": N_CUT, "shifts_per_step": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": "port",
                             "sample": "%d shifts per step split over %d processes (oracle/xcorr_oracle.fastXcorr, "
                                       "numpy+scipy.fft, the reference's own per-shift loop)" % (sample, procs)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from pydsproutines_b200 import _lib, dist as xdist
    from pydsproutines_b200 import xcorrRoutines as xr

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; pydsproutines_b200 has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    _lib.check(lib.xc_set_device(local_rank))

    if world == 1:
        k_rank, seed, p0, k0 = K_C2, 1002, 600001, 1234
        workload = "C2 fastXcorr freqsearch: 8192-sample cutout, 2^20 shifts, 8192-bin FFT per shift"
    else:
        k_rank, seed, k0 = K_RANK_MULTI, 1005, 4321
        p0 = int(50000017 * world / 8)
        workload = ("C5 freqsearch CAF sharded by shift range: 8192-sample cutout, %d x 2^23 shifts "
                    "(block + 8191-sample halo per rank), peak allgather" % world)
    cutout, rx, first = make_shard(seed, rank, world, k_rank, N_CUT, p0, k0)
    cells_rank = float(k_rank) * N_CUT
    cutout_d = torch.from_numpy(cutout).to(dev)
    rx_d = torch.from_numpy(rx).to(dev)
    cutout_pin = torch.from_numpy(cutout).pin_memory()
    rx_pin = torch.from_numpy(rx).pin_memory()
    cutout_h, rx_h = cutout_pin.numpy(), rx_pin.numpy()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2

    def step_device():
        qf2, fidx, peak = xr.fastXcorrPeak(cutout_d, rx_d, freqsearch=True)
        row = xdist.peak_to_row(peak.qf2, peak.fidx, first + peak.shift, 0, rank)
        return xdist.global_peak(xdist.allgather_peak(row, dev))

    def step_host():
        qf2, fidx, peak = xr.fastXcorrPeak(cutout_h, rx_h, freqsearch=True)
        row = xdist.peak_to_row(peak.qf2, peak.fidx, first + peak.shift, 0, rank)
        return xdist.global_peak(xdist.allgather_peak(row, dev)), qf2.nbytes + fidx.nbytes

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (also checks the answer) ----
    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        gp = step_device()
    assert gp["shift"] == p0 and gp["fidx"] == k0, "wrong peak: %r (expected shift %d bin %d)" % (gp, p0, k0)

    # ---- timed region: device-resident inputs ----
    lib.xc_profile_enable(1)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    barrier()
    launches0 = lib.xc_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kern_ms = []
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        flush.zero_()
        ev[i][0].record()
        step_device()
        ev[i][1].record()
        kern_ms.append(lib.xc_profile_last_ms())
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = lib.xc_launch_count() - launches0
    clocks = sampler.stop() if sampler else None
    lib.xc_profile_enable(0)
    ms_total = sum(a.elapsed_time(b) for a, b in ev)
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = cells_rank * world / (ms_step * 1e-3)

    # ---- end to end through the public API with pinned host buffers ----
    for _ in range(2):
        step_host()
    barrier()
    e2e_ms = []
    d2h = 0
    for i in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _, d2h = step_host()
        e2e_ms.append((time.perf_counter() - t0) * 1e3)
    barrier()
    e2e_step = float(np.mean(e2e_ms))
    if world > 1:
        t = torch.tensor([e2e_step], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_step = float(t.item())
    e2e_value = cells_rank * world / (e2e_step * 1e-3)

    if rank == 0:
        hbm_peak, peak_src = measured_peaks()
        kms = float(np.mean([k for k in kern_ms if k > 0])) if any(k > 0 for k in kern_ms) else ms_step
        algo_bytes = ALGO_BYTES_PER_SHIFT * k_rank
        achieved = algo_bytes / (kms * 1e-3) / 1e9
        tf = C_double()
        fp32_peak = None
        if lib.xc_measure_fp32_tflops(300.0, tf) == 0:
            fp32_peak = tf.value
        fp32_ach = FLOPS_PER_CELL * cells_rank / (kms * 1e-3) / 1e12
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload, "cutout": N_CUT, "shifts_per_rank": k_rank, "bins": N_CUT,
                           "l2": "256 MiB device memset between timed steps (outside the per-step CUDA events)",
                           "timing": "sum of per-step CUDA-event times on the launching stream, max over ranks",
                           "wall_s_timed_loop": t_wall},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                             "frac": achieved / hbm_peak, "traffic": NCU_DRAM_BYTES_C2 if world == 1 else None,
                             "traffic_source": "profiles/r1_c2_staged_ncu.md (ncu --set full, per launch)" if world == 1 else None,
                             "peak_source": peak_src,
                             "kernel": "fused_fft_staged_kernel<8192>", "kernel_ms": kms,
                             "algorithmic_bytes_per_launch": algo_bytes,
                             "note": "the path is FP32/issue bound, not HBM bound (SURVEY.md 8d); see fp32"},
                "fp32": {"achieved_tflops": fp32_ach, "peak_tflops_measured_ffma": fp32_peak,
                         "frac": (fp32_ach / fp32_peak) if fp32_peak else None,
                         "flops_per_cell": FLOPS_PER_CELL},
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_step,
                        "h2d_bytes_per_step": int(rx.nbytes + cutout.nbytes), "d2h_bytes_per_step": int(d2h + 32)},
                "gpu_launches": int(launches), "clocks": clocks}
        if world == 1 and not args.no_cpu:
            v, n = cpu_time_sample(1 << 14, 1)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": "%d of the 2^20 shifts, single thread (oracle/xcorr_oracle.fastXcorr)" % n}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def C_double():
    import ctypes
    return ctypes.c_double(0.0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if args.gpus > 1 and world == 1:
            raise SystemExit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d"
                             % (args.gpus, args.gpus))
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

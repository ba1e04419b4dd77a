#!/usr/bin/env python
"""bench.py -- CAF cells/s of the fused frequency-search fastXcorr on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c4|c6|bx|bx10]

A "step" is one pass of the hot path over one synthetic batch:
  N = 1 : BASELINE.json configs[1] (C2): 8192-sample cutout, 2^20 shifts, 8192-bin FFT per shift (the headline).
          The line also carries `configs`: device-timed C1, C3, C4 (and C6) with their own rooflines, and
          `config.strong`: the 2^26-sample C5 record on this one GPU.
  N > 1 : configs[4] (C5) sharded by shift range.  `value` is the weak line (every rank searches 2^23 shifts of a
          shard = block + 8191-sample halo of one N*2^23-sample record, so N = 8 is exactly the 2^26-sample C5 record);
          `config.strong` is the FIXED 2^26-sample record split N ways.  A half-amplitude decoy sits in rank 0's
          block, the only exchange is the allgather of the per-rank peak tuples (NCCL), every rank checks one 256-shift
          window of its results against the oracle outside the timed region.
`value` = cells (shifts x bins) of all ranks / device time with inputs resident in HBM (CUDA events, max over ranks);
`e2e` = the same through the public API with pinned HOST buffers (H2D + D2H inside the timed region).
`--workload` makes another config the headline of a single-GPU run (its own metric unit stays cells/s).
The reference arm (`--impl reference`) times the CPU restatement of the reference's numpy/scipy path (oracle/) on all
host cores on a bounded sample of the same workload.
"""
import argparse
import ctypes
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
K_RANK_WEAK = 1 << 23
LOG2_C5 = 26
ALGO_BYTES_PER_SHIFT = 16.0        # one new rx sample (8 B) + one result (f32 qf2 + u32 bin)
FLOPS_PER_CELL = 75.0              # 5*log2(8192) + 10
# dram__bytes_read.sum + dram__bytes_write.sum of one C2 launch of fused_fft_pp_kernel, from this round's
# `ncu --set full` capture of this very command (profiles/r2_c2_pp_ncu.md): 12.78 MB read, 0 written (the 8 MB of
# results stay in the 126 MB L2)
NCU_DRAM_BYTES_C2 = 12.78e6
NCU_TRAFFIC_SOURCE = "profiles/r2_c2_pp_ncu.md (ncu --set full of this command, per launch)"
FP32_NOMINAL_TFLOPS = 74.4         # 148 SMs x 128 lanes x 2 x 1.965 GHz


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        d["_source"] = "measured"
        return d
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "_source": "fallback"}


# ------------------------------------------------------------------------------------------------
# synthetic workload (seeded; block-keyed so that shards of different ranks agree on the halo)
# ------------------------------------------------------------------------------------------------
def noise_block(seed, b, n):
    rng = np.random.default_rng([seed, b])
    w = rng.standard_normal(2 * n, dtype=np.float32)
    return (np.float32(np.sqrt(0.5)) * w).view(np.complex64)


NOISE_BLOCK = 1 << 22      # samples per seeded noise block of the sharded record


def record_slice(seed, first, n, N, plants):
    """rx[first : first + n] of the global record (noise in seeded 2^22-sample blocks + the planted copies of the
    cutout: (position, amplitude, bin) each) and the QPSK cutout.  Every rank builds only its own slice."""
    from pydsproutines_b200 import signals
    cutout = signals.qpsk(np.random.default_rng([seed, 10 ** 6]), N)
    rx = np.empty(n, np.complex64)
    b0, b1 = first // NOISE_BLOCK, (first + n - 1) // NOISE_BLOCK
    for b in range(b0, b1 + 1):
        blk = noise_block(seed, b, NOISE_BLOCK)
        lo, hi = max(first, b * NOISE_BLOCK), min(first + n, (b + 1) * NOISE_BLOCK)
        rx[lo - first:hi - first] = blk[lo - b * NOISE_BLOCK:hi - b * NOISE_BLOCK]
    for pos, amp, k in plants:
        lo, hi = max(pos, first), min(pos + N, first + n)
        if hi > lo:
            planted = np.complex64(amp) * cutout * signals.tone(N, k, N)
            rx[lo - first:hi - first] += planted[lo - pos:hi - pos]
    return cutout, rx


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


CPU_NOTE = ("oracle/xcorr_oracle.fastXcorr = the reference's per-shift numpy + scipy.fft loop (xcorrRoutines.py:511-531), "
            "bit-identical to it on the golden vectors; conj(cutout) is hoisted out of the loop (the reference "
            "recomputes it per shift), which favours the CPU arm")


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
                                   "(bounded sample of the 2^20 shifts)", "cutout": N_CUT, "shifts_per_step": sample,
                       "host_cores": cores},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": "port",
                             "sample": "%d shifts per step split over %d processes of %d host cores; %s"
                                       % (sample, procs, cores, CPU_NOTE)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def event_times(fn, steps, warmup, flush=None):
    """Per-step CUDA-event times (ms) of fn() on torch's current stream (the stream the library enqueues on)."""
    import torch
    for _ in range(warmup):
        if flush is not None:
            flush.zero_()
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in ev]


def fp32_peak(lib):
    tf = ctypes.c_double(0.0)
    if lib.xc_measure_fp32_tflops(300.0, tf) == 0 and tf.value > 0:
        return tf.value, "measured (register-resident FFMA loop, xc_measure_fp32_tflops)"
    return FP32_NOMINAL_TFLOPS, "nominal 148 x 128 x 2 x 1.965 GHz"


def fp32_roofline(flops, ms, peak, src, kernel, hbm_bytes, peaks, traffic=None, traffic_source=None):
    ach = flops / (ms * 1e-3) / 1e12
    hbm = hbm_bytes / (ms * 1e-3) / 1e9
    return {"bound": "fp32", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "peak_source": src,
            "frac_of_nominal_74.4": ach / FP32_NOMINAL_TFLOPS, "kernel": kernel, "kernel_ms": ms,
            "algorithmic_flops_per_launch": flops, "traffic": traffic, "traffic_source": traffic_source,
            "hbm": {"achieved": hbm, "peak": float(peaks["hbm_gbs"]), "unit": "GB/s", "frac": hbm / float(peaks["hbm_gbs"]),
                    "algorithmic_bytes_per_launch": hbm_bytes, "peak_source": peaks["_source"],
                    "note": "SURVEY.md 8d: the path is FP32 bound, the HBM fraction is ~1e-4 by construction"}}


def config_line(which, lib, peaks, p32, p32src, steps):
    """Device-timed line of one non-headline config (inputs resident in HBM, answer checked)."""
    import torch
    from pydsproutines_b200 import signals
    from pydsproutines_b200 import xcorrRoutines as xr
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    lib.xc_profile_enable(0)
    if which == "c1":
        c = signals.config_c1()
        cu, rx = dev(c["cutout"]), dev(c["rx"])
        K = c["rx"].size - 4096 + 1
        sh = xr._Shifts(None, rx.shape[0], 4096, rx)
        q = xr._fastxcorr_call(cu, rx, sh, False)[0]
        assert int(torch.argmax(q)) == c["p0"], "C1 lost the planted peak"
        ms = float(np.median(event_times(lambda: xr._fastxcorr_call(cu, rx, sh, False), steps, 2, flush)))
        flops = K * (8.0 * 4096 + 10)                          # the sliding dot product
        # executed: fast correlation through 8192-point blocks (ovs.cuh): two transforms + the spectrum product per block of 4097 shifts
        nblk = -(-K // (8192 - 4096 + 1))
        executed = nblk * (2 * 8192 * (5.0 * 13 + 10) + 8.0 * 8192)
        r = fp32_roofline(executed, ms, p32, p32src, "ovs_kernel<8192> (overlap-save fast correlation, one wave of 128 CTAs)", 16.0 * K, peaks)
        r["algorithmic_flops_sliding_dot_product"] = flops
        r["frac_in_sliding_dot_product_flops"] = flops / (ms * 1e-3) / 1e12 / p32
        r["note"] = ("executed flops of the fast correlation (110 x fewer than the 8 N + 10 per shift of SURVEY 8d's accounting, which "
                     "this time would put at several times the FP32 peak); the call is one wave of 128 CTAs plus two small kernels: latency bound")
        return {"workload": "C1 fastXcorr no freqsearch: 4096-sample cutout, 2^20-sample rx, step 1", "ms": ms,
                "cells_per_s": K / (ms * 1e-3), "peak_checked": [c["p0"]], "roofline": r}
    if which == "c3":
        c = signals.config_c3()
        obj = xr.GroupXcorrFFT(c["ygroups"], c["starts"], 1, fftlen=c["F"])
        rx = dev(c["rx"])
        shl = xr._group_shifts(np.arange(c["K"]), rx)
        q, f = obj._handle.run(rx, shl, True, False)[:2]
        p = int(torch.argmax(q))
        assert p == c["p0"] and int(f[p]) == c["k0"], "C3 lost the planted peak"
        ms = float(np.median(event_times(lambda: obj._handle.run(rx, shl, True, False), steps, 2, flush)))
        flops = c["K"] * c["F"] * 80.0
        return {"workload": "C3 GroupXcorrFFT: 64 groups x 128, fftlen 4096, 2^22 shifts", "ms": ms,
                "cells_per_s": c["K"] * c["F"] / (ms * 1e-3), "peak_checked": [c["p0"], c["k0"]],
                "roofline": fp32_roofline(flops, ms, p32, p32src, "group_pp_kernel<4096> (staged group windows, contributor lists in tensor memory, pruned transform)", 16.0 * c["K"], peaks)}
    if which in ("c3r", "g128"):
        # C3's groups with a spacing that is not a multiple of the transform's 128-cell rows (starts 1000 g: the rotated-row
        # kernel), and the reference's default fftlen = group length (dense-fold kernel); 2^20 shifts each
        if which == "c3r":
            c = signals.config_c3(log2_shifts=20, spacing=1000)
            kern, name = "group_pp_kernel<4096, ROT> (rotated-row fold)", "GroupXcorrFFT: 64 groups x 128 at starts 1000 g, fftlen 4096, 2^20 shifts"
        else:
            c = signals.config_c3(log2_shifts=20, F=128)
            kern, name = "group_dense_kernel<128> (register-tiled fold, sub-warp FFT)", "GroupXcorrFFT: 64 groups x 128, fftlen 128 (default fftlen = group length), 2^20 shifts"
        obj = xr.GroupXcorrFFT(c["ygroups"], c["starts"], 1, fftlen=c["F"])
        rx = dev(c["rx"])
        shl = xr._group_shifts(np.arange(c["K"]), rx)
        q, f = obj._handle.run(rx, shl, True, False)[:2]
        p = int(torch.argmax(q))
        assert p == c["p0"] and int(f[p]) == c["k0"], "%s lost the planted peak" % which
        ms = float(np.median(event_times(lambda: obj._handle.run(rx, shl, True, False), steps, 2, flush)))
        flops = c["K"] * (c["F"] * (5.0 * np.log2(c["F"]) + 10.0) + 8.0 * 64 * 128)      # transform + fold
        return {"workload": name, "ms": ms, "cells_per_s": c["K"] * c["F"] / (ms * 1e-3), "peak_checked": [c["p0"], c["k0"]],
                "roofline": fp32_roofline(flops, ms, p32, p32src, kern, 16.0 * c["K"], peaks)}
    if which == "b1000":
        # a cutout length that is not a power of two (the default of the reference's benchmark_xcorrs.py): Bluestein, M = 2048
        N, K = 1000, 1 << 20
        cutout, rxa = signals.make_fastxcorr_case(1007, N, N + K - 1, 600001, 123)
        cu, rx = dev(cutout), dev(rxa)
        q, f, peak = xr.fastXcorrPeak(cu, rx, freqsearch=True)
        assert (peak.shift, peak.fidx) == (600001, 123), "lost the planted peak at N = 1000"
        lib.xc_profile_enable(1)
        kms = []
        for i in range(2 + steps):
            flush.zero_()
            xr.fastXcorrPeak(cu, rx, freqsearch=True)
            kms.append(lib.xc_profile_last_ms())
        lib.xc_profile_enable(0)
        ms = float(np.median(kms[2:]))
        cells = float(K) * N
        flops = cells * (5.0 * np.log2(N) + 10.0)
        r = fp32_roofline(flops, ms, p32, p32src, "fused_blue_pp_kernel<2048> (Bluestein, both transforms in the ping-pong layout)", 16.0 * K, peaks)
        r["executed_flops_per_launch"] = float(K) * 2 * 2048 * (5.0 * 11 + 10.0)
        r["note"] = "flops counted for the true length; Bluestein executes two 2048-point transforms per shift (4.5 x)"
        return {"workload": "fastXcorr freqsearch: 1000-sample cutout (not a power of two), 2^20 shifts", "ms": ms,
                "cells_per_s": cells / (ms * 1e-3), "peak_checked": [600001, 123], "roofline": r}
    if which in ("c4", "c6"):
        if which == "c4":
            c = signals.config_c4()
            h = xr._MtHandle(c["templates"].conj(), False)
            x = dev(c["x"])
            best, tidx = h.run(x, 0, c["K"], True, False, False)[:2]
            for pos, t, _ in c["plants"]:
                assert int(tidx[pos]) == t, "C4 lost a planted template"
            ms = float(np.median(event_times(lambda: h.run(x, 0, c["K"], True, False, False), steps, 1, flush)))
            cells, flops, passes = c["K"] * 256, c["K"] * 256 * 8.0 * 1024, 2
            name = "C4 multi-template: 256 templates x 1024 vs 2^24 samples"
            checked = [list(p[:2]) for p in c["plants"]]
            tcores = bool(h.uses_tensor_cores)
        else:
            c = signals.config_c3(log2_shifts=20)
            NFQ = 256
            freqs = (np.arange(NFQ) - NFQ // 2) * (1.0 / 65536.0)
            y = np.zeros(int(c["starts"][-1]) + 128, np.complex64)
            for g in range(64):
                y[int(c["starts"][g]):int(c["starts"][g]) + 128] = c["ygroups"][g]
            h = xr.GroupXcorr(y, c["starts"], np.full(64, 128), freqs, 1.0)._handle
            x = dev(c["rx"])
            ms = float(np.median(event_times(lambda: h.run(x, 0, c["K"], True, False), steps, 1, flush)))
            cells, flops, passes = c["K"] * NFQ, c["K"] * NFQ * 8.0 * 64 * 128, 3
            name = "C6 GroupXcorr (frequency list, SURVEY 8f): 64 groups x 128, 256 frequencies, 2^20 shifts"
            checked, tcores = None, True
        if which == "c4":
            # the shipped path for (best, template) at this template length is the fast correlation (mt_ovs.cuh, FP32); the
            # tensor-core contraction it replaces is timed beside it (XC_MT_DIRECT)
            hd = xr._MtHandle(c["templates"].conj(), False, direct=True)
            ms_tc = float(np.median(event_times(lambda: hd.run(x, 0, c["K"], True, False, False), steps, 1, flush)))
            S = 8192 - 1024 + 1
            nblk = -(-c["K"] // S)
            executed = nblk * ((1 + 256) * 8192 * (5.0 * 13 + 10) + 256 * 8.0 * 8192)
            r = fp32_roofline(executed, ms, p32, p32src, "mt_ovs2_kernel<8192> (overlap-save: one transform of the x block for all templates; X and the running maxima in tensor memory)", 16.0 * c["K"], peaks)
            r["note"] = ("executed flops of the fast correlation; SURVEY 8d counts the dense contraction (8 L flops per template and "
                         "shift = %.3g flop), which the tensor-core kernel runs in %.1f ms" % (flops, ms_tc))
            peak = float(peaks.get("bf16_tflops", 1590.0))
            return {"workload": name, "ms": ms, "cells_per_s": cells / (ms * 1e-3), "peak_checked": checked, "tensor_cores": False,
                    "roofline": r,
                    "tensor_core_contraction": {"ms": ms_tc, "bound": "tensor", "achieved": passes * flops / (ms_tc * 1e-3) / 1e12, "peak": peak,
                                                "frac": passes * flops / (ms_tc * 1e-3) / 1e12 / peak, "precision_passes": passes,
                                                "kernel": "mt_tc_kernel (XC_MT_DIRECT)"}}
        peak = float(peaks.get("bf16_tflops", 1590.0))
        ach = passes * flops / (ms * 1e-3) / 1e12
        return {"workload": name, "ms": ms, "cells_per_s": cells / (ms * 1e-3), "peak_checked": checked,
                "tensor_cores": tcores,
                "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                             "peak_source": peaks["_source"] + " (cuBLAS bf16 burst)", "precision_passes": passes,
                             "algorithmic_frac": flops / (ms * 1e-3) / 1e12 / peak, "kernel": "mt_tc_kernel",
                             "kernel_ms": ms, "traffic": None}}
    if which in ("n256", "n512", "n1024", "n2048", "n4096", "n16384"):
        # the frequency search at the other power-of-two lengths (2^20 shifts each, planted peak checked)
        N = int(which[1:])
        c = signals.config_c2(log2_shifts=20, N=N)
        cu, rx = dev(c["cutout"]), dev(c["rx"])
        q, f, peak = xr.fastXcorrPeak(cu, rx, freqsearch=True)
        assert (peak.shift, peak.fidx) == (c["p0"], c["k0"]), "lost the planted peak at N = %d" % N
        lib.xc_profile_enable(1)
        kms = []
        for i in range(2 + steps):
            flush.zero_()
            xr.fastXcorrPeak(cu, rx, freqsearch=True)
            kms.append(lib.xc_profile_last_ms())
        lib.xc_profile_enable(0)
        ms = float(np.median(kms[2:]))
        cells = float(c["K"]) * N
        flops = cells * (5.0 * np.log2(N) + 10.0)
        layout = ("fused_fft_kernel (one group of 512 threads, conj(cutout) in tensor memory)" if N == 16384 else
                  "fused_fft_subwarp_kernel (%d lanes per FFT, %d FFTs per warp, staged window, constants in tensor memory)" % (N // 32, 1024 // N) if N < 1024 else
                  "fused_fft_warp_kernel (one warp per FFT, staged window, constants in tensor memory)" if N == 1024 else
                  "fused_fft_pp_kernel (ping-pong layout)" if N in (2048, 4096) else "fused_fft_kernel (B groups per CTA, round 1)")
        return {"workload": "fastXcorr freqsearch: %d-sample cutout, 2^20 shifts, %d-bin FFT per shift" % (N, N), "ms": ms,
                "cells_per_s": cells / (ms * 1e-3), "peak_checked": [c["p0"], c["k0"]],
                "roofline": fp32_roofline(flops, ms, p32, p32src, layout, 16.0 * c["K"], peaks)}
    if which in ("bx", "bx10"):
        # the reference's own published benchmark: `python benchmarks/benchmark_xcorrs.py 1000000 128` (README.md:3-23):
        # cutout = x[10000 : 10000 + N], shifts 0..127, cp_fastXcorr on device arrays, results copied to the host
        from oracle import xcorr_oracle as orc
        N, K = (1000000, 128) if which == "bx" else (10000000, 128)
        rng = np.random.default_rng(7)
        x = (rng.standard_normal(10000 + N + K, dtype=np.float32) + 1j * rng.standard_normal(10000 + N + K, dtype=np.float32)).astype(np.complex64)
        cutout = x[10000:10000 + N].copy()
        shifts = np.arange(K)
        d_c, d_x = dev(cutout), dev(x)
        q, f = xr.cp_fastXcorr(d_c, d_x, freqsearch=True, shifts=shifts, BATCH=128)
        for i in (0, K - 1):                                   # two shifts against the oracle (numpy, float32 FFT)
            qo, fo = orc.fastXcorr(cutout, x[i:i + N], freqsearch=True)
            assert abs(q[i] - qo[0]) <= 1e-4 * qo[0] + 1e-9, "benchmark_xcorrs workload: QF2 differs from the oracle"
            if f[i] != fo[0]:
                X = np.abs(np.fft.fft(x[i:i + N].astype(np.complex128) * np.conj(cutout).astype(np.complex128))) ** 2
                assert abs(X[f[i]] - X[fo[0]]) <= 1e-5 * X[fo[0]], "benchmark_xcorrs workload: bin differs from the oracle"
        lib.xc_profile_enable(0)
        wall = []
        for i in range(2 + steps):
            flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            xr.cp_fastXcorr(d_c, d_x, freqsearch=True, shifts=shifts, BATCH=128)
            wall.append((time.perf_counter() - t0) * 1e3)
        ms = float(np.median(wall[2:]))
        log2l = int(np.ceil(np.log2(2 * N - 1)))
        L = 1 << log2l
        # window + conj(cutout) chirp read, column pass writes 8 L, the fused middle step reads 8 L + FFT_L(chirp) and writes 8 L,
        # the last row pass reads 8 L (round 2, first half: six passes over the work array)
        form_bytes = K * (2 * 8.0 * N + 5 * 8.0 * L)
        pub = {"bx": {"numpy_s": 15.9, "cython_ipp_4thr_s": 3.9, "cupy_s": 0.29}, "bx10": {"numpy_s": 192.0, "cython_ipp_4thr_s": 62.0, "cupy_s": 3.0}}[which]
        ach = form_bytes / (ms * 1e-3) / 1e9
        return {"workload": "benchmark_xcorrs.py %d 128 (reference's published benchmark): %d-sample cutout, 128 shifts, cp_fastXcorr "
                            "on device arrays, results on the host" % (N, N), "ms": ms, "cells_per_s": K * float(N) / (ms * 1e-3),
                "peak_checked": None, "published_reference": dict(pub, source="benchmarks/README.md:3-23, unnamed hardware"),
                "speedup_vs_published_cupy": pub["cupy_s"] * 1e3 / ms, "vs_baseline": K * float(N) / (ms * 1e-3) / (K * float(N) / pub["cupy_s"]),
                "roofline": {"bound": "hbm", "achieved": ach, "peak": float(peaks["hbm_gbs"]), "unit": "GB/s", "frac": ach / float(peaks["hbm_gbs"]),
                             "peak_source": peaks["_source"], "kernel": "big_cols_kernel / big_rows_kernel (four-step Bluestein, L = 2^%d)" % log2l,
                             "kernel_ms": ms, "formulation_bytes_per_call": form_bytes, "traffic": None,
                             "note": "bytes of the four-step formulation (window and chirped cutout read + 5 passes over 8 L bytes per shift, the Bluestein middle step fused with the second transform's column pass); the "
                                     "work arrays of a batch (about 1 GB) do not stay in the 126 MB L2 -- an L2-sized batch measured "
                                     "slower (12.5 vs 9.9 ms at N = 1e6), see bigfft.cu"}}
    raise SystemExit("unknown workload %r" % which)


def sharded_search(world, rank, dev, k_rank, seed, plants, expect, steps, warmup, lib, flush, host_pass):
    """One sharded run: every rank searches k_rank shifts of its slice (block + halo) of the record; peak allgather.
    Returns dict(ms (max over ranks), e2e_ms, launches, kern_ms, h2d, d2h, oracle window result)."""
    import torch
    import torch.distributed as dist
    from pydsproutines_b200 import dist as xdist
    from pydsproutines_b200 import xcorrRoutines as xr
    first = rank * k_rank
    cutout, rx = record_slice(seed, first, k_rank + N_CUT - 1, N_CUT, plants)
    cutout_d, rx_d = torch.from_numpy(cutout).to(dev), torch.from_numpy(rx).to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        qf2, fidx, peak = xr.fastXcorrPeak(cutout_d, rx_d, freqsearch=True)
        row = xdist.peak_to_row(peak.qf2, peak.fidx, first + peak.shift, 0, rank)
        return qf2, fidx, peak, xdist.global_peak(xdist.allgather_peak(row, dev))

    for _ in range(max(warmup, 1)):
        flush.zero_()
        qf2, fidx, lpeak, gp = step_device()
    assert gp["shift"] == expect[0] and gp["fidx"] == expect[1], "wrong global peak: %r (expected %r)" % (gp, expect)
    out = {"global_peak": [gp["shift"], gp["fidx"], gp["rank"]]}
    # a decoy planted in this rank's block must be this rank's LOCAL peak unless the true peak is here too
    found = -1
    for pos, amp, k in plants[1:]:
        if first <= pos < first + k_rank and not (first <= expect[0] < first + k_rank):
            assert first + lpeak.shift == pos and lpeak.fidx == k, "rank %d did not find its decoy" % rank
            found = rank
    # one 256-shift window of this rank's results against the oracle (outside the timed region)
    from oracle import xcorr_oracle as orc
    w0 = (k_rank // 3) & ~255
    pk = expect[0] - first
    if 0 <= pk < k_rank:
        w0 = max(0, min(k_rank - 256, pk - 128))
    qo, fo = orc.fastXcorr(cutout, rx[w0:w0 + 256 + N_CUT - 1], freqsearch=True)
    qg, fg = qf2[w0:w0 + 256].cpu().numpy(), fidx[w0:w0 + 256].cpu().numpy().astype(np.uint32)
    assert np.allclose(qg, qo, rtol=1e-4, atol=1e-7), "rank %d: QF2 window differs from the oracle" % rank
    bad = np.nonzero(fg != fo)[0]
    for i in bad:       # a differing bin must be a float32 near-tie of the float64 spectrum
        X = np.fft.fft(rx[w0 + i:w0 + i + N_CUT].astype(np.complex128) * np.conj(cutout).astype(np.complex128))
        a, b = abs(X[fg[i]]) ** 2, abs(X[fo[i]]) ** 2
        assert abs(a - b) <= 1e-5 * max(a, b), "rank %d: bin differs from the oracle at shift %d" % (rank, first + w0 + i)
    out["oracle_window"] = {"rank": rank, "first_shift": first + w0, "shifts": 256, "bins_differing_near_ties": int(bad.size)}
    if world > 1:       # which rank saw the decoy as its local peak (max over ranks; -1 = none)
        t = torch.tensor([found], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        found = int(t.item())
    out["decoy_found_by_rank"] = found if found >= 0 else None

    lib.xc_profile_enable(1)
    barrier()
    launches0 = lib.xc_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    kern_ms = []
    for i in range(steps):
        flush.zero_()
        ev[i][0].record()
        step_device()
        ev[i][1].record()
        kern_ms.append(lib.xc_profile_last_ms())
    barrier()
    out["launches"] = int(lib.xc_launch_count() - launches0)
    lib.xc_profile_enable(0)
    ms_total = sum(a.elapsed_time(b) for a, b in ev)
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    out["ms"] = ms_total / steps
    out["kern_ms"] = float(np.mean([k for k in kern_ms if k > 0])) if any(k > 0 for k in kern_ms) else out["ms"]
    if host_pass:
        cutout_h = torch.from_numpy(cutout).pin_memory().numpy()
        rx_h = torch.from_numpy(rx).pin_memory().numpy()

        def step_host():
            q, f, peak = xr.fastXcorrPeak(cutout_h, rx_h, freqsearch=True)
            row = xdist.peak_to_row(peak.qf2, peak.fidx, first + peak.shift, 0, rank)
            return xdist.global_peak(xdist.allgather_peak(row, dev)), q.nbytes + f.nbytes
        for _ in range(2):
            step_host()
        barrier()
        e2e_ms, d2h = [], 0
        for i in range(steps):
            flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            gph, d2h = step_host()
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
        barrier()
        assert gph["shift"] == expect[0] and gph["fidx"] == expect[1]
        e2e = float(np.mean(e2e_ms))
        if world > 1:
            t = torch.tensor([e2e], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e = float(t.item())
        out.update(e2e_ms=e2e, h2d=int(rx.nbytes + cutout.nbytes), d2h=int(d2h + 32))
    del rx_d, cutout_d
    torch.cuda.empty_cache()
    return out


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from pydsproutines_b200 import _lib

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; pydsproutines_b200 has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    _lib.check(lib.xc_set_device(local_rank))
    peaks = measured_peaks()
    warmup = max(args.warmup, 3)

    if world == 1 and args.workload != "c2":
        p32, p32src = fp32_peak(lib)
        cl = config_line(args.workload, lib, peaks, p32, p32src, args.steps)
        line = {"metric": METRIC, "value": cl["cells_per_s"], "unit": UNIT, "n_gpus": 1, "steps": args.steps,
                "warmup": 2, "ms_per_step": cl["ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": cl.get("vs_baseline"),
                "dtype": "bf16 hi/lo split, f32 accumulate" if cl["roofline"]["bound"] == "tensor" else "f32",
                "data": "synthetic", "config": {"workload": cl["workload"], "l2": "256 MiB device memset between timed steps",
                                                "peak_checked": cl["peak_checked"]},
                "roofline": cl["roofline"], "gpu_launches": None}
        print(json.dumps(line), flush=True)
        return

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if world == 1:
        k_rank, seed, plants, expect = K_C2, 1002, [(600001, 1.0, 1234)], (600001, 1234)
        workload = "C2 fastXcorr freqsearch: 8192-sample cutout, 2^20 shifts, 8192-bin FFT per shift"
    else:
        k_rank, seed = K_RANK_WEAK, 1005
        p0 = int(50000017 * world / 8)
        plants, expect = [(p0, 1.0, 4321), (1000003, 0.5, (4321 * 3) % N_CUT)], (p0, 4321)
        workload = ("C5 freqsearch CAF sharded by shift range (weak): 8192-sample cutout, %d x 2^23 shifts "
                    "(block + 8191-sample halo per rank), half-amplitude decoy in rank 0, peak allgather" % world)
    main = sharded_search(world, rank, dev, k_rank, seed, plants, expect, args.steps, warmup, lib, flush, True)
    clocks = sampler.stop() if sampler else None
    cells_rank = float(k_rank) * N_CUT
    value = cells_rank * world / (main["ms"] * 1e-3)

    # ---- strong scaling: the fixed 2^26-sample C5 record split `world` ways ----
    strong = None
    if not args.no_strong:
        k_strong = (1 << LOG2_C5) // world
        p0s = 50000017
        strong = sharded_search(world, rank, dev, k_strong, 1005, [(p0s, 1.0, 4321), (1000003, 0.5, (4321 * 3) % N_CUT)],
                                (p0s, 4321), max(1, min(args.steps, 3)), 1, lib, flush, False)

    if rank == 0:
        p32, p32src = fp32_peak(lib)
        roof = fp32_roofline(FLOPS_PER_CELL * cells_rank, main["kern_ms"], p32, p32src,
                             "fused_fft_pp_kernel<8192> (two thread groups per SM, constants in tensor memory)",
                             ALGO_BYTES_PER_SHIFT * k_rank, peaks,
                             NCU_DRAM_BYTES_C2 if world == 1 else None, NCU_TRAFFIC_SOURCE if world == 1 else None)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": warmup, "ms_per_step": main["ms"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload, "cutout": N_CUT, "shifts_per_rank": k_rank, "bins": N_CUT,
                           "l2": "256 MiB device memset between timed steps (outside the per-step CUDA events)",
                           "timing": "sum of per-step CUDA-event times on the launching stream, max over ranks",
                           "global_peak": main["global_peak"], "oracle_window": main["oracle_window"],
                           "decoy_found_by_rank": main.get("decoy_found_by_rank")},
                "roofline": roof,
                "e2e": {"value": cells_rank * world / (main["e2e_ms"] * 1e-3), "unit": UNIT, "ms_per_step": main["e2e_ms"],
                        "h2d_bytes_per_step": main["h2d"], "d2h_bytes_per_step": main["d2h"],
                        "path": "xcorrRoutines.fastXcorrPeak(numpy, numpy): pinned host buffers, chunked H2D | search | D2H pipeline"},
                "gpu_launches": main["launches"], "clocks": clocks}
        if strong is not None:
            line["config"]["strong"] = {
                "workload": "C5: fixed 2^26-sample record, %d shifts per rank" % k_strong, "ms_per_step": strong["ms"],
                "value": float(k_strong) * world * N_CUT / (strong["ms"] * 1e-3), "unit": UNIT,
                "global_peak": strong["global_peak"], "decoy_found_by_rank": strong.get("decoy_found_by_rank"),
                "oracle_window": strong["oracle_window"]}
        if world == 1 and not args.no_cpu:
            v, n = cpu_time_sample(1 << 14, 1)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": "%d of the 2^20 shifts, single thread of %d host cores; %s"
                                              % (n, os.cpu_count() or 1, CPU_NOTE)}
        if world == 1 and not args.no_configs:
            line["configs"] = []
            for w in ("c1", "c3", "c3r", "g128", "c4", "c6", "bx", "b1000", "n256", "n512", "n1024", "n2048", "n4096", "n16384"):
                try:
                    line["configs"].append(config_line(w, lib, peaks, p32, p32src, 3))
                except Exception as e:                      # a config failure must not hide the headline
                    line["configs"].append({"workload": w, "error": repr(e)})
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c1", "c3", "c3r", "g128", "c4", "c6", "bx", "bx10", "b1000", "n256", "n512", "n1024", "n2048", "n4096", "n16384"],
                    help="single-GPU headline config (default C2, BASELINE.json's metric config)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C3 / C4 / C6 lines of the N = 1 run")
    ap.add_argument("--no-strong", action="store_true", help="skip the fixed-size C5 (strong scaling) leg")
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

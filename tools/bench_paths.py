#!/usr/bin/env python
"""Device-timed survey of the secondary paths of the frequency search (everything bench.py's headline lines do not cover):
CAF rows, complex outputs, stepped shift lists, Bluestein lengths, chirp-z grids, short lengths, GroupXcorrFFT at other
fftlen.  One JSON line per path with the fraction of the roof that bounds it (FP32 for the transforms, HBM for the CAF
writes).   python tools/bench_paths.py [names...] [--steps 5]"""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, steps, warmup=2):
    import torch
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))


def main():
    import torch
    from pydsproutines_b200 import signals, _lib
    from pydsproutines_b200 import xcorrRoutines as xr
    names = [a for a in sys.argv[1:] if not a.startswith("-") and not a.isdigit()]
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 5
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    lib = _lib.load()
    tf = ctypes.c_double(0)
    lib.xc_measure_fp32_tflops(200.0, tf)
    p32 = tf.value
    hbm = 6458.4
    try:
        hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", hbm))
    except Exception:
        pass

    def line(name, ms, cells, flops=None, bytes_=None, **kw):
        d = {"path": name, "ms": round(ms, 4), "cells_per_s": cells / (ms * 1e-3)}
        if flops is not None:
            d["fp32_frac"] = round(flops / (ms * 1e-3) / 1e12 / p32, 4)
        if bytes_ is not None:
            d["hbm_frac"] = round(bytes_ / (ms * 1e-3) / 1e9 / hbm, 4)
        d.update(kw)
        print(json.dumps(d), flush=True)

    def want(n):
        return not names or n in names

    def fft_flops(N):
        return 5.0 * np.log2(N) + 10.0

    # ---- power-of-two searches with the outputs that leave the fast layouts ----
    for N in (8192, 4096, 1024):
        K = 1 << 18
        c = signals.config_c2(log2_shifts=18, N=N)
        cu, rx = dev(c["cutout"]), dev(c["rx"])
        if want("plain%d" % N):
            ms = timed(lambda: xr._fastxcorr_call(cu, rx, xr._Shifts(None, rx.shape[0], N, rx), True, want_fidx=True), steps)
            line("freqsearch N=%d contiguous (max, arg max), 2^18 shifts" % N, ms, K * N, K * N * fft_flops(N))
        if want("cplx%d" % N):
            ms = timed(lambda: xr._fastxcorr_call(cu, rx, xr._Shifts(None, rx.shape[0], N, rx), True, want_fidx=True, want_cplx=True), steps)
            line("freqsearch N=%d complex QF (absResult=False), 2^18 shifts" % N, ms, K * N, K * N * fft_flops(N))
        if want("step%d" % N):
            sh = np.arange(0, K, 3)
            dsh = xr._Shifts(sh, rx.shape[0], N, rx)
            ms = timed(lambda: xr._fastxcorr_call(cu, rx, dsh, True, want_fidx=True), steps)
            line("freqsearch N=%d shifts 0:K:3 (stepped list)" % N, ms, sh.size * N, sh.size * N * fft_flops(N))
        if want("caf%d" % N):
            Kc = (1 << 28) // (4 * N)            # 256 MiB of CAF rows
            dsh = xr._Shifts(np.arange(Kc), rx.shape[0], N, rx)
            ms = timed(lambda: xr._caf_call(cu, rx, dsh, False), steps)
            line("CAF rows N=%d (float32 [K, N], %d shifts)" % (N, Kc), ms, Kc * N, Kc * N * fft_flops(N), bytes_=4.0 * Kc * N)
            ms = timed(lambda: xr._caf_call(cu, rx, dsh, True), steps)
            line("CAF rows N=%d complex64 (%d shifts)" % (N, Kc), ms, Kc * N, Kc * N * fft_flops(N), bytes_=8.0 * Kc * N)
    # ---- shorter powers of two ----
    for N in (512, 256, 128, 64):
        if want("plain%d" % N):
            K = 1 << 20
            c = signals.config_c2(log2_shifts=20, N=N)
            cu, rx = dev(c["cutout"]), dev(c["rx"])
            ms = timed(lambda: xr._fastxcorr_call(cu, rx, xr._Shifts(None, rx.shape[0], N, rx), True, want_fidx=True), steps)
            line("freqsearch N=%d contiguous, 2^20 shifts" % N, ms, K * N, K * N * fft_flops(N))
    # ---- lengths that are not a power of two (Bluestein inside one SM); flops counted for the true length ----
    for N in (1000, 2000, 3000, 5000, 8191):
        if want("blue%d" % N):
            K = 1 << 17
            cutout, rxa = signals.make_fastxcorr_case(5, N, N + K - 1, 4321, 77)
            cu, rx = dev(cutout), dev(rxa)
            ms = timed(lambda: xr._fastxcorr_call(cu, rx, xr._Shifts(None, rx.shape[0], N, rx), True, want_fidx=True), steps)
            line("freqsearch N=%d (Bluestein), 2^17 shifts" % N, ms, K * N, K * N * fft_flops(N))
    # ---- chirp-z zoom ----
    if want("czt"):
        N, K = 4096, 1 << 17
        cutout, rxa = signals.make_fastxcorr_case(6, N, N + K - 1, 4321, 0)
        cu, rx = dev(cutout), dev(rxa)
        t0 = timed(lambda: xr.cztXcorr(cu, rx, -1000.0, 1000.0, 1e6, cztStep=2.0), steps)
        line("cztXcorr N=4096, 1001 bins, 2^17 shifts", t0, K * 1001)
    # ---- GroupXcorrFFT at other transform lengths ----
    for F, G, L, sp in ((2048, 16, 64, 0), (8192, 64, 128, 0), (4096, 8, 512, 0), (4096, 64, 128, 1000), (4096, 64, 128, -1),
                        (512, 16, 128, 1000), (128, 64, 128, 1024), (128, 64, 128, -1), (256, 32, 256, -1), (64, 32, 64, 1000),
                        (1024, 16, 128, 2000), (4096, 64, 128, 1024)):
        if want("group%d" % F):
            K = 1 << 18
            rng = np.random.default_rng(3)
            y = (rng.standard_normal((G, L)) + 1j * rng.standard_normal((G, L))).astype(np.complex64)
            if sp == 0:
                starts = np.arange(G) * (F // G) if L * G <= F else np.arange(G) * L
            elif sp > 0:
                starts = np.arange(G) * sp
            else:
                starts = np.cumsum(rng.integers(L, 2000, G)) - L     # ragged spacing, disjoint groups
                starts -= starts[0]
            rxa = (rng.standard_normal(K + int(starts[-1]) + L) + 1j * rng.standard_normal(K + int(starts[-1]) + L)).astype(np.complex64)
            obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F)
            rx = dev(rxa)
            sh = np.arange(K)
            ms = timed(lambda: obj._handle.run(rx, xr._group_shifts(sh, rx), True, False), steps)
            line("GroupXcorrFFT %d groups x %d, fftlen %d, starts %s, 2^18 shifts" % (G, L, F, "ragged" if sp < 0 else "%d g" % int(starts[1])), ms, K * F, K * F * (fft_flops(F) + 8.0 * G * L / F))


if __name__ == "__main__":
    main()

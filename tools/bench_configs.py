#!/usr/bin/env python
"""Device-timed throughput of the non-headline BASELINE.json configs (C1, C3, C4) -- one JSON line each.
bench.py stays the contract for the headline C2 / C5 numbers; this script exists so that every §8 row has a
measured figure under profiles/.   python tools/bench_configs.py [c1] [c3] [c4] [--steps 5]"""
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
    from pydsproutines_b200 import cupyExtensions as ce
    which = [a for a in sys.argv[1:] if not a.startswith("-")] or ["c1", "c3", "c4"]
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 5
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    lib = _lib.load()
    tf = __import__("ctypes").c_double(0)
    lib.xc_measure_fp32_tflops(200.0, tf)
    for w in which:
        if w == "c1":
            c = signals.config_c1()
            cu, rx = dev(c["cutout"]), dev(c["rx"])
            K = c["rx"].size - 4096 + 1
            ms = timed(lambda: xr._fastxcorr_call(cu, rx, xr._Shifts(None, rx.shape[0], 4096, rx), False), steps)
            cells, flops = K, K * (8 * 4096 + 10)
            name = "C1 fastXcorr no freqsearch: 4096-sample cutout, 2^20-sample rx, step 1"
        elif w == "c3":
            c = signals.config_c3()
            obj = xr.GroupXcorrFFT(c["ygroups"], c["starts"], 1, fftlen=c["F"])
            rx = dev(c["rx"])
            sh = np.arange(c["K"])
            ms = timed(lambda: obj._handle.run(rx, xr._group_shifts(sh, rx), True, False), steps)
            cells, flops = c["K"] * c["F"], c["K"] * c["F"] * 80.0
            name = "C3 GroupXcorrFFT: 64 groups x 128, fftlen 4096, 2^22 shifts"
        elif w == "c4":
            c = signals.config_c4()
            h = xr._MtHandle(c["templates"].conj(), False)
            x = dev(c["x"])
            ms = timed(lambda: h.run(x, 0, c["K"], True, False, False), steps, warmup=1)
            cells, flops = c["K"] * 256, c["K"] * 256 * 8.0 * 1024
            name = "C4 multi-template: 256 templates x 1024 vs 2^24 samples"
        elif w == "c6":
            # 8(f) rows: GroupXcorr over a frequency list, C3's groups (64 x 128 at 1024 g), 256 frequencies, 2^20 shifts
            c = signals.config_c3(log2_shifts=20)
            NFQ = 256
            freqs = (np.arange(NFQ) - NFQ // 2) * (1.0 / 65536.0)
            y = np.zeros(int(c["starts"][-1]) + 128, np.complex64)
            for g in range(64):
                y[int(c["starts"][g]):int(c["starts"][g]) + 128] = c["ygroups"][g]
            h = xr.GroupXcorr(y, c["starts"], np.full(64, 128), freqs, 1.0)._handle
            rx = dev(c["rx"])
            ms = timed(lambda: h.run(rx, 0, c["K"], True, False), steps, warmup=1)
            cells, flops = c["K"] * NFQ, c["K"] * NFQ * 8.0 * 64 * 128
            name = "C6 GroupXcorr (frequency list): 64 groups x 128, 256 frequencies, 2^20 shifts"
        else:
            continue
        extra = {}
        if w == "c6":
            peaks = {}
            pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
            if os.path.isfile(pk):
                peaks = json.load(open(pk))
            peak = float(peaks.get("bf16_tflops", 1590.0))
            ach = 3.0 * flops / (ms * 1e-3) / 1e12          # x hi/lo against template hi, x hi against template lo
            extra = {"roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                                  "peak_source": "measured" if peaks else "fallback", "precision_passes": 3}}
        if w == "c4":
            # tensor roofline: executed BF16 flops = 2 precision passes (hi/lo split of x) x algorithmic flops
            peaks = {}
            pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
            if os.path.isfile(pk):
                peaks = json.load(open(pk))
            peak = float(peaks.get("bf16_tflops", 1590.0))
            ach = 2.0 * flops / (ms * 1e-3) / 1e12
            extra = {"tensor_cores": bool(h.uses_tensor_cores),
                     "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                                  "peak_source": "measured" if peaks else "fallback", "precision_passes": 2}}
        print(json.dumps({"workload": name, "ms": ms, "cells_per_s": cells / (ms * 1e-3), **extra,
                          "algorithmic_tflops": flops / (ms * 1e-3) / 1e12, "fp32_peak_tflops_measured": tf.value,
                          "frac_of_fp32_peak": flops / (ms * 1e-3) / 1e12 / tf.value if tf.value else None}), flush=True)


if __name__ == "__main__":
    main()

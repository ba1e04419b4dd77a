#!/usr/bin/env python
"""Randomised differential check of the dispatch: the same search through the long-range kernels (one call) and through short
calls (the round-1 layouts), plus a few oracle samples.  python tools/fuzz_paths.py [cases] [seed]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from oracle import xcorr_oracle as orc
    from pydsproutines_b200 import signals
    from pydsproutines_b200 import xcorrRoutines as xr
    ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    bad = 0
    for case in range(ncases):
        kind = rng.integers(0, 4)
        if kind == 0:
            N = int(2 ** rng.integers(5, 14))
        elif kind == 1:
            N = int(rng.integers(100, 4096))
        else:
            N = int(rng.choice([128, 256, 512, 1024, 2048, 4096, 8192, 1000, 1500, 3000]))
        step = int(rng.choice([1, 1, 1, 2, 3, 5, 16, 17]))
        Kraw = int(rng.integers(2400, 9000))
        a = int(rng.integers(-40, 40))
        b = a + Kraw * step
        rxlen = max(b + N - int(rng.integers(0, 60)), N + 10)
        freq = bool(rng.integers(0, 4) > 0)
        cutout, rx = signals.make_fastxcorr_case(int(rng.integers(1, 1 << 30)), N, rxlen, int(rng.integers(50, max(60, rxlen - N - 50))),
                                                 int(rng.integers(0, N)), amp=0.7)
        obj = xr.CyIppXcorrFFT(cutout)
        try:
            if freq:
                q, f = obj.xcorr(rx, a, b, step)                      # one call
                qs, fs = [], []
                for lo in range(a, b, 1700 * step):
                    qq, ff = obj.xcorr(rx, lo, min(lo + 1700 * step, b), step)
                    qs.append(qq)
                    fs.append(ff)
                q2, f2 = np.concatenate(qs), np.concatenate(fs)
                ok = q.shape == q2.shape and np.allclose(q, q2, rtol=1e-5, atol=1e-8) and np.mean(f != f2) < 0.01
                sh = np.arange(a, b, step)
                inside = np.nonzero((sh >= 0) & (sh + N <= rxlen))[0]
                pick = inside[rng.integers(0, inside.size, 3)]
                rq, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=sh[pick])
                ok = ok and np.allclose(q[pick], rq, rtol=1e-4, atol=1e-7)
                outside = np.nonzero(~((sh >= 0) & (sh + N <= rxlen)))[0]
                ok = ok and not q[outside].any()
            else:
                sh = np.arange(max(a, 0), min(b, rxlen - N + 1), step)
                q = xr.fastXcorr(dev(cutout), dev(rx), shifts=sh).cpu().numpy()
                pick = rng.integers(0, sh.size, 4)
                rq = orc.fastXcorr(cutout, rx, shifts=sh[pick])
                q2 = np.concatenate([xr.fastXcorr(dev(cutout), dev(rx), shifts=sh[lo:lo + 1500]).cpu().numpy() for lo in range(0, sh.size, 1500)])
                ok = np.allclose(q[pick], rq, rtol=1e-4, atol=1e-7) and np.allclose(q, q2, rtol=2e-5, atol=1e-8)
        except Exception as e:      # noqa: BLE001
            ok = False
            print("case %d raised %r" % (case, e))
        if not ok:
            bad += 1
            print("MISMATCH case %d: N=%d step=%d a=%d b=%d rxlen=%d freq=%s" % (case, N, step, a, b, rxlen, freq))
    print("fuzz: %d cases, %d bad" % (ncases, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())

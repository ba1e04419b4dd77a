#!/usr/bin/env python
"""Randomised differential check of the GroupXcorrFFT dispatch: one long call (per-cell lists / rotated rows / dense fold /
warp kernels) against short calls (gather lists of group.cu) and oracle samples.  python tools/fuzz_groups.py [cases] [seed]"""
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
    ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    bad = 0
    for case in range(ncases):
        F = int(2 ** rng.integers(5, 14))
        L = int(rng.integers(8, min(F, 600) + 1))
        G = int(rng.integers(1, 70))
        mode = rng.integers(0, 3)
        if mode == 0:
            starts = np.arange(G) * int(rng.integers(L, 3000))
        elif mode == 1:
            starts = np.cumsum(rng.integers(L, L + 1500, G))
        else:
            starts = np.sort(rng.integers(0, 6000, G))
        starts = (starts - starts[0]).astype(np.int64)
        K = int(rng.integers(2400, 7000))
        y = signals.qpsk(rng, G * L).reshape(G, L)
        rx = signals.cnoise(rng, K + int(starts[-1]) + F + 3)
        obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F)
        try:
            q, f = obj.xcorr(dev(rx), np.arange(K))
            q, f = q.cpu().numpy(), f.cpu().numpy()
            qs, fs = [], []
            for lo in range(0, K, 1900):
                qq, ff = obj.xcorr(dev(rx), np.arange(lo, min(lo + 1900, K)))
                qs.append(qq.cpu().numpy())
                fs.append(ff.cpu().numpy())
            q2, f2 = np.concatenate(qs), np.concatenate(fs)
            ok = np.allclose(q, q2, rtol=3e-5, atol=1e-7 * q2.max()) and np.mean(f != f2) < 0.02
            pick = rng.integers(0, K, 2)
            qr, fr = orc.GroupXcorrFFT(y, starts, 1, fftlen=F).xcorr(rx, pick)
            ok = ok and np.allclose(q[pick], qr, rtol=1e-4, atol=1e-6 * qr.max())
        except Exception as e:      # noqa: BLE001
            ok = False
            print("case %d raised %r" % (case, e))
        if not ok:
            bad += 1
            print("MISMATCH case %d: F=%d G=%d L=%d mode=%d K=%d" % (case, F, G, L, mode, K))
    print("fuzz groups: %d cases, %d bad" % (ncases, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())

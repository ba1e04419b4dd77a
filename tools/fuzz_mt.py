#!/usr/bin/env python
"""Randomised differential check of the multi-template correlator: the fast-correlation path (templates of 512 ... 4096 samples
over long records) and the tensor-core contraction against the FP32 SIMT kernel.  python tools/fuzz_mt.py [cases] [seed]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from pydsproutines_b200 import signals
    from pydsproutines_b200 import xcorrRoutines as xr
    ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 12
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    bad = 0
    for case in range(ncases):
        T = int(rng.integers(1, 24))
        L = int(rng.choice([512, 513, 777, 1024, 2000, 4096, 300, 4097]))
        S = 8192 - L + 1 if L <= 4096 else 5000
        K = int(4 * S + rng.integers(-900, 900)) if rng.integers(0, 3) else int(rng.integers(3000, 9000))
        start = int(rng.integers(0, 50))
        templates = (signals.qpsk(rng, T * L).reshape(T, L) * rng.uniform(0.5, 1.5, (T, 1))).astype(np.complex64)
        x = signals.cnoise(rng, start + K + L - 1 + int(rng.integers(0, 40)))
        pos, tp = int(rng.integers(0, K)), int(rng.integers(0, T))
        x[start + pos:start + pos + L] += np.complex64(1.5) * templates[tp].conj()
        try:
            b, ti, _, _ = xr._MtHandle(templates, False).run(dev(x), start, K, True, False, False)
            b32, t32, _, _ = xr._MtHandle(templates, False, fp32=True).run(dev(x), start, K, True, False, False)
            bd, td, _, _ = xr._MtHandle(templates, False, direct=True).run(dev(x), start, K, True, False, False)
            b, ti, b32, t32, bd, td = (a.cpu().numpy() for a in (b, ti, b32, t32, bd, td))
            ok = np.allclose(b, b32, rtol=5e-5, atol=1e-7 * b32.max()) and np.allclose(bd, b32, rtol=1e-4, atol=1e-7 * b32.max())
            ok = ok and ti[pos] == tp and td[pos] == tp and np.mean(ti != t32) < 0.01 and np.mean(td != t32) < 0.01
        except Exception as e:      # noqa: BLE001
            ok = False
            print("case %d raised %r" % (case, e))
        if not ok:
            bad += 1
            print("MISMATCH case %d: T=%d L=%d K=%d start=%d" % (case, T, L, K, start))
    print("fuzz mt: %d cases, %d bad" % (ncases, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())

"""TEST INFRASTRUCTURE ONLY -- seeded input recipes shared by oracle/make_golden.py (which runs the real reference on them)
and the tests (which regenerate the same inputs on the GPU box, where /root/reference does not exist)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pydsproutines_b200 import signals  # noqa: E402


def big_inputs(recipe):
    """Regenerates the inputs of a `bigxcorr_*` / `bigczt_*` vector from its recipe (the arrays themselves would be
    megabytes; the seeded generator is deterministic, so only the recipe and the reference's outputs are stored)."""
    seed, N, rxlen, p0 = int(recipe["seed"]), int(recipe["N"]), int(recipe["rxlen"]), int(recipe["p0"])
    if "foff" in recipe:
        rng = np.random.default_rng(seed)
        cutout = signals.qpsk(rng, N)
        rx = signals.cnoise(rng, rxlen)
        n = np.arange(N)
        rx[p0:p0 + N] += np.complex64(0.8) * cutout * np.exp(2j * np.pi * float(recipe["foff"]) * n / float(recipe["fs"])).astype(np.complex64)
        return cutout, rx
    return signals.make_fastxcorr_case(seed, N, rxlen, p0, int(recipe["k0"]), amp=0.7)

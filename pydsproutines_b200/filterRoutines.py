"""Mirror of the one filterRoutines.py function on the detection hot path: cupyMovingAverage
(filterRoutines.py:1167-1203 -> filter.cu:291-347), the sliding energy used by the template correlator."""
import numpy as np

from . import _lib
from ._lib import check
from .xcorrRoutines import _select_device


def cupyMovingAverage(x, avgLength, NUM_PER_THREAD=33, THREADS_PER_BLK=32, sumInstead=False):
    """Causal moving average (or sum) of a float32 array, zero padded at the front, output length ==
    input length.  NUM_PER_THREAD / THREADS_PER_BLK are accepted for compatibility; the reference's
    argument checks are kept (TypeError for non-float32, ValueError for an even NUM_PER_THREAD)."""
    if _lib.np_dtype_of(x) != np.dtype(np.float32):
        raise TypeError("Must be float32, found %s" % _lib.np_dtype_of(x))      # cupyRequireDtype
    if NUM_PER_THREAD % 2 == 0:
        raise ValueError("NUM_PER_THREAD must be odd.")                          # :1180-1181
    x = _lib.contiguous(x)
    _select_device(x)
    out = _lib.empty_like_space(x, tuple(x.shape), np.float32)
    check(_lib.load().xc_moving_average_f32(_lib.ptr(x), int(np.prod(x.shape)), int(avgLength), 1 if sumInstead else 0,
                                            _lib.ptr(out), _lib.where_of(x), _lib.current_stream(x)))
    return out


def slidingEnergy(x, L):
    """valid-mode windowed energy of a complex64 array: E[s] = sum_{n<L} |x[s+n]|^2 (what
    xcorrRoutines.py:342-355 builds from cupyComplexMagnSq + cupyMovingAverage(sumInstead=True)[L-1:])."""
    if _lib.np_dtype_of(x) != np.dtype(np.complex64):
        raise TypeError("Must be complex64, found %s" % _lib.np_dtype_of(x))
    x = _lib.contiguous(x)
    _select_device(x)
    n = int(x.shape[0])
    if L < 1 or L > n:
        raise ValueError("need 1 <= L <= len(x)")
    out = _lib.empty_like_space(x, n - L + 1, np.float32)
    check(_lib.load().xc_sliding_energy_c64(_lib.ptr(x), n, int(L), _lib.ptr(out), _lib.where_of(x),
                                            _lib.current_stream(x)))
    return out


_lib.adopt_module(globals(), ['cupyMovingAverage', 'slidingEnergy'])

"""The part of the reference's spectralRoutines.py the detection path touches: the cached chirp-z transform object
(spectralRoutines.py:239-391) that cztXcorr and cp_fastXcorr_v2(cztObj=...) take.  It describes the frequency grid; the
transform itself runs inside the fused correlation kernel (xc_cztxcorr_c64), and run / runMany call the same kernel
stand-alone (xc_czt_c64)."""
import numpy as np

from . import _lib
from ._lib import check


class CZTCached:
    """Same constructor and attributes as the reference's CZTCached / CZTCachedGPU: k = int((f2 - f1) / binWidth + 1)
    bins of width (f2 - f1 + binWidth) / k from f1 (spectralRoutines.py:244-259), m = input length."""

    def __init__(self, xlength, f1, f2, binWidth, fs, convertTo32fc=False):
        self.binWidth = binWidth
        self.f1 = f1
        self.f2 = f2
        self.fs = fs
        self.k = int((f2 - f1) / binWidth + 1)
        self.m = int(xlength)
        self.convertTo32fc = convertTo32fc
        if self.k < 1:
            raise ValueError("empty frequency range")

    def getFreq(self):
        return np.arange(self.k) * self.binWidth + self.f1                    # spectralRoutines.py:308-309

    # start and step of the grid in cycles per sample (what the C-ABI takes)
    @property
    def f0_cycles(self):
        return float(self.f1) / float(self.fs)

    @property
    def df_cycles(self):
        return float(self.f2 - self.f1 + self.binWidth) / self.k / float(self.fs)

    def runMany(self, xmany, out=None):
        """CZT of every row of xmany (rows, m) -> (rows, k); complex64 in, complex128 out like the reference (complex64
        with convertTo32fc / for CUDA tensors, spectralRoutines.py:272-306, 361-391)."""
        if _lib.np_dtype_of(xmany) != np.dtype(np.complex64):
            raise TypeError("x must be complex64, got %s" % _lib.np_dtype_of(xmany))
        x = _lib.contiguous(xmany)
        if x.ndim != 2 or x.shape[1] != self.m:
            raise ValueError("expected rows of %d samples" % self.m)
        if _lib.is_device(x):
            check(_lib.load().xc_set_device(x.device.index or 0))
        g = _lib.empty_like_space(x, (x.shape[0], self.k), np.complex64)
        check(_lib.load().xc_czt_c64(_lib.ptr(x), self.m, x.shape[0], self.f0_cycles, self.df_cycles, self.k, _lib.ptr(g),
                                     _lib.where_of(x), _lib.current_stream(x)))
        if not _lib.is_torch(g) and not self.convertTo32fc:
            g = g.astype(np.complex128)
        if out is None:
            return g
        out[...] = g

    def run(self, x):
        return self.runMany(x.reshape(1, -1))[0]


CZTCachedGPU = CZTCached


_lib.adopt_module(globals(), ['CZTCached'])

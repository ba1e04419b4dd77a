"""The part of the reference's spectralRoutines.py the detection path touches: the cached chirp-z transform object
(spectralRoutines.py:239-391) that cztXcorr and cp_fastXcorr_v2(cztObj=...) take.  Here it is only the description
of the frequency grid -- the transform itself runs inside the fused correlation kernel (xc_cztxcorr_c64)."""
import numpy as np


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

    def run(self, x):
        raise NotImplementedError("the stand-alone transform is not part of the detection path; use cztXcorr or "
                                  "cp_fastXcorr_v2(cztObj=...)")

    runMany = run


CZTCachedGPU = CZTCached

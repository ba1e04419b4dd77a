"""TEST INFRASTRUCTURE ONLY -- CPU (numpy/scipy) restatement of the reference's cross-correlation
detection hot path.  It is the parity checker for the CUDA library and the timed `cpu_baseline`;
the product (pydsproutines_b200) never imports it.

Every function cites the reference file:line (relative to icyveins7/pydsproutines @ e005737) whose
arithmetic it follows.  Parity status: PINNED -- oracle/make_golden.py runs the *real* reference
(`oracle/ref_import.py`) in the build container and tests/test_oracle_golden.py checks this
restatement against those committed vectors (tests/golden/*.npz) bit for bit where the reference
has a CPU implementation (fastXcorr, GroupXcorrFFT); the GPU-only reference pieces
(TemplateCrossCorrelator, multiTemplateSlidingDotProduct, movingAverage, findLocalMaxima, row argmax)
are restated from their kernels and pinned by the reference's own known-answer unit tests
(xcorrRoutines.py:2125-2245, filterRoutines.py:1326-1335, cupyExtensions.py:697-712).

Float semantics are those of numpy >= 2 (complex64 stays complex64 through scipy.fft, float32
scalars stay float32 under `** 2.0`), i.e. what the reference computes in this image.
"""
import numpy as np
import scipy.fft as sfft
import scipy.signal as sps


# --------------------------------------------------------------------------------------------
# helpers used by the reference
# --------------------------------------------------------------------------------------------
def makeFreq(length, fs):
    """signalCreationRoutines.py:380-386 -- FFT bin frequencies, upper half wrapped negative."""
    freq = np.arange(length, dtype=np.float64) / length * fs
    freq[freq >= fs / 2] -= fs
    return freq


# --------------------------------------------------------------------------------------------
# a1/a2/a3: fastXcorr  (xcorrRoutines.py:460-580)
# --------------------------------------------------------------------------------------------
def fastXcorr(cutout, rx, freqsearch=False, outputCAF=False, shifts=None, absResult=True):
    """Same signature and return tuples as xcorrRoutines.fastXcorr (460-580).

    no freqsearch  : QF2[s] = |vdot(rx[s:s+N], cutout)|^2 / ||cutout||^2 / ||rx[s:s+N]||^2   (483-495)
    freqsearch     : per shift p = rx[s:s+N]*conj(cutout); X = fft(p); bin = argmax |X^2|       (511-531)
    outputCAF      : full (K, N) surface                                                        (550-580)
    absResult=False: complex QF scaled by the norms (not squared)                               (497-509, 533-548, 568-580)
    """
    N = len(cutout)
    if shifts is None:
        shifts = np.arange(len(rx) - N + 1)            # :473-474
    K = len(shifts)
    cnorm = np.linalg.norm(cutout)                      # :477
    cnormsq = cnorm ** 2.0                              # :478
    cconj = cutout.conj()                               # recomputed per shift in the reference (:521)

    if not freqsearch:
        if absResult is True:
            out = np.zeros(K, dtype=np.float64)
            for i in range(K):
                win = rx[shifts[i]: shifts[i] + N]
                v = np.abs(np.vdot(win, cutout)) ** 2.0             # :489-491
                out[i] = v                                           # stored to float64 first (:489)
                out[i] = out[i] / cnormsq / (np.linalg.norm(win) ** 2.0)   # :492-493
            return out
        out = np.zeros(K, dtype=np.complex128)
        for i in range(K):
            win = rx[shifts[i]: shifts[i] + N]
            out[i] = np.vdot(win, cutout)                            # :503-505
            out[i] = out[i] / cnorm / np.linalg.norm(win)            # :506-507
        return out

    if not outputCAF:
        freqlist = np.zeros(K, dtype=np.uint32)
        if absResult is True:
            out = np.zeros(K, dtype=np.float64)
            for i in range(K):
                win = rx[shifts[i]: shifts[i] + N]
                spec = sfft.fft(win * cconj)                         # :521-522
                specsq = spec ** 2.0                                 # complex square (:523)
                mag = np.abs(specsq)
                b = np.argmax(mag)                                   # first occurrence (:524)
                freqlist[i] = b
                out[i] = mag[b] / cnormsq / (np.linalg.norm(win) ** 2.0)   # :526-529
            return out, freqlist
        out = np.zeros(K, dtype=np.complex128)
        for i in range(K):
            win = rx[shifts[i]: shifts[i] + N]
            spec = sfft.fft(win * cconj)                             # :539-540
            b = np.argmax(np.abs(spec))                              # :541
            freqlist[i] = b
            out[i] = spec[b] / cnorm / np.linalg.norm(win)           # :543-546
        return out, freqlist

    if absResult is True:
        out = np.zeros((K, N), dtype=np.float64)
        for i in range(K):
            win = rx[shifts[i]: shifts[i] + N]
            spec = sfft.fft(win * cconj)                             # :559-560
            out[i] = np.abs(spec ** 2.0) / cnormsq / (np.linalg.norm(win) ** 2.0)   # :561-564
        return out
    out = np.zeros((K, N), dtype=np.complex128)
    for i in range(K):
        win = rx[shifts[i]: shifts[i] + N]
        spec = np.fft.fft(win * cconj)                               # numpy fft here (:575)
        out[i] = spec / cnorm / np.linalg.norm(win)                  # :577-578
    return out


# --------------------------------------------------------------------------------------------
# a6: CyIppXcorrFFT conventions on top of a2  (CyIppXcorrFFT.pyx:25-80, IppXcorrFFT.cpp:94-178)
# --------------------------------------------------------------------------------------------
def cyIppXcorrFFT(cutout, rx, startIdx, endIdx, step):
    """float32 peaks / int32 bins; shifts with i<0 or i+N>len(rx) give 0.0 / 0 (IppXcorrFFT.cpp:125-130)."""
    N = len(cutout)
    idx = np.arange(startIdx, endIdx, step)
    peaks = np.zeros(idx.size, dtype=np.float32)
    bins = np.zeros(idx.size, dtype=np.int32)
    ok = (idx >= 0) & (idx + N <= len(rx))
    if ok.any():
        r, f = fastXcorr(cutout, rx, freqsearch=True, shifts=idx[ok])
        peaks[ok] = r.astype(np.float32)
        bins[ok] = f.astype(np.int32)
    return peaks, bins


# --------------------------------------------------------------------------------------------
# a7/a8: GroupXcorrFFT  (xcorrRoutines.py:1047-1189; GroupXcorrFFT.cpp:53-199)
# --------------------------------------------------------------------------------------------
class GroupXcorrFFT:
    """Restates the numpy methods of xcorrRoutines.GroupXcorrFFT (1047-1189)."""

    def __init__(self, ygroups, starts, fs, autoConj=True, fftlen=None, autoZeroStarts=True):
        assert starts.size == ygroups.shape[0]                       # :1057
        self.starts = starts - starts[0] if autoZeroStarts else starts   # :1058-1060
        self.numGroups = self.starts.size
        self.fs = fs
        self.ygroupLen = ygroups.shape[1]
        self.fftlen = self.ygroupLen if fftlen is None else fftlen   # :1065-1068
        self.ygroupNormSq = np.linalg.norm(ygroups.flatten()) ** 2   # :1070
        self.ygroups = ygroups.conj() if autoConj else ygroups       # :1072-1073
        self.fftfreq = makeFreq(self.fftlen, self.fs)                # :1076
        self.groupPhases = np.exp(                                   # :1077-1079 (complex128)
            -1j * 2 * np.pi * self.fftfreq * self.starts.reshape((-1, 1)) / self.fs)

    def _surface(self, rx, shift):
        """One row of the CAF: xcorrRoutines.py:1152-1184."""
        G, L, F = self.numGroups, self.ygroupLen, self.fftlen
        prod = np.zeros((G, F), rx.dtype)                            # :1152
        esum = np.zeros(G)                                           # float64 collector (:1153)
        for g in range(G):
            seg = rx[shift + self.starts[g]: shift + self.starts[g] + L]   # :1157-1159
            esum[g] = np.linalg.norm(seg) ** 2                       # :1160
            np.multiply(self.ygroups[g, :], seg, out=prod[g, :L])    # :1163-1165
        spec = np.fft.fft(prod, n=F, axis=1)                         # :1168
        np.multiply(spec, self.groupPhases, out=spec)                # :1170
        comb = np.sum(spec, axis=0)                                  # :1172
        return np.abs(comb) ** 2 / np.sum(esum) / self.ygroupNormSq  # :1176 / :1183

    def default_shifts(self, rx):
        return np.arange(len(rx) - (self.starts[-1] + self.fftlen) + 1)   # :1141 (uses fftlen)

    def xcorr(self, rx, shifts=None, flattenToTime=True):
        if shifts is None:
            shifts = self.default_shifts(rx)
        else:
            assert shifts[-1] + self.starts[-1] + self.fftlen < rx.size   # :1143
        if flattenToTime:
            xc = np.zeros(shifts.size)
            fi = np.zeros(shifts.size, dtype=np.uint32)
            for i, s in enumerate(shifts):
                row = self._surface(rx, s)
                b = np.argmax(row)                                   # :1177
                xc[i] = row[b]
                fi[i] = b
            return xc, fi
        xc = np.zeros((shifts.size, self.fftlen))
        for i, s in enumerate(shifts):
            xc[i, :] = self._surface(rx, s)
        return xc


def cyGroupXcorrFFT(ygroups, offsets, fs, rx, shifts, fftlen=-1, autoConj=True):
    """CyGroupXcorrFFT.pyx:13-66 -- CAF only, float32 (K, F); offsets used as given (no auto-zero,
    GroupXcorrFFT.cpp:53-75); fftlen<L raises (GroupXcorrFFT.cpp:11-14)."""
    L = ygroups.shape[1]
    F = L if fftlen < 0 else fftlen
    if F < L:
        raise ValueError("INVALID_FFTLEN")
    o = GroupXcorrFFT(ygroups, np.asarray(offsets), fs, autoConj=autoConj, fftlen=F,
                      autoZeroStarts=False)
    out = np.zeros((len(shifts), F), dtype=np.float32)
    for i, s in enumerate(shifts):
        out[i] = o._surface(rx, int(s)).astype(np.float32)
    return out


# --------------------------------------------------------------------------------------------
# a11: sliding energy primitives  (complex_magn.cu:8-20, filter.cu:291-347, filterRoutines.py:1167-1203)
# --------------------------------------------------------------------------------------------
def complexMagnSq(x, out_dtype=np.float64):
    """complex_magn.cu:8-20 -- |x|^2 evaluated in the OUTPUT type (re*re + im*im)."""
    re = x.real.astype(out_dtype)
    im = x.imag.astype(out_dtype)
    return re * re + im * im


def movingAverage(x, avgLength, sumInstead=False):
    """filter.cu:291-347 -- causal length-L window, zero padded at the front, output length ==
    input length, accumulated in double, rounded to float32 (pinned by filterRoutines.py:1326-1335:
    equals lfilter(ones(L), 1, x) to 5 places)."""
    c = np.concatenate(([0.0], np.cumsum(x.astype(np.float64))))
    idx = np.arange(x.size)
    lo = np.maximum(idx + 1 - avgLength, 0)
    s = c[idx + 1] - c[lo]
    if not sumInstead:
        s = s / float(avgLength)
    return s.astype(np.float32)


def slidingEnergy(x, L):
    """valid-mode windowed energy E[s] = sum_{n<L} |x[s+n]|^2 (what xcorrRoutines.py:343-355 slices
    out of the zero padded moving sum as [L-1:])."""
    return movingAverage(complexMagnSq(x, np.float32), L, sumInstead=True)[L - 1:]


# --------------------------------------------------------------------------------------------
# a9: TemplateCrossCorrelator  (xcorrRoutines.py:277-371) with numpy.fft standing in for cupy.fft
# --------------------------------------------------------------------------------------------
class TemplateCrossCorrelator:
    def __init__(self, templates, inputSize):
        if templates.ndim != 2:
            raise ValueError("Templates must be a 2D array; 1 row for 1 template.")   # :295-296
        self._inputSize = inputSize
        self._L = templates.shape[1]
        self._templateNorms = np.linalg.norm(templates, axis=1)                       # :302
        padded = np.pad(templates, ((0, 0), (0, inputSize - self._L)))                # :306-309
        self._tfftconj = np.fft.fft(padded, axis=1).conj()                            # :310

    def correlate(self, x, returnMax=False):
        if x.ndim != 1 or x.size != self._inputSize:
            raise ValueError("x must be 1D of length %d" % self._inputSize)           # :335-336
        L = self._L
        xfft = np.fft.fft(x)                                                          # :339
        norms = movingAverage(complexMagnSq(x, np.float32), L, sumInstead=True) ** 0.5   # :342-348
        out = np.fft.ifft(xfft * self._tfftconj, axis=1)                              # :352
        # NB: for L == 1 the reference's `:-L+1` slice is empty; restated for L >= 2 only
        nout = out[:, :-L + 1] / norms[L - 1:]                                        # :355
        nout = nout / self._templateNorms.reshape((-1, 1))                            # :357
        if not returnMax:
            return nout
        mag = np.abs(nout)                                                            # :366
        tidx = np.argmax(mag, axis=0)                                                 # :367
        return np.take_along_axis(mag, tidx[None, :], 0).reshape(-1), tidx            # :369-371


# --------------------------------------------------------------------------------------------
# a10: multiTemplateSlidingDotProduct  (cupyExtensions.py:563-642 -> multiplySlices.cu:251-399)
# --------------------------------------------------------------------------------------------
def multiTemplateSlidingDotProduct(x, templates, startIdx, idxlen, templateEnergies=None):
    """QF2[s] = max_t |sum_n t[n] x[s+n]|^2 / E_t / E_x[s]; templates are NOT conjugated here
    (multiplySlices.cu:352-356); ties keep the lowest template index and a slide whose best value
    is not > 0 keeps index 0 (strict '>' against a zero-initialised output, 313-316, 382-386)."""
    if templates.ndim != 2:
        raise ValueError("Templates should be 2D; each row is an individual template.")
    T, L = templates.shape
    if startIdx < 0:
        raise ValueError("startIdx should be >= 0.")
    if startIdx + idxlen - 1 + L - 1 >= x.size:
        raise ValueError("final slide index should be within the bounds of d_x.")
    if templateEnergies is None:
        templateEnergies = np.sum(np.abs(templates) ** 2, axis=1).astype(np.float32)   # :598
    seg = x[startIdx: startIdx + idxlen + L - 1]
    ex = slidingEnergy(seg, L)
    best = np.zeros(idxlen, np.float32)
    bidx = np.zeros(idxlen, np.int32)
    for t in range(T):
        # sum_n t[n] x[s+n]  == correlate(x, conj(t)) in 'valid' mode
        d = np.correlate(seg, templates[t].conj(), mode="valid").astype(np.complex64)
        q = (complexMagnSq(d, np.float32) / templateEnergies[t] / ex).astype(np.float32)
        upd = q > best
        best[upd] = q[upd]
        bidx[upd] = t
    return bidx, best


# --------------------------------------------------------------------------------------------
# a12: row argmax (argmax.cu:93-152) and a13: local maxima (peakfinding.cu:14-58)
# --------------------------------------------------------------------------------------------
def argmaxAbsRows(x, useNormSqInstead=False):
    """argmax.cu:93-152 -- per row argmax of abs(x) (or abs(x)^2) and its value.  The reference kernel
    does not fix the tie order; the restatement (and the new kernel) take the LOWEST index, which is
    what numpy argmax and fastXcorr (xcorrRoutines.py:524) do."""
    mag = complexMagnSq(x, np.float32)
    if not useNormSqInstead:
        mag = np.abs(x).astype(np.float32)
    idx = np.argmax(mag, axis=1).astype(np.uint32)
    return idx, np.take_along_axis(mag, idx[:, None].astype(np.int64), 1).reshape(-1)


def findLocalMaxima(x, minHeight):
    """peakfinding.cu:14-58 -- indices with x[i] > minHeight and x[i] > both neighbours (neighbours
    outside the array count as 0).  Returned SORTED (the reference's atomic append is unordered)."""
    p = np.concatenate(([0.0], x.astype(np.float32), [0.0])).astype(np.float32)
    mid = p[1:-1]
    ok = (mid > np.float32(minHeight)) & (mid > p[:-2]) & (mid > p[2:])
    return np.nonzero(ok)[0].astype(np.int32)


# ---------------------------------------------------------------------------------------------
# a14: gather / copy helpers (restated from the kernels; the reference has no CPU version)
# ---------------------------------------------------------------------------------------------
def cupyCopyGroups32fc(x, y, xStarts, yStarts, lengths):
    """copying.cu copy_groups_kernel32fc via cupyExtensions.py:41-83 (in place on y)."""
    for xs, ys, n in zip(xStarts, yStarts, lengths):
        y[ys:ys + n] = x[xs:xs + n]


def cupyCopyIncrementalEqualSlicesToMatrix_32fc(x, startIdx, increment, rowLength, numRows):
    """copying.cu:83-137: rows are slices start + r*increment, zero past the end of x."""
    out = np.zeros((numRows, rowLength), np.complex64)
    for r in range(numRows):
        a = startIdx + r * increment
        seg = x[a:a + rowLength]
        out[r, :seg.size] = seg
    return out


def multiplySlidesNormalised(x, y, startIdx, idxlen, coefficient=None):
    """multiplySlices.cu:112-216: x[t] * y[s + t] / float32(sqrt(window energy in double) * coefficient)."""
    x = np.asarray(x, np.complex64)
    coef = float(np.linalg.norm(x.astype(np.complex128))) if coefficient is None else float(np.asarray(coefficient).reshape(-1)[0])
    out = np.zeros((idxlen, x.size), np.complex64)
    for r in range(idxlen):
        seg = np.zeros(x.size, np.complex64)
        w = y[startIdx + r:startIdx + r + x.size]
        seg[:w.size] = w
        e = float(np.sum(np.abs(seg.astype(np.complex128)) ** 2))
        out[r] = (x * seg) / np.float32(np.sqrt(e) * coef)
    return out


# ---------------------------------------------------------------------------------------------
# 8f-1: chirp-z frequency zoom
# ---------------------------------------------------------------------------------------------
def cztXcorr(cutout, rx, f_searchMin, f_searchMax, fs, cztStep=0.1, outputCAF=False, shifts=None):
    """xcorrRoutines.py:413-457 on the grid of CZTCached (spectralRoutines.py:239-309), evaluated directly:
    X[k] = sum_n pdt[n] exp(-2j pi (f1 + k d) n / fs), d = (f2 - f1 + step) / k  (what the chirp-z transform computes)."""
    cutout = np.asarray(cutout, np.complex64)
    rx = np.asarray(rx, np.complex64)
    N = cutout.size
    k = int((f_searchMax - f_searchMin) / cztStep + 1)
    f_search = np.arange(k) * cztStep + f_searchMin
    d = (f_searchMax - f_searchMin + cztStep) / k
    if shifts is None:
        shifts = np.arange(len(rx) - N + 1)
    n = np.arange(N)
    E = np.exp(-2j * np.pi * np.outer(f_searchMin + np.arange(k) * d, n) / fs)            # (k, N) complex128
    cn = np.linalg.norm(cutout)
    cc = cutout.conj()
    if outputCAF:
        out = np.zeros((len(shifts), k))
        for i, s in enumerate(shifts):
            sl = rx[s:s + N]
            X = E @ (sl * cc).astype(np.complex128)
            out[i] = np.abs(X) ** 2 / np.linalg.norm(sl) ** 2 / cn ** 2
        return out, f_search
    res = np.zeros(len(shifts), rx.dtype)
    freqs = np.zeros(len(shifts))
    for i, s in enumerate(shifts):
        sl = rx[s:s + N]
        X = E @ (sl * cc).astype(np.complex128)
        mi = int(np.argmax(np.abs(X)))
        res[i] = X[mi] / np.linalg.norm(sl) / cn
        freqs[i] = f_search[mi]
    return res, freqs


class GroupXcorr:
    """xcorrRoutines.py:852-954: coherent sum over groups of y against an arbitrary frequency list, evaluated as the
    reference does (freqMat @ (rxconcat * yconcat) per shift), complex128."""

    def __init__(self, y, starts, lengths, freqs, fs, autoConj=True, autoZeroStarts=True):
        y = np.asarray(y)
        starts = np.asarray(starts)
        lengths = np.asarray(lengths)
        self.starts = starts - starts[0] if autoZeroStarts else starts                      # :885-887
        self.lengths = lengths
        self.freqs = np.asarray(freqs, np.float64)
        self.fs = fs
        yy = y.conj() if autoConj else y
        G = starts.size
        self.yconcat = np.hstack([yy[starts[i]:starts[i] + lengths[i]] for i in range(G)])  # :894-905
        self.yconcatNormSq = np.linalg.norm(self.yconcat) ** 2
        idx = np.hstack([np.arange(starts[i], starts[i] + lengths[i]) for i in range(G)])
        self.freqMat = np.exp(-2j * np.pi * self.freqs.reshape(-1, 1) * idx / fs)          # :909-918

    def xcorr(self, rx, shifts=None):
        rx = np.asarray(rx)
        if shifts is None:
            shifts = np.arange(len(rx) - (self.starts[-1] + self.lengths[-1]) + 1)
        xc = np.zeros(len(shifts))
        fp = np.zeros(len(shifts))
        for i, s in enumerate(shifts):
            rc = np.hstack([rx[s + self.starts[g]:s + self.starts[g] + self.lengths[g]] for g in range(self.starts.size)])
            pf = np.abs(self.freqMat @ (rc * self.yconcat))
            m = int(np.argmax(pf))
            xc[i] = pf[m] ** 2 / np.linalg.norm(rc) ** 2 / self.yconcatNormSq                # :949
            fp[i] = self.freqs[m]
        return xc, fp


class GroupXcorrCZT:
    """xcorrRoutines.py:957-1039 with the CZT of every group evaluated directly on the bins CZTCached produces
    (f1 + k (f2 - f1 + binWidth) / nbins, spectralRoutines.py:239-262) and the group phases on the nominal grid
    np.arange(f1, f2 + binWidth / 2, binWidth) (:1007-1010)."""

    def __init__(self, y, starts, lengths, f1, f2, binWidth, fs, autoConj=True, autoZeroStarts=True):
        y = np.asarray(y)
        starts = np.asarray(starts)
        lengths = np.asarray(lengths)
        self.starts = starts - starts[0] if autoZeroStarts else starts
        self.lengths = lengths
        self.fs = fs
        self.groups = [(y[starts[i]:starts[i] + lengths[i]].conj() if autoConj else y[starts[i]:starts[i] + lengths[i]])
                       for i in range(starts.size)]
        self.ystackNormSq = sum(np.linalg.norm(g) ** 2 for g in self.groups)
        self.k = int((f2 - f1) / binWidth + 1)
        self.bins = f1 + np.arange(self.k) * (f2 - f1 + binWidth) / self.k
        self.cztFreq = np.arange(f1, f2 + binWidth / 2, binWidth)

    def xcorr(self, rx, shifts=None):
        rx = np.asarray(rx)
        if shifts is None:
            shifts = np.arange(len(rx) - (self.starts[-1] + self.lengths[-1]) + 1)
        phases = np.exp(-2j * np.pi * self.cztFreq * self.starts.reshape(-1, 1) / self.fs)
        E = [np.exp(-2j * np.pi * np.outer(self.bins, np.arange(len(g))) / self.fs) for g in self.groups]
        xc = np.zeros((len(shifts), self.k))
        for i, s in enumerate(shifts):
            tot = np.zeros(self.k, np.complex128)
            en = 0.0
            for g, yg in enumerate(self.groups):
                sl = rx[s + self.starts[g]:s + self.starts[g] + len(yg)]
                en += np.linalg.norm(sl) ** 2
                tot += (E[g] @ (yg * sl).astype(np.complex128)) * phases[g]
            xc[i] = np.abs(tot) ** 2 / en / self.ystackNormSq
        return xc, self.cztFreq


def timeScan(x, y, td_scan_range, fs, freqBounds=None, steeringvec=None):
    """The time-alignment half of fineFreqTimeSearch (xcorrRoutines.py:637-664) = GenXcorr.xcorr (697-719):
    cost[k] = sum_m X[m] conj(Y[m]) conj(S[k, m]) / ||X|| / ||Y||, S = exp(+2j pi makeFreq(N, fs) td_k)."""
    X = np.fft.fft(np.asarray(x, np.complex128))
    Y = np.fft.fft(np.asarray(y, np.complex128))
    r = X * Y.conj()
    f = makeFreq(X.size, fs)
    if freqBounds is not None:
        r[(f < freqBounds[0]) | (f >= freqBounds[1])] = 0
    if steeringvec is None:
        steeringvec = np.exp(2j * np.pi * f * np.asarray(td_scan_range).reshape(-1, 1))
    cost = (steeringvec.conj() @ r) / np.linalg.norm(X) / np.linalg.norm(Y)
    return np.asarray(td_scan_range)[int(np.argmax(np.abs(cost)))], cost


def fineFreqTimeSearch(x_aligned, y_aligned, fineRes, freqfound, freqRes, fs, td_scan_range, steeringvec=None,
                       td_scan_freqBounds=None):
    """xcorrRoutines.py:583-667: successive fine frequency scans pp[j] = vdot(conj(y) x, exp(-2j pi f_j n / fs)),
    then the time scan on the frequency-corrected x."""
    x = np.asarray(x_aligned, np.complex128)
    y = np.asarray(y_aligned, np.complex128)
    n = np.arange(x.size)
    finefreq = None
    for res in fineRes:
        grid = np.arange(freqfound - freqRes, freqfound + freqRes, res)
        pp = np.exp(-2j * np.pi * np.outer(grid, n) / fs) @ (y * x.conj())
        freqfound = grid[int(np.argmax(np.abs(pp)))]
        finefreq = freqfound
    if finefreq is not None:
        x = x * np.exp(-2j * np.pi * finefreq * n / fs)
    td, cost = timeScan(x, y, td_scan_range, fs, td_scan_freqBounds, steeringvec)
    return finefreq, td, cost


def calcQF2(x, y):
    """xcorrRoutines.py:833-848 (1-D and row-wise 2-D)."""
    x = np.atleast_2d(np.asarray(x, np.complex128))
    y = np.atleast_2d(np.asarray(y, np.complex128))
    q = np.abs(np.sum(x.conj() * y, axis=1)) ** 2 / np.sum(np.abs(x) ** 2, axis=1) / np.sum(np.abs(y) ** 2, axis=1)
    return q if q.size > 1 else q[0]


class GroupXcorrCZT_Permutations:
    """xcorrRoutines.py:1264-1690 (CPU methods xcorr / getCAF): per-template complex surfaces
    CZT(y_k rx[shift + start_g : + L]) * exp(-2j pi cztFreq start_g / fs), combined per permutation as
    |sum_g surface|^2 / sum_g ||rx_g||^2 / sum_g ||y_k||^2.  The CZT is evaluated directly on CZTCached's bins."""

    def __init__(self, ygroups, ygroupIdxs, groupStarts, f1, f2, binWidth, fs, autoConj=True):
        self.ygroups = np.asarray(ygroups).conj() if autoConj else np.asarray(ygroups)
        self.ygroupIdxs = np.asarray(ygroupIdxs)
        self.groupStarts = np.asarray(groupStarts)
        self.fs = fs
        self.length = self.ygroups.shape[1]
        self.ygroupsEnergy = np.linalg.norm(self.ygroups, axis=1) ** 2
        k = int((f2 - f1) / binWidth + 1)
        self.bins = f1 + np.arange(k) * (f2 - f1 + binWidth) / k
        self.cztFreq = np.arange(f1, f2 + binWidth / 2, binWidth)

    def xcorr(self, rx, shifts):
        rx = np.asarray(rx)
        E = np.exp(-2j * np.pi * np.outer(self.bins, np.arange(self.length)) / self.fs)
        ph = np.exp(-2j * np.pi * self.cztFreq * self.groupStarts.reshape(-1, 1) / self.fs)
        T, G = self.ygroups.shape[0], self.groupStarts.size
        self.xcTemplates = np.zeros((T, len(shifts), self.bins.size), np.complex128)
        self.rxgroupNormSq = np.zeros((G, len(shifts)))
        for i, s in enumerate(shifts):
            for g in range(G):
                sl = rx[s + self.groupStarts[g]:s + self.groupStarts[g] + self.length]
                self.rxgroupNormSq[g, i] = np.linalg.norm(sl) ** 2
            for k in range(T):
                g = self.ygroupIdxs[k]
                sl = rx[s + self.groupStarts[g]:s + self.groupStarts[g] + self.length]
                self.xcTemplates[k, i] = (E @ (self.ygroups[k] * sl).astype(np.complex128)) * ph[g]
        return self.cztFreq

    def getCAF(self, templateIdx):
        tot = np.zeros(self.xcTemplates.shape[1:], np.complex128)
        rn = np.zeros(self.rxgroupNormSq.shape[1])
        yn = 0.0
        for g in range(self.groupStarts.size):
            k = np.argwhere(self.ygroupIdxs == g)[templateIdx[g]][0]
            tot += self.xcTemplates[k]
            rn += self.rxgroupNormSq[g]
            yn += self.ygroupsEnergy[k]
        return np.abs(tot) ** 2 / rn.reshape(-1, 1) / yn


def cztRun(xmany, f1, f2, binWidth, fs):
    """CZTCached.runMany (spectralRoutines.py:239-306) evaluated directly: X[r, k] = sum_n x[r, n] exp(-2j pi f_k n / fs)
    on the bins f_k = f1 + k (f2 - f1 + binWidth) / K, K = int((f2 - f1) / binWidth + 1)."""
    x = np.atleast_2d(np.asarray(xmany)).astype(np.complex128)
    K = int((f2 - f1) / binWidth + 1)
    fk = f1 + np.arange(K) * (f2 - f1 + binWidth) / K
    return x @ np.exp(-2j * np.pi * np.outer(np.arange(x.shape[1]), fk) / fs)

"""Host-side mirror of the reference's cross-correlation API (xcorrRoutines.py of pydsproutines @ e005737)
on top of libxcorr_b200: same names, argument meaning, return tuples and error behaviour, so a caller
only swaps the import line.  numpy arrays are host buffers (staged by the library); torch CUDA tensors
take the place of the reference's cupy arrays and stay on the device.

Nothing here computes on the CPU beyond argument bookkeeping and dtype conversion of the results.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import XC_DEVICE, XC_FLAG_ZERO_OOB, XC_HOST, check, xc_peak


# ------------------------------------------------------------------------------------------------
# helpers
# ------------------------------------------------------------------------------------------------
def _require_c64(name, x):
    """cupyRequireDtype (cupyHelpers.py:50-69): complex64 only, TypeError otherwise."""
    dt = _lib.np_dtype_of(x)
    if dt != np.dtype(np.complex64):
        raise TypeError("%s must be complex64, got %s" % (name, dt))
    if x.ndim != 1:
        raise ValueError("%s must be 1-D" % name)
    return _lib.contiguous(x)


def _select_device(ref):
    if _lib.is_device(ref):
        check(_lib.load().xc_set_device(ref.device.index or 0))


class _Shifts:
    """A shift list as the kernels take it: arithmetic progression (shift0, step, K) or explicit int64 list."""

    def __init__(self, shifts, rxlen, N, like):
        self.array = None
        if shifts is None:
            self.shift0, self.step, self.K = 0, 1, max(rxlen - N + 1, 0)
            lo, hi = 0, self.K - 1
        else:
            if _lib.is_torch(shifts):
                shifts = shifts.detach().cpu().numpy()
            s = np.asarray(shifts).astype(np.int64).reshape(-1)
            self.K = int(s.size)
            lo, hi = (int(s.min()), int(s.max())) if s.size else (0, -1)
            if s.size <= 1:
                self.shift0, self.step = (int(s[0]) if s.size else 0), 1
            else:
                d = np.diff(s)
                if np.all(d == d[0]):
                    self.shift0, self.step = int(s[0]), int(d[0])
                else:
                    self.shift0, self.step = 0, 0
                    self.array = np.ascontiguousarray(s)
                    if _lib.is_device(like):
                        import torch
                        self.array = torch.from_numpy(self.array).to(like.device)
        if self.K and (lo < 0 or hi + N > rxlen):
            # numpy slicing would silently shorten the window and then fail to broadcast
            raise ValueError("shifts [%d, %d] with a cutout of %d samples leave rx (%d samples)" % (lo, hi, N, rxlen))


def _fastxcorr_call(cutout, rx, sh, freqsearch, want_qf2=True, want_fidx=False, want_cplx=False, want_peak=False,
                    flags=0):
    lib = _lib.load()
    where = _lib.where_of(cutout, rx)
    _select_device(rx)
    K = sh.K
    qf2 = _lib.empty_like_space(rx, K, np.float32) if want_qf2 else None
    fidx = _lib.empty_like_space(rx, K, np.uint32) if want_fidx else None
    cplx = _lib.empty_like_space(rx, K, np.complex64) if want_cplx else None
    peak = xc_peak() if want_peak else None
    check(lib.xc_fastxcorr_c64(_lib.ptr(cutout), cutout.shape[0], _lib.ptr(rx), rx.shape[0], sh.shift0, sh.step, K,
                               _lib.ptr(sh.array), 1 if freqsearch else 0, flags, _lib.ptr(qf2), _lib.ptr(fidx),
                               _lib.ptr(cplx), C.byref(peak) if peak is not None else None, where,
                               _lib.current_stream(rx)))
    return qf2, fidx, cplx, peak


def _caf_call(cutout, rx, sh, complex_out):
    lib = _lib.load()
    where = _lib.where_of(cutout, rx)
    _select_device(rx)
    N = cutout.shape[0]
    caf = _lib.empty_like_space(rx, (sh.K, N), np.complex64 if complex_out else np.float32)
    check(lib.xc_fastxcorr_caf_c64(_lib.ptr(cutout), N, _lib.ptr(rx), rx.shape[0], sh.shift0, sh.step, sh.K,
                                   _lib.ptr(sh.array), _lib.ptr(caf), 1 if complex_out else 0, where,
                                   _lib.current_stream(rx)))
    return caf


def _as(x, np_dtype):
    """Result dtype conversion (the reference returns float64 / uint32 / complex128 from fastXcorr)."""
    if x is None:
        return None
    if _lib.is_torch(x):
        import torch
        target = {np.dtype(np.float64): torch.float64, np.dtype(np.float32): torch.float32,
                  np.dtype(np.complex128): torch.complex128, np.dtype(np.complex64): torch.complex64}
        # device indices stay int32 (torch has no first-class uint32)
        return x.to(target[np.dtype(np_dtype)]) if np.dtype(np_dtype) in target else x
    return x.astype(np_dtype, copy=False)


# ------------------------------------------------------------------------------------------------
# a1/a2/a3: fastXcorr  (xcorrRoutines.py:460-580)
# ------------------------------------------------------------------------------------------------
def fastXcorr(cutout, rx, freqsearch=False, outputCAF=False, shifts=None, absResult=True):
    """Same signature and return values as the reference's fastXcorr (xcorrRoutines.py:460-580):

    freqsearch=False                    -> QF^2 float64[K]                 (complex128[K] if absResult=False)
    freqsearch=True                     -> (QF^2 float64[K], freqlist uint32[K])   ((complex128[K], uint32[K]))
    freqsearch=True, outputCAF=True     -> float64 (K, N)                  (complex128 (K, N))
    `shifts` defaults to every valid shift (473-474).
    """
    cutout = _require_c64("cutout", cutout)
    rx = _require_c64("rx", rx)
    sh = _Shifts(shifts, rx.shape[0], cutout.shape[0], rx)
    if not freqsearch:
        if absResult:
            qf2, _, _, _ = _fastxcorr_call(cutout, rx, sh, False)
            return _as(qf2, np.float64)
        _, _, z, _ = _fastxcorr_call(cutout, rx, sh, False, want_qf2=False, want_cplx=True)
        return _as(z, np.complex128)
    if not outputCAF:
        if absResult:
            qf2, fidx, _, _ = _fastxcorr_call(cutout, rx, sh, True, want_fidx=True)
            return _as(qf2, np.float64), fidx
        _, fidx, z, _ = _fastxcorr_call(cutout, rx, sh, True, want_qf2=False, want_fidx=True, want_cplx=True)
        return _as(z, np.complex128), fidx
    caf = _caf_call(cutout, rx, sh, not absResult)
    return _as(caf, np.float64 if absResult else np.complex128)


def _czt_call(cutout, rx, sh, cztobj, want_qf2=False, want_fidx=False, want_cplx=False, want_caf=False):
    """xc_cztxcorr_c64 on the grid of a CZTCached object; returns (qf2, fidx, cplx, caf) in the memory space of rx."""
    lib = _lib.load()
    _select_device(rx)
    where = _lib.where_of(cutout, rx)
    K = sh.K
    qf2 = _lib.empty_like_space(rx, K, np.float32) if want_qf2 else None
    fidx = _lib.empty_like_space(rx, K, np.uint32) if want_fidx else None
    z = _lib.empty_like_space(rx, K, np.complex64) if want_cplx else None
    caf = _lib.empty_like_space(rx, (K, cztobj.k), np.float32) if want_caf else None
    check(lib.xc_cztxcorr_c64(_lib.ptr(cutout), cutout.shape[0], _lib.ptr(rx), rx.shape[0], sh.shift0, sh.step, K,
                              _lib.ptr(sh.array), cztobj.f0_cycles, cztobj.df_cycles, cztobj.k, _lib.ptr(qf2),
                              _lib.ptr(fidx), _lib.ptr(z), _lib.ptr(caf), where, _lib.current_stream(rx)))
    return qf2, fidx, z, caf


def cztXcorr(cutout, rx, f_searchMin, f_searchMax, fs, cztStep=0.1, outputCAF=False, shifts=None):
    """Same signature and return values as the reference's cztXcorr (xcorrRoutines.py:413-457): the frequency search of
    fastXcorr on the chirp-z grid of CZTCached (spectralRoutines.py:239-309), k = int((f2 - f1) / step + 1) bins of
    width (f2 - f1 + step) / k starting at f1, labelled np.arange(k) * step + f1.

    outputCAF=False -> (complex QF at the best bin, complex64[K]; its frequency label, float64[K])
    outputCAF=True  -> (QF^2 float64 (K, k), frequency labels float64[k])
    """
    cutout = _require_c64("cutout", cutout)
    rx = _require_c64("rx", rx)
    from .spectralRoutines import CZTCached
    cztobj = CZTCached(cutout.shape[0], f_searchMin, f_searchMax, cztStep, fs)
    f_search = cztobj.getFreq()
    sh = _Shifts(shifts, rx.shape[0], cutout.shape[0], rx)
    _, fidx, z, caf = _czt_call(cutout, rx, sh, cztobj, want_cplx=not outputCAF, want_fidx=not outputCAF, want_caf=outputCAF)
    if outputCAF:
        return _as(caf, np.float64), f_search
    if _lib.is_torch(fidx):
        import torch
        freqs = torch.from_numpy(f_search).to(fidx.device)[fidx.to(torch.int64)]
    else:
        freqs = f_search[fidx.astype(np.int64)]
    return z, freqs


def fastXcorrPeak(cutout, rx, freqsearch=True, shifts=None):
    """(QF^2, frequency index, peak shift) of the whole search in one call -- the tuple ranks exchange in
    the shift-sharded multi-GPU run (pydsproutines_b200.dist)."""
    cutout = _require_c64("cutout", cutout)
    rx = _require_c64("rx", rx)
    sh = _Shifts(shifts, rx.shape[0], cutout.shape[0], rx)
    qf2, fidx, _, peak = _fastxcorr_call(cutout, rx, sh, freqsearch, want_fidx=freqsearch, want_peak=True)
    return qf2, fidx, peak


# ------------------------------------------------------------------------------------------------
# a4: cp_fastXcorr  (xcorrRoutines.py:29-167)
# ------------------------------------------------------------------------------------------------
def cp_fastXcorr(cutout, rx, freqsearch=True, outputCAF=False, shifts=None, absResult=True, BATCH=1024,
                 copyToCpu=True):
    """GPU fastXcorr of the reference.  Accepts numpy arrays or device tensors; BATCH is accepted and
    ignored (no (BATCH, N) intermediate exists any more).  Returns (float64[K], uint32[K]) as numpy
    when copyToCpu, else device tensors.  Like the reference, only the flattened frequency search is
    implemented (67, 164, 167 print "Not implemented.")."""
    if _lib.np_dtype_of(cutout) != _lib.np_dtype_of(rx):
        raise Exception("Cutout and Rx must be same type, please cast one of them manually.")   # :54-57
    if not freqsearch or outputCAF or not absResult:
        raise NotImplementedError("cp_fastXcorr implements freqsearch=True, outputCAF=False, absResult=True only "
                                  "(as the reference); use fastXcorr for the other variants")
    result, freqlist = fastXcorr(cutout, rx, freqsearch=True, shifts=shifts)
    if copyToCpu and _lib.is_torch(result):
        return result.cpu().numpy(), freqlist.cpu().numpy().astype(np.uint32)
    return result, freqlist


# ------------------------------------------------------------------------------------------------
# a5: cp_fastXcorr_v2  (xcorrRoutines.py:169-274)
# ------------------------------------------------------------------------------------------------
def cp_fastXcorr_v2(cutout, rx, startIdx=0, idxlen=None, THREADS_PER_BLOCK=32, numSlidesPerBlk=None, cztObj=None,
                    flattenCAF=False, BATCH=None):
    """Step-1 search over rx[startIdx : startIdx+idxlen].  Returns (freqIdx uint32, qf2 float32) -- in THAT
    order (xcorrRoutines.py:272) -- when flattenCAF, else the float32 (idxlen, N) ambiguity surface.
    THREADS_PER_BLOCK / numSlidesPerBlk / BATCH are accepted and ignored.  With cztObj (a spectralRoutines.CZTCached /
    CZTCachedGPU) the frequency axis is its chirp-z grid (xcorrRoutines.py:247-249): indices into cztObj.getFreq(),
    surface of shape (idxlen, cztObj.k)."""
    cutout = _require_c64("cutout", cutout)
    rx = _require_c64("rx", rx)
    if cztObj is not None and cztObj.m != cutout.shape[0]:
        raise ValueError("CZT object input length doesn't match the cutout array size")   # :212-215
    if idxlen is None:
        idxlen = rx.shape[0] - cutout.shape[0] - startIdx + 1       # :209-210
    if startIdx < 0 or idxlen < 0 or startIdx + idxlen - 1 + cutout.shape[0] > rx.shape[0]:
        raise ValueError("Search range exceeds the input array.")                 # cupyExtensions.py:512-515
    sh = _Shifts(np.arange(startIdx, startIdx + idxlen), rx.shape[0], cutout.shape[0], rx)
    if cztObj is not None:
        if flattenCAF:
            qf2, fidx, _, _ = _czt_call(cutout, rx, sh, cztObj, want_qf2=True, want_fidx=True)
            return fidx, qf2
        return _czt_call(cutout, rx, sh, cztObj, want_caf=True)[3]
    if flattenCAF:
        qf2, fidx, _, _ = _fastxcorr_call(cutout, rx, sh, True, want_fidx=True)
        return fidx, qf2
    return _caf_call(cutout, rx, sh, False)


# ------------------------------------------------------------------------------------------------
# a6: CyIppXcorrFFT  (CyIppXcorrFFT.pyx:25-80, IppXcorrFFT.cpp:94-194)
# ------------------------------------------------------------------------------------------------
class CyIppXcorrFFT:
    """Drop-in for cython_ext.CyIppXcorrFFT: float32 peaks, int32 bins, out-of-range shifts give 0 / 0."""

    def __init__(self, cutout, num_threads=1, autoConj=True):
        cutout = np.ascontiguousarray(cutout)
        if cutout.dtype != np.complex64 or cutout.ndim != 1:
            raise ValueError("Buffer dtype mismatch, expected 'complex64_t' 1-D")     # Cython's own check
        # the kernels conjugate the cutout themselves; autoConj=False means "already conjugated"
        self._cutout = cutout if autoConj else np.ascontiguousarray(cutout.conj())
        self._num_threads = num_threads
        self.m_productpeaks = None
        self.m_freqlistinds = None

    def xcorr(self, rx, startIdx, endIdx, step):
        if _lib.np_dtype_of(rx) != np.dtype(np.complex64) or rx.ndim != 1:
            raise ValueError("Buffer dtype mismatch, expected 'complex64_t' 1-D")
        rx = _lib.contiguous(rx)
        K = len(range(startIdx, endIdx, step))
        sh = _Shifts.__new__(_Shifts)
        sh.shift0, sh.step, sh.K, sh.array = int(startIdx), int(step), K, None
        cutout = self._cutout
        if _lib.is_device(rx):
            import torch
            cutout = torch.from_numpy(cutout).to(rx.device)
        qf2, fidx, _, _ = _fastxcorr_call(cutout, rx, sh, True, want_fidx=True, flags=XC_FLAG_ZERO_OOB)
        if not _lib.is_torch(fidx):
            fidx = fidx.view(np.int32)
        self.m_productpeaks, self.m_freqlistinds = qf2, fidx
        return qf2, fidx

    def results(self):
        return self.m_productpeaks, self.m_freqlistinds


# ------------------------------------------------------------------------------------------------
# a7: GroupXcorrFFT  (xcorrRoutines.py:1047-1262)
# ------------------------------------------------------------------------------------------------
def makeFreq(length, fs):
    """signalCreationRoutines.py:380-386."""
    f = np.arange(length, dtype=np.float64) / length * fs
    f[f >= fs / 2] -= fs
    return f


class _GroupHandle:
    """Owns an xc_group* (folded contributor lists + ||ygroups||^2 on the device)."""

    def __init__(self, ygroups, starts, fftlen, autoConj):
        lib = _lib.load()
        y = np.ascontiguousarray(ygroups, dtype=np.complex64)
        st = np.ascontiguousarray(starts, dtype=np.int64)
        self._h = C.c_void_p()
        check(lib.xc_group_create(_lib.ptr(y), y.shape[0], y.shape[1], _lib.ptr(st), int(fftlen), 1 if autoConj else 0,
                                  C.byref(self._h)))

    def run(self, rx, sh, want_flat, want_caf, want_peak=False):
        lib = _lib.load()
        where = _lib.where_of(rx)
        _select_device(rx)
        K = sh.K
        qf2 = _lib.empty_like_space(rx, K, np.float32) if want_flat else None
        fidx = _lib.empty_like_space(rx, K, np.uint32) if want_flat else None
        caf = _lib.empty_like_space(rx, (K, self.F), np.float32) if want_caf else None
        peak = xc_peak() if want_peak else None
        check(lib.xc_group_xcorr(self._h, _lib.ptr(rx), rx.shape[0], sh.shift0, sh.step, K, _lib.ptr(sh.array),
                                 _lib.ptr(qf2), _lib.ptr(fidx), _lib.ptr(caf),
                                 C.byref(peak) if peak is not None else None, where, _lib.current_stream(rx)))
        return qf2, fidx, caf, peak

    def __del__(self):
        try:
            if self._h:
                _lib.load().xc_group_destroy(self._h)
                self._h = None
        except Exception:
            pass


def _group_shifts(shifts, like):
    sh = _Shifts.__new__(_Shifts)
    sh.array = None
    if _lib.is_torch(shifts):
        shifts = shifts.detach().cpu().numpy()
    s = np.asarray(shifts).astype(np.int64).reshape(-1)
    sh.K = int(s.size)
    if s.size <= 1:
        sh.shift0, sh.step = (int(s[0]) if s.size else 0), 1
    else:
        d = np.diff(s)
        if np.all(d == d[0]):
            sh.shift0, sh.step = int(s[0]), int(d[0])
        else:
            sh.shift0, sh.step = 0, 0
            sh.array = np.ascontiguousarray(s)
            if _lib.is_device(like):
                import torch
                sh.array = torch.from_numpy(sh.array).to(like.device)
    return sh


class GroupXcorrFFT:
    """Drop-in for xcorrRoutines.GroupXcorrFFT (1047-1262): same constructor, attributes and
    xcorr / xcorrThreads / xcorrGPU methods.  The G zero-padded FFTs + phase rotation + sum of the
    reference are computed as ONE folded F-point FFT per shift (see csrc/group.cu)."""

    def __init__(self, ygroups, starts, fs, autoConj=True, fftlen=None, autoZeroStarts=True):
        ygroups = np.asarray(ygroups)
        starts = np.asarray(starts)
        assert starts.size == ygroups.shape[0]                                  # :1057
        self.starts = starts - starts[0] if autoZeroStarts else starts         # :1058-1060
        self.numGroups = self.starts.size
        self.fs = fs
        self.ygroupLen = ygroups.shape[1]
        self.fftlen = self.ygroupLen if fftlen is None else fftlen             # :1065-1068
        self.ygroupNormSq = np.linalg.norm(ygroups.flatten()) ** 2             # :1070
        self.ygroups = ygroups.conj() if autoConj else ygroups                 # :1072-1073
        self.fftfreq = makeFreq(self.fftlen, self.fs)                          # :1076
        self._handle = _GroupHandle(self.ygroups, self.starts, self.fftlen, False)
        self._handle.F = int(self.fftlen)

    @property
    def groupPhases(self):
        """The (G, F) complex128 phase table of the reference (1077-1079); not used by the kernels."""
        return np.exp(-1j * 2 * np.pi * self.fftfreq * self.starts.reshape((-1, 1)) / self.fs)

    def _shifts(self, rx, shifts):
        if shifts is None:
            shifts = np.arange(rx.shape[0] - (int(self.starts[-1]) + self.fftlen) + 1)    # :1141
        else:
            assert shifts[-1] + self.starts[-1] + self.fftlen < rx.shape[0]               # :1143
        sh = _group_shifts(shifts, rx)
        return sh

    def xcorr(self, rx, shifts=None, flattenToTime=True):
        """(xc float64[K], fi uint32[K]) when flattenToTime, else the CAF float64 (K, F)."""
        if _lib.np_dtype_of(rx) != np.dtype(np.complex64):
            raise TypeError("rx must be complex64")
        rx = _lib.contiguous(rx)
        sh = self._shifts(rx, shifts)
        if flattenToTime:
            qf2, fidx, _, _ = self._handle.run(rx, sh, True, False)
            return _as(qf2, np.float64), fidx
        _, _, caf, _ = self._handle.run(rx, sh, False, True)
        return _as(caf, np.float64)

    def xcorrThreads(self, rx, shifts=None, NUM_THREADS=4):
        """CAF only, like the reference (1114-1135); NUM_THREADS is accepted and ignored."""
        return self.xcorr(rx, shifts, flattenToTime=False)

    def xcorrGPU(self, rx, shifts=None, flattenToTime=True):
        """Device-resident variant (1191-1262): rx is a torch CUDA tensor, results stay on the device."""
        if not _lib.is_device(rx):
            raise TypeError("rx must be a CUDA tensor (the reference requires a cupy array here)")
        return self.xcorr(rx, shifts, flattenToTime)

    def xcorrPeak(self, rx, shifts=None):
        """(QF^2[K], fi[K], xc_peak) -- the per-call peak tuple used by the sharded multi-GPU run."""
        rx = _lib.contiguous(rx)
        sh = self._shifts(rx, shifts)
        qf2, fidx, _, peak = self._handle.run(rx, sh, True, False, want_peak=True)
        return qf2, fidx, peak


# ------------------------------------------------------------------------------------------------
# 8(f) rank 2: refinement around the found peak  (xcorrRoutines.py:583-772)
# ------------------------------------------------------------------------------------------------
def _c128(a):
    """complex128 view of an aligned pair of arrays (the reference's arithmetic promotes to complex128)."""
    if _lib.is_torch(a):
        import torch
        return a.to(torch.complex128).contiguous()
    return np.ascontiguousarray(a, dtype=np.complex128)


def _fine_freq(x, y, freq_cycles):
    lib = _lib.load()
    where = _lib.where_of(x, y)
    _select_device(x)
    fc = np.ascontiguousarray(freq_cycles, dtype=np.float64)
    pp = _lib.empty_like_space(x, fc.size, np.complex128)
    best = C.c_int64(-1)
    bdev = _lib.empty_like_space(x, 1, np.int64) if where == XC_DEVICE else None
    check(lib.xc_fine_freq_c128(_lib.ptr(x), _lib.ptr(y), x.shape[0], _lib.ptr(fc), fc.size, _lib.ptr(pp),
                                _lib.ptr(bdev) if bdev is not None else C.cast(C.byref(best), C.c_void_p), where,
                                _lib.current_stream(x)))
    return pp, (int(bdev.item()) if bdev is not None else int(best.value))


def _time_scan(x, y, shift_cycles, td_scan_range, fs, freqBounds, steeringvec):
    lib = _lib.load()
    where = _lib.where_of(x, y, steeringvec)
    _select_device(x)
    N = x.shape[0]
    td = None if td_scan_range is None else np.ascontiguousarray(td_scan_range, dtype=np.float64)
    if steeringvec is not None:
        steeringvec = _c128(steeringvec)
        assert steeringvec.ndim == 2 and steeringvec.shape[1] == N
        NT = steeringvec.shape[0]
    else:
        NT = td.size
    fb = None if freqBounds is None else np.ascontiguousarray(freqBounds, dtype=np.float64)
    cost = _lib.empty_like_space(x, NT, np.complex128)
    best = C.c_int64(-1)
    bdev = _lib.empty_like_space(x, 1, np.int64) if where == XC_DEVICE else None
    check(lib.xc_time_scan_c128(_lib.ptr(x), _lib.ptr(y), N, float(shift_cycles), _lib.ptr(td), NT, float(fs),
                                _lib.ptr(fb), _lib.ptr(steeringvec), _lib.ptr(cost),
                                _lib.ptr(bdev) if bdev is not None else C.cast(C.byref(best), C.c_void_p), where,
                                _lib.current_stream(x)))
    return cost, (int(bdev.item()) if bdev is not None else int(best.value))


def makeTimeScanSteervec(td_scan_range, fs, siglen):
    """xcorrRoutines.py:669-675 (only needed by callers that want the matrix itself; the scans generate it on the fly)."""
    return np.exp(1j * 2 * np.pi * makeFreq(siglen, fs) * np.asarray(td_scan_range).reshape((-1, 1)))


def fineFreqTimeSearch(x_aligned, y_aligned, fineRes, freqfound, freqRes, fs, td_scan_range, steeringvec=None,
                       td_scan_freqBounds=None):
    """Drop-in for xcorrRoutines.fineFreqTimeSearch (583-667): returns (finefreqfound, timediff, cost_vec complex128).
    Each frequency pass and the time scan are FP64 kernels (csrc/fine.cu); the spectra are exact dense DFTs, so any
    length works."""
    x = _c128(x_aligned)
    y = _c128(y_aligned)
    assert x.ndim == 1 and x.shape == y.shape
    shift_cycles = 0.0
    if len(fineRes) > 0:
        for i in range(len(fineRes)):
            fineFreq = np.arange(freqfound - freqRes, freqfound + freqRes, fineRes[i])           # :618
            _, ind = _fine_freq(x, y, fineFreq / float(fs))
            freqfound = fineFreq[ind]                                                             # :628-629
        finefreqfound = freqfound
        shift_cycles = float(freqfound) / float(fs)                                              # :632-633
    else:
        finefreqfound = None
    td = np.asarray(td_scan_range)
    cost, idx_td = _time_scan(x, y, shift_cycles, td, fs, td_scan_freqBounds, steeringvec)
    return finefreqfound, td[idx_td], cost


class GenXcorr:
    """Drop-in for xcorrRoutines.GenXcorr (679-719)."""

    def __init__(self, td_scan_range, fs, siglen):
        self.td_scan_range = np.asarray(td_scan_range)
        self.fs = fs
        self.siglen = siglen
        self.sigFreq = makeFreq(siglen, fs)
        self.td_scan_freqBounds = None
        self._steer = None

    @property
    def steeringvec(self):
        if self._steer is None:
            self._steer = makeTimeScanSteervec(self.td_scan_range, self.fs, self.siglen)
        return self._steer

    def setTDscan_freqBounds(self, td_scan_freqBounds):
        self.td_scan_freqBounds = td_scan_freqBounds

    def xcorr(self, x, y):
        cost, idx = _time_scan(_c128(x), _c128(y), 0.0, self.td_scan_range, self.fs, self.td_scan_freqBounds, None)
        return self.td_scan_range[idx], cost


def calcQF2(x, y):
    """Drop-in for xcorrRoutines.calcQF2 (833-848): QF^2 of two aligned arrays (1-D, or row-wise for 2-D) = the
    no-frequency-search correlator at the single shift 0."""
    if x.ndim == 1 and y.ndim == 1:
        q, _, _, _ = _fastxcorr_call(_require_c64("x", x), _require_c64("y", y), _Shifts(np.array([0]), y.shape[0], x.shape[0], y), False)
        return _as(q, np.float64)[0]
    if x.ndim == 2 and y.ndim == 2:
        rows = [calcQF2(x[i], y[i]) for i in range(x.shape[0])]
        if _lib.is_torch(rows[0]):
            import torch
            return torch.stack(rows)
        return np.array(rows)


def convertQF2toSNR(qf2):
    """xcorrRoutines.py:723-725."""
    return qf2 / (1.0 - qf2)


def convertQF2toEffSNR(qf2):
    """xcorrRoutines.py:728-730."""
    return 2.0 * qf2 / (1.0 - qf2)


def convertEffSNRtoQF2(effSNR):
    """xcorrRoutines.py:733-735."""
    return effSNR / (2 + effSNR)


def expectedEffSNR(snr1, snr2=np.inf, OSR=1):
    """xcorrRoutines.py:738-755."""
    y = 1.0 / (0.5 * (1 / snr1 + 1 / snr2 + 1 / snr1 / snr2))
    return y / OSR


def sigmaDTO(signalBW, noiseBW, integTime, effSNR):
    """xcorrRoutines.py:758-764."""
    beta = np.pi / np.sqrt(3) * signalBW
    return 1.0 / beta / np.sqrt(noiseBW * integTime * effSNR)


def sigmaDFO(noiseBW, integTime, effSNR):
    """xcorrRoutines.py:767-772."""
    return 0.55 / integTime / np.sqrt(noiseBW * integTime * effSNR)


# ------------------------------------------------------------------------------------------------
# 8(f): GroupXcorr / GroupXcorrGPU / GroupXcorrCZT  (xcorrRoutines.py:852-1039, 1896-2051)
# ------------------------------------------------------------------------------------------------
class _GroupFreqHandle:
    """Owns an xc_groupfreq*: the group samples with the frequency list folded in, packed for the tensor cores."""

    def __init__(self, ycat, starts, lengths, freq_cycles, start_freq_cycles=None):
        lib = _lib.load()
        y = np.ascontiguousarray(ycat, dtype=np.complex64)
        st = np.ascontiguousarray(starts, dtype=np.int64)
        ln = np.ascontiguousarray(lengths, dtype=np.int32)
        fc = np.ascontiguousarray(freq_cycles, dtype=np.float64)
        sfc = None if start_freq_cycles is None else np.ascontiguousarray(start_freq_cycles, dtype=np.float64)
        assert y.size == int(ln.sum()) and (sfc is None or sfc.size == fc.size)
        self.NF = int(fc.size)
        self._h = C.c_void_p()
        check(lib.xc_groupfreq_create(_lib.ptr(y), _lib.ptr(st), _lib.ptr(ln), int(st.size), _lib.ptr(fc), _lib.ptr(sfc),
                                      self.NF, 0, C.byref(self._h)))

    def run(self, rx, shift0, K, want_flat, want_caf):
        lib = _lib.load()
        where = _lib.where_of(rx)
        _select_device(rx)
        qf2 = _lib.empty_like_space(rx, K, np.float32) if want_flat else None
        fidx = _lib.empty_like_space(rx, K, np.uint32) if want_flat else None
        caf = _lib.empty_like_space(rx, (K, self.NF), np.float32) if want_caf else None
        check(lib.xc_groupfreq_xcorr(self._h, _lib.ptr(rx), rx.shape[0], int(shift0), int(K), _lib.ptr(qf2),
                                     _lib.ptr(fidx), _lib.ptr(caf), None, None, where, _lib.current_stream(rx)))
        return qf2, fidx, caf

    def run_complex(self, rx, shift0, K, out):
        """raw complex surface d[s, k] into `out` (torch CUDA complex64 [K, NF]); rx must be a CUDA tensor"""
        _select_device(rx)
        check(_lib.load().xc_groupfreq_xcorr(self._h, _lib.ptr(rx), rx.shape[0], int(shift0), int(K), None, None, None,
                                             _lib.ptr(out), None, XC_DEVICE, _lib.current_stream(rx)))

    def run_shifts(self, rx, shifts, want_flat, want_caf):
        """Any shift list: the kernel works on the covering contiguous range (a tile is 128 consecutive shifts) and
        the requested rows are picked from it."""
        if _lib.is_torch(shifts):
            shifts = shifts.detach().cpu().numpy()
        s = np.asarray(shifts).astype(np.int64).reshape(-1)
        if s.size == 0:
            return self.run(rx, 0, 0, want_flat, want_caf)
        lo, hi = int(s.min()), int(s.max())
        out = self.run(rx, lo, hi - lo + 1, want_flat, want_caf)
        if s.size == hi - lo + 1 and np.all(np.diff(s) == 1):
            return out
        pick = s - lo
        if _lib.is_device(rx):
            import torch
            pick = torch.from_numpy(pick).to(rx.device)
        return tuple(None if o is None else o[pick] for o in out)

    def __del__(self):
        try:
            if self._h:
                _lib.load().xc_groupfreq_destroy(self._h)
                self._h = None
        except Exception:
            pass


class GroupXcorr:
    """Drop-in for xcorrRoutines.GroupXcorr (852-954): groups y[starts[i] : starts[i] + lengths[i]] correlated
    coherently over an arbitrary frequency list; xcorr returns (xc float64[K], freqpeaks float64[K]).  The reference's
    freqMat @ (rxconcat * yconcat) per shift runs as one segmented tensor-core contraction (csrc/groupfreq.cu)."""

    def __init__(self, y, starts, lengths, freqs, fs, autoConj=True, autoZeroStarts=True):
        y = np.asarray(y)
        starts = np.asarray(starts)
        lengths = np.asarray(lengths)
        assert starts.size == lengths.size                                           # :883
        self.starts = starts - starts[0] if autoZeroStarts else starts              # :885-887
        self.lengths = lengths
        self.numGroups = self.starts.size
        self.freqs = np.asarray(freqs, dtype=np.float64)
        self.fs = fs
        yy = y.conj() if autoConj else y                                             # :894-905
        self.yconcat = np.hstack([yy[starts[i]:starts[i] + lengths[i]] for i in range(self.numGroups)])
        self.yconcatNormSq = np.linalg.norm(self.yconcat) ** 2                       # :906
        # |freqMat @ p| does not depend on a common start offset, so the (zeroed) starts serve as phase origin
        self._handle = _GroupFreqHandle(self.yconcat, self.starts - self.starts.min(), self.lengths,
                                        self.freqs / float(fs))

    def _shifts(self, rx, shifts):
        if shifts is None:
            return np.arange(rx.shape[0] - (int(self.starts[-1]) + int(self.lengths[-1])) + 1)    # :921
        assert shifts[-1] + self.starts[-1] + self.lengths[-1] < rx.shape[0]                      # :923-925
        return shifts

    def _rx(self, rx):
        if _lib.np_dtype_of(rx) != np.dtype(np.complex64):
            raise TypeError("rx must be complex64")
        return _lib.contiguous(rx)

    def _offset(self):
        return int(self.starts.min())

    def xcorr(self, rx, shifts=None):
        rx = self._rx(rx)
        shifts = self._shifts(rx, shifts)
        qf2, fidx, _ = self._handle.run_shifts(rx, np.asarray(shifts) + self._offset(), True, False)
        if _lib.is_torch(fidx):
            import torch
            freqpeaks = torch.from_numpy(self.freqs).to(fidx.device)[fidx.long()]
        else:
            freqpeaks = self.freqs[fidx]
        return _as(qf2, np.float64), freqpeaks

    def xcorrIndices(self, rx, shifts=None):
        """(xc float32[K], freqinds int32[K]) -- what GroupXcorrGPU.xcorrKernel returns (2000-2037)."""
        rx = self._rx(rx)
        shifts = self._shifts(rx, shifts)
        qf2, fidx, _ = self._handle.run_shifts(rx, np.asarray(shifts) + self._offset(), True, False)
        return qf2, (fidx.astype(np.int32) if not _lib.is_torch(fidx) else fidx)


class GroupXcorrGPU(GroupXcorr):
    """Drop-in for xcorrRoutines.GroupXcorrGPU (1896-2051); autoConj / autoZeroStarts at their defaults like the
    reference's constructor (1905)."""

    def __init__(self, y, starts, lengths, freqs, fs):
        super().__init__(y, starts, lengths, freqs, fs)

    def xcorrKernel(self, rx, shifts, numShiftsPerBlk=2, verbTiming=False):
        assert np.asarray(shifts).size % numShiftsPerBlk == 0                        # :1970
        return self.xcorrIndices(rx, shifts)


class GroupXcorrCZT:
    """Drop-in for xcorrRoutines.GroupXcorrCZT (957-1039): xcorr returns (CAF float64 (K, nbins), cztFreq).  The
    per-group cached CZT + group phase + coherent sum of the reference is the same segmented contraction with the
    bins of CZTCached (f1 + k (f2 - f1 + binWidth) / nbins) inside a group and the nominal grid for the group phase."""

    def __init__(self, y, starts, lengths, f1, f2, binWidth, fs, autoConj=True, autoZeroStarts=True):
        y = np.asarray(y)
        starts = np.asarray(starts)
        lengths = np.asarray(lengths)
        assert starts.size == lengths.size                                           # :971
        self.starts = starts - starts[0] if autoZeroStarts else starts              # :972-974
        self.lengths = lengths
        self.numGroups = self.starts.size
        self.fs = fs
        self.f1, self.f2, self.binWidth = f1, f2, binWidth
        self.maxLength = int(np.max(self.lengths))
        self.ystack = np.zeros((self.numGroups, self.maxLength), y.dtype)           # :985-990
        for i in range(self.numGroups):
            self.ystack[i, :lengths[i]] = y[starts[i]:starts[i] + lengths[i]]
        if autoConj:
            self.ystack = self.ystack.conj()
        self.ystackNormSq = np.linalg.norm(self.ystack.flatten()) ** 2              # :992
        from .spectralRoutines import CZTCached
        self.cztc = CZTCached(self.maxLength, f1, f2, binWidth, fs)                 # :995
        self.cztFreq = np.arange(self.f1, self.f2 + self.binWidth / 2, self.binWidth)   # :1007
        nb = self.cztc.k
        if self.cztFreq.size != nb:
            raise ValueError("operands could not be broadcast together: %d CZT bins against %d group phases"
                             % (nb, self.cztFreq.size))      # the reference fails the same way at :1033
        ycat = np.hstack([self.ystack[i, :lengths[i]] for i in range(self.numGroups)])
        bins = self.cztc.f0_cycles + np.arange(nb) * self.cztc.df_cycles
        self._off = int(self.starts.min())
        self._handle = _GroupFreqHandle(ycat, self.starts - self._off, self.lengths, bins,
                                        self.cztFreq / float(fs))

    def xcorr(self, rx, shifts=None):
        if _lib.np_dtype_of(rx) != np.dtype(np.complex64):
            raise TypeError("rx must be complex64")
        rx = _lib.contiguous(rx)
        if shifts is None:
            shifts = np.arange(rx.shape[0] - (int(self.starts[-1]) + int(self.lengths[-1])) + 1)   # :1000
        else:
            assert shifts[-1] + self.starts[-1] + self.lengths[-1] < rx.shape[0]                   # :1002-1004
        _, _, caf = self._handle.run_shifts(rx, np.asarray(shifts) + self._off, False, True)
        return _as(caf, np.float64), self.cztFreq


class pbIppGroupXcorrCZT:
    """Drop-in for the reference's pybind class pbIppGroupXcorrCZT (pybinds/ippGroupXcorrCZT/pbGroupXcorrCZT.cpp:8-33,
    GroupXcorrCZT.cpp): groups are added one by one (or sliced from a long array), xcorr(x, shiftStart, shiftStep,
    numShifts) returns the float32 (numShifts, k) CAF with k = int((f2 - f1) / fstep + 1).  Same checks and messages
    (pybind11 turns std::range_error / std::invalid_argument into ValueError).  The segmented tensor-core contraction of
    GroupXcorrCZT is rebuilt lazily whenever the set of groups changed."""

    def __init__(self, maxlen, f1, f2, fstep, fs, NUM_THREADS=1):
        if NUM_THREADS < 1:
            raise ValueError("Number of threads must be greater than 0")          # GroupXcorrCZT.cpp:32
        self._N, self._f1, self._f2, self._fstep, self._fs = int(maxlen), float(f1), float(f2), float(fstep), float(fs)
        self._k = int((self._f2 - self._f1) / self._fstep + 1)                    # GroupXcorrCZT.cpp:13
        self._threads = int(NUM_THREADS)
        self._starts, self._groups, self._energies = [], [], []
        self._handle = None

    def getNumThreads(self):
        return self._threads

    def addGroup(self, start, group, autoConj=True):
        group = np.asarray(group)
        if group.ndim != 1:
            raise ValueError("Input must be 1d")
        start, length = int(start), int(group.shape[0])
        for gs, g in zip(self._starts, self._groups):                              # GroupXcorrCZT.cpp:45-59
            ge = gs + g.size
            if gs <= start < ge:
                raise ValueError("Group start overlaps with existing group! [%d,%d)" % (gs, ge))
            if gs <= start + length < ge:
                raise ValueError("Group end overlaps with existing group! [%d,%d)" % (gs, ge))
        if length > self._N:
            raise ValueError("Length of group exceeds maximum length")
        g = np.ascontiguousarray(group, dtype=np.complex64)
        self._starts.append(start)
        self._groups.append(g.conj() if autoConj else g.copy())
        self._energies.append(float(np.linalg.norm(g) ** 2))
        self._handle = None

    def addGroupsFromArray(self, starts, lengths, arr, autoConj=True):
        starts, lengths, arr = np.asarray(starts), np.asarray(lengths), np.asarray(arr)
        if arr.ndim != 1:
            raise ValueError("Input must be 1d")
        if starts.ndim != 1:
            raise ValueError("Starts must be 1d")
        if lengths.ndim != 1:
            raise ValueError("Lengths must be 1d")
        if lengths.shape[0] != starts.shape[0]:
            raise ValueError("Lengths must be same size as starts")
        m = int(starts.min())                                                       # GroupXcorrCZT.cpp:84-95
        for s0, n in zip(starts, lengths):
            self.addGroup(int(s0) - m, arr[int(s0):int(s0) + int(n)], autoConj)

    def printGroups(self):
        for i, (gs, g, e) in enumerate(zip(self._starts, self._groups, self._energies)):
            print("Group %d: [%d, %d)" % (i, gs, gs + g.size))
            print("Energy = %f" % e)

    def resetGroups(self):
        self._starts, self._groups, self._energies = [], [], []
        self._handle = None

    def _build(self):
        from .spectralRoutines import CZTCached
        order = np.argsort(np.asarray(self._starts, dtype=np.int64), kind="stable")
        starts = np.asarray(self._starts, dtype=np.int64)[order]
        lengths = np.array([self._groups[i].size for i in order], dtype=np.int32)
        ycat = np.hstack([self._groups[i] for i in order])
        cz = CZTCached(self._N, self._f1, self._f2, self._fstep, self._fs)
        bins = cz.f0_cycles + np.arange(cz.k) * cz.df_cycles
        nominal = (self._f1 + np.arange(self._k) * self._fstep) / self._fs         # ippe Slope(f1, fstep), GroupXcorrCZT.cpp:340-345
        self._off = int(starts.min())
        self._handle = _GroupFreqHandle(ycat, starts - self._off, lengths, bins, nominal)

    def xcorr(self, x, shiftStart, shiftStep, numShifts):
        if shiftStep < 0:
            raise ValueError("shiftStep cannot be negative")                       # GroupXcorrCZT.cpp:111-112
        if not self._starts:
            raise ValueError("No groups have been defined!")
        if _lib.np_dtype_of(x) != np.dtype(np.complex64):
            raise TypeError("x must be complex64")
        x = _lib.contiguous(x)
        for gs, g in zip(self._starts, self._groups):                              # GroupXcorrCZT.cpp:153-163
            xi = int(shiftStart) + gs
            if xi < 0:
                raise ValueError("Shifts accesses negative indices!")
            if xi + int(numShifts) * int(shiftStep) + g.size >= x.shape[0]:
                raise ValueError("Input length is insufficient for search range!")
        if self._handle is None:
            self._build()
        shifts = int(shiftStart) + np.arange(int(numShifts), dtype=np.int64) * max(int(shiftStep), 0) + self._off
        if int(shiftStep) == 0:
            shifts = np.full(int(numShifts), int(shiftStart) + self._off, dtype=np.int64)
        _, _, caf = self._handle.run_shifts(x, shifts, False, True)
        return caf


class GroupXcorrCZT_Permutations:
    """Drop-in for xcorrRoutines.GroupXcorrCZT_Permutations (1264-1690): every template of every group is correlated
    once over the CZT grid (complex surfaces kept on the device), then any permutation of one template per group is
    combined with getCAF / getCAF_GPU.  Each template is one tensor-core contraction (csrc/groupfreq.cu, complex
    epilogue); the combination is one kernel (xc_groupperm_combine)."""

    def __init__(self, ygroups, ygroupIdxs, groupStarts, f1, f2, binWidth, fs, autoConj=True):
        ygroups = np.asarray(ygroups)
        ygroupIdxs = np.asarray(ygroupIdxs)
        groupStarts = np.asarray(groupStarts)
        assert ygroups.shape[0] == ygroupIdxs.size                                       # :1298-1300
        assert np.unique(ygroupIdxs).size == groupStarts.size                            # :1301-1303
        self.numTemplates = ygroupIdxs.size
        self.numGroups = groupStarts.size
        assert np.all(np.sort(np.unique(ygroupIdxs)) == np.arange(self.numGroups))       # :1309-1311
        self.groupStarts = groupStarts
        self.ygroupIdxs = ygroupIdxs
        self.fs = fs
        self.length = ygroups.shape[1]
        self.f1, self.f2, self.binWidth = f1, f2, binWidth
        self.ygroups = ygroups.conj() if autoConj else ygroups                           # :1323-1324
        self.ygroupsEnergy = np.linalg.norm(self.ygroups, axis=1) ** 2                   # :1325
        from .spectralRoutines import CZTCached
        cz = CZTCached(self.length, f1, f2, binWidth, fs)
        self._cztFreq = np.arange(self.f1, self.f2 + self.binWidth / 2, self.binWidth)   # :1353
        self._nb = cz.k
        if self._cztFreq.size != self._nb:
            raise ValueError("operands could not be broadcast together: %d CZT bins against %d group phases"
                             % (self._nb, self._cztFreq.size))
        bins = cz.f0_cycles + np.arange(self._nb) * cz.df_cycles
        self._handles = [_GroupFreqHandle(self.ygroups[k], np.array([int(groupStarts[ygroupIdxs[k]])]),
                                          np.array([self.length]), bins, self._cztFreq / float(fs))
                         for k in range(self.numTemplates)]
        self.d_xcTemplates = None
        self.d_rxgroupNormSq = None

    def _run(self, d_rx, shifts):
        import torch
        from .filterRoutines import slidingEnergy
        s = np.asarray(shifts).astype(np.int64).reshape(-1)
        lo, hi = int(s.min()), int(s.max())
        Kc = hi - lo + 1
        contiguous = s.size == Kc and np.all(np.diff(s) == 1)
        pick = None if contiguous else torch.from_numpy(s - lo).to(d_rx.device)
        surf = torch.empty((self.numTemplates, s.size, self._nb), dtype=torch.complex64, device=d_rx.device)
        tmp = None if contiguous else torch.empty((Kc, self._nb), dtype=torch.complex64, device=d_rx.device)
        for k, h in enumerate(self._handles):
            if contiguous:
                h.run_complex(d_rx, lo, Kc, surf[k])
            else:
                h.run_complex(d_rx, lo, Kc, tmp)
                surf[k] = tmp[pick]
        en = torch.empty((self.numGroups, s.size), dtype=torch.float32, device=d_rx.device)
        for g in range(self.numGroups):
            a = lo + int(self.groupStarts[g])
            e = slidingEnergy(d_rx[a:a + Kc + self.length - 1], self.length)
            en[g] = e if contiguous else e[pick]
        self.d_xcTemplates, self.d_rxgroupNormSq = surf, en

    def xcorrGPU(self, rx, shifts, batchSz=32):
        """rx: CUDA tensor (complex64).  Fills d_xcTemplates (T, K, nbins) complex64 and d_rxgroupNormSq (G, K); returns
        the CZT frequencies as numpy (:1327-1450)."""
        if not _lib.is_device(rx):
            raise TypeError("rx must be a CUDA tensor (the reference requires a cupy array here)")
        if _lib.np_dtype_of(rx) != np.dtype(np.complex64):
            raise TypeError("rx must be complex64")
        assert shifts[-1] + self.groupStarts[-1] + self.length < rx.shape[0]             # :1337
        self._run(_lib.contiguous(rx), shifts)
        return self._cztFreq

    def _combine(self, templateIdx):
        import torch
        templateIdx = np.asarray(templateIdx)
        assert templateIdx.size == self.numGroups                                         # :1453
        tsel = np.array([np.argwhere(self.ygroupIdxs == g)[templateIdx[g]][0] for g in range(self.numGroups)], np.int32)
        gsel = np.arange(self.numGroups, dtype=np.int32)
        ynormsq = float(np.sum(self.ygroupsEnergy[tsel]))
        surf = self.d_xcTemplates
        K = surf.shape[1]
        caf = torch.empty((K, self._nb), dtype=torch.float32, device=surf.device)
        check(_lib.load().xc_groupperm_combine(_lib.ptr(surf), K, self._nb, _lib.ptr(tsel), _lib.ptr(gsel), self.numGroups,
                                               _lib.ptr(self.d_rxgroupNormSq), ynormsq, _lib.ptr(caf),
                                               _lib.current_stream(surf)))
        return caf

    def getCAF_GPU(self, templateIdx):
        """(K, nbins) float64 CUDA tensor (:1452-1483)."""
        return _as(self._combine(templateIdx), np.float64)

    def xcorr(self, rx, shifts=None, numThreads=1):
        """Host-array variant (:1485-1547): same kernels, rx staged to the device once; returns the CZT frequencies."""
        import torch
        if _lib.np_dtype_of(rx) != np.dtype(np.complex64):
            raise TypeError("rx must be complex64")
        if shifts is None:
            shifts = np.arange(rx.shape[0] - (int(self.groupStarts[-1]) + self.length) + 1)   # :1487
        else:
            assert shifts[-1] + self.groupStarts[-1] + self.length < rx.shape[0]              # :1489-1491
        d_rx = rx if _lib.is_device(rx) else torch.from_numpy(np.ascontiguousarray(rx)).cuda()
        self._run(d_rx.contiguous(), shifts)
        return self._cztFreq

    @property
    def xcTemplates(self):
        """(T, K, nbins) complex128 numpy like the reference's attribute (:1493-1500)."""
        return self.d_xcTemplates.cpu().numpy().astype(np.complex128)

    @property
    def rxgroupNormSq(self):
        return self.d_rxgroupNormSq.cpu().numpy().astype(np.float64)

    def getCAF(self, templateIdx, numThreads=4):
        """(K, nbins) float64 numpy (:1549-1585)."""
        return self._combine(templateIdx).cpu().numpy().astype(np.float64)


# ------------------------------------------------------------------------------------------------
# a8: CyGroupXcorrFFT  (CyGroupXcorrFFT.pyx:13-66, GroupXcorrFFT.cpp:3-199)
# ------------------------------------------------------------------------------------------------
class CyGroupXcorrFFT:
    """Drop-in for cython_ext.CyGroupXcorrFFT: offsets used as given, CAF float32 (K, F) only."""

    def __init__(self, ygroups, offsets, fs, fftlen=-1, autoConj=True):
        ygroups = np.ascontiguousarray(ygroups)
        offsets = np.ascontiguousarray(offsets)
        if ygroups.dtype != np.complex64 or ygroups.ndim != 2:
            raise ValueError("Buffer dtype mismatch, expected 'complex64_t' 2-D")
        if offsets.dtype != np.int32:
            raise ValueError("Buffer dtype mismatch, expected 'int32_t'")
        L = ygroups.shape[1]
        self.fftlen = L if fftlen < 0 else int(fftlen)
        if self.fftlen < L:
            raise RuntimeError("INVALID_FFTLEN")                   # GroupXcorrFFT.cpp:11-14 throws int 1
        self._handle = _GroupHandle(ygroups, offsets.astype(np.int64), self.fftlen, autoConj)
        self._handle.F = self.fftlen

    def xcorr(self, rx, shifts, NUM_THREADS=1):
        if _lib.np_dtype_of(rx) != np.dtype(np.complex64):
            raise ValueError("Buffer dtype mismatch, expected 'complex64_t'")
        rx = _lib.contiguous(rx)
        sh = _group_shifts(shifts, rx)
        _, _, caf, _ = self._handle.run(rx, sh, False, True)
        return caf


# ------------------------------------------------------------------------------------------------
# a9: TemplateCrossCorrelator  (xcorrRoutines.py:277-371)
# ------------------------------------------------------------------------------------------------
class _MtHandle:
    def __init__(self, templates, conj, energies=None, fp32=False, direct=False):
        lib = _lib.load()
        if _lib.is_torch(templates):
            templates = templates.detach().cpu().numpy()
        t = np.ascontiguousarray(templates, dtype=np.complex64)
        e = None
        if energies is not None:
            if _lib.is_torch(energies):
                energies = energies.detach().cpu().numpy()
            e = np.ascontiguousarray(energies, dtype=np.float32)
        self.T, self.L = t.shape
        self._h = C.c_void_p()
        flags = (_lib.XC_MT_CONJ if conj else 0) | (_lib.XC_MT_FP32 if fp32 else 0) | (_lib.XC_MT_DIRECT if direct else 0)
        check(lib.xc_mt_create(_lib.ptr(t), self.T, self.L, flags, _lib.ptr(e), C.byref(self._h)))

    @property
    def uses_tensor_cores(self):
        """True when this template set runs on tcgen05 (every output: max / arg max and the full (T, K) complex QF)."""
        return bool(_lib.load().xc_mt_uses_tensor_cores(self._h))

    def run(self, x, start, K, want_best, want_full, sqrt_out, want_peak=False):
        lib = _lib.load()
        where = _lib.where_of(x)
        _select_device(x)
        best = _lib.empty_like_space(x, K, np.float32) if want_best else None
        tidx = _lib.empty_like_space(x, K, np.int32) if want_best else None
        full = _lib.empty_like_space(x, (self.T, K), np.complex64) if want_full else None
        peak = xc_peak() if want_peak else None
        check(lib.xc_mt_correlate(self._h, _lib.ptr(x), x.shape[0], int(start), int(K), _lib.ptr(best), _lib.ptr(tidx),
                                  _lib.ptr(full), 1 if sqrt_out else 0, C.byref(peak) if peak is not None else None,
                                  where, _lib.current_stream(x)))
        return best, tidx, full, peak

    def __del__(self):
        try:
            if self._h:
                _lib.load().xc_mt_destroy(self._h)
                self._h = None
        except Exception:
            pass


class TemplateCrossCorrelator:
    """Drop-in for xcorrRoutines.TemplateCrossCorrelator (277-371).  `templates` is (T, L) complex64
    (numpy or device tensor); correlate(x) returns the complex QF (T, inputSize-L+1), or with
    returnMax=True the tuple (QF float32[K], templateIdx int64[K])."""

    def __init__(self, templates, inputSize):
        if templates.ndim != 2:
            raise ValueError("Templates must be a 2D array; 1 row for 1 template.")     # :295-296
        if _lib.np_dtype_of(templates) != np.dtype(np.complex64):
            raise TypeError("templates must be complex64")
        self._inputSize = int(inputSize)
        self._templateLength = templates.shape[1]
        self._handle = _MtHandle(templates, True)

    def correlate(self, x, returnMax=False):
        if _lib.np_dtype_of(x) != np.dtype(np.complex64):
            raise TypeError("x must be complex64")
        if x.ndim != 1 or x.shape[0] != self._inputSize:
            raise ValueError("x must be 1D with length %d" % self._inputSize)            # :335-336
        x = _lib.contiguous(x)
        K = self._inputSize - self._templateLength + 1
        if not returnMax:
            _, _, full, _ = self._handle.run(x, 0, K, False, True, False)
            return full
        best, tidx, _, _ = self._handle.run(x, 0, K, True, False, True)
        if _lib.is_torch(tidx):
            import torch
            return best, tidx.to(torch.int64)
        return best, tidx.astype(np.int64)                                               # cp.argmax dtype (:367)


_lib.adopt_module(globals(), ['fastXcorr', 'cztXcorr', 'fastXcorrPeak', 'cp_fastXcorr', 'cp_fastXcorr_v2', 'CyIppXcorrFFT', 'GroupXcorrFFT', 'GroupXcorr', 'GroupXcorrGPU', 'GroupXcorrCZT', 'GroupXcorrCZT_Permutations', 'pbIppGroupXcorrCZT', 'CyGroupXcorrFFT', 'TemplateCrossCorrelator', 'fineFreqTimeSearch', 'GenXcorr', 'calcQF2'])

"""Mirror of the reference's cupyExtensions.py launch wrappers that sit on the detection hot path
(cupyExtensions.py:268-319, 337-387, 563-642, 651-686), on top of libxcorr_b200.  Device arrays are
torch CUDA tensors (the reference takes cupy arrays); numpy arrays are accepted too and staged."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check
from .xcorrRoutines import _MtHandle, _select_device


def cupyRequireDtype(dtype, var):
    """cupyHelpers.py:50-69."""
    if _lib.np_dtype_of(var) != np.dtype(dtype):
        raise TypeError("Must be %s, found %s" % (np.dtype(dtype), _lib.np_dtype_of(var)))


def cupyComplexMagnSq(d_x, out_dtype=np.float64, THREADS_PER_BLOCK=128):
    """cupyExtensions.py:337-387 -> complex_magn.cu:8-20: |x|^2 in float32 or float64 (complex64 input)."""
    cupyRequireDtype(np.complex64, d_x)
    out_dtype = np.dtype(out_dtype)
    if out_dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
        raise TypeError("out_dtype must be float32 or float64")
    d_x = _lib.contiguous(d_x)
    _select_device(d_x)
    n = int(np.prod(d_x.shape))
    out = _lib.empty_like_space(d_x, tuple(d_x.shape), out_dtype)
    check(_lib.load().xc_complex_magnsq_c64(_lib.ptr(d_x), n, _lib.ptr(out), 1 if out_dtype == np.float64 else 0,
                                            _lib.where_of(d_x), _lib.current_stream(d_x)))
    return out


def cupyArgmaxAbsRows_complex64(d_x, d_argmax=None, d_max=None, returnMaxValues=False, THREADS_PER_BLOCK=128,
                                useNormSqInstead=False):
    """cupyExtensions.py:268-319 -> argmax.cu:93-152.  Ties go to the LOWEST index (numpy's rule)."""
    cupyRequireDtype(np.complex64, d_x)
    if d_x.ndim != 2:
        raise ValueError("d_x must be 2D")
    d_x = _lib.contiguous(d_x)
    _select_device(d_x)
    rows, cols = d_x.shape
    if d_argmax is None:
        d_argmax = _lib.empty_like_space(d_x, rows, np.uint32)
    elif tuple(d_argmax.shape) != (rows,):
        raise ValueError("d_argmax shape must be 1D of length %d" % rows)
    if returnMaxValues:
        if d_max is None:
            d_max = _lib.empty_like_space(d_x, rows, np.float32)
        else:
            cupyRequireDtype(np.float32, d_max)
            if tuple(d_max.shape) != (rows,):
                raise ValueError("d_max shape must be 1D of length %d" % rows)
    check(_lib.load().xc_row_argmax_c64(_lib.ptr(d_x), rows, cols, 1 if useNormSqInstead else 0, _lib.ptr(d_argmax),
                                        _lib.ptr(d_max) if returnMaxValues else None, _lib.where_of(d_x),
                                        _lib.current_stream(d_x)))
    if returnMaxValues:
        return d_argmax, d_max
    return d_argmax


def multiTemplateSlidingDotProduct(d_x, d_templates, startIdx, idxlen, d_templateEnergies=None, numSlidesPerBlk=None,
                                   THREADS_PER_BLOCK=128):
    """cupyExtensions.py:563-642 -> multiplySlices.cu:251-399.  Templates are used as given (the caller
    pre-conjugates).  Returns (d_templateIdx int32[idxlen], d_qf2 float32[idxlen])."""
    cupyRequireDtype(np.complex64, d_x)
    cupyRequireDtype(np.complex64, d_templates)
    if d_templates.ndim != 2:
        raise ValueError("Templates should be 2D; each row is an individual template.")
    T, L = d_templates.shape
    if startIdx < 0:
        raise ValueError("startIdx should be >= 0.")
    endIdx = startIdx + idxlen - 1
    if endIdx + L - 1 >= d_x.shape[0]:
        raise ValueError("final slide index (%d) should be within the bounds of d_x (%d)."
                         % (endIdx + L - 1, d_x.shape[0]))
    if d_templateEnergies is not None:
        cupyRequireDtype(np.float32, d_templateEnergies)
        if int(np.prod(d_templateEnergies.shape)) != T:
            raise ValueError("d_templateEnergies should have 1 value per template.")
    h = _mt_handle_for(d_templates, d_templateEnergies)
    best, tidx, _, _ = h.run(_lib.contiguous(d_x), startIdx, idxlen, True, False, False)
    return tidx, best


# The reference calls multiTemplateSlidingDotProduct once per batch with the same device templates; building a handle
# costs a device-to-host copy of the templates, the BF16 exactness check and the repack.  The last few handles are kept,
# keyed on the template buffer (address, shape, torch version counter) and the energies buffer.
_MT_CACHE = []
_MT_CACHE_MAX = 4


def _buffer_key(a):
    if a is None:
        return None
    if _lib.is_torch(a):
        return ("t", a.data_ptr(), tuple(a.shape), a._version, str(a.device))
    return ("n", a.__array_interface__["data"][0], a.shape, hash(a.tobytes()) if a.size <= 1 << 16 else None)


def _mt_handle_for(templates, energies):
    key = (_buffer_key(templates), _buffer_key(energies))
    cacheable = all(k is None or k[-1] is not None or k[0] == "t" for k in key)
    if cacheable:
        for k, h, _, _ in _MT_CACHE:
            if k == key:
                return h
    h = _MtHandle(templates, False, energies)
    if cacheable:
        _MT_CACHE.append((key, h, templates, energies))     # the references keep the buffers (and their addresses) alive
        del _MT_CACHE[:-_MT_CACHE_MAX]
    return h


def cupyFindLocalMaxima(x, minHeight, numOutputPerBlk=32, THREADS_PER_BLK=32, maxNumPeaks=10000):
    """cupyExtensions.py:651-686 -> peakfinding.cu:14-58.  Returns (peakIndex int32[maxNumPeaks],
    numPeaksFound int32[1]); unlike the reference's atomic append the indices come out ascending, and
    numPeaksFound may exceed maxNumPeaks (only the first maxNumPeaks are stored)."""
    cupyRequireDtype(np.float32, x)
    x = _lib.contiguous(x)
    _select_device(x)
    peakIndex = _lib.empty_like_space(x, maxNumPeaks, np.int32)
    if _lib.is_torch(peakIndex):
        peakIndex.zero_()
    else:
        peakIndex[:] = 0
    numPeaksFound = _lib.empty_like_space(x, 1, np.int32)
    check(_lib.load().xc_find_local_maxima_f32(_lib.ptr(x), int(np.prod(x.shape)), float(minHeight), _lib.ptr(peakIndex),
                                               int(maxNumPeaks), _lib.ptr(numPeaksFound), 1, _lib.where_of(x),
                                               _lib.current_stream(x)))
    return peakIndex, numPeaksFound


# ------------------------------------------------------------------------------------------------
# gather / copy helpers: compatibility functions, not used by the fused paths
# ------------------------------------------------------------------------------------------------
def cupyCopyGroups32fc(x, y, xStarts, yStarts, lengths, threads_per_blk=256):
    """cupyExtensions.py:41-83: y[yStarts[g] : yStarts[g] + lengths[g]] = x[xStarts[g] : ...] for every group,
    in place; like the reference, no dtype, bounds or overlap checks beyond the equal-length assert."""
    assert len(xStarts) == len(yStarts) and len(xStarts) == len(lengths)
    where = _lib.where_of(x, y, xStarts, yStarts, lengths)
    _select_device(x)
    if where == _lib.XC_HOST:
        xStarts, yStarts, lengths = (np.ascontiguousarray(a, dtype=np.int32) for a in (xStarts, yStarts, lengths))
    check(_lib.load().xc_copy_groups_c64(_lib.ptr(x), int(np.prod(x.shape)), _lib.ptr(y), int(np.prod(y.shape)),
                                         _lib.ptr(xStarts), _lib.ptr(yStarts), _lib.ptr(lengths), len(xStarts), where,
                                         _lib.current_stream(x)))


def cupyCopyIncrementalEqualSlicesToMatrix_32fc(d_x, startIdx, increment, rowLength, numRows, d_out=None):
    """cupyExtensions.py:161-215: d_out[r, :] = d_x[startIdx + r * increment : ... + rowLength] (0 past the end)."""
    cupyRequireDtype(np.complex64, d_x)
    d_x = _lib.contiguous(d_x)
    if d_out is None:
        d_out = _lib.empty_like_space(d_x, (numRows, rowLength), np.complex64)
    else:
        cupyRequireDtype(np.complex64, d_out)
        if tuple(d_out.shape) != (numRows, rowLength):
            raise ValueError("d_out must have the shape %d, %d" % (numRows, rowLength))
    _select_device(d_x)
    check(_lib.load().xc_copy_slices_to_matrix_c64(_lib.ptr(d_x), int(d_x.shape[0]), int(startIdx), int(increment),
                                                   int(rowLength), int(numRows), _lib.ptr(d_out),
                                                   _lib.where_of(d_x, d_out), _lib.current_stream(d_x)))
    return d_out


def multiplySlidesNormalised(d_x, d_y, startIdx, idxlen, THREADS_PER_BLOCK=128, numSlidesPerBlk=None, coefficient=None):
    """cupyExtensions.py:491-560: (idxlen, len(d_x)) complex64 matrix of d_x[t] * d_y[startIdx + r + t] divided by
    the norm of the d_y window times `coefficient` (a one-element float64 array; default the norm of d_x)."""
    cupyRequireDtype(np.complex64, d_x)
    cupyRequireDtype(np.complex64, d_y)
    if startIdx < 0 or startIdx + idxlen > d_y.shape[0]:
        raise ValueError("startIdx and idxlen should be within the bounds of d_y.")
    coef = 0.0
    if coefficient is not None:
        cupyRequireDtype(np.float64, coefficient)
        if int(np.prod(coefficient.shape)) != 1:
            raise ValueError("coefficient should be a single element 1D array.")
        coef = float(coefficient.reshape(-1)[0])
    d_x, d_y = _lib.contiguous(d_x), _lib.contiguous(d_y)
    _select_device(d_y)
    out = _lib.empty_like_space(d_y, (int(idxlen), int(d_x.shape[0])), np.complex64)
    check(_lib.load().xc_multiply_slides_normalised_c64(_lib.ptr(d_x), int(d_x.shape[0]), _lib.ptr(d_y), int(d_y.shape[0]),
                                                        int(startIdx), int(idxlen), coef, _lib.ptr(out),
                                                        _lib.where_of(d_x, d_y), _lib.current_stream(d_y)))
    return out


_lib.adopt_module(globals(), ['cupyComplexMagnSq', 'cupyArgmaxAbsRows_complex64', 'multiTemplateSlidingDotProduct', 'cupyFindLocalMaxima', 'cupyCopyGroups32fc', 'cupyCopyIncrementalEqualSlicesToMatrix_32fc', 'multiplySlidesNormalised'])

"""The input side of the detection path: SDR capture files hold raw interleaved int16 I/Q
(usrpRoutines.py:51-66 of the reference).  The reference widens them to complex64 on the CPU and then copies 8 bytes per
sample to the GPU; here the int16 pairs are copied (4 bytes per sample) and widened by a CUDA kernel
(xc_convert_i16_c64), following the reference's own benchmark finding (benchmarks/benchmark_cupyCopyAndConvert.py:15-26).
The result-sink helper produces the BLOB pair of the reference's type-1 results table (xcorrDatabase/_core.py:94-107)."""
import numpy as np

from . import _lib
from ._lib import XC_DEVICE, XC_HOST, check


def convertInt16ToComplex64(iq, device=None):
    """iq: interleaved int16 (I, Q, I, Q, ...) as a numpy array or a torch CUDA tensor.  Returns complex64: a torch CUDA
    tensor when `device` is given (or iq already lives on a device), otherwise a numpy array."""
    lib = _lib.load()
    if _lib.np_dtype_of(iq) != np.dtype(np.int16):
        raise TypeError("iq must be int16, got %s" % _lib.np_dtype_of(iq))
    iq = _lib.contiguous(iq).reshape(-1)
    if iq.shape[0] % 2:
        raise ValueError("an interleaved I/Q buffer needs an even number of int16 values")
    n = iq.shape[0] // 2
    if _lib.is_device(iq) or device is not None:
        import torch
        dev = iq.device if _lib.is_device(iq) else torch.device(device)
        check(lib.xc_set_device(dev.index or 0))
        out = torch.empty(n, dtype=torch.complex64, device=dev)
        if n == 0:
            return out
        check(lib.xc_convert_i16_c64(_lib.ptr(iq), n, _lib.ptr(out), XC_DEVICE if _lib.is_device(iq) else XC_HOST, XC_DEVICE,
                                     _lib.current_stream(out)))
        return out
    out = np.empty(n, dtype=np.complex64)
    if n == 0:
        return out
    check(lib.xc_convert_i16_c64(_lib.ptr(iq), n, _lib.ptr(out), XC_HOST, XC_HOST, None))
    return out


def simpleBinRead(filename, numSamps=-1, in_dtype=np.int16, out_dtype=np.complex64, offset=0, device=None):
    """Drop-in for usrpRoutines.simpleBinRead (51-66); numSamps counts complex samples.  `device` (e.g. "cuda:0") is the
    one addition: the samples are returned as a torch CUDA tensor, widened on the device."""
    if in_dtype == np.complex64 or in_dtype == np.complex128:
        raise TypeError("in_dtype must be a real type. You likely want float32 or float64 instead.")      # :57-60
    raw = np.fromfile(filename, dtype=in_dtype, count=numSamps * 2 if numSamps >= 0 else -1, offset=offset)
    if np.dtype(in_dtype) == np.dtype(np.int16) and np.dtype(out_dtype) == np.dtype(np.complex64):
        return convertInt16ToComplex64(raw, device)
    if np.dtype(in_dtype) == np.dtype(np.float32) and np.dtype(out_dtype) == np.dtype(np.complex64):
        data = raw.view(np.complex64)             # already the right bits, no conversion
        if device is not None:
            import torch
            return torch.from_numpy(data).to(device)
        return data
    raise NotImplementedError("only int16 and float32 capture files are part of the detection path")


def multiBinRead(filenames, numSamps, in_dtype=np.int16, out_dtype=np.complex64, offset=0, device=None):
    """usrpRoutines.multiBinRead (69-83): the files back to back."""
    parts = [simpleBinRead(f, numSamps, in_dtype, out_dtype, offset, device) for f in filenames]
    if device is not None:
        import torch
        return torch.cat(parts)
    return np.concatenate(parts)


def type1ResultBlobs(qf2, freqIdx):
    """(qf2 BLOB, freqIdx BLOB) of the reference's type-1 results table: float64 and uint32, C order
    (xcorrDatabase/_core.py:94-107)."""
    if _lib.is_torch(qf2):
        qf2 = qf2.detach().cpu().numpy()
    if _lib.is_torch(freqIdx):
        freqIdx = freqIdx.detach().cpu().numpy()
    return (np.ascontiguousarray(qf2, dtype=np.float64).tobytes(),
            np.ascontiguousarray(freqIdx).astype(np.uint32).tobytes())

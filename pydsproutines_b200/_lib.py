"""ctypes binding of libxcorr_b200.so (include/xcorr_b200.h).

Follows the reference's own ctypes-DLL convention (cpuWola.py:38-66): the library sits next to this
file, outputs are allocated by the caller, sizes travel as fixed-width ints and every call returns an
int32 status.  Status codes are mapped to the exception types the reference raises for the same
conditions (cupyHelpers.py:50-77, cupyExtensions.py:512-515, xcorrRoutines.py:48-57).

There is NO CPU fallback: if the shared library is missing, or no CUDA device is visible, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libxcorr_b200.so")

XC_HOST, XC_DEVICE = 0, 1
XC_FLAG_ZERO_OOB = 1
XC_MT_CONJ, XC_MT_FP32, XC_MT_DIRECT = 1, 2, 4
XC_NOALLOC_MAX_SHIFTS = 1 << 20
XC_OK, XC_ERR_ARG, XC_ERR_RANGE, XC_ERR_UNSUPPORTED, XC_ERR_NOMEM, XC_ERR_CUDA = 0, -1, -2, -3, -4, -5


class xc_peak(C.Structure):
    _fields_ = [("qf2", C.c_float), ("fidx", C.c_uint32), ("index", C.c_int64), ("shift", C.c_int64),
                ("aux", C.c_int32), ("rank", C.c_int32)]


_vp, _i32, _i64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float

# name -> (restype, argtypes): exactly the declarations of include/xcorr_b200.h
SIGNATURES = {
    "xc_version": (_i32, []),
    "xc_last_error": (C.c_char_p, []),
    "xc_device_count": (_i32, []),
    "xc_set_device": (_i32, [_i32]),
    "xc_pointer_device": (_i32, [_vp]),
    "xc_release_workspace": (_i32, []),
    "xc_device_info": (_i32, [C.POINTER(_i32)]),
    "xc_launch_count": (_i64, []),
    "xc_profile_enable": (_i32, [_i32]),
    "xc_profile_last_ms": (C.c_double, []),
    "xc_measure_fp32_tflops": (_i32, [C.c_double, C.POINTER(C.c_double)]),
    "xc_sliding_energy_c64": (_i32, [_vp, _i64, _i32, _vp, _i32, _vp]),
    "xc_moving_average_f32": (_i32, [_vp, _i64, _i32, _i32, _vp, _i32, _vp]),
    "xc_complex_magnsq_c64": (_i32, [_vp, _i64, _vp, _i32, _i32, _vp]),
    "xc_fastxcorr_c64": (_i32, [_vp, _i32, _vp, _i64, _i64, _i64, _i64, _vp, _i32, _i32, _vp, _vp, _vp,
                                C.POINTER(xc_peak), _i32, _vp]),
    "xc_fastxcorr_caf_c64": (_i32, [_vp, _i32, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _i32, _i32, _vp]),
    "xc_fastxcorr_supported": (_i32, [_i32]),
    "xc_fastxcorr_nofreq_noalloc_c64": (_i32, [_vp, _i32, _vp, _i64, _i64, _i64, _vp, _vp]),
    "xc_cztxcorr_c64": (_i32, [_vp, _i32, _vp, _i64, _i64, _i64, _i64, _vp, C.c_double, C.c_double, _i32, _vp, _vp, _vp, _vp,
                               _i32, _vp]),
    "xc_group_create": (_i32, [_vp, _i32, _i32, _vp, _i32, _i32, C.POINTER(_vp)]),
    "xc_group_xcorr": (_i32, [_vp, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp, C.POINTER(xc_peak), _i32, _vp]),
    "xc_group_destroy": (_i32, [_vp]),
    "xc_groupfreq_create": (_i32, [_vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, C.POINTER(_vp)]),
    "xc_groupfreq_xcorr": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, C.POINTER(xc_peak), _i32, _vp]),
    "xc_groupperm_combine": (_i32, [_vp, _i64, _i32, _vp, _vp, _i32, _vp, C.c_double, _vp, _vp]),
    "xc_groupfreq_destroy": (_i32, [_vp]),
    "xc_fine_freq_c128": (_i32, [_vp, _vp, _i64, _vp, _i64, _vp, _vp, _i32, _vp]),
    "xc_time_scan_c128": (_i32, [_vp, _vp, _i64, C.c_double, _vp, _i64, C.c_double, _vp, _vp, _vp, _vp, _i32, _vp]),
    "xc_mt_create": (_i32, [_vp, _i32, _i32, _i32, _vp, C.POINTER(_vp)]),
    "xc_mt_correlate": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _i32, C.POINTER(xc_peak), _i32, _vp]),
    "xc_mt_destroy": (_i32, [_vp]),
    "xc_mt_uses_tensor_cores": (_i32, [_vp]),
    "xc_czt_c64": (_i32, [_vp, _i32, _i64, C.c_double, C.c_double, _i32, _vp, _i32, _vp]),
    "xc_convert_i16_c64": (_i32, [_vp, _i64, _vp, _i32, _i32, _vp]),
    "xc_copy_groups_c64": (_i32, [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i32, _i32, _vp]),
    "xc_copy_slices_to_matrix_c64": (_i32, [_vp, _i64, _i64, _i64, _i64, _i64, _vp, _i32, _vp]),
    "xc_multiply_slides_normalised_c64": (_i32, [_vp, _i32, _vp, _i64, _i64, _i64, C.c_double, _vp, _i32, _vp]),
    "xc_row_argmax_c64": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _i32, _vp]),
    "xc_dist_unique_id": (_i32, [_vp]),
    "xc_dist_init": (_i32, [_i32, _i32, _vp]),
    "xc_dist_allgather_peak": (_i32, [C.POINTER(xc_peak), C.POINTER(xc_peak), C.POINTER(_i32), _vp]),
    "xc_dist_finalize": (_i32, []),
    "xc_peak_argmax": (_i32, [C.POINTER(xc_peak), _i32]),
    "xc_find_local_maxima_f32": (_i32, [_vp, _i64, _f32, _vp, _i32, _vp, _i32, _i32, _vp]),
}

_lib = None


def load():
    """Loads the shared library (once) and declares every prototype.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError("%s is missing: run `make` (or __graft_entry__.build()) first; "
                          "pydsproutines_b200 has no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)       # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class XcorrB200Error(RuntimeError):
    pass


def check(status):
    """Maps a status code to the reference's exception types."""
    if status == XC_OK:
        return
    msg = load().xc_last_error().decode("utf-8", "replace")
    if status in (XC_ERR_ARG, XC_ERR_RANGE):
        raise ValueError(msg)
    if status == XC_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    if status == XC_ERR_NOMEM:
        raise MemoryError(msg)
    raise XcorrB200Error(msg)


# ------------------------------------------------------------------------------------------------
# buffers: numpy arrays are host memory, torch CUDA tensors are device memory (the stand-in for the
# reference's cupy arrays)
# ------------------------------------------------------------------------------------------------
def is_torch(x):
    return type(x).__module__.split(".")[0] == "torch"


def is_device(x):
    return is_torch(x) and x.is_cuda


_TORCH_DTYPES = {}


def torch_dtype(np_dtype):
    import torch
    if not _TORCH_DTYPES:
        _TORCH_DTYPES.update({np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64,
                              np.dtype(np.complex64): torch.complex64, np.dtype(np.complex128): torch.complex128,
                              np.dtype(np.int32): torch.int32, np.dtype(np.int16): torch.int16,
                              np.dtype(np.int64): torch.int64, np.dtype(np.uint32): torch.int32})
    return _TORCH_DTYPES[np.dtype(np_dtype)]


def np_dtype_of(x):
    if is_torch(x):
        import torch
        return {torch.float32: np.dtype(np.float32), torch.float64: np.dtype(np.float64),
                torch.complex64: np.dtype(np.complex64), torch.complex128: np.dtype(np.complex128),
                torch.int32: np.dtype(np.int32), torch.int64: np.dtype(np.int64),
                torch.int16: np.dtype(np.int16)}.get(x.dtype)
    return np.asarray(x).dtype


def ptr(x):
    """Address of the first element (None -> NULL)."""
    if x is None:
        return None
    if is_torch(x):
        return C.c_void_p(x.data_ptr())
    return C.c_void_p(x.ctypes.data)


def where_of(*arrays):
    """XC_DEVICE if all given buffers are CUDA tensors, XC_HOST if all are host; mixing raises."""
    dev = [is_device(a) for a in arrays if a is not None]
    if all(dev) and dev:
        return XC_DEVICE
    if any(dev):
        raise TypeError("all arrays of a call must live in the same memory space (all numpy or all torch.cuda)")
    return XC_HOST


def contiguous(x):
    if is_torch(x):
        return x.contiguous()
    return np.ascontiguousarray(x)


PINNED_MIN_BYTES = 1 << 16


def host_empty(shape, np_dtype):
    """Host output buffer.  Large results come from torch's caching pinned-memory allocator (when a CUDA device
    is there), so that the library's device-to-host copies run at PCIe speed and overlap the kernels of the
    pipelined host path; small ones are plain numpy arrays."""
    shape = tuple(int(v) for v in np.atleast_1d(shape))
    nbytes = int(np.prod(shape, dtype=np.int64)) * np.dtype(np_dtype).itemsize
    if nbytes >= PINNED_MIN_BYTES:
        try:
            import torch
            if torch.cuda.is_available():
                t = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
                return t.numpy().view(np_dtype).reshape(shape)      # keeps `t` alive through .base
        except Exception:
            pass
    return np.empty(shape, dtype=np_dtype)


def empty_like_space(ref, shape, np_dtype):
    """Output buffer in the memory space of `ref`."""
    if is_device(ref):
        import torch
        return torch.empty(shape, dtype=torch_dtype(np_dtype), device=ref.device)
    return host_empty(shape, np_dtype)


# ------------------------------------------------------------------------------------------------
# foreign device arrays: anything that exposes __cuda_array_interface__ (CuPy, Numba, ...) is adopted
# zero-copy as a torch tensor on the device that owns the pointer (cupyHelpers.py:80-83 requireCupyArray is
# the reference's gate; the protocol, not a framework, is what a drop-in has to accept).  Results come back
# as torch CUDA tensors, which expose the same protocol (cupy.asarray(t) / numba.cuda.as_cuda_array(t)).
# ------------------------------------------------------------------------------------------------
def adopt(x):
    if x is None or isinstance(x, (np.ndarray, np.generic, int, float, complex, str, bool, list, tuple)) or is_torch(x):
        return x
    cai = getattr(x, "__cuda_array_interface__", None)
    if cai is None:
        return x
    import torch
    dev = load().xc_pointer_device(C.c_void_p(int(cai["data"][0]))) if cai["data"][0] else 0
    if dev < 0:
        raise TypeError("__cuda_array_interface__ object does not point to device memory")
    return torch.as_tensor(x, device=torch.device("cuda", dev))


def adopting(fn):
    """Decorator of the public entry points: adopts foreign device arrays among the arguments."""
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        return fn(*[adopt(a) for a in args], **{k: adopt(v) for k, v in kwargs.items()})
    return wrapper


def adopt_module(namespace, names):
    """Wraps the listed functions / the public methods of the listed classes of a module in `adopting`."""
    for name in names:
        obj = namespace[name]
        if isinstance(obj, type):
            for attr, val in list(vars(obj).items()):
                if callable(val) and (not attr.startswith("_") or attr == "__init__"):
                    setattr(obj, attr, adopting(val))
        else:
            namespace[name] = adopting(obj)


def current_stream(ref):
    """cudaStream_t the library enqueues on: torch's current stream for device buffers."""
    if is_device(ref):
        import torch
        return C.c_void_p(torch.cuda.current_stream(ref.device).cuda_stream)
    return None


def device_available():
    return load().xc_device_count() > 0

"""GPU tests of the runtime around the kernels (run on the B200 box: `pytest -m gpu`): the pipelined XC_HOST path, device
arrays that are not torch tensors (__cuda_array_interface__), two devices in one process and the NCCL peak exchange of
the C-ABI (the last two need two visible devices and are skipped otherwise)."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

from pydsproutines_b200 import signals

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def xr():
    import torch
    assert torch.cuda.is_available()
    import pydsproutines_b200.xcorrRoutines as m
    return m


def _dev(a, index=0):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(torch.device("cuda", index))


def test_host_pipeline_matches_device_path(xr):
    """K >= 2^17 contiguous shifts from numpy buffers take the chunked H2D | search | D2H pipeline (api.cu); the
    results must be bit-identical to the device-resident call, the peak tuple too."""
    N, K = 1024, (1 << 17) + 77
    cutout, rx = signals.make_fastxcorr_case(77, N, K + N - 1, 90001, 300)
    qh, fh, ph = xr.fastXcorrPeak(cutout, rx, freqsearch=True)
    qd, fd, pd = xr.fastXcorrPeak(_dev(cutout), _dev(rx), freqsearch=True)
    assert isinstance(qh, np.ndarray) and qh.dtype == np.float32 and fh.dtype == np.uint32
    assert np.array_equal(qh, qd.cpu().numpy())
    assert np.array_equal(fh, fd.cpu().numpy().astype(np.uint32))
    assert (ph.shift, ph.fidx) == (pd.shift, pd.fidx) == (90001, 300)
    # pageable (not pinned) inputs and the reference-shaped entry point
    q2, f2 = xr.fastXcorr(cutout.copy(), rx.copy(), freqsearch=True)
    assert q2.dtype == np.float64 and np.array_equal(q2.astype(np.float32), qh) and np.array_equal(f2, fh)


class _ForeignDeviceArray:
    """Stand-in for a CuPy / Numba device array: exposes only the CUDA array interface (v3)."""

    def __init__(self, tensor, typestr):
        self._keep = tensor
        self.__cuda_array_interface__ = {"shape": tuple(tensor.shape), "typestr": typestr,
                                         "data": (tensor.data_ptr(), False), "version": 3, "strides": None}


def test_cuda_array_interface_objects_are_device_buffers(xr):
    """cupyHelpers.py:80-83 gates on cupy arrays; the drop-in takes anything that speaks __cuda_array_interface__."""
    cutout, rx = signals.make_fastxcorr_case(5, 256, 256 + 999, 400, 17)
    tc, tr = _dev(cutout), _dev(rx)
    q0, f0 = xr.fastXcorr(tc, tr, freqsearch=True)
    q1, f1 = xr.fastXcorr(_ForeignDeviceArray(tc, "<c8"), _ForeignDeviceArray(tr, "<c8"), freqsearch=True)
    assert q1.is_cuda and np.array_equal(q1.cpu().numpy(), q0.cpu().numpy()) and np.array_equal(f1.cpu().numpy(), f0.cpu().numpy())
    assert hasattr(q1, "__cuda_array_interface__")          # what the caller wraps back into its own framework
    from pydsproutines_b200 import cupyExtensions as ce
    m = ce.cupyComplexMagnSq(_ForeignDeviceArray(tr, "<c8"), out_dtype=np.float32)
    assert np.allclose(m.cpu().numpy(), np.abs(rx) ** 2, rtol=1e-6)


def test_two_devices_in_one_process(xr):
    """Per-device workspace and per-device kernel attributes (227 KB dynamic shared memory opt-in): the same search on
    cuda:0 then cuda:1 then cuda:0 again from one process."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible devices")
    N, K = 8192, 148 * 16 + 5                    # takes the ping-pong layout (>= 16 shifts per SM)
    cutout, rx = signals.make_fastxcorr_case(9, N, K + N - 1, 1000, 4000)
    res = []
    for index in (0, 1, 0):
        q, f = xr.fastXcorr(_dev(cutout, index), _dev(rx, index), freqsearch=True)
        assert q.device.index == index
        res.append((q.cpu().numpy(), f.cpu().numpy()))
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])
    assert np.array_equal(res[0][0], res[2][0])
    assert int(np.argmax(res[1][0])) == 1000 and int(res[1][1][1000]) == 4000
    # host buffers follow the current device of the library
    from pydsproutines_b200 import _lib
    _lib.check(_lib.load().xc_set_device(1))
    qh, fh = xr.fastXcorr(cutout, rx, freqsearch=True)
    _lib.check(_lib.load().xc_set_device(0))
    assert np.array_equal(qh.astype(np.float32), res[0][0].astype(np.float32))


_DIST_CHILD = r'''
import ctypes, os, sys, time
import numpy as np
sys.path.insert(0, %r)
from pydsproutines_b200 import _lib
rank, world, idfile = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
lib = _lib.load()
_lib.check(lib.xc_set_device(rank))
uid = ctypes.create_string_buffer(128)
if rank == 0:
    _lib.check(lib.xc_dist_unique_id(uid))
    with open(idfile + ".tmp", "wb") as f:
        f.write(uid.raw)
    os.replace(idfile + ".tmp", idfile)
else:
    for _ in range(600):
        if os.path.exists(idfile):
            break
        time.sleep(0.05)
    uid = ctypes.create_string_buffer(open(idfile, "rb").read(), 128)
_lib.check(lib.xc_dist_init(rank, world, uid))
mine = _lib.xc_peak(qf2=0.25 + 0.5 * (rank == 1), fidx=10 + rank, index=5, shift=1000 * (rank + 1), aux=0, rank=-1)
allp = (_lib.xc_peak * world)()
best = ctypes.c_int32(-1)
_lib.check(lib.xc_dist_allgather_peak(ctypes.byref(mine), allp, ctypes.byref(best), None))
assert [p.rank for p in allp] == list(range(world)), [p.rank for p in allp]
assert [p.shift for p in allp] == [1000 * (r + 1) for r in range(world)]
assert best.value == 1 and allp[best.value].fidx == 11
_lib.check(lib.xc_dist_finalize())
print("rank %%d ok" %% rank)
''' % ROOT


def test_c_abi_nccl_peak_exchange(tmp_path):
    """xc_dist_unique_id / init / allgather_peak / finalize between two processes, one GPU each."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible devices")
    idfile = str(tmp_path / "nccl_id")
    procs = [subprocess.Popen([sys.executable, "-c", _DIST_CHILD, str(r), "2", idfile], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and ("rank %d ok" % r) in o, o[-2000:]

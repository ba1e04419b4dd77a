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


def test_pbIppGroupXcorrCZT_call_surface(xr):
    """The incremental add-group class of the reference's pybind module (pbGroupXcorrCZT.cpp:8-33) against golden vectors
    of the real GroupXcorrCZT (same mathematics, GroupXcorrCZT.cpp:106-330), plus its argument checks."""
    import glob
    golden = os.path.join(ROOT, "tests", "golden")
    for path in sorted(glob.glob(os.path.join(golden, "groupfreq_*.npz"))):
        d = np.load(path)
        y, starts, lengths, rx = d["y"], d["starts"], d["lengths"], d["rx"]
        K = int(d["shifts"].size)
        pb = xr.pbIppGroupXcorrCZT(int(lengths.max()), float(d["f1"]), float(d["f2"]), float(d["bw"]), float(d["fs"]), 4)
        assert pb.getNumThreads() == 4
        pb.addGroupsFromArray(starts.astype(np.int32), lengths.astype(np.int32), y)
        out = pb.xcorr(rx, 0, 1, K)
        assert out.dtype == np.float32 and out.shape == d["caf"].shape
        np.testing.assert_allclose(out, d["caf"], rtol=1e-4, atol=2e-6 * d["caf"].max())
        # groups added one by one in another order, a stepped shift list, device input
        pb.resetGroups()
        z = starts - starts.min()
        for g in reversed(range(starts.size)):
            pb.addGroup(int(z[g]), y[starts[g]:starts[g] + lengths[g]])
        out2 = pb.xcorr(_dev(rx), 1, 3, (K - 1) // 3)
        np.testing.assert_allclose(out2.cpu().numpy(), d["caf"][1:1 + 3 * ((K - 1) // 3):3], rtol=1e-4, atol=2e-6 * d["caf"].max())
    pb = xr.pbIppGroupXcorrCZT(12, -0.1, 0.1, 0.1, 100.0)
    data = np.arange(200).astype(np.float32).view(np.complex64)               # the reference's own test.py
    with pytest.raises(ValueError, match="No groups"):
        pb.xcorr(data, 0, 1, 3)
    pb.addGroup(0, data[10:20])
    pb.addGroup(60, data[70:82])
    with pytest.raises(ValueError, match="overlaps"):
        pb.addGroup(65, data[75:85])
    with pytest.raises(ValueError, match="maximum length"):
        pb.addGroup(20, data[0:13])
    with pytest.raises(ValueError, match="negative"):
        pb.xcorr(data, 9, -1, 3)
    with pytest.raises(ValueError, match="insufficient"):
        pb.xcorr(data, 9, 1, 30)
    with pytest.raises(ValueError, match="greater than 0"):
        xr.pbIppGroupXcorrCZT(12, -0.1, 0.1, 0.1, 100.0, 0)
    assert pb.xcorr(data, 9, 1, 3).shape == (3, 3)


def test_sharded_group_and_template_drivers_on_the_device(xr):
    """dist.sharded_group_peak / sharded_mt_peak with the real searches (world size 1: the allgather is the identity)."""
    from oracle import xcorr_oracle as orc
    from pydsproutines_b200 import dist as xd
    c = signals.config_c3(log2_shifts=13)
    obj = xr.GroupXcorrFFT(c["ygroups"], c["starts"], 1, fftlen=c["F"])
    window = xd.group_window(obj.starts, obj.ygroupLen)
    rx = c["rx"][:c["K"] + window - 1]
    q, f, g = xd.sharded_group_peak(obj, _dev(rx), 1000, 0)
    assert q.shape[0] == c["K"] and g["shift"] == 1000 + c["p0"] and g["fidx"] == c["k0"] and g["rank"] == 0
    rng = np.random.default_rng(12)
    templates = signals.qpsk(rng, 5 * 64).reshape(5, 64)
    x = signals.cnoise(rng, 3000)
    x[1234:1234 + 64] += 2.0 * templates[3]
    tcc = xr.TemplateCrossCorrelator(templates, x.size)
    best, tidx, g = xd.sharded_mt_peak(tcc, _dev(x), 50, 0)
    assert best.shape[0] == 3000 - 63 and g["shift"] == 50 + 1234 and g["fidx"] == 3
    ti_ref, q_ref = orc.multiTemplateSlidingDotProduct(x, templates.conj(), 0, 3000 - 63)
    np.testing.assert_allclose(best.cpu().numpy(), q_ref, rtol=1e-4, atol=1e-6 * q_ref.max())


@pytest.mark.parametrize("N", [1 << 16, 50000, 1 << 20])
def test_fine_search_long_arrays_fft(xr, N):
    """fineFreqTimeSearch above 2^14 samples transforms with the FP64 Stockham / Bluestein kernels (the exact O(N^2) DFT
    below): same time-offset pick as the oracle, cost within the single-precision noise of the reference's np.fft.fft on
    complex64 inputs, and N = 2^20 well inside 50 ms on the device."""
    import time
    import torch
    from oracle import xcorr_oracle as orc
    rng = np.random.default_rng(N % 1000)
    base = rng.standard_normal(N) + 1j * rng.standard_normal(N)
    Bf = np.fft.fft(base)
    fvec = np.fft.fftfreq(N)
    Bf[np.abs(fvec) > 0.25] = 0
    x = np.fft.ifft(Bf)
    y = np.fft.ifft(Bf * np.exp(-2j * np.pi * fvec * 0.3))
    xn = (x + 0.05 * (rng.standard_normal(N) + 1j * rng.standard_normal(N))).astype(np.complex64)
    yn = (y + 0.05 * (rng.standard_normal(N) + 1j * rng.standard_normal(N))).astype(np.complex64)
    td = np.arange(-1.0, 1.0, 0.05)
    ff, tdiff, cost = xr.fineFreqTimeSearch(xn, yn, [], 0.0, 0.0, 1.0, td)
    if N <= (1 << 16):
        ffo, tdo, costo = orc.fineFreqTimeSearch(xn, yn, [], 0.0, 0.0, 1.0, td)
        assert tdiff == tdo
        np.testing.assert_allclose(np.asarray(cost), costo, rtol=0, atol=5e-6)
    assert abs(tdiff - 0.3) < 0.051
    if N == (1 << 20):
        dx, dy = _dev(xn), _dev(yn)
        xr.fineFreqTimeSearch(dx, dy, [], 0.0, 0.0, 1.0, td)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        xr.fineFreqTimeSearch(dx, dy, [], 0.0, 0.0, 1.0, td)
        torch.cuda.synchronize()
        assert (time.perf_counter() - t0) < 0.05


def test_handles_are_bound_to_their_device(xr):
    """A handle used on another device than the one it was created on is an argument error, not a fault."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible devices")
    c = signals.config_c3(log2_shifts=8)
    obj = xr.GroupXcorrFFT(c["ygroups"], c["starts"], 1, fftlen=c["F"])          # created on the current device (0)
    with pytest.raises(ValueError, match="created on device 0"):
        obj.xcorr(_dev(c["rx"], 1), np.arange(8))
    q, f = obj.xcorr(_dev(c["rx"], 0), np.arange(8))
    assert q.shape[0] == 8

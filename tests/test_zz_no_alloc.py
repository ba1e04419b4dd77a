"""xc_fastxcorr_nofreq_noalloc_c64: the allocation-free variant of the no-frequency-search correlator (static device
scratch, include/xcorr_b200.h).  No reference counterpart; checked through the C-ABI against the existing entry
point xc_fastxcorr_c64(freqsearch = 0) on two calls in a row.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from pydsproutines_b200 import _lib, signals

pytestmark = pytest.mark.gpu

RTOL = 2e-5      # the QF^2 tolerance of tests/test_gpu_parity.py (same kernels: the outputs are in fact bit-identical)


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _existing(lib, torch, cutout, rx, shift0, K):
    q = torch.empty(K, dtype=torch.float32, device="cuda")
    st = lib.xc_fastxcorr_c64(_vp(cutout), cutout.numel(), _vp(rx), rx.numel(), shift0, 1, K, None, 0, 0, _vp(q), None,
                              None, None, _lib.XC_DEVICE, None)
    _lib.check(st)
    torch.cuda.synchronize()
    return q.cpu().numpy()


def _noalloc(lib, torch, cutout, rx, shift0, K, poison=False):
    q = torch.full((K,), -1.0, dtype=torch.float32, device="cuda") if poison else torch.empty(K, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    st = lib.xc_fastxcorr_nofreq_noalloc_c64(_vp(cutout), cutout.numel(), _vp(rx), rx.numel(), shift0, K, _vp(q), None)
    _lib.check(st)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return q.cpu().numpy(), free0 - free1


def test_noalloc_matches_existing_entry_point_twice():
    import torch
    lib = _lib.load()
    torch.cuda.set_device(0)
    _lib.check(lib.xc_set_device(0))
    # call 1: 1000-sample cutout, 200 000 shifts from shift 123; call 2: another cutout length, fewer shifts, other data --
    # the scratch still holds call 1's energies beyond call 2's range
    cases = [(11, 1000, 200000 + 1000 + 122, 50123, 123, 200000), (12, 4096, 70000 + 4095, 31000, 0, 70000)]
    grew = []
    for seed, N, rxlen, p0, shift0, K in cases:
        cutout, rx = signals.make_fastxcorr_case(seed, N, rxlen, p0)
        dc, drx = torch.from_numpy(cutout).cuda(), torch.from_numpy(rx).cuda()
        want = _existing(lib, torch, dc, drx, shift0, K)
        got, used = _noalloc(lib, torch, dc, drx, shift0, K, poison=True)
        grew.append(used)
        assert got.shape == want.shape and got.min() >= 0.0
        np.testing.assert_allclose(got, want, rtol=RTOL, atol=1e-6 * float(want.max()))
        assert int(np.argmax(got)) == p0 - shift0 == int(np.argmax(want))
    # the second call found the scratch address cached and the module loaded: device memory in use did not move
    assert grew[1] == 0, "the allocation-free call changed the free device memory by %d bytes" % grew[1]
    # same call again gives the same bits (nothing depends on what the previous call left in the scratch)
    again, _ = _noalloc(lib, torch, dc, drx, shift0, K)
    assert np.array_equal(again, got)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    lib = _lib.load()
    torch.cuda.set_device(0)
    _lib.check(lib.xc_set_device(0))
    N, K = 16, _lib.XC_NOALLOC_MAX_SHIFTS + 1
    dc = torch.ones(N, dtype=torch.complex64, device="cuda")
    drx = torch.ones(K + N - 1, dtype=torch.complex64, device="cuda")
    q = torch.zeros(K, dtype=torch.float32, device="cuda")
    st = lib.xc_fastxcorr_nofreq_noalloc_c64(_vp(dc), N, _vp(drx), drx.numel(), 0, K, _vp(q), None)
    assert st == _lib.XC_ERR_UNSUPPORTED
    with pytest.raises(NotImplementedError, match="at most 1048576 shifts"):
        _lib.check(st)
    # the maximum itself is accepted; shifts leaving rx are a range error, NULL outputs an argument error
    st = lib.xc_fastxcorr_nofreq_noalloc_c64(_vp(dc), N, _vp(drx), drx.numel(), 0, K - 1, _vp(q), None)
    _lib.check(st)
    torch.cuda.synchronize()
    np.testing.assert_allclose(q[:K - 1].cpu().numpy(), 1.0, rtol=RTOL)
    assert lib.xc_fastxcorr_nofreq_noalloc_c64(_vp(dc), N, _vp(drx), drx.numel(), 2, K - 1, _vp(q), None) == _lib.XC_ERR_RANGE
    assert lib.xc_fastxcorr_nofreq_noalloc_c64(_vp(dc), N, _vp(drx), drx.numel(), 0, 8, None, None) == _lib.XC_ERR_ARG

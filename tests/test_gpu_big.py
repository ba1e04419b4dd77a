"""GPU parity of the four-step frequency search (bigfft.cu: cutouts beyond one block FFT, any length) against golden
vectors of the REAL reference (tests/golden/bigxcorr_*.npz, bigczt_*.npz; inputs regenerated from their recipes) and
the oracle.  Tolerances as everywhere: peak shift and bin identical, QF^2 within 1e-4 relative, a differing per-shift
bin only as a float64-adjudicated near-tie."""
import glob
import os

import numpy as np
import pytest

from oracle import xcorr_oracle as orc
from oracle.recipes import big_inputs
from pydsproutines_b200 import signals

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL = 1e-4


def _load(pattern):
    files = sorted(glob.glob(os.path.join(GOLDEN, pattern)))
    assert files, "golden vectors missing: %s" % pattern
    return [(os.path.basename(f)[:-4], np.load(f)) for f in files]


@pytest.fixture(scope="module")
def xr():
    import torch
    assert torch.cuda.is_available()
    import pydsproutines_b200.xcorrRoutines as m
    return m


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _spec64(cutout, rx, s):
    w = rx[s:s + len(cutout)].astype(np.complex128) * cutout.astype(np.complex128).conj()
    return np.abs(np.fft.fft(w)) ** 2


def _check_bins(got, ref, spectrum_of):
    got, ref = np.asarray(got).astype(np.int64), np.asarray(ref).astype(np.int64)
    diff = np.nonzero(got != ref)[0]
    assert diff.size <= max(1, got.size // 200), "too many bin mismatches: %d of %d" % (diff.size, got.size)
    for i in diff:
        p = spectrum_of(int(i))
        assert abs(p[got[i]] - p[ref[i]]) <= 1e-5 * p[ref[i]], "bin mismatch at %d is not a near-tie" % i


@pytest.mark.parametrize("name,d", _load("bigxcorr_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("space", ["host", "device"])
def test_large_cutout_golden(xr, name, d, space):
    cutout, rx = big_inputs(d)
    shifts = d["shifts"] if "shifts" in d.files else None
    sh = np.arange(len(rx) - len(cutout) + 1) if shifts is None else shifts
    conv = _dev if space == "device" else (lambda a: a)
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    c, r = conv(cutout), conv(rx)
    q, f = xr.fastXcorr(c, r, freqsearch=True, shifts=shifts)
    q, f = back(q), back(f)
    assert q.dtype == np.float64 and q.shape == d["qf2_freq"].shape
    np.testing.assert_allclose(q, d["qf2_freq"], rtol=RTOL, atol=1e-6 * d["qf2_freq"].max())
    _check_bins(f, d["fidx_freq"], lambda i: _spec64(cutout, rx, int(sh[i])))
    k = int(np.argmax(q))
    assert sh[k] == int(d["p0"]) and int(f[k]) == int(d["k0"])
    z, fz = xr.fastXcorr(c, r, freqsearch=True, shifts=shifts, absResult=False)
    z, fz = back(z), back(fz)
    same = fz.astype(np.int64) == d["cplx_fidx"].astype(np.int64)
    assert same[k] and same.sum() >= same.size - max(1, same.size // 200)
    np.testing.assert_allclose(z[same], d["cplx_freq"][same], rtol=0, atol=5e-5 * np.abs(d["cplx_freq"]).max())
    if "caf" in d.files:
        caf = back(xr.fastXcorr(c, r, freqsearch=True, outputCAF=True, shifts=shifts))
        assert caf.shape == d["caf"].shape
        np.testing.assert_allclose(caf, d["caf"], rtol=RTOL, atol=2e-6 * d["caf"].max())


@pytest.mark.parametrize("name,d", _load("bigczt_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_large_chirpz_golden(xr, name, d):
    cutout, rx = big_inputs(d)
    args = (cutout, rx, float(d["f1"]), float(d["f2"]), float(d["fs"]), float(d["step"]))
    res, freqs = xr.cztXcorr(*args, shifts=d["shifts"])
    same = freqs == d["freqs"]
    for i in np.nonzero(~same)[0]:                 # a differing pick must be a near-tie of the reference's own surface
        row = d["caf"][i].astype(np.float64)
        ka, kb = int(np.argmin(np.abs(d["fsearch"] - freqs[i]))), int(np.argmin(np.abs(d["fsearch"] - d["freqs"][i])))
        assert abs(row[ka] - row[kb]) <= 1e-4 * row[kb], "frequency pick at %d is not a near-tie" % i
    assert same.sum() >= same.size - max(1, same.size // 50)
    np.testing.assert_allclose(res[same], d["res"][same], rtol=0, atol=5e-5 * np.abs(d["res"]).max())
    i = int(np.argmax(np.abs(res)))
    assert d["shifts"][i] == int(d["p0"]) and freqs[i] == d["freqs"][i]
    caf, fsearch = xr.cztXcorr(*args, outputCAF=True, shifts=d["shifts"])
    assert caf.shape == d["caf"].shape and np.array_equal(fsearch, d["fsearch"])
    np.testing.assert_allclose(caf, d["caf"], rtol=RTOL, atol=2e-6 * d["caf"].max())


@pytest.mark.parametrize("N", [1 << 16, 1 << 20, 50000, 1000003])
def test_large_cutout_vs_oracle(xr, N):
    """Sizes up to the reference's own benchmark (benchmark_xcorrs.py: a 1e6-sample cutout), a few shifts each, windows at
    both ends of rx, the peak on a high bin."""
    K = 6
    p0, k0 = 4, N - 3
    cutout, rx = signals.make_fastxcorr_case(900 + (N % 97), N, N + K - 1, p0, k0, amp=0.6)
    qr, fr = orc.fastXcorr(cutout, rx, freqsearch=True)
    q, f = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True)
    q, f = q.cpu().numpy(), f.cpu().numpy()
    np.testing.assert_allclose(q, qr, rtol=RTOL, atol=1e-6 * qr.max())
    _check_bins(f, fr, lambda i: _spec64(cutout, rx, i))
    assert int(np.argmax(q)) == p0 and int(f[p0]) == k0


def test_large_cutout_cyipp_zero_out_of_range(xr):
    """CyIppXcorrFFT semantics on the four-step path: shifts that leave rx give 0.0 / 0 (IppXcorrFFT.cpp:125-130)."""
    N = 20000
    cutout, rx = signals.make_fastxcorr_case(5, N, N + 40, 10, 777, amp=0.7)
    obj = xr.CyIppXcorrFFT(cutout)
    pp, fi = obj.xcorr(rx, -3, 50, 1)
    assert pp.dtype == np.float32 and fi.dtype == np.int32 and pp.size == 53
    assert np.all(pp[:3] == 0) and np.all(fi[:3] == 0) and np.all(pp[3 + 41:] == 0)
    qr, fr = orc.fastXcorr(cutout, rx, freqsearch=True)
    np.testing.assert_allclose(pp[3:3 + 41], qr, rtol=RTOL, atol=1e-6 * qr.max())
    assert int(np.argmax(pp)) == 13 and int(fi[13]) == 777

"""GPU parity tests (run on the B200 box: `pytest -m gpu`): the CUDA library, called through the
reference-shaped Python API over the ctypes C-ABI, against (1) the golden vectors generated from the
REAL reference (tests/golden, oracle/make_golden.py), (2) the numpy oracle on seeded inputs and
(3) the reference's own known-answer tests.

Tolerances (BASELINE.json north_star): peak time index and frequency bin identical; QF^2 within
1e-4 relative (plus an absolute floor of 1e-6 of the largest value for near-zero off-peak values).
Per-shift bins on noise-only shifts must be identical too unless the two candidate bins are a
float32 near-tie, which is adjudicated with a float64 recomputation."""
import glob
import os

import numpy as np
import pytest

from oracle import xcorr_oracle as orc
from pydsproutines_b200 import signals

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RTOL = 1e-4


def _load(pattern):
    files = sorted(glob.glob(os.path.join(GOLDEN, pattern)))
    assert files, "golden vectors missing: %s" % pattern
    return [(os.path.basename(f)[:-4], np.load(f)) for f in files]


def assert_qf2(got, ref, rtol=RTOL):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape
    atol = 1e-6 * float(np.max(np.abs(ref))) if ref.size else 0.0
    bad = np.abs(got - ref) > rtol * np.abs(ref) + atol
    assert not bad.any(), "QF^2 mismatch at %s: got %s ref %s" % (np.nonzero(bad)[0][:5], got[bad][:5], ref[bad][:5])


def assert_bins(got, ref, spectrum_of, what="bin"):
    """Bins must be equal; a differing bin is accepted only if a float64 recomputation of the
    spectrum shows the two candidates within 1e-5 relative (a float32 near-tie)."""
    got = np.asarray(got).astype(np.int64)
    ref = np.asarray(ref).astype(np.int64)
    assert got.shape == ref.shape
    diff = np.nonzero(got != ref)[0]
    assert diff.size <= max(1, got.size // 200), "too many %s mismatches: %d of %d" % (what, diff.size, got.size)
    for i in diff:
        p = spectrum_of(int(i))
        assert abs(p[got[i]] - p[ref[i]]) <= 1e-5 * p[ref[i]], "%s mismatch at %d is not a near-tie" % (what, i)


def _spec64(cutout, rx, s):
    w = rx[s:s + len(cutout)].astype(np.complex128) * cutout.astype(np.complex128).conj()
    return np.abs(np.fft.fft(w)) ** 2


def _group_spec64(y, starts, rx, F, s):
    """float64 |sum_g phase_g FFT_F(conj(y_g) rx_g)|^2 of GroupXcorrFFT at shift s (xcorrRoutines.py:1151-1184), via the fold"""
    z = np.zeros(F, np.complex128)
    L = y.shape[1]
    st = np.asarray(starts, np.int64) - int(np.min(starts)) * 0
    for g in range(y.shape[0]):
        a = int(st[g])
        m = (a + np.arange(L)) % F
        np.add.at(z, m, rx[s + a:s + a + L].astype(np.complex128) * y[g].astype(np.complex128).conj())
    return np.abs(np.fft.fft(z)) ** 2


def assert_templates(got, ref, templates, x, first_shift, conj=False):
    """Template indices must be equal; a differing index is accepted only if the float64 QF^2 of the two candidate templates
    at that shift agree to 1e-5 relative (a float32 near-tie), and for at most 0.5 % of the shifts."""
    got, ref = np.asarray(got).astype(np.int64), np.asarray(ref).astype(np.int64)
    assert got.shape == ref.shape
    diff = np.nonzero(got != ref)[0]
    assert diff.size <= max(1, got.size // 200), "too many template mismatches: %d of %d" % (diff.size, got.size)
    L = templates.shape[1]
    for i in diff:
        seg = x[first_shift + i:first_shift + i + L].astype(np.complex128)
        q = []
        for k in (got[i], ref[i]):
            tk = templates[k].astype(np.complex128)
            tk = tk.conj() if conj else tk
            q.append(abs(np.dot(tk, seg)) ** 2 / np.sum(np.abs(tk) ** 2))
        assert abs(q[0] - q[1]) <= 1e-5 * max(q), "template mismatch at %d is not a near-tie" % i


def _groupfreq_surface64(y, starts, lengths, freqs, fs, rx, s):
    """float64 |sum_g sum_t conj-free y_g[t] rx[s + z_g + t] exp(-j 2 pi f (z_g + t) / fs)|^2 over the frequency list
    (GroupXcorr.xcorr, xcorrRoutines.py:852-954; y enters conjugated by autoConj)"""
    z = np.asarray(starts, np.int64) - int(starts[0])
    acc = np.zeros(len(freqs), np.complex128)
    for g in range(len(starts)):
        n = np.arange(int(lengths[g])) + int(z[g])
        prod = rx[s + n].astype(np.complex128) * y[int(starts[g]):int(starts[g]) + int(lengths[g])].astype(np.complex128).conj()
        acc += np.exp(-2j * np.pi * np.outer(np.asarray(freqs, np.float64), n) / fs) @ prod
    return np.abs(acc) ** 2


@pytest.fixture(scope="module")
def xr():
    import torch
    assert torch.cuda.is_available()
    from pydsproutines_b200 import _lib
    assert _lib.load().xc_device_count() > 0
    import pydsproutines_b200.xcorrRoutines as m
    return m


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ---------------------------------------------------------------------------------------------
# fastXcorr vs golden vectors of the real reference
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,d", _load("fastxcorr_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("space", ["host", "device"])
def test_fastxcorr_golden(xr, name, d, space):
    cutout, rx = d["cutout"], d["rx"]
    shifts = d["shifts"] if "shifts" in d.files else None
    sh = np.arange(len(rx) - len(cutout) + 1) if shifts is None else shifts
    conv = _dev if space == "device" else (lambda a: a)
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    c, r = conv(cutout), conv(rx)
    q = back(xr.fastXcorr(c, r, shifts=shifts))
    assert q.dtype == np.float64
    assert_qf2(q, d["qf2_nofreq"])
    assert np.argmax(q) == np.argmax(d["qf2_nofreq"])
    q, f = xr.fastXcorr(c, r, freqsearch=True, shifts=shifts)        # any length: Bluestein when not a power of two
    q, f = back(q), back(f)
    assert_qf2(q, d["qf2_freq"])
    assert_bins(f, d["fidx_freq"], lambda i: _spec64(cutout, rx, int(sh[i])))
    k = int(np.argmax(q))
    assert sh[k] == int(d["p0"]) and int(f[k]) == int(d["k0"])
    if "caf" in d.files:
        z = back(xr.fastXcorr(c, r, shifts=shifts, absResult=False))
        np.testing.assert_allclose(z, d["cplx_nofreq"], rtol=0, atol=2e-5 * np.abs(d["cplx_nofreq"]).max())
        z, f2 = xr.fastXcorr(c, r, freqsearch=True, shifts=shifts, absResult=False)
        z, f2 = back(z), back(f2)
        same = f2.astype(np.int64) == d["cplx_fidx"].astype(np.int64)
        assert_bins(f2, d["cplx_fidx"], lambda i: _spec64(cutout, rx, int(sh[i])))
        np.testing.assert_allclose(z[same], d["cplx_freq"][same], rtol=0, atol=2e-5 * np.abs(d["cplx_freq"]).max())
        caf = back(xr.fastXcorr(c, r, freqsearch=True, outputCAF=True, shifts=shifts))
        assert caf.shape == d["caf"].shape
        np.testing.assert_allclose(caf, d["caf"], rtol=RTOL, atol=1e-6 * d["caf"].max())
        cafc = back(xr.fastXcorr(c, r, freqsearch=True, outputCAF=True, shifts=shifts, absResult=False))
        np.testing.assert_allclose(cafc, d["caf_cplx"], rtol=0, atol=2e-5 * np.abs(d["caf_cplx"]).max())


@pytest.mark.parametrize("N", [3, 5, 17, 31, 33, 100, 257, 480, 1000, 1023, 2049, 3000, 5000, 8191])
def test_fastxcorr_any_length_bluestein_vs_oracle(xr, N):
    """Cutout lengths that are not a power of two (the reference takes any length, xcorrRoutines.py:511-531):
    chirp-z inside the fused kernel, every padded plan from 64 to 16384 points."""
    K = 64 if N <= 1100 else 16
    p0, k0 = K // 3, (3 * N) // 8 + 1
    cutout, rx = signals.make_fastxcorr_case(300 + N, N, N + K - 1, p0, k0 % N, amp=0.9)
    qr, fr = orc.fastXcorr(cutout, rx, freqsearch=True)
    q, f = xr.fastXcorr(cutout, rx, freqsearch=True)
    assert_qf2(q, qr)
    assert_bins(f, fr, lambda i: _spec64(cutout, rx, i))
    if N >= 100:      # shorter cutouts do not stand out of the noise
        assert int(np.argmax(q)) == p0 and int(f[p0]) == k0 % N
    # CAF rows and complex QF through the same Bluestein kernel
    sub = np.arange(0, min(K, 12))
    caf_ref = orc.fastXcorr(cutout, rx, freqsearch=True, outputCAF=True, shifts=sub)
    caf = xr.fastXcorr(cutout, rx, freqsearch=True, outputCAF=True, shifts=sub)
    assert caf.shape == caf_ref.shape
    np.testing.assert_allclose(caf, caf_ref, rtol=RTOL, atol=2e-6 * caf_ref.max())
    z_ref, f_ref = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=sub, absResult=False)
    z, fz = xr.fastXcorr(cutout, rx, freqsearch=True, shifts=sub, absResult=False)
    same = np.asarray(fz).astype(np.int64) == np.asarray(f_ref).astype(np.int64)
    assert_bins(fz, f_ref, lambda i: _spec64(cutout, rx, int(sub[i])))
    np.testing.assert_allclose(np.asarray(z)[same], np.asarray(z_ref)[same], rtol=0, atol=3e-5 * np.abs(z_ref).max())
    cafc_ref = orc.fastXcorr(cutout, rx, freqsearch=True, outputCAF=True, shifts=sub, absResult=False)
    cafc = xr.fastXcorr(cutout, rx, freqsearch=True, outputCAF=True, shifts=sub, absResult=False)
    np.testing.assert_allclose(cafc, cafc_ref, rtol=0, atol=3e-5 * np.abs(cafc_ref).max())
    # device buffers and a strided shift list
    shifts = np.arange(1, K, 3)
    q2, f2 = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, shifts=shifts)
    q2, f2 = q2.cpu().numpy(), f2.cpu().numpy()
    np.testing.assert_allclose(q2, q[shifts], rtol=1e-6)
    assert np.array_equal(f2, f[shifts])


@pytest.mark.parametrize("log2n", list(range(1, 15)))
def test_fastxcorr_every_power_of_two_vs_oracle(xr, log2n):
    """Every block-FFT plan (N = 2 .. 16384) against the numpy oracle on a seeded planted case."""
    N = 1 << log2n
    K = 96 if N <= 2048 else 24
    p0, k0 = K // 3, (3 * N) // 8 + 1 if N > 2 else 1
    cutout, rx = signals.make_fastxcorr_case(100 + log2n, N, N + K - 1, p0, k0 % N, amp=0.9)
    qr, fr = orc.fastXcorr(cutout, rx, freqsearch=True)
    q, f = xr.fastXcorr(cutout, rx, freqsearch=True)
    assert_qf2(q, qr)
    assert_bins(f, fr, lambda i: _spec64(cutout, rx, i))
    assert int(np.argmax(q)) == int(np.argmax(qr))
    if N >= 64:      # below that a noise window can beat the planted copy
        assert int(np.argmax(q)) == p0 and int(f[p0]) == k0 % N
    assert_qf2(xr.fastXcorr(cutout, rx), orc.fastXcorr(cutout, rx))


def test_fastxcorr_shift_lists_and_errors(xr):
    cutout, rx = signals.make_fastxcorr_case(31, 128, 3000, 777, 5)
    ragged = np.array([5, 6, 100, 777, 778, 2000, 13, 2872], dtype=np.int64)      # not an arithmetic progression
    for fs in (False, True):
        ref = orc.fastXcorr(cutout, rx, freqsearch=fs, shifts=ragged)
        got = xr.fastXcorr(cutout, rx, freqsearch=fs, shifts=ragged)
        if fs:
            assert_qf2(got[0], ref[0])
            assert np.array_equal(got[1], ref[1])
        else:
            assert_qf2(got, ref)
    step = np.arange(2871, 0, -7)                                                  # negative step progression
    assert_qf2(xr.fastXcorr(cutout, rx, shifts=step), orc.fastXcorr(cutout, rx, shifts=step))
    q, f = xr.fastXcorr(cutout, rx, freqsearch=True, shifts=step)
    qr, fr = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=step)
    assert_qf2(q, qr)
    assert np.array_equal(f, fr)
    with pytest.raises(ValueError):
        xr.fastXcorr(cutout, rx, shifts=np.array([0, 2873]))
    with pytest.raises(TypeError):
        xr.fastXcorr(cutout.astype(np.complex128), rx)
    e = xr.fastXcorr(cutout, rx[:128], freqsearch=True)                           # a single shift
    assert e[0].shape == (1,)
    empty = xr.fastXcorr(cutout, rx, shifts=np.array([], dtype=np.int64))
    assert empty.shape == (0,)


def test_cp_fastxcorr_variants_and_cyipp(xr):
    cutout, rx = signals.make_fastxcorr_case(32, 256, 256 + 500, 123, 40)
    qr, fr = orc.fastXcorr(cutout, rx, freqsearch=True)
    q, f = xr.cp_fastXcorr(cutout, rx)
    assert q.dtype == np.float64 and f.dtype == np.uint32
    assert_qf2(q, qr)
    assert np.array_equal(f, fr)
    q, f = xr.cp_fastXcorr(_dev(cutout), _dev(rx))                                 # device in, copyToCpu
    assert isinstance(q, np.ndarray) and np.array_equal(f, fr)
    fi, q2 = xr.cp_fastXcorr_v2(_dev(cutout), _dev(rx), startIdx=100, idxlen=50, flattenCAF=True)   # (freqIdx, qf2) order
    assert q2.dtype.is_floating_point and np.array_equal(fi.cpu().numpy(), fr[100:150])
    assert_qf2(q2.cpu().numpy(), qr[100:150])
    caf = xr.cp_fastXcorr_v2(_dev(cutout), _dev(rx), startIdx=120, idxlen=8)
    np.testing.assert_allclose(caf.cpu().numpy(), orc.fastXcorr(cutout, rx, True, True, np.arange(120, 128)),
                               rtol=RTOL, atol=1e-6)
    with pytest.raises(ValueError):
        xr.cp_fastXcorr_v2(_dev(cutout), _dev(rx), startIdx=400, idxlen=200)
    with pytest.raises(Exception):
        xr.cp_fastXcorr(cutout, rx.astype(np.complex128))
    # CyIppXcorrFFT conventions: float32 / int32, out-of-range shifts are zero (IppXcorrFFT.cpp:125-130)
    pr, br = orc.cyIppXcorrFFT(cutout, rx, -6, 520, 3)
    p, b = xr.CyIppXcorrFFT(cutout).xcorr(rx, -6, 520, 3)
    assert p.dtype == np.float32 and b.dtype == np.int32 and p.size == pr.size
    assert_qf2(p, pr)
    assert np.array_equal(b, br)
    assert p[0] == 0 and p[1] == 0 and p[-1] == 0 and b[-1] == 0


# ---------------------------------------------------------------------------------------------
# GroupXcorrFFT vs golden vectors of the real reference
# ---------------------------------------------------------------------------------------------
def _group_spec64(d, s):
    y, starts, F = d["ygroups"].astype(np.complex128), d["starts"] - d["starts"][0], int(d["F"])
    L = y.shape[1]
    z = np.zeros(F, np.complex128)
    for g in range(y.shape[0]):
        seg = d["rx"][s + starts[g]: s + starts[g] + L].astype(np.complex128)
        np.add.at(z, (starts[g] + np.arange(L)) % F, y[g].conj() * seg)
    return np.abs(np.fft.fft(z)) ** 2


@pytest.mark.parametrize("name,d", _load("groupxcorr_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("space", ["host", "device"])
def test_groupxcorr_golden(xr, name, d, space):
    F = int(d["F"])
    obj = xr.GroupXcorrFFT(d["ygroups"], d["starts"], 1, fftlen=F)
    rx = _dev(d["rx"]) if space == "device" else d["rx"]
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    xc, fi = obj.xcorr(rx, d["shifts"])
    xc, fi = back(xc), back(fi)
    assert xc.dtype == np.float64
    assert_qf2(xc, d["xc"])
    assert_bins(fi, d["fi"], lambda i: _group_spec64(d, int(d["shifts"][i])))
    k = int(np.argmax(xc))
    assert d["shifts"][k] == int(d["p0"]) and int(fi[k]) == int(d["k0"])
    caf = back(obj.xcorr(rx, d["shifts"], flattenToTime=False))
    assert caf.shape == d["caf"].shape
    np.testing.assert_allclose(caf, d["caf"], rtol=RTOL, atol=1e-6 * d["caf"].max())
    if space == "host":
        np.testing.assert_allclose(obj.xcorrThreads(rx, d["shifts"]), d["caf"], rtol=RTOL, atol=1e-6 * d["caf"].max())
        # CyGroupXcorrFFT: offsets as given, float32 CAF
        cy = xr.CyGroupXcorrFFT(d["ygroups"], (d["starts"] - d["starts"][0]).astype(np.int32), 1, fftlen=F)
        out = cy.xcorr(d["rx"], d["shifts"].astype(np.int32))
        assert out.dtype == np.float32 and out.shape == d["caf"].shape
        np.testing.assert_allclose(out, d["caf"], rtol=RTOL, atol=1e-6 * d["caf"].max())
    else:
        xc2, fi2 = obj.xcorrGPU(rx, d["shifts"])
        assert np.array_equal(fi2.cpu().numpy(), fi)


def test_groupxcorr_errors(xr):
    y = signals.qpsk(np.random.default_rng(1), 64).reshape(4, 16)
    with pytest.raises(ValueError):
        xr.GroupXcorrFFT(y, np.array([0, 20, 40, 60]), 1, fftlen=8)              # fftlen < group length
    with pytest.raises(RuntimeError):
        xr.CyGroupXcorrFFT(y, np.array([0, 20, 40, 60], np.int32), 1, fftlen=8)   # INVALID_FFTLEN
    with pytest.raises(NotImplementedError):
        xr.GroupXcorrFFT(y, np.array([0, 20, 40, 60]), 1, fftlen=9000)           # beyond the Bluestein plans


@pytest.mark.parametrize("G,L,F,ragged", [(4, 16, 48, False), (5, 20, 20, True), (8, 100, 100, False),
                                          (6, 33, 1000, True), (3, 1200, 3000, False)])
def test_groupxcorr_any_fftlen_bluestein_vs_oracle(xr, G, L, F, ragged):
    """fftlen that is not a power of two (the default fftlen is the group length, xcorrRoutines.py:1065-1068):
    folded gather with chirped weights + the two-FFT Bluestein kernel."""
    rng = np.random.default_rng(1000 + F)
    y = signals.qpsk(rng, G * L).reshape(G, L)
    gap = L + (7 if ragged else 0)
    starts = np.arange(G) * gap + (rng.integers(0, 5, G) if ragged else 0)
    starts = (starts - starts[0]).astype(np.int64)
    K = 40
    span = int(starts[-1]) + F
    rx = signals.cnoise(rng, K + span + 8)
    p0, k0 = 17, (F // 3) + 1
    for g in range(G):
        n = np.arange(L)
        rx[p0 + starts[g]:p0 + starts[g] + L] += 2.0 * y[g] * np.exp(2j * np.pi * k0 * (starts[g] + n) / F).astype(np.complex64)
    shifts = np.arange(K)
    ref = orc.GroupXcorrFFT(y, starts, 1, fftlen=F)
    xc_ref, fi_ref = ref.xcorr(rx, shifts)
    obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F)
    xc, fi = obj.xcorr(rx, shifts)
    assert_qf2(xc, xc_ref)
    assert_bins(fi, fi_ref, lambda i: _group_spec64(y, starts, rx, F, int(shifts[i])), what="group bin")
    assert int(np.argmax(xc)) == p0 and int(fi[p0]) == k0
    xc2, fi2 = obj.xcorrGPU(_dev(rx), shifts)
    np.testing.assert_allclose(xc2.cpu().numpy(), xc, rtol=1e-6)
    # CAF rows (flattenToTime=False) and the CAF-only CyGroupXcorrFFT
    caf_ref = ref.xcorr(rx, shifts[:12], flattenToTime=False)
    caf = obj.xcorr(rx, shifts[:12], flattenToTime=False)
    assert caf.shape == caf_ref.shape
    np.testing.assert_allclose(caf, caf_ref, rtol=RTOL, atol=2e-6 * caf_ref.max())
    cy = xr.CyGroupXcorrFFT(y, starts.astype(np.int32), 1, fftlen=F)
    np.testing.assert_allclose(cy.xcorr(rx, shifts[:12].astype(np.int32)), caf_ref, rtol=RTOL, atol=2e-6 * caf_ref.max())


# ---------------------------------------------------------------------------------------------
# template correlators: the reference's known-answer tests (xcorrRoutines.py:2130-2241)
# ---------------------------------------------------------------------------------------------
def _qpsk_templates():
    rng = np.random.default_rng(7)
    x = signals.qpsk(rng, 100).astype(np.complex64)
    return x, np.vstack((x[20:40] * 1.234, x[40:60] * 2.345)).astype(np.complex64)


@pytest.mark.parametrize("space", ["host", "device"])
def test_template_correlator_known_answers(xr, space):
    x, templates = _qpsk_templates()
    conv = _dev if space == "device" else (lambda a: a)
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    tc = xr.TemplateCrossCorrelator(conv(templates), x.size)
    out = back(tc.correlate(conv(x)))
    assert out.shape == (2, 81) and out.dtype == np.complex64
    np.testing.assert_almost_equal(np.abs(out[0, 20]), 1.0, 5)
    np.testing.assert_almost_equal(np.abs(out[1, 40]), 1.0, 5)
    qf, tidx = tc.correlate(conv(x), returnMax=True)
    qf, tidx = back(qf), back(tidx)
    assert tidx.dtype == np.int64 and tidx[20] == 0 and tidx[40] == 1
    np.testing.assert_allclose(qf, np.abs(out).max(axis=0), rtol=1e-5)
    for i in range(2):
        np.testing.assert_almost_equal(np.abs(out[i]), orc.fastXcorr(templates[i], x) ** 0.5, 5)   # :2180-2200
        np.testing.assert_allclose(np.abs(out[i]), np.abs(orc.fastXcorr(templates[i], x, absResult=False)), atol=2e-6)
    ref = orc.TemplateCrossCorrelator(templates, x.size).correlate(x)
    np.testing.assert_allclose(out, ref, atol=3e-6)
    with pytest.raises(ValueError):
        xr.TemplateCrossCorrelator(templates[0], 100)
    with pytest.raises(ValueError):
        tc.correlate(conv(x[:50]))


def test_multitemplate_dot_vs_oracle():
    from pydsproutines_b200 import cupyExtensions as ce
    rng = np.random.default_rng(8)
    for T, L, n in ((5, 24, 300), (3, 1, 40), (17, 100, 5000), (2, 1030, 4000)):
        templates = signals.qpsk(rng, T * L).reshape(T, L)
        x = signals.cnoise(rng, n)
        x[n // 3:n // 3 + L] += templates[T - 1]
        K = n - L + 1
        ti_ref, q_ref = orc.multiTemplateSlidingDotProduct(x, templates.conj(), 0, K)
        ti, q = ce.multiTemplateSlidingDotProduct(x, templates.conj().copy(), 0, K)
        assert ti.dtype == np.int32 and q.dtype == np.float32
        assert_qf2(q, q_ref)
        if L > 1:        # with L == 1 every template gives QF^2 == 1: the index is decided by rounding
            bad = np.nonzero(ti != ti_ref)[0]
            assert bad.size <= K // 100
        if L > 1:
            assert int(np.argmax(q)) == n // 3 and ti[n // 3] == T - 1
        ti2, q2 = ce.multiTemplateSlidingDotProduct(_dev(x), _dev(templates.conj()), 7, K - 11)
        assert np.array_equal(ti2.cpu().numpy(), ti[7:K - 4]) and np.allclose(q2.cpu().numpy(), q[7:K - 4], rtol=1e-6)
    with pytest.raises(ValueError):
        ce.multiTemplateSlidingDotProduct(x, templates, -1, 10)
    with pytest.raises(ValueError):
        ce.multiTemplateSlidingDotProduct(x, templates, 0, x.size)
    with pytest.raises(ValueError):
        ce.multiTemplateSlidingDotProduct(x, templates[0], 0, 10)
    with pytest.raises(TypeError):
        ce.multiTemplateSlidingDotProduct(x.astype(np.complex128), templates, 0, 10)


# ---------------------------------------------------------------------------------------------
# primitives: sliding energy / moving average / |x|^2 / row argmax / local maxima
# ---------------------------------------------------------------------------------------------
def test_moving_average_known_answers():
    import scipy.signal as sps
    from pydsproutines_b200 import filterRoutines as fr
    x = np.random.default_rng(9).random(100).astype(np.float32)               # filterRoutines.py:1326-1335
    np.testing.assert_almost_equal(fr.cupyMovingAverage(x, 20, sumInstead=True), sps.lfilter(np.ones(20), 1, x), 5)
    np.testing.assert_almost_equal(fr.cupyMovingAverage(x, 20), sps.lfilter(np.ones(20) / 20, 1, x), 5)
    with pytest.raises(ValueError):
        fr.cupyMovingAverage(x, 20, NUM_PER_THREAD=32)                          # :1308-1313
    with pytest.raises(TypeError):
        fr.cupyMovingAverage(x.astype(np.float64), 20)
    rng = np.random.default_rng(10)
    for n, L in ((5000, 1), (5000, 33), (70000, 4096), (20000, 20000), (3, 7)):
        x = rng.random(n).astype(np.float32)
        got = fr.cupyMovingAverage(_dev(x), L, sumInstead=True).cpu().numpy()
        np.testing.assert_allclose(got, orc.movingAverage(x, L, sumInstead=True), rtol=2e-6)
    z = signals.cnoise(rng, 50000)
    for L in (1, 100, 8192, 50000):
        np.testing.assert_allclose(fr.slidingEnergy(z, L), orc.slidingEnergy(z, L), rtol=2e-6)
        ref = np.array([np.linalg.norm(z[s:s + L].astype(np.complex128)) ** 2 for s in (0, (50000 - L) // 2, 50000 - L)])
        np.testing.assert_allclose(fr.slidingEnergy(z, L)[[0, (50000 - L) // 2, 50000 - L]], ref, rtol=1e-6)
    # very long windows take the first window of a tile from cumulative chunk sums (csrc/energy.cu): several tiles, a window
    # that starts on / off a chunk boundary, the multi-block cutout norm of the no-frequency-search correlator
    z = signals.cnoise(rng, 400000)
    for L in (16384, 100001, 393217):
        got = fr.slidingEnergy(_dev(z), L).cpu().numpy()
        c = np.concatenate([[0.0], np.cumsum(np.abs(z.astype(np.complex128)) ** 2)])
        np.testing.assert_allclose(got, c[L:] - c[:-L], rtol=2e-6)
    from pydsproutines_b200 import xcorrRoutines as xr
    cut = z[1000:1000 + 70001].copy()
    q = xr.fastXcorr(_dev(cut), _dev(z[:80000])).cpu().numpy()
    assert int(np.argmax(q)) == 1000 and abs(q[1000] - 1.0) < 2e-5
    np.testing.assert_allclose(q[[0, 5000]], orc.fastXcorr(cut, z[:80000], shifts=np.array([0, 5000])), rtol=1e-4, atol=1e-9)


def test_magnsq_argmax_local_maxima():
    from pydsproutines_b200 import cupyExtensions as ce
    rng = np.random.default_rng(11)
    z = signals.cnoise(rng, 10001)
    np.testing.assert_allclose(ce.cupyComplexMagnSq(z), orc.complexMagnSq(z, np.float64), rtol=1e-12)
    m32 = ce.cupyComplexMagnSq(_dev(z), np.float32).cpu().numpy()
    assert m32.dtype == np.float32
    np.testing.assert_allclose(m32, orc.complexMagnSq(z, np.float32), rtol=3e-7)
    X = signals.cnoise(rng, 37 * 1000).reshape(37, 1000)
    X[5, 10] = X[5, 900] = 9 + 0j                                                # exact tie -> lowest index
    for ns in (False, True):
        idx_ref, val_ref = orc.argmaxAbsRows(X, useNormSqInstead=ns)
        idx, val = ce.cupyArgmaxAbsRows_complex64(X, returnMaxValues=True, useNormSqInstead=ns)
        assert idx.dtype == np.uint32 and np.array_equal(idx, idx_ref) and idx[5] == 10
        np.testing.assert_allclose(val, val_ref, rtol=1e-6)
    idx = ce.cupyArgmaxAbsRows_complex64(_dev(X))
    assert np.array_equal(idx.cpu().numpy().astype(np.uint32), idx_ref)
    with pytest.raises(TypeError):
        ce.cupyArgmaxAbsRows_complex64(X.astype(np.complex128))
    x = np.zeros(50, np.float32)                                                  # cupyExtensions.py:697-712
    x[5], x[8], x[9], x[37], x[41] = 1.0, 0.5, 0.6, 0.9, 0.7
    pk, cnt = ce.cupyFindLocalMaxima(x, 0.4)
    assert cnt.dtype == np.int32 and int(cnt[0]) == 4 and list(pk[:4]) == [5, 9, 37, 41]
    big = rng.random(300000).astype(np.float32)
    ref = orc.findLocalMaxima(big, 0.9)
    pk, cnt = ce.cupyFindLocalMaxima(_dev(big), 0.9, maxNumPeaks=ref.size + 10)
    assert int(cnt[0]) == ref.size and np.array_equal(pk.cpu().numpy()[:ref.size], ref)
    pk, cnt = ce.cupyFindLocalMaxima(big, 0.9, maxNumPeaks=100)                   # capped output
    assert int(cnt[0]) == ref.size and np.array_equal(pk, ref[:100])
    with pytest.raises(TypeError):
        ce.cupyFindLocalMaxima(big.astype(np.float64), 0.5)


# ---------------------------------------------------------------------------------------------
# tensor-core (tcgen05) multi-template kernel: tile edges, template blocks, padding, dispatch
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,L,n,start", [
    (1, 32, 700, 0),        # one template: N padded to 16
    (16, 33, 1000, 3),      # L not a multiple of the 32-sample stage, unaligned start
    (40, 250, 4000, 129),   # N = 48, several tiles, start past one tile
    (256, 64, 3000, 0),     # full-width block
    (300, 48, 1500, 5),     # two template blocks (256 + 44 padded)
    (8, 1500, 6000, 0),     # long templates: fewer pipeline stages
])
def test_mt_tensor_core_vs_oracle(xr, T, L, n, start):
    rng = np.random.default_rng(100 + T + L)
    alphabet = np.array([1, 1j, -1, -1j, 0.5 + 0.5j, -2.0, 0.25j], dtype=np.complex64)   # all exact in BF16
    templates = alphabet[rng.integers(0, alphabet.size, (T, L))]
    x = signals.cnoise(rng, n)
    K = n - L + 1 - start
    p = start + K // 2
    x[p:p + L] += 3.0 * templates[T - 1].conj()
    h = xr._MtHandle(templates, False)
    assert h.uses_tensor_cores
    best, tidx, _, _ = h.run(_dev(x), start, K, True, False, False)
    best, tidx = best.cpu().numpy(), tidx.cpu().numpy()
    ti_ref, q_ref = orc.multiTemplateSlidingDotProduct(x, templates, start, K)
    assert_qf2(best, q_ref)
    assert int(np.argmax(best)) == K // 2 and tidx[K // 2] == T - 1
    assert_templates(tidx, ti_ref, templates, x, start)
    # host-buffer call and the sqrt output agree with the device-buffer call
    best_h, tidx_h, _, _ = h.run(x, start, K, True, False, True)
    np.testing.assert_allclose(best_h, np.sqrt(best), rtol=1e-6)
    assert np.array_equal(tidx_h, tidx)


def test_mt_inexact_templates_split_hi_lo(xr):
    """Templates that are not exact in BF16 (frequency-shifted copies, scaled symbols): hi + lo split, 12 MMAs per
    K step; the FP32 kernel (XC_MT_FP32 flag of xc_mt_create) must agree, also on the full (T, K) complex output."""
    rng = np.random.default_rng(5)
    base = signals.qpsk(rng, 4 * 200).reshape(4, 200)
    tones = np.exp(2j * np.pi * np.outer(np.array([0.0, 0.013, -0.02, 0.11]), np.arange(200))).astype(np.complex64)
    templates = (base * tones * np.float32(1.234)).astype(np.complex64)
    x = signals.cnoise(rng, 3000)
    x[700:900] += 2.0 * templates[2].conj()
    K = 3000 - 200 + 1
    h = xr._MtHandle(templates, False)
    assert h.uses_tensor_cores
    best, tidx, _, _ = h.run(_dev(x), 0, K, True, False, False)
    best, tidx = best.cpu().numpy(), tidx.cpu().numpy()
    ti_ref, q_ref = orc.multiTemplateSlidingDotProduct(x, templates, 0, K)
    assert_qf2(best, q_ref)
    assert int(np.argmax(best)) == 700 and tidx[700] == 2
    assert_templates(tidx, ti_ref, templates, x, 0)
    h32 = xr._MtHandle(templates, False, fp32=True)
    assert not h32.uses_tensor_cores
    b32, t32, _, _ = h32.run(_dev(x), 0, K, True, False, False)
    b32, t32 = b32.cpu().numpy(), t32.cpu().numpy()
    np.testing.assert_allclose(best, b32, rtol=1e-4, atol=1e-9)
    for i in np.nonzero(tidx != t32)[0]:            # a differing template must be a near-tie of the float64 products
        seg = x[i:i + 200].astype(np.complex128)
        qa, qb = (abs(np.dot(templates[k].astype(np.complex128), seg)) ** 2 / np.sum(np.abs(templates[k].astype(np.float64)) ** 2)
                  for k in (tidx[i], t32[i]))
        assert abs(qa - qb) <= 1e-5 * max(qa, qb)
    # the full (T, K) complex QF: tensor cores (complex epilogue, template-major) against the FP32 kernel
    full_tc = h.run(_dev(x), 0, K, False, True, True)[2].cpu().numpy()
    full_32 = h32.run(_dev(x), 0, K, False, True, True)[2].cpu().numpy()
    assert full_tc.shape == (4, K) and full_tc.dtype == np.complex64
    np.testing.assert_allclose(full_tc, full_32, rtol=0, atol=2e-5 * np.abs(full_32).max())


@pytest.mark.parametrize("T,L", [(6, 700), (3, 4096), (20, 512)])
def test_mt_fast_correlation_vs_fp32_and_oracle(xr, T, L):
    """Templates of 512 ... 4096 samples over a long record: (best, template) come from the overlap-save path
    (csrc/mt_ovs.cuh: one transform of the x block for all templates); against the FP32 SIMT kernel (XC_MT_FP32) on every
    shift, against the oracle on windows, start offset and the sqrt output included."""
    rng = np.random.default_rng(T * 10 + L)
    templates = signals.qpsk(rng, T * L).reshape(T, L) * np.linspace(0.7, 1.3, T, dtype=np.float32)[:, None]
    templates = templates.astype(np.complex64)
    K, start = 4 * (8192 - L + 1) + 777, 13
    x = signals.cnoise(rng, start + K + L - 1 + 50)
    plants = [(5000, 1 % T), (start + K - 3000, (T - 1))]
    for pos, t in plants:
        x[pos:pos + L] += np.complex64(1.5) * templates[t].conj()
    h = xr._MtHandle(templates, False)
    best, tidx, _, _ = h.run(_dev(x), start, K, True, False, False)
    best, tidx = best.cpu().numpy(), tidx.cpu().numpy()
    for pos, t in plants:
        assert tidx[pos - start] == t and best[pos - start] > 0.5
    h32 = xr._MtHandle(templates, False, fp32=True)
    b32, t32, _, _ = h32.run(_dev(x), start, K, True, False, False)
    b32, t32 = b32.cpu().numpy(), t32.cpu().numpy()
    np.testing.assert_allclose(best, b32, rtol=3e-5, atol=1e-7 * float(b32.max()))
    assert_templates(tidx, t32, templates, x, start)
    for a in (0, plants[0][0] - start - 8, K - 16):
        ti_ref, q_ref = orc.multiTemplateSlidingDotProduct(x, templates, start + a, 16)
        assert_qf2(best[a:a + 16], q_ref)
        assert_templates(tidx[a:a + 16], ti_ref, templates, x, start + a)
    bs, ts, _, _ = h.run(_dev(x), start, K, True, False, True)
    np.testing.assert_allclose(bs.cpu().numpy(), np.sqrt(best), rtol=1e-6)
    assert np.array_equal(ts.cpu().numpy(), tidx)


# ---------------------------------------------------------------------------------------------
# a14: gather / copy helpers
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("space", ["host", "device"])
def test_copy_helpers_vs_oracle(xr, space):
    from pydsproutines_b200 import cupyExtensions as ce
    rng = np.random.default_rng(14)
    conv = _dev if space == "device" else (lambda a: a.copy())
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    x = signals.cnoise(rng, 5000)
    xs = np.array([0, 700, 4100, 33], np.int32)
    ys = np.array([10, 1000, 2500, 3500], np.int32)
    ln = np.array([300, 1, 900, 257], np.int32)
    y0 = signals.cnoise(rng, 4000)
    y_ref = y0.copy()
    orc.cupyCopyGroups32fc(x, y_ref, xs, ys, ln)
    y = conv(y0)
    ce.cupyCopyGroups32fc(conv(x), y, conv(xs), conv(ys), conv(ln))
    assert np.array_equal(back(y), y_ref)
    for start, inc, rowlen, rows in ((0, 1, 64, 10), (17, 300, 1000, 16), (4000, 7, 500, 40)):    # the last one runs off the end
        m = back(ce.cupyCopyIncrementalEqualSlicesToMatrix_32fc(conv(x), start, inc, rowlen, rows))
        assert np.array_equal(m, orc.cupyCopyIncrementalEqualSlicesToMatrix_32fc(x, start, inc, rowlen, rows))
    with pytest.raises(ValueError):
        ce.cupyCopyIncrementalEqualSlicesToMatrix_32fc(conv(x), 0, 1, 8, 4, conv(np.zeros((4, 9), np.complex64)))
    t = signals.qpsk(rng, 100)
    z = back(ce.multiplySlidesNormalised(conv(t), conv(x), 11, 300))
    np.testing.assert_allclose(z, orc.multiplySlidesNormalised(t, x, 11, 300), rtol=2e-6, atol=1e-7)
    coef = np.array([2.5], np.float64)
    z = back(ce.multiplySlidesNormalised(conv(t), conv(x), 4950, 50, coefficient=conv(coef)))             # windows past the end
    np.testing.assert_allclose(z, orc.multiplySlidesNormalised(t, x, 4950, 50, coef), rtol=2e-6, atol=1e-7)
    # a row of products summed = the complex QF of fastXcorr at that shift (how cp_fastXcorr_v2 used the matrix)
    qf = xr.fastXcorr(t, x, absResult=False, shifts=np.arange(11, 311))
    zz = back(ce.multiplySlidesNormalised(conv(t.conj().copy()), conv(x), 11, 300))
    np.testing.assert_allclose(np.abs(zz.sum(axis=1)), np.abs(qf), rtol=0, atol=3e-5)
    with pytest.raises(ValueError):
        ce.multiplySlidesNormalised(conv(t), conv(x), 4990, 20)
    with pytest.raises(TypeError):
        ce.multiplySlidesNormalised(conv(t.astype(np.complex128)) if space == "host" else conv(t).to(dtype=__import__("torch").complex128), conv(x), 0, 5)


# ---------------------------------------------------------------------------------------------
# chirp-z frequency zoom (cztXcorr) vs golden vectors of the real reference
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,d", _load("cztxcorr_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("space", ["host", "device"])
def test_cztxcorr_golden(xr, name, d, space):
    conv = _dev if space == "device" else (lambda a: a)
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    args = (conv(d["cutout"]), conv(d["rx"]), float(d["f1"]), float(d["f2"]), float(d["fs"]), float(d["step"]))
    res, freqs = xr.cztXcorr(*args, shifts=d["shifts"])
    res, freqs = back(res), back(freqs)
    assert res.dtype == np.complex64 and freqs.dtype == np.float64
    same = freqs == d["freqs"]
    for i in np.nonzero(~same)[0]:                 # a differing pick must be a near-tie of the reference's own surface
        row = d["caf"][i].astype(np.float64)
        ka, kb = int(np.argmin(np.abs(d["fsearch"] - freqs[i]))), int(np.argmin(np.abs(d["fsearch"] - d["freqs"][i])))
        assert abs(row[ka] - row[kb]) <= 1e-4 * row[kb], "frequency pick at %d is not a near-tie" % i
    assert same.sum() >= same.size - max(1, same.size // 50)
    np.testing.assert_allclose(res[same], d["res"][same], rtol=0, atol=3e-5 * np.abs(d["res"]).max())
    i = int(np.argmax(np.abs(res)))
    assert d["shifts"][i] == int(d["p0"]) and freqs[i] == d["freqs"][i]
    caf, fsearch = xr.cztXcorr(*args, outputCAF=True, shifts=d["shifts"])
    caf = back(caf)
    assert caf.dtype == np.float64 and caf.shape == d["caf"].shape and np.array_equal(fsearch, d["fsearch"])
    np.testing.assert_allclose(caf, d["caf"], rtol=RTOL, atol=2e-6 * d["caf"].max())


def test_cztxcorr_limits(xr):
    rng = np.random.default_rng(3)
    c, rx = signals.qpsk(rng, 9000), signals.cnoise(rng, 9100)
    # N + bins - 1 > 16384 takes the four-step path (tests/test_gpu_big.py pins it to the reference)
    caf, fsearch = xr.cztXcorr(c, rx, -1.0, 1.0, 100.0, cztStep=0.0002, outputCAF=True, shifts=np.arange(3))
    assert caf.shape == (3, fsearch.size) and fsearch.size == 10001 and np.isfinite(caf).all()
    with pytest.raises(TypeError):
        xr.cztXcorr(c.astype(np.complex128), rx, -1.0, 1.0, 100.0)


def test_cp_fastxcorr_v2_czt_branch(xr):
    """cp_fastXcorr_v2(cztObj=...) (xcorrRoutines.py:247-249): indices into the chirp-z grid, (freqIdx, qf2) order."""
    from pydsproutines_b200.spectralRoutines import CZTCachedGPU
    d = dict(np.load(os.path.join(GOLDEN, "cztxcorr_n1000_hz.npz")))
    cutout, rx = d["cutout"], d["rx"]
    obj = CZTCachedGPU(cutout.size, float(d["f1"]), float(d["f2"]), float(d["step"]), float(d["fs"]))
    assert obj.k == d["caf"].shape[1] and np.array_equal(obj.getFreq(), d["fsearch"])
    s0, n = int(d["shifts"][0]), int(d["shifts"].size)
    fi, q = xr.cp_fastXcorr_v2(_dev(cutout), _dev(rx), s0, n, cztObj=obj, flattenCAF=True)
    fi, q = fi.cpu().numpy(), q.cpu().numpy()
    assert q.dtype == np.float32
    assert_qf2(q, d["caf"].max(axis=1))
    am = d["caf"].argmax(axis=1)
    for i in np.nonzero(fi != am)[0]:
        assert abs(d["caf"][i, fi[i]] - d["caf"][i, am[i]]) <= 1e-4 * d["caf"][i, am[i]]
    assert (fi != am).sum() <= max(1, fi.size // 50)
    surf = xr.cp_fastXcorr_v2(_dev(cutout), _dev(rx), s0, n, cztObj=obj).cpu().numpy()
    assert surf.dtype == np.float32 and surf.shape == d["caf"].shape
    np.testing.assert_allclose(surf, d["caf"], rtol=RTOL, atol=2e-6 * d["caf"].max())
    with pytest.raises(ValueError):
        xr.cp_fastXcorr_v2(_dev(cutout[:500].copy()), _dev(rx), 0, 10, cztObj=obj)


# ---------------------------------------------------------------------------------------------
# grouped correlators over a frequency list (GroupXcorr / GroupXcorrGPU / GroupXcorrCZT): tensor-core segmented
# contraction vs golden vectors of the real reference and vs the oracle
# ---------------------------------------------------------------------------------------------
def _groupfreq_bins_ok(got_f, ref_f, surface_row, freqs):
    """frequency picks must be equal; a differing pick is accepted only as a near-tie of the float64 surface"""
    diff = np.nonzero(got_f != ref_f)[0]
    assert diff.size <= max(1, got_f.size // 50), "too many frequency mismatches: %d of %d" % (diff.size, got_f.size)
    for i in diff:
        row = surface_row(int(i))
        a = row[int(np.nonzero(freqs == got_f[i])[0][0])]
        b = row[int(np.nonzero(freqs == ref_f[i])[0][0])]
        assert abs(a - b) <= 1e-4 * b, "frequency mismatch at %d is not a near-tie" % i


@pytest.mark.parametrize("name,d", _load("groupfreq_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("space", ["host", "device"])
def test_groupfreq_golden(xr, name, d, space):
    conv = _dev if space == "device" else (lambda a: a)
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    rx = conv(d["rx"])
    fs = float(d["fs"])
    a = xr.GroupXcorr(d["y"], d["starts"], d["lengths"], d["freqs"], fs)
    xc, fp = a.xcorr(rx, d["shifts"])
    xc, fp = back(xc), back(fp)
    assert xc.dtype == np.float64 and fp.dtype == np.float64
    assert_qf2(xc, d["xc"])
    assert int(np.argmax(xc)) == int(d["p0"]) and fp[int(d["p0"])] == d["freqpeaks"][int(d["p0"])]
    o = orc.GroupXcorr(d["y"], d["starts"], d["lengths"], d["freqs"], fs)

    def row(i):
        s = int(d["shifts"][i])
        rc = np.hstack([d["rx"][s + o.starts[g]:s + o.starts[g] + o.lengths[g]] for g in range(o.starts.size)])
        return np.abs(o.freqMat @ (rc * o.yconcat).astype(np.complex128)) ** 2
    _groupfreq_bins_ok(fp, d["freqpeaks"], row, np.asarray(d["freqs"], np.float64))
    b = xr.GroupXcorrCZT(d["y"], d["starts"], d["lengths"], float(d["f1"]), float(d["f2"]), float(d["bw"]), fs)
    caf, cf = b.xcorr(rx, d["shifts"])
    caf = back(caf)
    assert caf.dtype == np.float64 and caf.shape == d["caf"].shape and np.array_equal(cf, d["cztFreq"])
    np.testing.assert_allclose(caf, d["caf"], rtol=RTOL, atol=2e-6 * d["caf"].max())


def test_groupfreq_vs_oracle_many_tiles(xr):
    """more shifts than one 128-shift tile and more frequencies than one 256-column block, strided and explicit
    shift lists, GroupXcorrGPU.xcorrKernel return types"""
    rng = np.random.default_rng(77)
    ylen, fs = 3000, 1.0
    starts = np.array([7, 500, 1333, 2100, 2600])
    lengths = np.array([100, 48, 129, 16, 250])
    y = signals.qpsk(rng, ylen)
    freqs = np.linspace(-0.01, 0.01, 300)
    K = 700
    z = starts - starts[0]
    rx = signals.cnoise(rng, K + int(z[-1]) + int(lengths[-1]) + 8)
    p0, foff = 411, freqs[123]
    for g in range(starts.size):
        n = np.arange(lengths[g]) + z[g]
        rx[p0 + z[g]:p0 + z[g] + lengths[g]] += (0.9 * y[starts[g]:starts[g] + lengths[g]] * np.exp(2j * np.pi * foff * n / fs)).astype(np.complex64)
    a = xr.GroupXcorrGPU(y, starts, lengths, freqs, fs)
    o = orc.GroupXcorr(y, starts, lengths, freqs, fs)
    for shifts in (np.arange(K), np.arange(3, K, 5), np.array([411, 2, 640, 128, 127])):
        xc, fp = a.xcorr(rx, shifts)
        rxc, rfp = o.xcorr(rx, shifts)
        assert_qf2(xc, rxc)
        _groupfreq_bins_ok(fp, rfp, lambda i: _groupfreq_surface64(y, starts, lengths, freqs, fs, rx, int(shifts[i])), freqs)
        if 411 in shifts:
            i = int(np.argmax(xc))
            assert shifts[i] == 411 and fp[i] == freqs[123]
    q, fi = a.xcorrKernel(rx, np.arange(K), numShiftsPerBlk=2)
    assert q.dtype == np.float32 and fi.dtype == np.int32 and int(fi[411]) == 123
    with pytest.raises(AssertionError):
        a.xcorr(rx, np.arange(K + 20))                       # reads past the end of rx (xcorrRoutines.py:923-925)


# ---------------------------------------------------------------------------------------------
# refinement around the peak (fineFreqTimeSearch / GenXcorr / calcQF2): FP64 kernels vs the real reference
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,d", _load("fine_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("space", ["host", "device"])
def test_fine_search_golden(xr, name, d, space):
    conv = _dev if space == "device" else (lambda a: a)
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    x, y = conv(d["x"]), conv(d["y"])
    bounds = d["bounds"] if "bounds" in d.files else None
    fs = float(d["fs"])
    ff, td, cost = xr.fineFreqTimeSearch(x, y, list(d["fineRes"]), float(d["f0"]), float(d["freqRes"]), fs, d["td"],
                                         td_scan_freqBounds=bounds)
    cost = back(cost)
    if np.isnan(d["finefreq"]):
        assert ff is None
    else:
        assert ff == float(d["finefreq"])                       # the same grid point, bit for bit
    assert td == float(d["timediff"])
    assert cost.dtype == np.complex128
    np.testing.assert_allclose(cost, d["cost"], rtol=0, atol=5e-8)
    g = xr.GenXcorr(d["td"], fs, d["x"].size)
    if bounds is not None:
        g.setTDscan_freqBounds(bounds)
    gtd, gcost = g.xcorr(x, y)
    assert gtd == float(d["gen_td"])
    np.testing.assert_allclose(back(gcost), d["gen_cost"], rtol=0, atol=5e-7)       # both reference FFTs in float32 (numpy >= 2)
    # a caller-supplied steering matrix (fineFreqTimeSearch's steeringvec argument) gives the same scan
    S = xr.makeTimeScanSteervec(d["td"], fs, d["x"].size)
    ff2, td2, cost2 = xr.fineFreqTimeSearch(x, y, list(d["fineRes"]), float(d["f0"]), float(d["freqRes"]), fs, d["td"],
                                            steeringvec=conv(S), td_scan_freqBounds=bounds)
    assert td2 == td and (ff2 == ff or (ff is None and ff2 is None))
    np.testing.assert_allclose(back(cost2), d["cost"], rtol=0, atol=5e-8)
    q = xr.calcQF2(x, y)
    assert abs(float(q) - float(d["qf2"])) <= RTOL * float(d["qf2"])
    q2 = back(xr.calcQF2(conv(np.stack([d["x"], d["y"]])), conv(np.stack([d["y"], d["y"]]))))
    assert abs(q2[0] - float(d["qf2"])) <= RTOL * float(d["qf2"]) and abs(q2[1] - 1.0) <= RTOL


def test_snr_helpers(xr):
    """xcorrRoutines.py:723-772 (closed-form helpers that callers apply to the QF^2 trace)."""
    assert xr.convertQF2toSNR(0.5) == 1.0 and xr.convertQF2toEffSNR(0.5) == 2.0
    assert abs(xr.convertEffSNRtoQF2(xr.convertQF2toEffSNR(0.3)) - 0.3) < 1e-15
    assert xr.expectedEffSNR(10.0, np.inf, 2) == 10.0                    # the docstring example of the reference (:744-750)
    assert abs(xr.sigmaDFO(1e4, 0.1, 10.0) - 0.55 / 0.1 / np.sqrt(1e4 * 0.1 * 10.0)) < 1e-15
    assert abs(xr.sigmaDTO(5e3, 1e4, 0.1, 10.0) - 1.0 / (np.pi / np.sqrt(3) * 5e3) / np.sqrt(1e4 * 0.1 * 10.0)) < 1e-18


# ---------------------------------------------------------------------------------------------
# capture ingestion: raw int16 I/Q -> complex64 on the device (usrpRoutines.simpleBinRead, bit-exact)
# ---------------------------------------------------------------------------------------------
def test_int16_ingestion(xr, tmp_path):
    import torch
    from pydsproutines_b200 import usrpRoutines as ur
    rng = np.random.default_rng(5)
    for n in (0, 1, 3, 4, 1001, 1 << 16):
        iq = rng.integers(-32768, 32768, size=2 * n, dtype=np.int16)
        ref = iq.astype(np.float32).view(np.complex64)            # usrpRoutines.py:61-64
        assert np.array_equal(ur.convertInt16ToComplex64(iq), ref)
        d = ur.convertInt16ToComplex64(iq, device="cuda:0")
        assert d.is_cuda and d.dtype == torch.complex64 and np.array_equal(d.cpu().numpy(), ref)
        if n:
            d2 = ur.convertInt16ToComplex64(torch.from_numpy(iq).cuda())
            assert np.array_equal(d2.cpu().numpy(), ref)
    # file reader: sample count, byte offset, device placement, straight into the correlator
    iq = rng.integers(-2000, 2000, size=2 * 5000, dtype=np.int16)
    path = os.path.join(str(tmp_path), "cap.bin")
    iq.tofile(path)
    ref = iq.astype(np.float32).view(np.complex64)
    assert np.array_equal(ur.simpleBinRead(path), ref)
    assert np.array_equal(ur.simpleBinRead(path, numSamps=100, offset=4 * 7), ref[7:107])
    rx = ur.simpleBinRead(path, device="cuda:0")
    cut = ref[1234:1234 + 256].copy()
    q = xr.fastXcorr(_dev(cut), rx).cpu().numpy()
    assert int(np.argmax(q)) == 1234 and abs(q[1234] - 1.0) < 1e-4
    both = ur.multiBinRead([path, path], 50)
    assert np.array_equal(both, np.concatenate([ref[:50], ref[:50]]))
    with pytest.raises(TypeError):
        ur.simpleBinRead(path, in_dtype=np.complex64)
    qb, fb = ur.type1ResultBlobs(np.array([0.25, 0.5], np.float32), np.array([3, 7], np.int32))
    assert np.array_equal(np.frombuffer(qb, np.float64), [0.25, 0.5]) and np.array_equal(np.frombuffer(fb, np.uint32), [3, 7])


@pytest.mark.parametrize("name,d", _load("groupperm_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_groupperm_golden(xr, name, d):
    """GroupXcorrCZT_Permutations (xcorrRoutines.py:1264-1690): per-template complex surfaces from the tensor-core
    contraction, permutations combined on the device; host (xcorr / getCAF) and device (xcorrGPU / getCAF_GPU) methods."""
    obj = xr.GroupXcorrCZT_Permutations(d["ygroups"], d["idxs"], d["starts"], float(d["f1"]), float(d["f2"]), float(d["bw"]),
                                        float(d["fs"]))
    cf = obj.xcorr(d["rx"], d["shifts"])
    assert np.array_equal(cf, d["cztFreq"])
    scale = np.abs(d["xcTemplates"]).max()
    assert obj.xcTemplates.dtype == np.complex128
    np.testing.assert_allclose(obj.xcTemplates, d["xcTemplates"], rtol=0, atol=3e-5 * scale)
    np.testing.assert_allclose(obj.rxgroupNormSq, d["rxgroupNormSq"], rtol=1e-5)
    for key, perm in (("caf_planted", tuple(d["planted"])), ("caf_other", (0, 0, 0))):
        caf = obj.getCAF(np.array(perm))
        assert caf.dtype == np.float64 and caf.shape == d[key].shape
        np.testing.assert_allclose(caf, d[key], rtol=RTOL, atol=2e-6 * d[key].max())
    i, b = np.unravel_index(int(np.argmax(caf := obj.getCAF(d["planted"]))), caf.shape)
    assert i == int(d["p0"]) and cf[b] == float(d["foff"])
    # device methods, strided shifts
    sh = d["shifts"][1::3]
    cf2 = obj.xcorrGPU(_dev(d["rx"]), sh)
    assert np.array_equal(cf2, d["cztFreq"])
    g = obj.getCAF_GPU(d["planted"])
    assert g.is_cuda and g.cpu().numpy().dtype == np.float64
    np.testing.assert_allclose(g.cpu().numpy(), d["caf_planted"][1::3], rtol=RTOL, atol=2e-6 * d["caf_planted"].max())
    with pytest.raises(TypeError):
        obj.xcorrGPU(d["rx"], sh)


def test_groupfreq_and_fine_limits(xr):
    """error behaviour of the 8(f) rows: dtype checks, empty shift lists, unsupported sizes fail loudly"""
    rng = np.random.default_rng(3)
    y = signals.qpsk(rng, 300)
    starts, lengths = np.array([0, 150]), np.array([64, 100])
    rx = signals.cnoise(rng, 600)
    a = xr.GroupXcorr(y, starts, lengths, np.linspace(-0.01, 0.01, 9), 1.0)
    with pytest.raises(TypeError):
        a.xcorr(rx.astype(np.complex128))
    with pytest.raises(IndexError):                 # the reference indexes shifts[-1] (xcorrRoutines.py:923)
        a.xcorr(rx, np.array([], dtype=np.int64))
    one, f1 = a.xcorr(rx, np.array([17]))
    ref, rf = orc.GroupXcorr(y, starts, lengths, np.linspace(-0.01, 0.01, 9), 1.0).xcorr(rx, np.array([17]))
    assert_qf2(one, ref) and f1[0] == rf[0]
    with pytest.raises(NotImplementedError):
        xr.GroupXcorr(y, starts, lengths, np.linspace(-0.1, 0.1, 3000), 1.0)          # > 2048 frequencies
    with pytest.raises(AssertionError):
        xr.fineFreqTimeSearch(rx[:100], rx[:99], [], 0.0, 0.0, 1.0, np.arange(-1, 1, 0.1))
    # a one-sample-long "signal" and a single time offset still work
    ff, td, cost = xr.fineFreqTimeSearch(rx[:1], rx[:1], [], 0.0, 0.0, 1.0, np.array([0.0]))
    assert ff is None and td == 0.0 and abs(cost[0] - 1.0) < 1e-12


@pytest.mark.parametrize("name,d", _load("cztrun_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
@pytest.mark.parametrize("space", ["host", "device"])
def test_czt_run_golden(xr, name, d, space):
    """stand-alone CZTCached.run / runMany (spectralRoutines.py:239-306) through the fused correlation kernel"""
    from pydsproutines_b200.spectralRoutines import CZTCached
    conv = _dev if space == "device" else (lambda a: a)
    back = (lambda t: t.cpu().numpy()) if space == "device" else (lambda a: a)
    x = d["x"]
    c = CZTCached(x.shape[1], float(d["f1"]), float(d["f2"]), float(d["bw"]), float(d["fs"]))
    assert np.array_equal(c.getFreq(), d["freq"])
    many = back(c.runMany(conv(x)))
    assert many.shape == d["many"].shape
    if space == "host":
        assert many.dtype == np.complex128
    # float32 transform of m samples: error relative to the row's spectrum scale
    scale = np.abs(d["many"]).max()
    np.testing.assert_allclose(many, d["many"], rtol=0, atol=2e-5 * scale)
    assert np.all(many[-1] == 0)
    one = back(c.run(conv(x[0])))
    np.testing.assert_allclose(one, d["one"], rtol=0, atol=2e-5 * scale)
    with pytest.raises(TypeError):
        c.runMany(x.astype(np.complex128))


@pytest.mark.parametrize("N", [64, 128, 256, 512, 1024, 2048, 4096])
def test_pingpong_layout_other_lengths(xr, N):
    """The ping-pong layout at 1024 points (sixteen warps per SM, one FFT per warp: csrc/fused_fft_warp.cuh), 2048 points
    (eight groups of 64 threads per SM) and 4096 points (four groups of 128): one contiguous call against calls of < 16 shifts per SM (B-groups-per-CTA layout) and against the oracle."""
    K = 40000
    cutout, rx = signals.make_fastxcorr_case(31 + N, N, N + K - 1, 23456, N - 5, amp=0.7)
    q, f = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True)
    q, f = q.cpu().numpy(), f.cpu().numpy()
    assert int(np.argmax(q)) == 23456 and int(f[23456]) == N - 5
    qs, fs = [], []
    for lo in range(0, K, 2000):
        qq, ff = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, shifts=np.arange(lo, min(lo + 2000, K)))
        qs.append(qq.cpu().numpy())
        fs.append(ff.cpu().numpy())
    q2, f2 = np.concatenate(qs), np.concatenate(fs)
    np.testing.assert_allclose(q, q2, rtol=2e-6, atol=1e-9)
    assert_bins(f, f2, lambda i: _spec64(cutout, rx, i))
    idx = np.r_[np.arange(0, 64), np.arange(23456 - 32, 23456 + 32), np.arange(K - 64, K)]
    rq, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=idx)
    assert_qf2(q[idx], rq)
    assert_bins(f[idx], rf, lambda i: _spec64(cutout, rx, int(idx[i])))


def test_pingpong_layout_matches_b2_layout(xr):
    """Contiguous N = 8192 searches of >= 16 shifts per SM run the ping-pong layout (csrc/fused_fft_pp.cuh: two thread
    groups per SM over one staged rx window, constants in tensor memory, exact twiddles folded into the butterflies);
    shorter calls and stepped shift lists run the B = 2 layout (csrc/fused_fft_kernel.cuh).  The same record through both:
    QF^2 within 2e-6 relative (different twiddle rounding), bins identical unless a float64 near-tie; windows that leave
    rx on both sides (CyIppXcorrFFT's zero rule, IppXcorrFFT.cpp:125-130), several window refills per CTA."""
    cutout, rx = signals.make_fastxcorr_case(21, 8192, 8192 + 300000, 123456, 4321, amp=0.7)
    obj = xr.CyIppXcorrFFT(cutout)
    for a, b in ((-50, 300101), (100, 2601 + 100)):
        q, f = obj.xcorr(rx, a, b, 1)                       # one call: ping-pong layout
        qs, fs = [], []
        for lo in range(a, b, 1500):                         # calls of < 16 shifts per SM: B = 2 layout
            qq, ff = obj.xcorr(rx, lo, min(lo + 1500, b), 1)
            qs.append(qq)
            fs.append(ff)
        q2, f2 = np.concatenate(qs), np.concatenate(fs)
        assert q.shape == q2.shape
        np.testing.assert_allclose(q, q2, rtol=2e-6, atol=1e-9)
        for i in np.nonzero(f != f2)[0]:
            p = _spec64(cutout, rx, a + int(i))
            assert abs(p[f[i]] - p[f2[i]]) <= 1e-5 * p[f2[i]]
        assert (f != f2).sum() <= max(1, f.size // 200)
    # and against the oracle: zeros outside, peak where planted
    q, f = xr.CyIppXcorrFFT(cutout).xcorr(rx, -50, 300101, 1)
    assert q.dtype == np.float32 and f.dtype == np.int32 and q.size == 300151
    assert not q[:50].any() and not f[:50].any() and not q[-100:].any() and q[50] > 0 and q[-101] > 0
    i = int(np.argmax(q))
    assert i == 123456 + 50 and f[i] == 4321
    idx = np.r_[np.arange(0, 40), np.arange(123456 - 20, 123456 + 21), np.arange(299960, 300001)]
    rq, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=idx)
    assert_qf2(q[idx + 50], rq)
    assert_bins(f[idx + 50], rf, lambda i: _spec64(cutout, rx, int(idx[i])))
    assert f[123456 + 50] == rf[60]


@pytest.mark.parametrize("N,step", [(64, 16), (128, 2), (256, 16), (512, 7), (1024, 3), (2048, 2), (4096, 16), (8192, 5)])
def test_staged_layouts_stepped_ranges(xr, N, step):
    """Shift ranges with a small step (CyIppXcorrFFT.xcorr(rx, start, end, step), fastXcorr(shifts=arange(a, b, step))) run
    the staged-window layouts too: against the contiguous search at the same shifts and against the oracle."""
    K = 60000
    cutout, rx = signals.make_fastxcorr_case(50 + N + step, N, N + K - 1, 30000 - 30000 % step + 0, N // 3, amp=0.7)
    p0 = 30000 - 30000 % step
    q1, f1 = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True)
    q1, f1 = q1.cpu().numpy(), f1.cpu().numpy()
    sh = np.arange(7, K, step)
    q, f = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, shifts=sh)
    q, f = q.cpu().numpy(), f.cpu().numpy()
    np.testing.assert_allclose(q, q1[sh], rtol=2e-6, atol=1e-9)
    assert_bins(f, f1[sh], lambda i: _spec64(cutout, rx, int(sh[i])))
    obj = xr.CyIppXcorrFFT(cutout)
    a, b = -3 * step, K + 2 * step                              # windows leaving rx on both sides
    qc, fc = obj.xcorr(rx, a, b, step)
    shc = np.arange(a, b, step)
    inside = (shc >= 0) & (shc + N <= rx.size)
    assert not qc[~inside].any() and not fc[~inside].any()
    np.testing.assert_allclose(qc[inside], q1[shc[inside]], rtol=2e-6, atol=1e-9)
    j = int(np.argmax(qc))
    assert shc[j] == p0 and fc[j] == N // 3
    idx = np.r_[0:16, sh.size - 16:sh.size]
    rq, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=sh[idx])
    assert_qf2(q[idx], rq)
    assert_bins(f[idx], rf, lambda i: _spec64(cutout, rx, int(sh[idx[i]])))


@pytest.mark.parametrize("N,step", [(64, 3), (128, 1), (256, 1), (512, 5), (1024, 1), (2048, 1), (4096, 3), (8192, 1)])
def test_staged_layouts_complex_qf(xr, N, step):
    """absResult=False (complex QF at the arg max, xcorrRoutines.py:533-548) from the staged-window layouts: against short
    calls (B-groups-per-CTA layout) and the oracle; the squared magnitude equals the QF^2 of the same call."""
    K = 20000
    cutout, rx = signals.make_fastxcorr_case(90 + N, N, N + K - 1, 12345 - 12345 % step, 17, amp=0.6)
    sh = np.arange(0, K, step)
    z, f = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, absResult=False, shifts=sh)
    z, f = z.cpu().numpy(), f.cpu().numpy()
    q, f1 = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, shifts=sh)
    np.testing.assert_allclose(np.abs(z) ** 2, q.cpu().numpy(), rtol=3e-6, atol=1e-10)
    assert np.array_equal(f, f1.cpu().numpy())
    zs, fs = [], []
    for lo in range(0, sh.size, 2000):
        zz, ff = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, absResult=False, shifts=sh[lo:lo + 2000])
        zs.append(zz.cpu().numpy())
        fs.append(ff.cpu().numpy())
    z2, f2 = np.concatenate(zs), np.concatenate(fs)
    assert_bins(f, f2, lambda i: _spec64(cutout, rx, int(sh[i])))
    same = f == f2
    np.testing.assert_allclose(z[same], z2[same], rtol=0, atol=3e-6 * float(np.abs(z2).max()))
    idx = np.r_[0:12, sh.size - 12:sh.size, (12345 - 12345 % step) // step]
    rz, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=sh[idx], absResult=False)
    ok = f[idx] == rf
    assert ok.sum() >= idx.size - 1
    np.testing.assert_allclose(z[idx][ok], rz[ok], rtol=0, atol=1e-5 * float(np.abs(rz).max()))


@pytest.mark.parametrize("N,step", [(1000, 1), (777, 4), (1500, 1), (2049, 2), (3000, 1), (4095, 3)])
def test_pingpong_bluestein_lengths(xr, N, step):
    """Cutout lengths that are not a power of two over shift ranges: Bluestein in the ping-pong layout
    (csrc/fused_blue_pp.cuh; padded lengths 2048 / 4096 / 8192) against short calls (csrc/bluestein.cuh) and the oracle;
    windows that leave rx through CyIppXcorrFFT's zero rule."""
    K = 9000
    p0 = 4321 - 4321 % step
    cutout, rx = signals.make_fastxcorr_case(7 * N + step, N, N + K - 1, p0, N // 5, amp=0.7)
    sh = np.arange(0, K, step)
    q, f = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, shifts=sh)
    q, f = q.cpu().numpy(), f.cpu().numpy()
    assert int(np.argmax(q)) == p0 // step and int(f[p0 // step]) == N // 5
    qs, fs = [], []
    for lo in range(0, sh.size, 1500):
        qq, ff = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, shifts=sh[lo:lo + 1500])
        qs.append(qq.cpu().numpy())
        fs.append(ff.cpu().numpy())
    q2, f2 = np.concatenate(qs), np.concatenate(fs)
    np.testing.assert_allclose(q, q2, rtol=5e-6, atol=1e-9)
    assert_bins(f, f2, lambda i: _spec64(cutout, rx, int(sh[i])))
    idx = np.r_[0:10, sh.size - 10:sh.size, p0 // step]
    rq, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=sh[idx])
    assert_qf2(q[idx], rq)
    assert_bins(f[idx], rf, lambda i: _spec64(cutout, rx, int(sh[idx[i]])))
    a, b = -5 * step, K + 3 * step
    qc, fc = xr.CyIppXcorrFFT(cutout).xcorr(rx, a, b, step)
    shc = np.arange(a, b, step)
    inside = (shc >= 0) & (shc + N <= rx.size)
    assert not qc[~inside].any() and not fc[~inside].any()
    np.testing.assert_allclose(qc[inside], q[shc[inside] // step], rtol=5e-6, atol=1e-9)


@pytest.mark.parametrize("N,nb", [(1500, 301), (3000, 1001), (700, 64)])
def test_pingpong_bluestein_complex_outputs(xr, N, nb):
    """The complex QF of the Bluestein searches from the ping-pong layout: cztXcorr's default output (complex QF at the best
    bin of a chirp-z grid, xcorrRoutines.py:413-457) and fastXcorr(absResult=False) at a length that is not a power of two,
    over shift ranges long enough for that layout, against short calls (csrc/bluestein.cuh) and the oracle."""
    K = 6000
    df = 2.0
    fs = 4.0 * N * df                         # grid step = a quarter of the resolution fs / N: adjacent bins differ clearly
    cutout, rx = signals.make_fastxcorr_case(3 * N + nb, N, N + K - 1, 2500, 0, amp=0.8)
    f1, f2 = -df * (nb // 2), df * (nb // 2) - (0 if nb % 2 else df)
    rx[2500:2500 + N] *= np.exp(2j * np.pi * (f1 + 7 * df) * np.arange(N) / fs).astype(np.complex64)      # a tone on grid point 7
    z, fr = xr.cztXcorr(_dev(cutout), _dev(rx), f1, f2, fs, cztStep=df)
    z = z.cpu().numpy()
    fr = fr.cpu().numpy() if hasattr(fr, "cpu") else np.asarray(fr)
    assert int(np.argmax(np.abs(z))) == 2500 and abs(fr[2500] - (f1 + 7 * df)) < 1e-9
    zs, fs_ = [], []
    for lo in range(0, K, 1500):
        zz, ff = xr.cztXcorr(_dev(cutout), _dev(rx), f1, f2, fs, cztStep=df, shifts=np.arange(lo, lo + 1500))
        zs.append(zz.cpu().numpy())
        fs_.append(ff.cpu().numpy() if hasattr(ff, "cpu") else np.asarray(ff))
    z2, f2_ = np.concatenate(zs), np.concatenate(fs_)
    same = fr == f2_
    assert same.sum() >= K - K // 200
    np.testing.assert_allclose(z[same], z2[same], rtol=0, atol=5e-6 * float(np.abs(z2).max()))
    idx = np.r_[0:6, 2497:2503, K - 6:K]
    rz, rf = orc.cztXcorr(cutout, rx, f1, f2, fs, cztStep=df, shifts=idx)
    ok = fr[idx] == rf
    assert ok.sum() >= idx.size - 1 and ok[8]
    np.testing.assert_allclose(z[idx][ok], rz[ok], rtol=0, atol=2e-5 * float(np.abs(rz).max()))
    # fastXcorr(absResult=False) at the same length (DFT grid)
    zc, fc = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True, absResult=False)
    qc, fq = xr.fastXcorr(_dev(cutout), _dev(rx), freqsearch=True)
    np.testing.assert_allclose(np.abs(zc.cpu().numpy()) ** 2, qc.cpu().numpy(), rtol=5e-6, atol=1e-10)
    assert np.array_equal(fc.cpu().numpy(), fq.cpu().numpy())
    rz, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=idx, absResult=False)
    okb = fc.cpu().numpy()[idx] == rf
    assert okb.sum() >= idx.size - 1
    np.testing.assert_allclose(zc.cpu().numpy()[idx][okb], rz[okb], rtol=0, atol=2e-5 * float(np.abs(rz).max()))


@pytest.mark.parametrize("N", [1024, 1500, 4096, 6000, 8192])
def test_nofreq_overlap_save_vs_direct_and_oracle(xr, N):
    """fastXcorr without frequency search over many shifts runs as a fast correlation through 8192- / 16384-point blocks
    (csrc/ovs.cuh) from 1024-sample cutouts on: against the sliding dot product (the allocation-free entry point still runs
    it) on every shift, and against the oracle, QF^2 and the complex result (absResult=False)."""
    import ctypes as C
    import torch
    from pydsproutines_b200 import _lib
    K = 30011
    p0 = 12345
    cutout, rx = signals.make_fastxcorr_case(400 + N, N, N + K - 1 + 100, p0, 0, amp=0.9)
    dc, drx = _dev(cutout), _dev(rx)
    sh = np.arange(5, 5 + K)
    q = xr.fastXcorr(dc, drx, shifts=sh).cpu().numpy()
    assert q.shape == (K,) and int(np.argmax(q)) == p0 - 5
    lib = _lib.load()
    qd = torch.empty(K, dtype=torch.float32, device="cuda")
    _lib.check(lib.xc_fastxcorr_nofreq_noalloc_c64(C.c_void_p(dc.data_ptr()), N, C.c_void_p(drx.data_ptr()), drx.numel(), 5, K,
                                                   C.c_void_p(qd.data_ptr()), None))
    torch.cuda.synchronize()
    np.testing.assert_allclose(q, qd.cpu().numpy(), rtol=2e-5, atol=1e-7 * float(q.max()))
    idx = np.r_[0:8, p0 - 5 - 4:p0 - 5 + 4, K - 8:K]
    assert_qf2(q[idx], orc.fastXcorr(cutout, rx, shifts=sh[idx]))
    z = xr.fastXcorr(dc, drx, shifts=sh, absResult=False).cpu().numpy()
    rz = orc.fastXcorr(cutout, rx, shifts=sh[idx], absResult=False)
    np.testing.assert_allclose(z[idx], rz, rtol=0, atol=2e-5 * float(np.abs(rz).max()))
    np.testing.assert_allclose(np.abs(z) ** 2, q, rtol=1e-5, atol=1e-9)


def test_warp_layout_windows_leaving_rx(xr):
    """N = 1024 (one warp per FFT over a staged window): windows that leave rx on either side give 0 / 0
    (IppXcorrFFT.cpp:125-130), the rest agrees with short calls (B-groups-per-CTA layout) and the oracle; several refills."""
    cutout, rx = signals.make_fastxcorr_case(77, 1024, 1024 + 90000, 54321, 1000, amp=0.8)
    obj = xr.CyIppXcorrFFT(cutout)
    a, b = -37, 90000 + 1 + 41
    q, f = obj.xcorr(rx, a, b, 1)
    assert q.size == b - a and not q[:37].any() and not f[:37].any() and not q[-41:].any() and not f[-41:].any()
    assert q[37] > 0 and q[-42] > 0
    i = int(np.argmax(q))
    assert i == 54321 + 37 and f[i] == 1000
    qs, fs = [], []
    for lo in range(a, b, 2000):
        qq, ff = obj.xcorr(rx, lo, min(lo + 2000, b), 1)
        qs.append(qq)
        fs.append(ff)
    q2, f2 = np.concatenate(qs), np.concatenate(fs)
    np.testing.assert_allclose(q, q2, rtol=2e-6, atol=1e-9)
    assert_bins(f[37:-41], f2[37:-41], lambda i: _spec64(cutout, rx, int(i)))
    idx = np.r_[np.arange(0, 48), np.arange(54321 - 24, 54321 + 24), np.arange(90000 - 47, 90001)]
    rq, rf = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=idx)
    assert_qf2(q[idx + 37], rq)
    assert_bins(f[idx + 37], rf, lambda i: _spec64(cutout, rx, int(idx[i])))


@pytest.mark.parametrize("G,L,kind,F", [(64, 128, "1000g", 4096), (64, 128, "ragged", 4096), (40, 100, "ragged", 4096),
                                         (24, 200, "ragged", 4096), (64, 128, "dense", 4096), (3, 700, "ragged", 4096),
                                         (16, 64, "128g", 2048), (50, 100, "ragged", 2048), (64, 128, "dense", 2048),
                                         (16, 128, "1000g", 1024), (30, 50, "ragged", 1024), (5, 1000, "ragged", 1024),
                                         (16, 128, "1000g", 512), (8, 300, "ragged", 512), (10, 20, "ragged", 32),
                                         (64, 128, "1000g", 8192), (20, 300, "ragged", 8192), (64, 128, "128g", 8192)])
def test_group_rotated_rows_vs_oracle_and_ell_layout(xr, G, L, kind, F):
    """GroupXcorrFFT, fftlen 4096 (and 2048: rows of 64 cells), contiguous shifts, group starts that are NOT aligned to the 128-cell rows of the transform
    (csrc/group_pp.cuh, ROT variant: sub-groups of 128 samples sorted by row, rotated thread mapping, carry into the next
    row): against the same search in short calls (per-cell gather lists, csrc/group.cu) and against the oracle; group
    lengths that are not a multiple of 128, folds that wrap around fftlen and overlap each other."""
    K = 1 << 14
    rng = np.random.default_rng(G * 1000 + L + F)
    if kind == "1000g":
        starts = np.arange(G) * 1000
    elif kind == "128g":
        starts = np.arange(G) * 128
    elif kind == "dense":
        starts = np.sort(rng.integers(0, 3000, G))             # heavily overlapping cells
    else:
        starts = np.cumsum(rng.integers(L, 3 * L + 777, G))
    starts = (starts - starts[0]).astype(np.int64)
    y = signals.qpsk(rng, G * L).reshape(G, L)
    span = int(starts[-1]) + F
    rx = signals.cnoise(rng, K + span)
    p0, k0 = 9876, 1501 % F
    for g in range(G):
        a = p0 + int(starts[g])
        rx[a:a + L] += np.complex64(0.8) * y[g] * signals.tone(L, k0, F, offset=int(starts[g]))
    obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F)
    q, f, peak = obj.xcorrPeak(_dev(rx), np.arange(K))
    q, f = q.cpu().numpy(), f.cpu().numpy()
    assert (peak.shift, peak.fidx) == (p0, k0) and int(np.argmax(q)) == p0 and int(f[p0]) == k0
    qs, fs = [], []
    for lo in range(0, K, 2048):                               # < 16 shifts per SM: the gather-list layout
        qq, ff = obj.xcorr(_dev(rx), np.arange(lo, lo + 2048))
        qs.append(qq.cpu().numpy())
        fs.append(ff.cpu().numpy())
    q2, f2 = np.concatenate(qs), np.concatenate(fs)
    np.testing.assert_allclose(q, q2, rtol=2e-5, atol=1e-7 * float(q.max()))
    assert_bins(f, f2, lambda i: _group_spec64(y, starts, rx, F, int(i)), what="group bin")
    idx = np.r_[np.arange(0, 12), np.arange(p0 - 6, p0 + 6), np.arange(K - 12, K)]
    qr, fr = orc.GroupXcorrFFT(y, starts, 1, fftlen=F).xcorr(rx, idx)
    assert_qf2(q[idx], qr)
    assert_bins(f[idx], fr, lambda i: _group_spec64(y, starts, rx, F, int(idx[i])), what="group bin")


def test_group_rotated_rows_nonzero_first_start(xr):
    """autoZeroStarts=False with a first start of 333 samples (xcorrRoutines.py:1058-1060): the fold is rotated by the first
    start as well; fftlen 4096 and 1024, against the oracle and the gather-list layout."""
    for F, G, L in ((4096, 20, 128), (1024, 12, 96)):
        rng = np.random.default_rng(F + G)
        starts = (np.cumsum(rng.integers(L, 900, G)) + 333).astype(np.int64)
        y = signals.qpsk(rng, G * L).reshape(G, L)
        K = 1 << 13
        rx = signals.cnoise(rng, K + int(starts[-1]) + F)
        p0, k0 = 4000, 77
        for g in range(G):
            a = p0 + int(starts[g])
            rx[a:a + L] += y[g] * signals.tone(L, k0, F, offset=int(starts[g]))
        obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F, autoZeroStarts=False)
        q, f = obj.xcorr(_dev(rx), np.arange(K))
        q, f = q.cpu().numpy(), f.cpu().numpy()
        assert int(np.argmax(q)) == p0 and int(f[p0]) == k0
        q2, f2 = obj.xcorr(_dev(rx), np.arange(1000, 3000))
        np.testing.assert_allclose(q[1000:3000], q2.cpu().numpy(), rtol=2e-5, atol=1e-7)
        idx = np.r_[np.arange(0, 8), np.arange(p0 - 4, p0 + 4), np.arange(K - 8, K)]
        qr, fr = orc.GroupXcorrFFT(y, starts, 1, fftlen=F, autoZeroStarts=False).xcorr(rx, idx)
        assert_qf2(q[idx], qr)
        assert_bins(f[idx], fr, lambda i: _group_spec64(y, starts, rx, F, int(idx[i])), what="group bin")


@pytest.mark.parametrize("G,L,F,kind", [(64, 128, 128, "1024g"), (20, 100, 128, "ragged"), (30, 256, 256, "ragged"),
                                         (50, 64, 64, "1000g"), (16, 40, 64, "ragged"), (7, 128, 256, "ragged"),
                                         (100, 128, 128, "dense")])
def test_group_dense_fold_short_fftlen(xr, G, L, F, kind):
    """GroupXcorrFFT at short transform lengths (fftlen 64 / 128 / 256, including the reference's default fftlen = group
    length): the dense-fold kernel (csrc/group_dense.cuh: register-tiled fold of G F products per shift, sub-warp FFT of
    the tile) against short calls (gather lists, csrc/group.cu) and the oracle; odd and even rotations, groups shorter than
    fftlen, blocks that straddle a group's wrap-around point."""
    K = 1 << 13
    rng = np.random.default_rng(G * 100 + L + F)
    if kind == "1024g":
        starts = np.arange(G) * 1024
    elif kind == "1000g":
        starts = np.arange(G) * 1000 + 3
    elif kind == "dense":
        starts = np.sort(rng.integers(0, 5000, G))
    else:
        starts = np.cumsum(rng.integers(L, 3 * L + 77, G))
    starts = (starts - starts[0]).astype(np.int64)
    y = signals.qpsk(rng, G * L).reshape(G, L)
    rx = signals.cnoise(rng, K + int(starts[-1]) + F)
    p0, k0 = 5005, (F // 3) | 1
    for g in range(G):
        a = p0 + int(starts[g])
        rx[a:a + L] += np.complex64(0.8) * y[g] * signals.tone(L, k0, F, offset=int(starts[g]))
    obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F)
    q, f, peak = obj.xcorrPeak(_dev(rx), np.arange(K))
    q, f = q.cpu().numpy(), f.cpu().numpy()
    assert (peak.shift, peak.fidx) == (p0, k0) and int(np.argmax(q)) == p0 and int(f[p0]) == k0
    qs, fs = [], []
    for lo in range(0, K, 2048):
        qq, ff = obj.xcorr(_dev(rx), np.arange(lo, lo + 2048))
        qs.append(qq.cpu().numpy())
        fs.append(ff.cpu().numpy())
    q2, f2 = np.concatenate(qs), np.concatenate(fs)
    np.testing.assert_allclose(q, q2, rtol=2e-5, atol=1e-7 * float(q.max()))
    assert_bins(f, f2, lambda i: _group_spec64(y, starts, rx, F, int(i)), what="group bin")
    idx = np.r_[np.arange(0, 12), np.arange(p0 - 6, p0 + 6), np.arange(K - 12, K)]
    qr, fr = orc.GroupXcorrFFT(y, starts, 1, fftlen=F).xcorr(rx, idx)
    assert_qf2(q[idx], qr)
    assert_bins(f[idx], fr, lambda i: _group_spec64(y, starts, rx, F, int(idx[i])), what="group bin")


def test_group_kernels_ragged_shift_counts(xr):
    """Shift counts that are not a multiple of the kernels' tiles (64 shifts per CTA tile in the dense-fold kernel, groups x
    chunk in the rotated-row kernels), a first shift that is not zero, and windows that run past the end of rx for the last
    shifts are not requested (the reference asserts them out): every fftlen family against the oracle at both ends."""
    for F, G, L in ((128, 30, 128), (256, 9, 200), (1024, 12, 128), (2048, 20, 64), (4096, 40, 128), (8192, 30, 128)):
        rng = np.random.default_rng(F + 7)
        starts = np.cumsum(rng.integers(L, 700, G)).astype(np.int64)
        starts -= starts[0]
        y = signals.qpsk(rng, G * L).reshape(G, L)
        K = 5000 + 37
        first = 11
        rx = signals.cnoise(rng, first + K + int(starts[-1]) + F + 5)
        p0, k0 = 3333, F // 2 + 3
        for g in range(G):
            a = p0 + int(starts[g])
            rx[a:a + L] += y[g] * signals.tone(L, k0, F, offset=int(starts[g]))
        obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F)
        sh = np.arange(first, first + K)
        q, f = obj.xcorr(_dev(rx), sh)
        q, f = q.cpu().numpy(), f.cpu().numpy()
        assert q.shape == (K,) and int(np.argmax(q)) == p0 - first and int(f[p0 - first]) == k0
        idx = np.r_[0:10, K - 10:K]
        qr, fr = orc.GroupXcorrFFT(y, starts, 1, fftlen=F).xcorr(rx, sh[idx])
        assert_qf2(q[idx], qr)
        assert_bins(f[idx], fr, lambda i: _group_spec64(y, starts, rx, F, int(sh[idx[i]])), what="group bin")


def test_groupfreq_large_planted_and_sampled(xr):
    """C3's groups (64 x 128 at 1024 g) against 256 frequencies over 2^17 shifts: the planted (shift, frequency) is the
    global peak, and 48 randomly sampled shifts agree with the oracle (size-independent checks at bench-like size)."""
    c = signals.config_c3(log2_shifts=17)
    F = c["F"]
    NFQ = 256
    freqs = (np.arange(NFQ) - 100) * (1.0 / F) + c["k0"] / F - 28.0 / F          # contains k0 / F exactly at index 128
    k_true = int(np.argmin(np.abs(freqs - c["k0"] / F)))
    y = np.zeros(int(c["starts"][-1]) + 128, np.complex64)
    for g in range(64):
        y[int(c["starts"][g]):int(c["starts"][g]) + 128] = c["ygroups"][g]
    lengths = np.full(64, 128)
    a = xr.GroupXcorr(y, c["starts"], lengths, freqs, 1.0)
    K = c["K"]
    xc, fp = a.xcorr(_dev(c["rx"]), np.arange(K))
    xc, fp = xc.cpu().numpy(), fp.cpu().numpy()
    i = int(np.argmax(xc))
    assert i == c["p0"] and fp[i] == freqs[k_true] and xc[i] > 0.3
    o = orc.GroupXcorr(y, c["starts"], lengths, freqs, 1.0)
    rng = np.random.default_rng(9)
    sample = np.unique(np.concatenate([rng.integers(0, K, 47), [c["p0"]]]))
    rxc, rfp = o.xcorr(c["rx"], sample)
    assert_qf2(xc[sample], rxc)
    _groupfreq_bins_ok(fp[sample], rfp, lambda i: _groupfreq_surface64(y, c["starts"], lengths, freqs, 1.0, c["rx"], int(sample[i])), freqs)

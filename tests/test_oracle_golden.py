"""Pins the numpy restatement (oracle/xcorr_oracle.py) to vectors produced by the REAL reference
(oracle/make_golden.py, run in the build container) and to the reference's own known-answer tests."""
import glob
import os

import numpy as np
import pytest
import scipy.signal as sps

from oracle import xcorr_oracle as orc
from pydsproutines_b200 import signals

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(pattern):
    files = sorted(glob.glob(os.path.join(GOLDEN, pattern)))
    assert files, "golden vectors missing: %s" % pattern
    return [(os.path.basename(f)[:-4], np.load(f)) for f in files]


@pytest.mark.parametrize("name,d", _load("fastxcorr_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_fastxcorr_restatement_is_bit_exact(name, d):
    cutout, rx = d["cutout"], d["rx"]
    shifts = d["shifts"] if "shifts" in d.files else None
    assert np.array_equal(orc.fastXcorr(cutout, rx, shifts=shifts), d["qf2_nofreq"])
    q, f = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=shifts)
    assert q.dtype == np.float64 and f.dtype == np.uint32
    assert np.array_equal(q, d["qf2_freq"]) and np.array_equal(f, d["fidx_freq"])
    # planted peak is where the generator put it
    sh = np.arange(len(rx) - len(cutout) + 1) if shifts is None else shifts
    assert sh[np.argmax(q)] == int(d["p0"]) and f[np.argmax(q)] == int(d["k0"])
    if "caf" in d.files:
        assert np.array_equal(orc.fastXcorr(cutout, rx, shifts=shifts, absResult=False), d["cplx_nofreq"])
        q, f = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=shifts, absResult=False)
        assert np.array_equal(q, d["cplx_freq"]) and np.array_equal(f, d["cplx_fidx"])
        assert np.array_equal(orc.fastXcorr(cutout, rx, True, True, shifts), d["caf"])
        assert np.array_equal(orc.fastXcorr(cutout, rx, True, True, shifts, False), d["caf_cplx"])


@pytest.mark.parametrize("name,d", _load("groupxcorr_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_groupxcorr_restatement_is_bit_exact(name, d):
    o = orc.GroupXcorrFFT(d["ygroups"], d["starts"], 1, fftlen=int(d["F"]))
    xc, fi = o.xcorr(d["rx"], d["shifts"])
    assert xc.dtype == np.float64 and fi.dtype == np.uint32
    assert np.array_equal(xc, d["xc"]) and np.array_equal(fi, d["fi"])
    assert np.array_equal(o.xcorr(d["rx"], d["shifts"], flattenToTime=False), d["caf"])
    assert d["shifts"][np.argmax(xc)] == int(d["p0"]) and fi[np.argmax(xc)] == int(d["k0"])


def test_cyipp_conventions():
    c, rx = signals.make_fastxcorr_case(5, 32, 200, 60, 3)
    p, b = orc.cyIppXcorrFFT(c, rx, -4, 190, 2)
    assert p.dtype == np.float32 and b.dtype == np.int32 and p.size == len(range(-4, 190, 2))
    assert p[0] == 0 and p[1] == 0 and p[-1] == 0 and b[-1] == 0      # out of range -> 0 (IppXcorrFFT.cpp:125-130)
    assert np.argmax(p) == (60 + 4) // 2 and b[np.argmax(p)] == 3


# ---- reference known-answer tests restated (xcorrRoutines.py:2130-2241) --------------------------
def _qpsk_templates():
    rng = np.random.default_rng(7)
    x = signals.qpsk(rng, 100).astype(np.complex64)
    templates = np.vstack((x[20:40] * 1.234, x[40:60] * 2.345)).astype(np.complex64)
    return x, templates


def test_template_correlator_known_answers():
    x, templates = _qpsk_templates()
    tc = orc.TemplateCrossCorrelator(templates, x.size)
    out = tc.correlate(x)
    assert out.shape == (2, 100 - 20 + 1)
    np.testing.assert_almost_equal(np.abs(out[0, 20]), 1.0, 5)       # :2237-2241
    np.testing.assert_almost_equal(np.abs(out[1, 40]), 1.0, 5)
    qf, tidx = tc.correlate(x, returnMax=True)
    assert tidx[20] == 0 and tidx[40] == 1
    assert np.array_equal(qf, np.abs(out).max(axis=0)) and np.array_equal(tidx, np.abs(out).argmax(axis=0))
    # single template equals sqrt(fastXcorr) to 6 dp (:2180-2200)
    for i in range(2):
        ref = orc.fastXcorr(templates[i], x) ** 0.5
        np.testing.assert_almost_equal(np.abs(out[i]), ref, 5)
    with pytest.raises(ValueError):
        orc.TemplateCrossCorrelator(templates[0], 100)


def test_multitemplate_dot_matches_template_correlator():
    rng = np.random.default_rng(8)
    T, L, n = 5, 24, 300
    templates = signals.qpsk(rng, T * L).reshape(T, L)
    x = signals.cnoise(rng, n)
    x[100:100 + L] += templates[3]
    tidx, qf2 = orc.multiTemplateSlidingDotProduct(x, templates.conj(), 0, n - L + 1)
    qf, tidx2 = orc.TemplateCrossCorrelator(templates, n).correlate(x, returnMax=True)
    assert np.array_equal(tidx, tidx2.astype(np.int32))
    np.testing.assert_allclose(qf2, qf ** 2, rtol=2e-4)
    assert np.argmax(qf2) == 100 and tidx[100] == 3


def test_moving_average_known_answer():
    x = np.random.default_rng(9).random(100).astype(np.float32)        # filterRoutines.py:1326-1335
    np.testing.assert_almost_equal(orc.movingAverage(x, 20, sumInstead=True), sps.lfilter(np.ones(20), 1, x), 5)
    np.testing.assert_almost_equal(orc.movingAverage(x, 20), sps.lfilter(np.ones(20) / 20, 1, x), 5)


def test_find_local_maxima_known_answer():
    x = np.zeros(50, np.float32)                                       # cupyExtensions.py:697-712
    x[5], x[8], x[9], x[37], x[41] = 1.0, 0.5, 0.6, 0.9, 0.7
    assert list(orc.findLocalMaxima(x, 0.4)) == [5, 9, 37, 41]
    idx, val = orc.argmaxAbsRows(np.array([[1, 3j, -3, 0], [0, 0, 0, 0]], np.complex64))
    assert list(idx) == [1, 0] and list(val) == [3.0, 0.0]


@pytest.mark.parametrize("name,d", _load("cztxcorr_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_cztxcorr_restatement_matches_reference(name, d):
    """The direct-evaluation restatement of cztXcorr against the reference's chirp-z implementation (the reference
    runs the transform in complex128, so the two agree to float32 rounding of the result, not bit for bit)."""
    args = (d["cutout"], d["rx"], float(d["f1"]), float(d["f2"]), float(d["fs"]), float(d["step"]))
    res, freqs = orc.cztXcorr(*args, shifts=d["shifts"])
    assert res.dtype == np.complex64 and freqs.dtype == np.float64
    np.testing.assert_allclose(res, d["res"], rtol=0, atol=1e-6)
    assert np.array_equal(freqs, d["freqs"])
    caf, fsearch = orc.cztXcorr(*args, outputCAF=True, shifts=d["shifts"])
    np.testing.assert_allclose(caf, d["caf"], rtol=1e-5, atol=1e-7 * d["caf"].max())
    assert np.array_equal(fsearch, d["fsearch"])
    # the planted offset is found at the planted shift, to within one frequency step
    i = int(np.argmax(np.abs(res)))
    assert d["shifts"][i] == int(d["p0"]) and abs(freqs[i] - float(d["foff"])) <= float(d["step"])


@pytest.mark.parametrize("name,d", _load("groupfreq_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_groupfreq_restatement_matches_reference(name, d):
    """GroupXcorr (xcorrRoutines.py:852-954) and GroupXcorrCZT (957-1039) restatements against the reference's
    outputs; the CZT variant is evaluated directly instead of through the chirp-z convolution, so the surfaces
    agree to rounding, not bit for bit."""
    a = orc.GroupXcorr(d["y"], d["starts"], d["lengths"], d["freqs"], float(d["fs"]))
    xc, fp = a.xcorr(d["rx"], d["shifts"])
    np.testing.assert_allclose(xc, d["xc"], rtol=1e-9, atol=0)
    assert np.array_equal(fp, d["freqpeaks"])
    b = orc.GroupXcorrCZT(d["y"], d["starts"], d["lengths"], float(d["f1"]), float(d["f2"]), float(d["bw"]), float(d["fs"]))
    caf, cf = b.xcorr(d["rx"], d["shifts"])
    assert np.array_equal(cf, d["cztFreq"]) and caf.shape == d["caf"].shape
    np.testing.assert_allclose(caf, d["caf"], rtol=1e-6, atol=1e-9 * d["caf"].max())
    # planted at p0, at the searched frequency next to the planted offset
    i = int(np.argmax(xc))
    assert i == int(d["p0"])
    assert abs(fp[i] - float(d["foff"])) <= np.min(np.abs(np.diff(np.sort(d["freqs"])))) + 1e-12


@pytest.mark.parametrize("name,d", _load("fine_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_fine_search_restatement_matches_reference(name, d):
    """fineFreqTimeSearch (xcorrRoutines.py:583-667), GenXcorr.xcorr (697-719) and calcQF2 (833-848).  The picks
    (fine frequency, time offset) must be identical; the cost vectors agree to 5e-8 absolute only, because numpy >= 2
    runs np.fft.fft of the complex64 inputs in single precision when the golden vectors are generated, while the
    restatement (like the reference under numpy 1.x) transforms in double."""
    bounds = d["bounds"] if "bounds" in d.files else None
    ff, td, cost = orc.fineFreqTimeSearch(d["x"], d["y"], list(d["fineRes"]), float(d["f0"]), float(d["freqRes"]),
                                          float(d["fs"]), d["td"], td_scan_freqBounds=bounds)
    if np.isnan(d["finefreq"]):
        assert ff is None
    else:
        assert ff == float(d["finefreq"])
    assert td == float(d["timediff"])
    np.testing.assert_allclose(cost, d["cost"], rtol=0, atol=5e-8)
    gtd, gcost = orc.timeScan(d["x"], d["y"], d["td"], float(d["fs"]), bounds)
    assert gtd == float(d["gen_td"])
    np.testing.assert_allclose(gcost, d["gen_cost"], rtol=0, atol=5e-7)       # both reference FFTs in float32 (numpy >= 2)
    np.testing.assert_allclose(orc.calcQF2(d["x"], d["y"]), float(d["qf2"]), rtol=1e-5)


@pytest.mark.parametrize("name,d", _load("groupperm_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_groupperm_restatement_matches_reference(name, d):
    """GroupXcorrCZT_Permutations.xcorr + getCAF (xcorrRoutines.py:1485-1585)."""
    o = orc.GroupXcorrCZT_Permutations(d["ygroups"], d["idxs"], d["starts"], float(d["f1"]), float(d["f2"]), float(d["bw"]),
                                       float(d["fs"]))
    cf = o.xcorr(d["rx"], d["shifts"])
    assert np.array_equal(cf, d["cztFreq"])
    scale = np.abs(d["xcTemplates"]).max()
    np.testing.assert_allclose(o.xcTemplates, d["xcTemplates"], rtol=0, atol=2e-6 * scale)    # golden stored as complex64
    np.testing.assert_allclose(o.rxgroupNormSq, d["rxgroupNormSq"], rtol=1e-6)
    for key, perm in (("caf_planted", tuple(d["planted"])), ("caf_other", (0, 0, 0))):
        np.testing.assert_allclose(o.getCAF(np.array(perm)), d[key], rtol=1e-6, atol=1e-9 * d[key].max())
    caf = o.getCAF(d["planted"])
    i, b = np.unravel_index(int(np.argmax(caf)), caf.shape)
    assert i == int(d["p0"]) and cf[b] == float(d["foff"]) and caf.max() > 3 * d["caf_other"].max()


@pytest.mark.parametrize("name,d", _load("cztrun_*.npz"), ids=lambda v: v if isinstance(v, str) else "")
def test_czt_run_restatement_matches_reference(name, d):
    """CZTCached.run / runMany (spectralRoutines.py:239-306): direct evaluation against the chirp-z convolution."""
    many = orc.cztRun(d["x"], float(d["f1"]), float(d["f2"]), float(d["bw"]), float(d["fs"]))
    scale = np.abs(d["many"]).max()
    np.testing.assert_allclose(many, d["many"], rtol=0, atol=1e-9 * scale)
    np.testing.assert_allclose(many[0], d["one"], rtol=0, atol=1e-9 * scale)
    assert np.all(many[-1] == 0)


# ---------------------------------------------------------------------------------------------
# cutouts beyond one block FFT (tests/golden/big*.npz hold recipes + the real reference's outputs)
# ---------------------------------------------------------------------------------------------
def _big(pattern):
    import glob
    return sorted(glob.glob(os.path.join(GOLDEN, pattern)))


@pytest.mark.parametrize("path", _big("bigxcorr_*.npz"), ids=os.path.basename)
def test_oracle_large_cutouts_bit_exact(path):
    from oracle.recipes import big_inputs
    d = np.load(path)
    cutout, rx = big_inputs(d)
    shifts = d["shifts"] if "shifts" in d.files else None
    q, f = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=shifts)
    assert np.array_equal(q, d["qf2_freq"]) and np.array_equal(f, d["fidx_freq"])
    k = int(np.argmax(q))
    sh = np.arange(q.size) if shifts is None else shifts
    assert sh[k] == int(d["p0"]) and int(f[k]) == int(d["k0"])


@pytest.mark.parametrize("path", _big("bigczt_*.npz"), ids=os.path.basename)
def test_oracle_large_chirpz(path):
    from oracle.recipes import big_inputs
    d = np.load(path)
    cutout, rx = big_inputs(d)
    res, fr = orc.cztXcorr(cutout, rx, float(d["f1"]), float(d["f2"]), float(d["fs"]), cztStep=float(d["step"]), shifts=d["shifts"])
    assert np.array_equal(fr, d["freqs"])
    np.testing.assert_allclose(res, d["res"], rtol=0, atol=1e-6 * np.abs(d["res"]).max())

"""GPU tests at BASELINE.json's full sizes (C1..C4; C5's sharding is covered by bench.py --gpus N and the
gloo test): the planted peak must come back with the exact shift / bin / template, the oracle is evaluated
on the peak window and on seeded random windows (it needs 0.24 ms .. 5 ms per shift), and size-independent
properties hold over the whole output (scale invariance of QF^2, shard == slice of the full search,
QF^2 <= 1)."""
import numpy as np
import pytest

from oracle import xcorr_oracle as orc
from pydsproutines_b200 import signals
from test_gpu_parity import _dev, assert_bins, assert_qf2, _spec64, _group_spec64, assert_templates

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def xr():
    import torch
    assert torch.cuda.is_available()
    import pydsproutines_b200.xcorrRoutines as m
    return m


def _windows(rng, K, p0, half, nrand, width):
    w = [np.arange(max(p0 - half, 0), min(p0 + half + 1, K))]
    for a in rng.integers(0, K - width, nrand):
        w.append(np.arange(a, a + width))
    return np.unique(np.concatenate(w))


def test_c1_fastxcorr_nofreq_full(xr):
    c = signals.config_c1()
    cutout, rx, p0 = c["cutout"], c["rx"], c["p0"]
    q = xr.fastXcorr(_dev(cutout), _dev(rx)).cpu().numpy()
    K = rx.size - cutout.size + 1
    assert q.shape == (K,) and int(np.argmax(q)) == p0
    assert q.max() <= 1.0 + 1e-5 and q.min() >= 0.0
    idx = _windows(np.random.default_rng(1), K, p0, 64, 8, 64)
    assert_qf2(q[idx], orc.fastXcorr(cutout, rx, shifts=idx))
    # host path gives the same numbers; scaling rx does not change QF^2
    assert np.array_equal(xr.fastXcorr(cutout, rx), q)
    q2 = xr.fastXcorr(_dev(cutout), _dev(rx * np.complex64(3.0))).cpu().numpy()
    np.testing.assert_allclose(q2, q, rtol=2e-5, atol=1e-9)


def test_c2_fastxcorr_freqsearch_full(xr):
    c = signals.config_c2()
    cutout, rx, p0, k0, K = c["cutout"], c["rx"], c["p0"], c["k0"], c["K"]
    q, f, peak = xr.fastXcorrPeak(_dev(cutout), _dev(rx))
    q, f = q.cpu().numpy(), f.cpu().numpy()
    assert q.shape == (K,) and int(np.argmax(q)) == p0 and int(f[p0]) == k0
    assert (peak.shift, peak.fidx, peak.index) == (p0, k0, p0) and abs(peak.qf2 - q[p0]) == 0
    assert q.max() <= 1.0 + 1e-5 and f.max() < 8192
    idx = _windows(np.random.default_rng(2), K, p0, 96, 12, 32)
    qr, fr = orc.fastXcorr(cutout, rx, freqsearch=True, shifts=idx)
    assert_qf2(q[idx], qr)
    assert_bins(f[idx], fr, lambda i: _spec64(cutout, rx, int(idx[i])))
    # a shard (block + halo) reproduces the matching slice of the full search bit for bit
    a, n = 300000, 5000
    qs, fs = xr.fastXcorr(_dev(cutout), _dev(rx[a:a + n + 8191]), freqsearch=True)
    assert np.array_equal(qs.cpu().numpy().astype(np.float32), q[a:a + n]) and np.array_equal(fs.cpu().numpy(), f[a:a + n])


def test_c3_groupxcorr_full(xr):
    c = signals.config_c3()
    y, starts, rx, F, p0, k0, K = c["ygroups"], c["starts"], c["rx"], c["F"], c["p0"], c["k0"], c["K"]
    obj = xr.GroupXcorrFFT(y, starts, 1, fftlen=F)
    q, f, peak = obj.xcorrPeak(_dev(rx), np.arange(K))
    q, f = q.cpu().numpy(), f.cpu().numpy()
    assert q.shape == (K,) and int(np.argmax(q)) == p0 and int(f[p0]) == k0
    assert (peak.shift, peak.fidx) == (p0, k0)
    idx = _windows(np.random.default_rng(3), K, p0, 8, 4, 8)
    ref = orc.GroupXcorrFFT(y, starts, 1, fftlen=F)
    qr, fr = ref.xcorr(rx, idx)
    assert_qf2(q[idx], qr)
    assert_bins(f[idx], fr, lambda i: _group_spec64(y, starts, rx, F, int(idx[i])), what="group bin")
    caf = obj.xcorr(_dev(rx), idx[:6], flattenToTime=False).cpu().numpy()
    np.testing.assert_allclose(caf, ref.xcorr(rx, idx[:6], flattenToTime=False), rtol=1e-4, atol=1e-6 * caf.max())


def test_c4_multitemplate_full():
    from pydsproutines_b200 import cupyExtensions as ce
    c = signals.config_c4()
    templates, x, plants, K = c["templates"], c["x"], c["plants"], c["K"]
    tidx, q = ce.multiTemplateSlidingDotProduct(_dev(x), _dev(templates.conj()), 0, K)
    tidx, q = tidx.cpu().numpy(), q.cpu().numpy()
    assert q.shape == (K,)
    for pos, t, a in plants:
        lo = max(pos - 2000, 0)
        assert lo + int(np.argmax(q[lo:pos + 2000])) == pos and int(tidx[pos]) == t
    assert int(np.argmax(q)) == plants[-1][0]
    # oracle on windows: every template on the window around each plant, 4 templates on random windows
    rng = np.random.default_rng(4)
    L = templates.shape[1]
    for pos, t, a in plants:
        seg = x[pos - 16:pos + 16 + L - 1]
        ti_ref, q_ref = orc.multiTemplateSlidingDotProduct(seg, templates.conj(), 0, 32)
        assert_qf2(q[pos - 16:pos + 16], q_ref)
        assert_templates(tidx[pos - 16:pos + 16], ti_ref, templates, x, pos - 16)
    for a in rng.integers(0, K - 64, 3):
        seg = x[a:a + 64 + L - 1]
        ti_ref, q_ref = orc.multiTemplateSlidingDotProduct(seg, templates.conj(), 0, 64)
        assert_qf2(q[a:a + 64], q_ref)
        assert_templates(tidx[a:a + 64], ti_ref, templates, x, int(a))

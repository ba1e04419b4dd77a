"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the REAL reference
(/root/reference, imported through oracle/ref_import.py) in the build container.

    PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden.py

The vectors hold inputs and the reference's outputs; tests/test_oracle_golden.py pins the numpy
restatement (oracle/xcorr_oracle.py) to them and the -m gpu tests pin the CUDA library to them.
Known-answer cases of the reference's own unit tests are restated in tests/ directly
(xcorrRoutines.py:2130-2241, filterRoutines.py:1326-1335, cupyExtensions.py:697-712).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_import  # noqa: E402
from recipes import big_inputs  # noqa: E402
from pydsproutines_b200 import signals  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def fastxcorr_cases(ref):
    cases = {}
    # (name, seed, N, rxlen, p0, k0, shifts or None)
    specs = [
        ("n64", 11, 64, 400, 123, 5, None),
        ("n100_nonpow2", 12, 100, 350, 77, 9, None),
        ("n256_step7", 13, 256, 2000, 704, 200, np.arange(4, 1500, 7)),
        ("n1024", 14, 1024, 1024 + 96, 40, 1000, None),
        ("n4096_win", 15, 4096, 4096 + 600, 300, 4000, np.arange(280, 320)),
        ("n8192_win", 16, 8192, 8192 + 700, 345, 1234, np.arange(337, 353)),
    ]
    for name, seed, N, rxlen, p0, k0, shifts in specs:
        cutout, rx = signals.make_fastxcorr_case(seed, N, rxlen, p0, k0, amp=0.7)
        d = dict(cutout=cutout, rx=rx, p0=p0, k0=k0)
        if shifts is not None:
            d["shifts"] = shifts
        d["qf2_nofreq"] = ref.fastXcorr(cutout, rx, shifts=shifts)
        q, f = ref.fastXcorr(cutout, rx, freqsearch=True, shifts=shifts)
        d["qf2_freq"], d["fidx_freq"] = q, f
        if N <= 256:
            d["cplx_nofreq"] = ref.fastXcorr(cutout, rx, shifts=shifts, absResult=False)
            q, f = ref.fastXcorr(cutout, rx, freqsearch=True, shifts=shifts, absResult=False)
            d["cplx_freq"], d["cplx_fidx"] = q, f
            d["caf"] = ref.fastXcorr(cutout, rx, freqsearch=True, outputCAF=True, shifts=shifts)
            d["caf_cplx"] = ref.fastXcorr(cutout, rx, freqsearch=True, outputCAF=True,
                                          shifts=shifts, absResult=False)
        cases[name] = d
    return cases


def group_cases(ref):
    cases = {}
    rng = np.random.default_rng(21)
    specs = [
        # name, G, L, F, starts, K, p0, k0
        ("g4_l16_f64", 4, 16, 64, np.array([0, 40, 80, 120]), 60, 17, 5),
        ("g8_l32_f256_ragged", 8, 32, 256, np.array([10, 47, 99, 140, 300, 333, 420, 511]), 48, 30, 200),
        ("g64_l128_f4096_win", 64, 128, 4096, np.arange(64) * 1024, 12, 5, 777),
        ("g6_l20_f32_wrap", 6, 20, 32, np.array([0, 25, 50, 77, 100, 131]), 40, 11, 3),
    ]
    for name, G, L, F, starts, K, p0, k0 in specs:
        ygroups = signals.qpsk(rng, G * L).reshape(G, L)
        z = starts - starts[0]
        rx = signals.cnoise(rng, K + int(z[-1]) + F + 1)
        for g in range(G):
            a = p0 + int(z[g])
            rx[a:a + L] += np.complex64(0.8) * ygroups[g] * signals.tone(L, k0, F, offset=int(z[g]))
        obj = ref.GroupXcorrFFT(ygroups, starts, 1, fftlen=F)
        shifts = np.arange(K)
        xc, fi = obj.xcorr(rx, shifts)
        caf = obj.xcorr(rx, shifts, flattenToTime=False)
        cases[name] = dict(ygroups=ygroups, starts=starts, rx=rx, F=F, shifts=shifts,
                           xc=xc, fi=fi, caf=caf, p0=p0, k0=k0)
    return cases


def czt_cases(ref):
    """cztXcorr (xcorrRoutines.py:413-457): a planted cutout with a fractional-bin frequency offset."""
    cases = {}
    # name, seed, N, rxlen, p0, f_off (cycles/sample * fs), fs, f1, f2, step
    specs = [
        ("n100_fine", 31, 100, 260, 60, 0.0123, 1.0, -0.02, 0.02, 0.0005),
        ("n1000_hz", 32, 1000, 1200, 90, 37.5, 10000.0, -100.0, 100.0, 2.5),
        ("n4096_narrow", 33, 4096, 4096 + 40, 17, -3.2, 1000.0, -5.0, 5.0, 0.1),
    ]
    for name, seed, N, rxlen, p0, foff, fs, f1, f2, step in specs:
        rng = np.random.default_rng(seed)
        cutout = signals.qpsk(rng, N)
        rx = signals.cnoise(rng, rxlen)
        rx[p0:p0 + N] += np.complex64(0.8) * cutout * np.exp(2j * np.pi * foff * np.arange(N) / fs).astype(np.complex64)
        shifts = np.arange(max(p0 - 12, 0), min(p0 + 12, rxlen - N + 1))
        res, freqs = ref.cztXcorr(cutout, rx, f1, f2, fs, cztStep=step, shifts=shifts)
        caf, fsearch = ref.cztXcorr(cutout, rx, f1, f2, fs, cztStep=step, outputCAF=True, shifts=shifts)
        cases[name] = dict(cutout=cutout, rx=rx, shifts=shifts, f1=f1, f2=f2, fs=fs, step=step, p0=p0, foff=foff,
                           res=res, freqs=freqs, caf=caf.astype(np.float32), fsearch=fsearch)
    return cases


def groupfreq_cases(ref):
    """GroupXcorr (xcorrRoutines.py:852-954) and GroupXcorrCZT (957-1039): ragged groups cut from a long template,
    planted in noise with a frequency offset that sits on / between the searched frequencies."""
    cases = {}
    # name, seed, ylen, starts, lengths, fs, freqs (GroupXcorr) , (f1, f2, bw) (GroupXcorrCZT), foff, p0, K
    specs = [
        ("g3_ragged", 41, 400, np.array([5, 120, 300]), np.array([40, 33, 64]), 1000.0,
         np.arange(-20.0, 20.0, 1.25), (-10.0, 10.0, 0.5), 6.25, 21, 50),
        ("g8_l128", 42, 9000, np.arange(8) * 1100 + 16, np.full(8, 128), 1.0,
         np.linspace(-0.002, 0.002, 41), (-0.001, 0.001, 0.00005), 0.0007, 9, 24),
        ("g2_long", 43, 1500, np.array([0, 800]), np.array([600, 530]), 100.0,
         np.array([-3.0, -1.0, 0.0, 0.4, 1.7, 2.0, 5.5]), (-1.0, 1.0, 0.125), 0.4, 3, 12),
    ]
    for name, seed, ylen, starts, lengths, fs, freqs, (f1, f2, bw), foff, p0, K in specs:
        rng = np.random.default_rng(seed)
        y = signals.qpsk(rng, ylen)
        z = starts - starts[0]
        rx = signals.cnoise(rng, K + int(z[-1]) + int(lengths[-1]) + 8)
        for g in range(starts.size):
            n = np.arange(lengths[g]) + z[g]
            seg = y[starts[g]:starts[g] + lengths[g]] * np.exp(2j * np.pi * foff * n / fs)
            rx[p0 + z[g]:p0 + z[g] + lengths[g]] += (0.8 * seg).astype(np.complex64)
        shifts = np.arange(K)
        a = ref.GroupXcorr(y, starts, lengths, freqs, fs)
        xc, fp = a.xcorr(rx, shifts)
        b = ref.GroupXcorrCZT(y, starts, lengths, f1, f2, bw, fs)
        caf, cf = b.xcorr(rx, shifts)
        cases[name] = dict(y=y, starts=starts, lengths=lengths, fs=fs, freqs=freqs, f1=f1, f2=f2, bw=bw, rx=rx,
                           shifts=shifts, xc=xc, freqpeaks=fp, caf=caf, cztFreq=cf, p0=p0, foff=foff)
    return cases


def fine_cases(ref):
    """fineFreqTimeSearch (xcorrRoutines.py:583-667) and GenXcorr (679-719): y = x delayed by a fraction of a sample
    and shifted by a fraction of a coarse frequency bin, both in noise."""
    cases = {}
    # name, seed, N, fs, frac delay (samples), freq offset, coarse freq found, freqRes, fineRes, td range, bounds
    specs = [
        ("n1000", 51, 1000, 1000.0, 0.37, 0.31, 0.0, 1.0, [0.1, 0.01], np.arange(-0.002, 0.002, 0.00005), None),
        ("n777_bounds", 52, 777, 48000.0, -0.21, -13.0, -61.776, 61.776, [5.0, 0.5, 0.05],
         np.arange(-2, 2, 0.05) / 48000.0, (-12000.0, 12000.0)),
        ("n4096_nofreq", 53, 4096, 1.0, 0.45, 0.0, 0.0, 0.0, [], np.arange(-1.0, 1.0, 0.01), None),
    ]
    for name, seed, N, fs, delay, foff, f0, fres, fineRes, td, bounds in specs:
        rng = np.random.default_rng(seed)
        base = (rng.standard_normal(N) + 1j * rng.standard_normal(N))
        # band-limit to half the band so that a fractional delay is well defined
        B = np.fft.fft(base)
        fvec = np.fft.fftfreq(N)
        B[np.abs(fvec) > 0.25] = 0
        x = np.fft.ifft(B)
        yd = np.fft.ifft(B * np.exp(-2j * np.pi * fvec * delay))
        # fineFreqTimeSearch removes +f from x to match y: y carries -foff relative to x
        yv = yd * np.exp(-2j * np.pi * foff * np.arange(N) / fs)
        xn = (x + 0.05 * (rng.standard_normal(N) + 1j * rng.standard_normal(N))).astype(np.complex64)
        yn = (yv + 0.05 * (rng.standard_normal(N) + 1j * rng.standard_normal(N))).astype(np.complex64)
        ff, tdiff, cost = ref.fineFreqTimeSearch(xn, yn, fineRes, f0, fres, fs, td, td_scan_freqBounds=bounds)
        g = ref.GenXcorr(td, fs, N)
        if bounds is not None:
            g.setTDscan_freqBounds(np.array(bounds))
        gtd, gcost = g.xcorr(xn, yn)
        d = dict(x=xn, y=yn, fs=fs, fineRes=np.array(fineRes, np.float64), f0=f0, freqRes=fres, td=td,
                 finefreq=np.nan if ff is None else ff, timediff=tdiff, cost=cost, gen_td=gtd, gen_cost=gcost,
                 delay=delay, foff=foff, qf2=ref.calcQF2(xn, yn))
        if bounds is not None:
            d["bounds"] = np.array(bounds)
        cases[name] = d
    return cases


def groupperm_cases(ref):
    """GroupXcorrCZT_Permutations (xcorrRoutines.py:1264-1690), CPU methods xcorr + getCAF: 3 groups with 2 / 3 / 1
    candidate templates; the permutation (1, 2, 0) is planted."""
    rng = np.random.default_rng(61)
    L, fs = 96, 2000.0
    idxs = np.array([0, 0, 1, 1, 1, 2])
    starts = np.array([0, 300, 777])
    ygroups = signals.qpsk(rng, idxs.size * L).reshape(idxs.size, L)
    f1, f2, bw = -20.0, 20.0, 1.0
    K, p0, foff = 40, 13, 7.0
    rx = signals.cnoise(rng, K + int(starts[-1]) + L + 4)
    planted = (1, 2, 0)
    for g in range(3):
        k = np.argwhere(idxs == g)[planted[g]][0]
        n = np.arange(L) + starts[g]
        rx[p0 + starts[g]:p0 + starts[g] + L] += (0.8 * ygroups[k] * np.exp(2j * np.pi * foff * n / fs)).astype(np.complex64)
    obj = ref.GroupXcorrCZT_Permutations(ygroups, idxs, starts, f1, f2, bw, fs)
    shifts = np.arange(K)
    cf = obj.xcorr(rx, shifts, numThreads=1)
    d = dict(ygroups=ygroups, idxs=idxs, starts=starts, f1=f1, f2=f2, bw=bw, fs=fs, rx=rx, shifts=shifts, cztFreq=cf,
             xcTemplates=obj.xcTemplates.astype(np.complex64), rxgroupNormSq=obj.rxgroupNormSq, p0=p0, foff=foff,
             planted=np.array(planted))
    for name, perm in (("caf_planted", planted), ("caf_other", (0, 0, 0))):
        d[name] = obj.getCAF(np.array(perm), numThreads=1)
    return {"g3": d}


def czt_run_cases(ref):
    """spectralRoutines.CZTCached.run / runMany (the reference's own chirp-z convolution, complex128)."""
    import spectralRoutines as sr            # the reference's module (sys.path set by ref_import.load())
    cases = {}
    for name, seed, m, rows, f1, f2, bw, fs in (("m100", 71, 100, 3, -5.0, 5.0, 0.25, 100.0),
                                                ("m1000", 72, 1000, 2, 100.0, 140.0, 0.5, 8000.0)):
        rng = np.random.default_rng(seed)
        x = signals.cnoise(rng, rows * m).reshape(rows, m)
        x[-1] = 0                              # an all-zero row transforms to zeros
        c = sr.CZTCached(m, f1, f2, bw, fs)
        cases[name] = dict(x=x, f1=f1, f2=f2, bw=bw, fs=fs, many=c.runMany(x), one=c.run(x[0]), freq=c.getFreq())
    return cases


def big_cases(ref):
    """fastXcorr(freqsearch=True) of the real reference for cutouts beyond one block FFT: a power of two (2^17), lengths
    that are not (9001, 20000 with a stepped shift list, 100000).  Inputs are stored as recipes (see big_inputs)."""
    cases = {}
    specs = [
        ("n9001", 81, 9001, 9001 + 6, 3, 4500, None, True),
        ("n20000_step", 82, 20000, 20000 + 300, 120, 19999, np.arange(0, 300, 12), False),
        ("n32768", 83, 32768, 32768 + 11, 7, 12345, None, False),
        ("n100000", 84, 100000, 100000 + 20, 13, 31337, None, False),
        ("n131072", 85, 131072, 131072 + 24, 9, 70001, None, False),
    ]
    for name, seed, N, rxlen, p0, k0, shifts, full in specs:
        recipe = dict(seed=seed, N=N, rxlen=rxlen, p0=p0, k0=k0)
        cutout, rx = big_inputs(recipe)
        d = dict(recipe)
        if shifts is not None:
            d["shifts"] = shifts
        q, f = ref.fastXcorr(cutout, rx, freqsearch=True, shifts=shifts)
        d["qf2_freq"], d["fidx_freq"] = q, f
        z, fz = ref.fastXcorr(cutout, rx, freqsearch=True, shifts=shifts, absResult=False)
        d["cplx_freq"], d["cplx_fidx"] = z.astype(np.complex64), fz
        if full:
            d["caf"] = ref.fastXcorr(cutout, rx, freqsearch=True, outputCAF=True, shifts=shifts).astype(np.float32)
        cases[name] = d
    return cases


def bigczt_cases(ref):
    """cztXcorr with cutout length + bins - 1 beyond 16384 points."""
    cases = {}
    for name, seed, N, rxlen, p0, foff, fs, f1, f2, step in (("n20000", 91, 20000, 20000 + 30, 11, 1.37, 1000.0, -5.0, 5.0, 0.05),):
        recipe = dict(seed=seed, N=N, rxlen=rxlen, p0=p0, foff=foff, fs=fs)
        cutout, rx = big_inputs(recipe)
        shifts = np.arange(max(p0 - 6, 0), min(p0 + 6, rxlen - N + 1))
        res, freqs = ref.cztXcorr(cutout, rx, f1, f2, fs, cztStep=step, shifts=shifts)
        caf, fsearch = ref.cztXcorr(cutout, rx, f1, f2, fs, cztStep=step, outputCAF=True, shifts=shifts)
        cases[name] = dict(recipe, shifts=shifts, f1=f1, f2=f2, step=step, res=res, freqs=freqs, caf=caf.astype(np.float32),
                           fsearch=fsearch)
    return cases


def main():
    ref = ref_import.load()
    os.makedirs(OUT, exist_ok=True)
    only = sys.argv[1:]
    for fam, cases in (("fastxcorr", fastxcorr_cases(ref)), ("groupxcorr", group_cases(ref)), ("cztxcorr", czt_cases(ref)),
                       ("groupfreq", groupfreq_cases(ref)), ("fine", fine_cases(ref)),
                       ("groupperm", groupperm_cases(ref)),
                       ("cztrun", czt_run_cases(ref)), ("bigxcorr", big_cases(ref)), ("bigczt", bigczt_cases(ref))):
        for name, d in cases.items():
            if only and fam not in only:
                continue
            path = os.path.join(OUT, "%s_%s.npz" % (fam, name))
            np.savez_compressed(path, **d)
            print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()

"""Seeded synthetic inputs for the five BASELINE.json configurations and the
three reference signal generators the hot path's tests need.

Reference: signalCreationRoutines.py:27,47-69 (`randPSKsyms`, QPSK alphabet {1, j, -1, -j}),
72-104 (`randnoise`, complex AWGN), 380-386 (`makeFreq`).  Everything is complex64, generated from
`np.random.default_rng(seed)`; the same seed gives the same arrays on every machine, which is what lets
the GPU box, the build container and the golden fixtures agree on the inputs.
"""
import numpy as np


def makeFreq(length, fs):
    """signalCreationRoutines.py:380-386."""
    f = np.arange(length, dtype=np.float64) / length * fs
    f[f >= fs / 2] -= fs
    return f


def qpsk(rng, n):
    """Unit-amplitude QPSK symbols from {1, j, -1, -j} (signalCreationRoutines.py:27,47-69)."""
    alphabet = np.array([1, 1j, -1, -1j], dtype=np.complex64)
    return alphabet[rng.integers(0, 4, size=n)]


def cnoise(rng, n, power=1.0):
    """Circular complex Gaussian noise, E|w|^2 = power (signalCreationRoutines.py:72-104)."""
    s = np.sqrt(power / 2.0)
    w = rng.standard_normal(n, dtype=np.float32) + 1j * rng.standard_normal(n, dtype=np.float32)
    return (s * w).astype(np.complex64)


def tone(n, k0, F, offset=0):
    """exp(+j 2 pi k0 (offset + n) / F) as complex64, evaluated in double."""
    ph = (np.arange(n, dtype=np.int64) + offset) * (k0 % F) % F
    return np.exp(2j * np.pi * ph / F).astype(np.complex64)


# ------------------------------------------------------------------------------------------------
# C1 / C2 / C5: single cutout (optionally frequency shifted) planted in noise
# ------------------------------------------------------------------------------------------------
def make_fastxcorr_case(seed, N, rxlen, p0, k0=0, amp=1.0, decoy=None):
    """cutout = N QPSK symbols; rx = CN(0,1) noise of `rxlen` samples with amp*cutout*tone(k0)
    added at p0 (expected peak: shift p0, bin k0).  `decoy` = (position, amplitude, bin) adds a
    second, weaker copy (C5 uses it to exercise the cross-rank reduction)."""
    rng = np.random.default_rng(seed)
    cutout = qpsk(rng, N)
    rx = cnoise(rng, rxlen)
    rx[p0:p0 + N] += np.complex64(amp) * cutout * tone(N, k0, N)
    if decoy is not None:
        dp, da, dk = decoy
        rx[dp:dp + N] += np.complex64(da) * cutout * tone(N, dk, N)
    return cutout, rx


def config_c1(scale=1.0):
    """C1: 4096-sample cutout vs 2^20 samples, no freqsearch, planted at 345 678 (seed 1001)."""
    rxlen = int((1 << 20) * scale)
    p0 = int(345678 * scale)
    c, rx = make_fastxcorr_case(1001, 4096, rxlen, p0)
    return dict(cutout=c, rx=rx, p0=p0, k0=0)


def config_c2(log2_shifts=20, N=8192):
    """C2: N=8192 cutout, 2^20 shifts, planted at 600 001 with a +1234-bin tone (seed 1002)."""
    K = 1 << log2_shifts
    p0 = int(600001 * K / (1 << 20))
    k0 = 1234 % N
    c, rx = make_fastxcorr_case(1002, N, K + N - 1, p0, k0)
    return dict(cutout=c, rx=rx, p0=p0, k0=k0, K=K)


def config_c5(log2_len=26, N=8192):
    """C5: as C2 with a 2^26-sample rx, peak at 50 000 017 (+4321 bins) and a half-amplitude decoy
    early in the record (seed 1005)."""
    rxlen = 1 << log2_len
    p0 = int(50000017 * rxlen / (1 << 26))
    k0 = 4321 % N
    dp = int(1000003 * rxlen / (1 << 26))
    c, rx = make_fastxcorr_case(1005, N, rxlen, p0, k0, decoy=(dp, 0.5, (k0 * 3) % N))
    return dict(cutout=c, rx=rx, p0=p0, k0=k0, K=rxlen - N + 1, decoy=dp)


# ------------------------------------------------------------------------------------------------
# C3: grouped (bursty) preamble
# ------------------------------------------------------------------------------------------------
def config_c3(log2_shifts=22, G=64, L=128, F=4096, spacing=1024):
    """C3: G QPSK groups of L samples at starts spacing*g, fftlen F, 2^22 shifts; planted at
    p0 = 2 000 003 (scaled) with tone exp(+j2pi k0 (s_g+n)/F), k0 = 777 (seed 1003); fs = 1."""
    rng = np.random.default_rng(1003)
    K = 1 << log2_shifts
    starts = (np.arange(G) * spacing).astype(np.int64)
    ygroups = qpsk(rng, G * L).reshape(G, L)
    span = int(starts[-1]) + F          # the reference's default-shift rule uses fftlen (xcorrRoutines.py:1141)
    rx = cnoise(rng, K + span)
    p0 = int(2000003 * K / (1 << 22))
    k0 = 777 % F
    for g in range(G):
        a = p0 + int(starts[g])
        rx[a:a + L] += ygroups[g] * tone(L, k0, F, offset=int(starts[g]))
    return dict(ygroups=ygroups, starts=starts, rx=rx, F=F, p0=p0, k0=k0, K=K)


# ------------------------------------------------------------------------------------------------
# C4: many templates
# ------------------------------------------------------------------------------------------------
def config_c4(log2_len=24, T=256, L=1024):
    """C4: T QPSK templates of L samples vs 2^24 noise samples; templates 3, 77, 200 (mod T) planted at
    1e6, 8e6, 15e6 (scaled) with amplitudes 1.0, 1.234, 2.345 (seed 1004)."""
    rng = np.random.default_rng(1004)
    n = 1 << log2_len
    templates = qpsk(rng, T * L).reshape(T, L)
    x = cnoise(rng, n)
    plants = []
    for t, pos, a in ((3, 1000000, 1.0), (77, 8000000, 1.234), (200, 15000000, 2.345)):
        t = t % T
        pos = int(pos * n / (1 << 24))
        x[pos:pos + L] += np.complex64(a) * templates[t]
        plants.append((pos, t, a))
    return dict(templates=templates, x=x, plants=plants, K=n - L + 1)

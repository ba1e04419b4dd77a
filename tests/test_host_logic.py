"""CPU-only tests: the C-ABI library loads and exports every symbol include/xcorr_b200.h declares, the
host-side mirror keeps the reference's argument checks, compute calls fail loudly without a GPU (no
CPU fallback), and the shift-range sharding + peak exchange works over gloo with world_size 2."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from pydsproutines_b200 import _lib
    if not os.path.isfile(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    from pydsproutines_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "xcorr_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(xc_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(raw, name), "libxcorr_b200.so does not export %s" % name
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.xc_version() >= 100
    assert ctypes.sizeof(_lib.xc_peak) == 32
    assert lib.xc_fastxcorr_supported(8192) == 1 and lib.xc_fastxcorr_supported(100) == 1      # any length (Bluestein)
    assert lib.xc_fastxcorr_supported(16384) == 1 and lib.xc_fastxcorr_supported(1) == 0
    # beyond one block FFT: the four-step path, powers of two to 2^28, any other length while 2 N - 1 <= 2^28
    assert lib.xc_fastxcorr_supported(8193) == 1 and lib.xc_fastxcorr_supported(1000000) == 1
    assert lib.xc_fastxcorr_supported(1 << 28) == 1 and lib.xc_fastxcorr_supported((1 << 28) + 1) == 0


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from pydsproutines_b200 import xcorrRoutines as xr
    from pydsproutines_b200 import _lib
    c = np.ones(8, np.complex64)
    rx = np.ones(64, np.complex64)
    with pytest.raises(_lib.XcorrB200Error, match="no CUDA device"):
        xr.fastXcorr(c, rx)
    with pytest.raises(_lib.XcorrB200Error, match="no CUDA device"):
        xr.fastXcorr(c, rx, freqsearch=True)
    with pytest.raises(_lib.XcorrB200Error):
        xr.GroupXcorrFFT(np.ones((2, 4), np.complex64), np.array([0, 8]), 1, fftlen=8)
    with pytest.raises(_lib.XcorrB200Error):
        xr.TemplateCrossCorrelator(np.ones((2, 4), np.complex64), 64)
    # the product never imports the oracle
    for mod in ("xcorrRoutines", "cupyExtensions", "filterRoutines", "dist", "_lib", "signals"):
        src = open(os.path.join(ROOT, "pydsproutines_b200", mod + ".py")).read()
        assert "oracle" not in src.replace("the oracle", "").replace("numpy oracle", ""), mod


def test_argument_checks_match_reference(lib):
    from pydsproutines_b200 import xcorrRoutines as xr
    from pydsproutines_b200 import cupyExtensions as ce
    from pydsproutines_b200 import filterRoutines as fr
    c = np.ones(8, np.complex64)
    rx = np.ones(64, np.complex64)
    with pytest.raises(TypeError):
        xr.fastXcorr(c.astype(np.complex128), rx)
    with pytest.raises(ValueError):
        xr.fastXcorr(c, rx, shifts=np.array([60]))                        # window leaves rx
    with pytest.raises(ValueError):
        xr.cp_fastXcorr_v2(c, rx, startIdx=50, idxlen=20)
    with pytest.raises(Exception):
        xr.cp_fastXcorr(c, rx.astype(np.complex128))
    with pytest.raises(ValueError):
        xr.TemplateCrossCorrelator(c, 64)                                   # templates must be 2-D
    with pytest.raises(ValueError):
        ce.multiTemplateSlidingDotProduct(rx, c.reshape(1, -1), -1, 4)
    with pytest.raises(ValueError):
        ce.multiTemplateSlidingDotProduct(rx, c.reshape(1, -1), 0, 64)
    with pytest.raises(TypeError):
        ce.cupyFindLocalMaxima(np.zeros(4, np.float64), 0.1)
    with pytest.raises(ValueError):
        fr.cupyMovingAverage(np.zeros(4, np.float32), 2, NUM_PER_THREAD=32)
    with pytest.raises(TypeError):
        fr.cupyMovingAverage(np.zeros(4, np.float64), 2)
    with pytest.raises(RuntimeError):
        xr.CyGroupXcorrFFT(np.ones((2, 16), np.complex64), np.array([0, 20], np.int32), 1, fftlen=8)
    sh = xr._Shifts(np.arange(3, 40, 4), 64, 8, rx)
    assert (sh.shift0, sh.step, sh.K, sh.array) == (3, 4, 10, None)
    sh = xr._Shifts(np.array([1, 5, 6]), 64, 8, rx)
    assert sh.array is not None and sh.array.dtype == np.int64 and sh.K == 3
    sh = xr._Shifts(None, 64, 8, rx)
    assert (sh.shift0, sh.step, sh.K) == (0, 1, 57)


def test_shard_ranges_and_global_peak():
    from pydsproutines_b200 import dist as xd
    for K, world in ((10, 3), (1 << 20, 8), (5, 8), (0, 2)):
        parts = [xd.shard_range(K, r, world) for r in range(world)]
        assert parts[0][0] == 0 and parts[-1][1] == K
        assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
        assert max(b - a for a, b in parts) - min(b - a for a, b in parts) <= 1
    first, n = xd.shard_samples(100, 16, 1, 4)
    assert (first, n) == (25, 25 + 15)
    rows = [xd.peak_to_row(0.5, 7, 900, 0, 0), xd.peak_to_row(0.9, 3, 5000, 0, 1),
            xd.peak_to_row(0.9, 4, 4000, 0, 2), xd.peak_to_row(0.0, 0, -1, 0, 3)]
    g = xd.global_peak(np.stack(rows))
    assert g == dict(qf2=0.9, fidx=4, shift=4000, aux=0, rank=2)             # tie -> lowest shift
    assert xd.global_peak(np.stack(rows[3:]))["rank"] == -1


_GLOO_WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, %(root)r)
import torch.distributed as dist
from pydsproutines_b200 import dist as xd, signals
from oracle import xcorr_oracle as orc     # the CPU checker stands in for the GPU search in this host-logic test
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
N, total = 64, 1000
cutout, rx = signals.make_fastxcorr_case(77, N, total, 700, 9, decoy=(100, 0.5, 20))
K = total - N + 1
first, n = xd.shard_samples(K, N, rank, world)
shard = rx[first:first + n]
q, f = orc.fastXcorr(cutout, shard, freqsearch=True)
k = int(np.argmax(q))
rows = xd.allgather_peak(xd.peak_to_row(q[k], f[k], first + k, 0, rank))
g = xd.global_peak(rows)
assert rows.shape == (world, 5)
assert g["shift"] == 700 and g["fidx"] == 9 and g["rank"] == 1, g
qa, fa = orc.fastXcorr(cutout, rx, freqsearch=True)
lo, hi = xd.shard_range(K, rank, world)
assert np.array_equal(q, qa[lo:hi]) and np.array_equal(f, fa[lo:hi])        # halo makes shards exact
dist.destroy_process_group()
sys.stdout.write("rank" + str(rank) + "-ok\n")
"""


def test_sharded_peak_exchange_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER % {"root": ROOT})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29517", str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "rank0-ok" in r.stdout and "rank1-ok" in r.stdout


_GLOO_WORKER_GROUP_MT = r"""
import os, sys, types
import numpy as np
sys.path.insert(0, %(root)r)
import torch.distributed as dist
from pydsproutines_b200 import dist as xd, signals
from oracle import xcorr_oracle as orc     # the CPU checker stands in for the GPU search in this host-logic test
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
rng = np.random.default_rng(5)
# ---- GroupXcorrFFT: 4 groups x 16, fftlen 64, halo = starts[-1] + L - 1; peak in rank 1, weaker decoy in rank 0 ----
G, L, F, K = 4, 16, 64, 240
starts = np.array([0, 40, 80, 120])
y = signals.qpsk(rng, G * L).reshape(G, L)
window = xd.group_window(starts, L)
rx = signals.cnoise(rng, K + window - 1)
for p0, k0, amp in ((170, 5, 1.0), (30, 11, 0.5)):
    for g in range(G):
        a = p0 + int(starts[g])
        rx[a:a + L] += np.complex64(amp) * y[g] * signals.tone(L, k0, F, offset=int(starts[g]))
first, n = xd.shard_samples(K, window, rank, world)
spec = types.SimpleNamespace(starts=starts, ygroupLen=L)
ref = orc.GroupXcorrFFT(y, starts, 1, fftlen=F)
pad = np.zeros(F + 1, np.complex64)          # the reference asserts fftlen samples of slack behind the last window (:1143)
q, f, g = xd.sharded_group_peak(spec, rx[first:first + n], first, rank, search=lambda r, sh: ref.xcorr(np.concatenate([r, pad]), sh))
assert g["shift"] == 170 and g["fidx"] == 5 and g["rank"] == 1, g
qa, fa = ref.xcorr(np.concatenate([rx, pad]), np.arange(K))
lo, hi = xd.shard_range(K, rank, world)
assert np.array_equal(q, qa[lo:hi]) and np.array_equal(f, fa[lo:hi])
if rank == 0:
    assert int(np.argmax(q)) == 30 and f[30] == 11                      # the decoy wins locally, loses globally
# ---- multi-template: 6 templates x 32, halo = L - 1; template 4 planted in rank 0, decoy template 1 in rank 1 ----
T, L, n_x = 6, 32, 400
templates = signals.qpsk(rng, T * L).reshape(T, L)
x = signals.cnoise(rng, n_x)
x[60:60 + L] += 2.0 * templates[4]
x[300:300 + L] += 1.0 * templates[1]
K = n_x - L + 1
first, n = xd.shard_samples(K, L, rank, world)
spec = types.SimpleNamespace(_templateLength=L)
def search(xs):
    ti, q = orc.multiTemplateSlidingDotProduct(xs, templates.conj(), 0, xs.size - L + 1)
    return q, ti
best, tidx, g = xd.sharded_mt_peak(spec, x[first:first + n], first, rank, search=search)
assert g["shift"] == 60 and g["fidx"] == 4 and g["rank"] == 0, g
if rank == 1:
    k = int(np.argmax(best)); assert first + k == 300 and tidx[k] == 1
# an empty shard never wins
_, _, g2 = xd.sharded_mt_peak(spec, x[:L - 1] if rank == 1 else x[first:first + n], first, rank, search=search)
assert g2["rank"] == 0 and g2["shift"] == 60
dist.destroy_process_group()
sys.stdout.write("rank" + str(rank) + "-ok\n")
"""


def test_sharded_group_and_template_search_gloo_world2(tmp_path):
    """SURVEY.md 8e halos of the other two correlators (starts[-1] + L - 1 for GroupXcorrFFT, L - 1 for the template
    correlator): shards reproduce the unsharded search exactly and the cross-rank reduction rejects a local decoy."""
    script = tmp_path / "worker2.py"
    script.write_text(_GLOO_WORKER_GROUP_MT % {"root": ROOT})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29519", str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "rank0-ok" in r.stdout and "rank1-ok" in r.stdout


def test_peak_argmax_rule(lib):
    """The deterministic cross-rank reduction of the C-ABI (largest QF^2, lowest shift on ties, negative shift = empty
    shard) agrees with pydsproutines_b200.dist.global_peak."""
    from pydsproutines_b200 import _lib, dist as xdist
    rows = [(0.5, 3, 900), (0.75, 4, 5000), (0.75, 9, 4000), (0.9, 1, -1)]
    arr = (_lib.xc_peak * len(rows))()
    for i, (q, f, s) in enumerate(rows):
        arr[i] = _lib.xc_peak(qf2=q, fidx=f, index=0, shift=s, aux=0, rank=i)
    assert lib.xc_peak_argmax(arr, len(rows)) == 2
    gp = xdist.global_peak([xdist.peak_to_row(q, f, s, 0, i) for i, (q, f, s) in enumerate(rows)])
    assert gp["rank"] == 2 and gp["shift"] == 4000
    assert lib.xc_peak_argmax(arr, 1) == 0 and lib.xc_peak_argmax(None, 0) == -1


def test_foreign_arrays_without_the_interface_are_left_alone():
    from pydsproutines_b200 import _lib
    a = np.ones(4, np.complex64)
    assert _lib.adopt(a) is a and _lib.adopt(None) is None and _lib.adopt(3) == 3

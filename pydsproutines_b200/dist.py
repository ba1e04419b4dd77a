"""Shift-range sharding of a search across the GPUs of one box (BASELINE.json north_star).

The reference has no multi-GPU code; every shift is independent, so rank r takes a contiguous block of
the shift list and the matching slice of rx plus an overlapped halo of (window - 1) samples.  Per-shift
outputs stay rank-local.  The ONE exchange step is an allgather of the per-rank peak tuple (32 bytes
each) over torch.distributed -- NCCL over NVLink on GPUs, gloo in the CPU tests -- followed by the same
deterministic arg max on every rank (largest QF^2, lowest shift on ties).
"""
import numpy as np

PEAK_FIELDS = ("qf2", "fidx", "shift", "aux", "rank")


def shard_range(K, rank, world):
    """Contiguous block [lo, hi) of the K shift-list entries owned by `rank` (sizes differ by at most 1)."""
    base, rem = divmod(int(K), int(world))
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    return lo, hi


def shard_samples(K, window, rank, world, shift0=0):
    """(first sample, number of samples) of rx the rank needs for its step-1 shifts: its block plus a halo of
    window-1 samples shared with the next rank."""
    lo, hi = shard_range(K, rank, world)
    if hi <= lo:
        return shift0 + lo, 0
    return shift0 + lo, (hi - lo) + window - 1


def peak_to_row(qf2, fidx, shift, aux, rank):
    """Packs a peak tuple into 4 float64 (exact for |values| < 2^53) + rank."""
    return np.array([float(qf2), float(fidx), float(shift), float(aux), float(rank)], dtype=np.float64)


def global_peak(rows):
    """Deterministic reduction of the gathered (world, 5) rows: largest QF^2, lowest shift on ties.
    Rows with a negative shift (empty shards) never win."""
    rows = np.asarray(rows, dtype=np.float64).reshape(-1, 5)
    best = None
    for r in rows:
        if r[2] < 0:
            continue
        if best is None or r[0] > best[0] or (r[0] == best[0] and r[2] < best[2]):
            best = r
    if best is None:
        return dict(qf2=0.0, fidx=0, shift=-1, aux=0, rank=-1)
    return dict(qf2=float(best[0]), fidx=int(best[1]), shift=int(best[2]), aux=int(best[3]), rank=int(best[4]))


def allgather_peak(row, device=None):
    """All ranks contribute their packed peak row; returns the (world, 5) float64 array on every rank.
    Uses the default process group (NCCL needs the tensor on the rank's GPU; gloo takes CPU tensors)."""
    import torch
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return np.asarray(row, dtype=np.float64).reshape(1, 5)
    world = dist.get_world_size()
    t = torch.from_numpy(np.asarray(row, dtype=np.float64))
    if device is not None:
        t = t.to(device)
    out = torch.empty(world * 5, dtype=torch.float64, device=t.device)
    dist.all_gather_into_tensor(out, t)
    return out.cpu().numpy().reshape(world, 5)


def sharded_fastxcorr_peak(cutout, rx_shard, shard_first_shift, rank, device=None, freqsearch=True):
    """One rank's part of a sharded fastXcorr: searches every shift of its shard (rx_shard holds the block
    plus halo), then exchanges peaks.  Returns (qf2_local, fidx_local, global_peak_dict)."""
    from . import xcorrRoutines as xr
    K_local = int(rx_shard.shape[0]) - int(cutout.shape[0]) + 1
    if K_local > 0:
        qf2, fidx, peak = xr.fastXcorrPeak(cutout, rx_shard, freqsearch=freqsearch)
        row = peak_to_row(peak.qf2, peak.fidx, shard_first_shift + peak.shift, 0, rank)
    else:
        qf2 = fidx = None
        row = peak_to_row(0.0, 0, -1, 0, rank)
    rows = allgather_peak(row, device)
    return qf2, fidx, global_peak(rows)


def sharded_peak(local, shard_first_shift, rank, device=None):
    """The exchange step shared by the three sharded correlators.  `local` = (QF^2, index (frequency bin or template),
    shift inside the shard) of the rank's best shift, or None for an empty shard."""
    if local is None:
        row = peak_to_row(0.0, 0, -1, 0, rank)
    else:
        q, k, s = local
        row = peak_to_row(q, k, shard_first_shift + int(s), 0, rank)
    return global_peak(allgather_peak(row, device))


def group_window(starts, L):
    """Samples one shift of a GroupXcorrFFT needs behind it: starts[-1] - starts[0] + L (the halo is one less)."""
    starts = np.asarray(starts, dtype=np.int64)
    return int(starts.max() - starts.min()) + int(L)


def sharded_group_peak(obj, rx_shard, shard_first_shift, rank, device=None, search=None):
    """One rank's part of a shift-sharded GroupXcorrFFT search (halo = starts[-1] + L - 1 samples):
    searches every shift of its shard (block + halo) and exchanges peaks.  `obj` is a GroupXcorrFFT; `search(rx, shifts)`
    -> (QF^2[K], bins[K]) replaces obj.xcorr in the CPU test.  Returns (qf2_local, fidx_local, global_peak_dict)."""
    window = group_window(obj.starts, obj.ygroupLen)
    K_local = int(rx_shard.shape[0]) - window + 1
    if K_local <= 0:
        return None, None, sharded_peak(None, shard_first_shift, rank, device)
    shifts = np.arange(K_local)
    if search is None:
        # the handle, not obj.xcorr: the reference's argument check wants fftlen samples of slack behind the last window
        # (xcorrRoutines.py:1143), which a halo of window - 1 samples does not have -- the kernels only read the windows
        from . import xcorrRoutines as xr
        from ._lib import contiguous
        rx_shard = contiguous(rx_shard)
        qf2, fidx, _, peak = obj._handle.run(rx_shard, xr._group_shifts(shifts, rx_shard), True, False, want_peak=True)
        local = (peak.qf2, peak.fidx, peak.shift)
    else:
        qf2, fidx = search(rx_shard, shifts)
        k = int(np.argmax(qf2))
        local = (qf2[k], fidx[k], k)
    return qf2, fidx, sharded_peak(local, shard_first_shift, rank, device)


def sharded_mt_peak(correlator, x_shard, shard_first_shift, rank, device=None, search=None):
    """One rank's part of a shift-sharded multi-template search (halo = L - 1 samples).  `correlator` is a
    TemplateCrossCorrelator (its handle gives QF^2 and the template index per shift); `search(x)` -> (QF^2[K], tidx[K])
    replaces it in the CPU test.  The peak tuple carries the template index in `fidx`."""
    L = int(correlator._templateLength)
    K_local = int(x_shard.shape[0]) - L + 1
    if K_local <= 0:
        return None, None, sharded_peak(None, shard_first_shift, rank, device)
    if search is None:
        best, tidx, _, peak = correlator._handle.run(x_shard, 0, K_local, True, False, False, want_peak=True)
        local = (peak.qf2, peak.aux, peak.shift)
    else:
        best, tidx = search(x_shard)
        k = int(np.argmax(best))
        local = (best[k], tidx[k], k)
    return best, tidx, sharded_peak(local, shard_first_shift, rank, device)

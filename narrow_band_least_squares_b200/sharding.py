"""Time-segment sharding of one long record across the GPUs of a box.

Band-windows are independent given the filtered samples, so rank g owns, in every band, the windows
whose START falls in [g*G/R, (g+1)*G/R).  It loads the samples those windows cover plus a filter
halo on both sides: the zero-phase IIR response of band b has decayed below 2^-80 after
K_b = ceil(80 ln2 / -ln r_b) samples (r_b = largest pole radius), so a segment filtered from zero
state agrees with the whole-record filter to below FP64 rounding inside the owned windows.  The
Hann taper is applied by GLOBAL sample index (Plan.global_t0 / global_npts).  No data-path
collective is needed; only the small per-window result arrays are gathered (NCCL under torchrun,
gloo in the CPU tests).

The reference has no counterpart (its only parallelism is joblib over bands,
narrow_band_least_squares.py:285); the sharded result must equal the unsharded one — tests check
lags and LTS flags bit-for-bit.
"""
from __future__ import annotations

import math

import numpy as np


def pole_radius(sos):
    """Largest pole modulus of a cascade of second-order sections (rows b0 b1 b2 1 a1 a2)."""
    r = 0.0
    for row in np.atleast_2d(sos):
        a1, a2 = float(row[4]), float(row[5])
        disc = a1 * a1 - 4.0 * a2
        if disc < 0:
            m = math.sqrt(abs(a2))
        else:
            m = max(abs((-a1 + math.sqrt(disc)) / 2), abs((-a1 - math.sqrt(disc)) / 2))
        r = max(r, m)
    return r


def filter_halo(sos_bands, bits=80):
    """Samples after which every band's impulse response is below 2^-bits."""
    k = 0
    for sos in sos_bands:
        r = pole_radius(sos)
        if r >= 1.0:
            raise ValueError("unstable or marginal filter section: no finite halo exists")
        if r > 0.0:
            k = max(k, int(math.ceil(bits * math.log(2.0) / -math.log(r))))
    return k


class Shard:
    """Rank ``rank`` of ``world``: owned windows per band and the sample segment to load."""

    def __init__(self, global_npts, win_len, win_hop, win_count, rank, world, halo):
        g = int(global_npts)
        self.rank, self.world, self.global_npts = rank, world, g
        lo_t, hi_t = (g * rank) // world, (g * (rank + 1)) // world
        nb = len(win_len)
        self.w_lo = np.zeros(nb, dtype=np.int64)
        self.w_hi = np.zeros(nb, dtype=np.int64)
        seg_lo, seg_hi = None, None
        for b in range(nb):
            n, h, w = int(win_len[b]), int(win_hop[b]), int(win_count[b])
            a = min(w, -(-lo_t // h))
            z = w if rank == world - 1 else min(w, -(-hi_t // h))
            self.w_lo[b], self.w_hi[b] = a, max(a, z)
            if z > a:
                s0, s1 = a * h, (z - 1) * h + n
                seg_lo = s0 if seg_lo is None else min(seg_lo, s0)
                seg_hi = s1 if seg_hi is None else max(seg_hi, s1)
        if seg_lo is None:
            seg_lo, seg_hi = lo_t, lo_t
        self.seg_lo = max(0, (seg_lo - halo) // 16 * 16)      # 128-B aligned start keeps the tiles aligned
        self.seg_hi = min(g, seg_hi + halo)
        if rank == 0:
            self.seg_lo = 0            # the record's true ends: the edge shards filter exactly like
        if rank == world - 1:
            self.seg_hi = g            # the unsharded run (zero state AT the record boundary)
        self.npts = self.seg_hi - self.seg_lo
        self.win_count = (self.w_hi - self.w_lo).astype(np.int32)
        self.win_first = np.array([int(self.w_lo[b]) * int(win_hop[b]) - self.seg_lo for b in range(nb)], dtype=np.int64)

    def plan(self, nchan, fs, sos, zerophase, win_len, win_hop, rij, **kw):
        from .engine import Plan
        return Plan(nchan, self.npts, fs, sos, zerophase, win_len, win_hop, self.win_count, rij,
                    win_first=self.win_first, global_t0=self.seg_lo, global_npts=self.global_npts, **kw)


def merge_band_major(parts, counts_per_rank):
    """Per-rank window-major arrays (rows = that rank's windows, band after band) -> the global
    window-major array (band after band, windows in time order inside a band).

    parts[r]: array with sum(counts_per_rank[r]) rows; counts_per_rank[r][b] = windows of band b on rank r.
    """
    nb = len(counts_per_rank[0])
    offs = [np.concatenate(([0], np.cumsum(c))) for c in counts_per_rank]
    rows = []
    for b in range(nb):
        for r, part in enumerate(parts):
            rows.append(part[offs[r][b]:offs[r][b + 1]])
    return np.concatenate(rows, axis=0) if rows else parts[0][:0]


def gather_results(local: dict, counts, group=None):
    """All-gather the per-window arrays of every rank with torch.distributed and merge them into
    global band-major order.  ``local``: name -> numpy array or torch tensor with one row per local
    window; ``counts``: this rank's windows per band.  Returns dict of numpy arrays on every rank.
    """
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dev_backend = dist.get_backend(group)
    counts_t = torch.as_tensor(np.asarray(counts, dtype=np.int64))
    if dev_backend == "nccl":
        counts_t = counts_t.cuda()
    all_counts = [torch.empty_like(counts_t) for _ in range(world)]
    dist.all_gather(all_counts, counts_t, group=group)
    all_counts = [c.cpu().numpy() for c in all_counts]
    totals = [int(c.sum()) for c in all_counts]
    mx = max(totals + [1])
    out = {}
    for name, arr in local.items():
        t = arr if isinstance(arr, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(arr))
        if dev_backend == "nccl" and not t.is_cuda:
            t = t.cuda()
        pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[:t.shape[0]] = t
        bufs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(bufs, pad, group=group)
        parts = [b[:totals[r]].cpu().numpy() for r, b in enumerate(bufs)]
        out[name] = merge_band_major(parts, all_counts)
    return out

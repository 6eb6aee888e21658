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


def all_shards(global_npts, win_len, win_hop, win_count, world, halo):
    """Every rank's Shard — pure arithmetic, so each rank knows every other rank's window counts without a collective."""
    return [Shard(global_npts, win_len, win_hop, win_count, r, world, halo) for r in range(world)]


class PackedGather:
    """The per-window estimates of every rank in ONE collective.

    Each rank's libnbls outputs (mdccm, mad, vel, baz, sigma_tau, slowness) are written by the kernels straight into
    one device block of 7 x e_max doubles (``outputs()`` hands libnbls the column pointers), so gathering is a single
    ``all_gather_into_tensor`` of that block — no padding copies, no per-array calls, no host round trip — followed by
    one index_select with a precomputed permutation into global band-major order (band after band, windows in time
    order).  Integer lags and LTS weights ride in a second / third block only when asked for.  Window counts per rank
    come from ``all_shards`` (no collective).  Works with NCCL (CUDA tensors) and gloo (CPU tensors; the CPU tests).
    """

    NCOL = 7      # mdccm, mad, vel, baz, sigma_tau, slowness[E][2]

    def __init__(self, counts_per_rank, rank, npairs=0, want_lag=False, want_weights=False, device="cuda", group=None):
        import torch
        self.group, self.rank = group, rank
        counts = np.asarray(counts_per_rank, dtype=np.int64)             # (world, B)
        self.world, self.nbands = counts.shape
        self.counts = counts
        totals = counts.sum(axis=1)
        self.e_local = int(totals[rank])
        self.e_max = int(max(totals.max(), 1))
        self.e_total = int(totals.sum())
        offs = np.concatenate((np.zeros((self.world, 1), np.int64), np.cumsum(counts, axis=1)), axis=1)
        idx = [r * self.e_max + offs[r, b] + np.arange(counts[r, b]) for b in range(self.nbands) for r in range(self.world)]
        self.index = torch.as_tensor(np.concatenate(idx) if idx else np.zeros(0, np.int64), device=device)
        self.local = torch.zeros((self.NCOL, self.e_max), dtype=torch.float64, device=device)
        self.all = torch.empty((self.world, self.NCOL, self.e_max), dtype=torch.float64, device=device)
        self.local_lag = self.all_lag = self.local_w = self.all_w = None
        if want_lag:
            self.local_lag = torch.zeros((self.e_max, npairs), dtype=torch.int32, device=device)
            self.all_lag = torch.empty((self.world, self.e_max, npairs), dtype=torch.int32, device=device)
        if want_weights:
            self.local_w = torch.zeros((self.e_max, npairs), dtype=torch.uint8, device=device)
            self.all_w = torch.empty((self.world, self.e_max, npairs), dtype=torch.uint8, device=device)

    def outputs(self, lag_ptr=None, weights_ptr=None):
        """_capi.Outputs whose members point into the packed block (device pointers for nbls_run_device)."""
        from . import _capi
        o = _capi.Outputs()
        base, step = self.local.data_ptr(), 8 * self.e_max
        o.mdccm, o.mad, o.vel, o.baz, o.sigma_tau, o.slowness = (base + i * step for i in range(6))
        o.lag = self.local_lag.data_ptr() if self.local_lag is not None else lag_ptr
        o.weights = self.local_w.data_ptr() if self.local_w is not None else weights_ptr
        o.cmax = None
        return o

    def gather(self):
        """-> dict of tensors in global band-major order on every rank (device-resident under NCCL)."""
        import torch.distributed as dist
        dist.all_gather_into_tensor(self.all.view(self.world * self.NCOL, self.e_max), self.local, group=self.group)
        flat = self.all.permute(1, 0, 2).reshape(self.NCOL, self.world * self.e_max)
        cols = flat[:5].index_select(1, self.index)
        out = dict(zip(("mdccm", "mad", "vel", "baz", "sigma_tau"), cols))
        slow = self.all[:, 5:7, :].reshape(self.world * self.e_max, 2)
        out["slowness"] = slow.index_select(0, self.index)
        if self.all_lag is not None:
            dist.all_gather_into_tensor(self.all_lag.view(self.world * self.e_max, -1), self.local_lag, group=self.group)
            out["lag"] = self.all_lag.reshape(self.world * self.e_max, -1).index_select(0, self.index)
        if self.all_w is not None:
            dist.all_gather_into_tensor(self.all_w.view(self.world * self.e_max, -1), self.local_w, group=self.group)
            out["weights"] = self.all_w.reshape(self.world * self.e_max, -1).index_select(0, self.index)
        return out

"""Seeded synthetic plane-wave records for tests, smoke and bench.

Host-side numpy only.  A record is C channels of  s(t - tau_i) + sigma * n_i(t)
with s unit-variance white noise delayed per element by linear interpolation
of an over-length realisation (sub-sample delays, residuals never vanish),
tau_i = -(s_x x_i + s_y y_i), (s_x, s_y) = (sin baz, cos baz) / v.
Every 10th minute is noise-only.  LTS cases flip one channel's polarity and
shift another by 0.5 s (mirrors lts_array/example.py:69-73 of the reference).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from .geometry import get_rij


@dataclass
class SynthRecord:
    data: np.ndarray            # (C, T) float64, channel-major
    fs: float
    lat: list
    lon: list
    rij: np.ndarray             # (2, C) km (from lat/lon through the Vincenty inverse)
    baz: float
    vel: float
    starttime: float = 0.0
    meta: dict = field(default_factory=dict)

    @property
    def nchans(self):
        return self.data.shape[0]

    @property
    def npts(self):
        return self.data.shape[1]


def disc_geometry(rng, nchans, aperture_km, lat0=64.0, lon0=-147.0):
    """C elements uniform in a disc of diameter ``aperture_km`` about (lat0, lon0)."""
    r = 0.5 * aperture_km * np.sqrt(rng.random(nchans))
    th = 2.0 * np.pi * rng.random(nchans)
    x, y = r * np.cos(th), r * np.sin(th)
    x -= x.mean()
    y -= y.mean()
    lat = lat0 + y / 111.19
    lon = lon0 + x / (111.19 * np.cos(np.radians(lat0)))
    return lat.tolist(), lon.tolist()


def plane_wave_record(seed, nchans, npts, fs, aperture_km, baz=135.0, vel=0.34, sigma=0.3,
                      noise_every_min=10, flip_channel=None, shift_channel=None, shift_s=0.5,
                      dead_channel=None, dtype=np.float64) -> SynthRecord:
    rng = np.random.default_rng(seed)
    lat, lon = disc_geometry(rng, nchans, aperture_km)
    rij = get_rij(lat, lon, nchans)
    sx, sy = np.sin(np.radians(baz)) / vel, np.cos(np.radians(baz)) / vel
    tau = -(sx * rij[0] + sy * rij[1])                      # seconds, per element
    if shift_channel is not None:
        tau = tau.copy()
        tau[shift_channel] += shift_s
    pad = int(np.ceil(np.max(np.abs(tau)) * fs)) + 2
    s = rng.standard_normal(npts + 2 * pad)
    if noise_every_min:
        # zero the coherent signal during every 10th minute (noise-only windows)
        t_min = (np.arange(npts + 2 * pad) - pad) / fs / 60.0
        s[(np.floor(t_min).astype(np.int64) % noise_every_min) == (noise_every_min - 1)] = 0.0
    data = np.empty((nchans, npts), dtype=dtype)
    base = np.arange(npts, dtype=np.float64) + pad
    for i in range(nchans):
        pos = base - tau[i] * fs
        k = np.floor(pos).astype(np.int64)
        frac = pos - k
        data[i] = (1.0 - frac) * s[k] + frac * s[k + 1] + sigma * rng.standard_normal(npts)
    if flip_channel is not None:
        data[flip_channel] *= -1.0
    if dead_channel is not None:
        data[dead_channel] = 0.0
    return SynthRecord(data=data, fs=float(fs), lat=lat, lon=lon, rij=rij, baz=baz, vel=vel,
                       meta=dict(seed=seed, sigma=sigma, aperture_km=aperture_km))


# BASELINE.json configs (SURVEY.md §8(a) header): name -> parameters
CONFIGS = {
    "c1": dict(seed=1001, nchans=4, hours=1.0, fs=40.0, nbands=8, alpha=1.0, aperture_km=0.2,
               winlen=("constant", 30.0), fmin=0.5, fmax=5.0),
    "c2": dict(seed=1002, nchans=6, hours=6.0, fs=100.0, nbands=12, alpha=0.75, aperture_km=0.5,
               winlen=("constant", 30.0), fmin=0.5, fmax=5.0, flip_channel=3, shift_channel=1),
    "c3": dict(seed=1003, nchans=25, hours=24.0, fs=100.0, nbands=20, alpha=1.0, aperture_km=1.0,
               winlen=("adaptive", 60.0, 30.0), fmin=0.5, fmax=5.0),
    "c4": dict(seed=1004, nchans=12, hours=720.0, fs=100.0, nbands=30, alpha=0.5, aperture_km=0.8,
               winlen=("constant", 30.0), fmin=0.5, fmax=5.0, flip_channel=3, shift_channel=1),
    "c5": dict(seed=1005, nchans=8, hours=24.0, fs=100.0, nbands=20, alpha=1.0, aperture_km=0.3,
               winlen=("constant", 30.0), fmin=0.5, fmax=5.0, narrays=64),
}


def config_record(name, hours=None, seed_offset=0) -> SynthRecord:
    c = CONFIGS[name]
    h = c["hours"] if hours is None else hours
    npts = int(round(h * 3600.0 * c["fs"]))
    return plane_wave_record(c["seed"] + seed_offset, c["nchans"], npts, c["fs"], c["aperture_km"],
                             flip_channel=c.get("flip_channel"), shift_channel=c.get("shift_channel"))


# ---- block-wise records: any segment of a long record can be synthesised on its own --------------------------------
# A long record is the concatenation of BLOCK-sample blocks, block k drawn from its own seed.  Ranks of a time-sharded
# run (sharding.py) therefore build exactly the samples of their segment without anybody holding the whole record, and
# a small record can be built whole AND in shards to check the multi-GPU path against the single-GPU one.
BLOCK = 360000      # one hour at 100 sps


def _block_seed(seed, k):
    return (int(seed) * 1000003 + 7919 * int(k) + 17) % (2 ** 63 - 1)


def record_geometry(seed, nchans, aperture_km):
    """(lat, lon, rij, tau0) of the array of a block-wise record: geometry from rng(seed), plane-wave delays in s."""
    rng = np.random.default_rng(seed)
    lat, lon = disc_geometry(rng, nchans, aperture_km)
    rij = get_rij(lat, lon, nchans)
    return lat, lon, rij


def _delays(rij, baz, vel, shift_channel, shift_s):
    sx, sy = np.sin(np.radians(baz)) / vel, np.cos(np.radians(baz)) / vel
    tau = -(sx * rij[0] + sy * rij[1])
    if shift_channel is not None:
        tau = tau.copy()
        tau[shift_channel] += shift_s
    return tau


def blocked_record(seed, nchans, t0, t1, fs, aperture_km, baz=135.0, vel=0.34, sigma=0.3, noise_every_min=10,
                   flip_channel=None, shift_channel=None, shift_s=0.5, block=BLOCK) -> SynthRecord:
    """Samples [t0, t1) of the block-wise record ``seed`` (numpy, host).  Same signal model as plane_wave_record."""
    lat, lon, rij = record_geometry(seed, nchans, aperture_km)
    tau = _delays(rij, baz, vel, shift_channel, shift_s)
    pad = int(np.ceil(np.max(np.abs(tau)) * fs)) + 2
    data = np.empty((nchans, t1 - t0), dtype=np.float64)
    for k in range(t0 // block, (t1 - 1) // block + 1):
        rng = np.random.default_rng(_block_seed(seed, k))
        s = rng.standard_normal(block + 2 * pad)
        if noise_every_min:
            t_min = (np.arange(block + 2 * pad) - pad + k * block) / fs / 60.0
            s[(np.floor(t_min).astype(np.int64) % noise_every_min) == (noise_every_min - 1)] = 0.0
        a, b = max(t0, k * block), min(t1, (k + 1) * block)
        base = np.arange(a - k * block, b - k * block, dtype=np.float64) + pad
        for i in range(nchans):
            noise = rng.standard_normal(block)[a - k * block:b - k * block]
            pos = base - tau[i] * fs
            kk = np.floor(pos).astype(np.int64)
            frac = pos - kk
            data[i, a - t0:b - t0] = (1.0 - frac) * s[kk] + frac * s[kk + 1] + sigma * noise
    if flip_channel is not None:
        data[flip_channel] *= -1.0
    return SynthRecord(data=data, fs=float(fs), lat=lat, lon=lon, rij=rij, baz=baz, vel=vel,
                       meta=dict(seed=seed, sigma=sigma, aperture_km=aperture_km, t0=t0, t1=t1, block=block))


def blocked_record_cuda(seed, nchans, t0, t1, fs, aperture_km, device, baz=135.0, vel=0.34, sigma=0.3,
                        noise_every_min=10, flip_channel=None, shift_channel=None, shift_s=0.5, block=BLOCK, ld=None):
    """The same record model synthesised on the GPU with torch (bench plumbing for records too long to draw with
    numpy in reasonable time: c4 is 3.1e9 samples).  Returns (tensor [nchans, ld] float64 on ``device``, rij).
    Deterministic for a given seed and GPU model; NOT bit-identical to ``blocked_record`` (different generator)."""
    import torch
    lat, lon, rij = record_geometry(seed, nchans, aperture_km)
    tau = _delays(rij, baz, vel, shift_channel, shift_s)
    pad = int(np.ceil(np.max(np.abs(tau)) * fs)) + 2
    n = t1 - t0
    ld = n if ld is None else ld
    out = torch.zeros((nchans, ld), dtype=torch.float64, device=device)
    gen = torch.Generator(device=device)
    for k in range(t0 // block, (t1 - 1) // block + 1):
        gen.manual_seed(_block_seed(seed, k))
        s = torch.randn(block + 2 * pad, dtype=torch.float64, device=device, generator=gen)
        if noise_every_min:
            idx = torch.arange(block + 2 * pad, device=device, dtype=torch.float64) - pad + k * block
            minute = torch.floor(idx / fs / 60.0).to(torch.int64)
            s = torch.where(minute % noise_every_min == noise_every_min - 1, torch.zeros_like(s), s)
        a, b = max(t0, k * block), min(t1, (k + 1) * block)
        base = torch.arange(a - k * block, b - k * block, device=device, dtype=torch.float64) + pad
        for i in range(nchans):
            noise = torch.randn(block, dtype=torch.float64, device=device, generator=gen)[a - k * block:b - k * block]
            pos = base - float(tau[i]) * fs
            kk = torch.floor(pos)
            frac = pos - kk
            kk = kk.to(torch.int64)
            v = (1.0 - frac) * s[kk] + frac * s[kk + 1] + sigma * noise
            out[i, a - t0:b - t0] = -v if flip_channel == i else v
    return out, rij, lat, lon

"""Array geometry on the host: lat/lon -> zero-mean km coordinates -> co-array.

Mirrors (does not copy) the reference's
  * ``DataBin.getrij``            lts_array/lts_array/classes/lts_data_class.py:82-124
  * ``helpers.get_rij``           helpers.py:239-284
  * ``LsBeam.calculate_co_array`` lts_array/lts_array/classes/lts_classes.py:836-850
The Vincenty inverse replaces ``obspy.geodetics.base.calc_vincenty_inverse``
(obspy is not a dependency of this package); it iterates each element to
|d lambda| < 1e-12 on WGS84.
"""
from __future__ import annotations

import math

import numpy as np

WGS84_A = 6378137.0
WGS84_F = 1.0 / 298.257223563


def _vincenty_one(lat0, lon0, lat1, lon1):
    a, f = WGS84_A, WGS84_F
    b = a * (1.0 - f)
    wrap = lambda x: (x + 180.0) % 360.0 - 180.0  # noqa: E731
    lon0, lon1 = wrap(lon0), wrap(lon1)
    if abs(lat0 - lat1) < 1e-8 and abs(lon0 - lon1) < 1e-8:
        return 0.0, 0.0
    dlon = math.radians(lon1 - lon0)
    u0 = math.atan((1.0 - f) * math.tan(math.radians(lat0)))
    u1 = math.atan((1.0 - f) * math.tan(math.radians(lat1)))
    s0, c0, s1, c1 = math.sin(u0), math.cos(u0), math.sin(u1), math.cos(u1)
    lam = dlon
    for _ in range(1000):
        sl, cl = math.sin(lam), math.cos(lam)
        sin_sig = math.hypot(c1 * sl, c0 * s1 - s0 * c1 * cl)
        if sin_sig == 0.0:
            return 0.0, 0.0
        cos_sig = s0 * s1 + c0 * c1 * cl
        sig = math.atan2(sin_sig, cos_sig)
        sin_al = c0 * c1 * sl / sin_sig
        cos2_al = 1.0 - sin_al * sin_al
        cos_2sm = cos_sig - 2.0 * s0 * s1 / cos2_al if cos2_al != 0.0 else 0.0
        cc = f / 16.0 * cos2_al * (4.0 + f * (4.0 - 3.0 * cos2_al))
        lam_new = dlon + (1.0 - cc) * f * sin_al * (
            sig + cc * sin_sig * (cos_2sm + cc * cos_sig * (2.0 * cos_2sm ** 2 - 1.0)))
        done = abs(lam_new - lam) < 1e-12
        lam = lam_new
        if done:
            break
    sl, cl = math.sin(lam), math.cos(lam)
    usq = cos2_al * (a * a - b * b) / (b * b)
    aa = 1.0 + usq / 16384.0 * (4096.0 + usq * (-768.0 + usq * (320.0 - 175.0 * usq)))
    bb = usq / 1024.0 * (256.0 + usq * (-128.0 + usq * (74.0 - 47.0 * usq)))
    dsig = bb * sin_sig * (cos_2sm + bb / 4.0 * (
        cos_sig * (2.0 * cos_2sm ** 2 - 1.0)
        - bb / 6.0 * cos_2sm * (4.0 * sin_sig ** 2 - 3.0) * (4.0 * cos_2sm ** 2 - 3.0)))
    return b * aa * (sig - dsig), math.degrees(math.atan2(c1 * sl, c0 * s1 - s0 * c1 * cl)) % 360.0


def vincenty_inverse(lat0: float, lon0: float, lat, lon):
    """Geodesic distance [m] and forward azimuth [deg, cw from N] from one point to many (WGS84)."""
    lat = np.atleast_1d(np.asarray(lat, dtype=np.float64))
    lon = np.atleast_1d(np.asarray(lon, dtype=np.float64))
    if np.any(np.abs(lat) > 90) or abs(lat0) > 90:
        raise ValueError("latitude out of range")
    res = [_vincenty_one(float(lat0), float(lon0), float(a), float(b)) for a, b in zip(lat, lon)]
    return np.array([r[0] for r in res]), np.array([r[1] for r in res])


def get_rij(latlist, lonlist, nchans=None):
    """Zero-mean Cartesian element coordinates [km], shape (2, C): row 0 = x (east), row 1 = y (north).

    Same contract and error as the reference's ``get_rij`` / ``DataBin.getrij``
    (ValueError when the list lengths differ from the channel count).
    """
    if nchans is None:
        nchans = len(latlist)
    if len(latlist) != nchans or len(lonlist) != nchans:
        raise ValueError('Mismatch between the number of stream channels and the latitude or longitude list length.')
    lat = np.asarray(latlist, dtype=np.float64)
    lon = np.asarray(lonlist, dtype=np.float64)
    x = np.zeros(nchans)
    y = np.zeros(nchans)
    if nchans > 1:
        dist, az = vincenty_inverse(float(lat[0]), float(lon[0]), lat[1:], lon[1:])
        ang = ((450.0 - az) % 360.0) * np.pi / 180.0
        x[1:] = dist / 1000.0 * np.cos(ang)
        y[1:] = dist / 1000.0 * np.sin(ang)
    x -= np.mean(x)
    y -= np.mean(y)
    return np.array([x.tolist(), y.tolist()])


def pair_index(nchans: int) -> np.ndarray:
    """Co-array pair list (P, 2) int32, i<j, row-major — lts_classes.py:840."""
    ii, jj = np.triu_indices(nchans, k=1)
    return np.stack([ii, jj], axis=1).astype(np.int32)


class CoArray:
    """Co-array geometry shared by every band and window (lts_classes.py:836-850)."""

    def __init__(self, rij: np.ndarray):
        rij = np.asarray(rij, dtype=np.float64)
        self.nchans = rij.shape[1]
        self.idx = pair_index(self.nchans)
        self.idx_pair = [(int(a), int(b)) for a, b in self.idx]
        self.xij = (rij[:, self.idx[:, 0]] - rij[:, self.idx[:, 1]]).T.copy()   # (P, 2)
        self.co_array_num = self.xij.shape[0]
        self.xij_mad = np.array([1.4826 * np.median(np.abs(self.xij[:, d])) for d in range(2)])
        self.xij_standardized = self.xij / self.xij_mad[None, :]

    def check_rank(self):
        if np.linalg.matrix_rank(self.xij) < 2:
            raise RuntimeError('Co-array is ill posed for the least squares problem. Check array coordinates. xij rank < 2')

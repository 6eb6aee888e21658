"""Host-side tables for the FAST-LTS kernel: the seeded subset stream, h, and the correction factors.

All of these depend only on (P, alpha) and are uploaded once per job.  They mirror, without
copying, lts_array/lts_array/classes/lts_classes.py of the reference:
  random_set            :16-36     (16-bit LCG, restarted at seed 0 for every window, :76)
  LTSEstimator.h_calc   :1013
  raw_corfactor_lts     :156-218   rew_corfactor_lts :260-323   (Pison et al. 2002)
  raw_consfactor_lts    :221-245   rew_consfactor_lts :326-344
"""
from __future__ import annotations

import functools
import math

import numpy as np
from scipy.special import erfinv

N_STARTS = 500        # lts_classes.py:991
N_KEEP = 10           # :993
CSTEPS1 = 4           # :995
CSTEPS2 = 100         # :997


@functools.lru_cache(maxsize=64)
def subset_stream(npairs: int, count: int = N_STARTS + N_KEEP) -> np.ndarray:
    """(count, 2) int32 pair indices drawn by the reference's LCG from seed 0.

    seed <- (seed*5761 + 999) mod 65536 ; index = floor(seed/65536 * P) ; the second index is
    redrawn while it equals the first.  seed*5761+999 < 2**53, so integer arithmetic reproduces the
    reference's float floor/divide exactly.
    """
    out = np.zeros((count, 2), dtype=np.int32)
    seed = 0
    for i in range(count):
        first = -1
        for k in range(2):
            while True:
                seed = (seed * 5761 + 999) % 65536
                num = int(math.floor((seed / 65536) * npairs))
                if k == 0 or num != first:
                    break
            if k == 0:
                first = num
            out[i, k] = num
    out.setflags(write=False)
    return out


def h_value(npairs: int, alpha: float, ndim: int = 2) -> int:
    half = np.floor((npairs + ndim + 1) / 2)
    return int(np.floor(2 * half - npairs + 2 * (npairs - half) * alpha))


def _pison_factor(c500, c875, p, n, alpha):
    def one(co):
        y1 = 1 + co[0][0] / np.power(p, co[1][0])
        y2 = 1 + co[0][1] / np.power(p, co[1][1])
        rhs = np.array([np.log(1 - y1), np.log(1 - y2)]).reshape(2, 1)
        lhs = np.array([[1.0, np.log(1.0 / (co[2][0] * p ** 2))], [1.0, np.log(1.0 / (co[2][1] * p ** 2))]])
        sol = np.linalg.lstsq(lhs, rhs, rcond=-1)[0]
        return 1.0 - np.exp(sol[0]) / np.power(n, sol[1])
    f5, f8 = one(c500), one(c875)
    if 0.5 <= alpha <= 0.875:
        f = f5 + (f8 - f5) / 0.375 * (alpha - 0.5)
    elif 0.875 < alpha < 1:
        f = f8 + (1 - f8) / 0.125 * (alpha - 0.875)
    else:
        raise ValueError("alpha must lie in [0.5, 1)")
    return float(np.ndarray.item(1.0 / f))


def raw_corfactor(p, n, alpha):
    return _pison_factor([[-0.487338281979106, -0.340762058011], [0.405511279418594, 0.37972360544988], [3, 5]],
                         [[-0.251778730491252, -0.146660023184295], [0.883966931611758, 0.86292940340761], [3, 5]],
                         p, n, alpha)


def rew_corfactor(p, n, alpha):
    return _pison_factor([[-0.417574780492848, -0.175753709374146], [1.83958876341367, 1.8313809497999], [3, 5]],
                         [[-0.267522855927958, -0.161200683014406], [1.17559984533974, 1.21675019853961], [3, 5]],
                         p, n, alpha)


def raw_consfactor(h, n):
    x = (h + n) / (2 * n)
    q = np.sqrt(2) * erfinv(2 * x - 1)
    phi = (1 / np.sqrt(2 * np.pi)) * np.exp(-0.5 * q ** 2)
    return float(1 / np.sqrt(1 - (2 * n / (h * (1 / q))) * phi))


def rew_consfactor_table(n):
    """rew_consfactor_lts as a function of the number of unit weights, entries 0..n (NaN where undefined)."""
    tab = np.full(n + 1, np.nan)
    with np.errstate(all="ignore"):
        for nw in range(1, n):
            q = erfinv(2 * ((nw + n) / (2 * n)) - 1) * np.sqrt(2)
            a = (1 / np.sqrt(2 * np.pi)) * np.exp(-0.5 * (1 / (1 / q)) ** 2)
            b = 1 / q
            tab[nw] = 1 / np.sqrt(1 - ((2 * n) / (nw * b)) * a)
    return tab


@functools.lru_cache(maxsize=64)
def lts_constants(npairs: int, alpha: float):
    """(h, raw_factor, rew_factor1, rew_cons table) for a co-array of ``npairs`` and trimming ``alpha``."""
    h = h_value(npairs, alpha)
    raw = raw_corfactor(2, npairs, alpha) * raw_consfactor(h, npairs)
    rew1 = rew_corfactor(2, npairs, alpha)
    return h, raw, rew1, rew_consfactor_table(npairs)

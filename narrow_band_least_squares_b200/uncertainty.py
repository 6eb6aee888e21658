"""Slowness-ellipse uncertainties ``conf_int_vel`` / ``conf_int_baz``.

Replaces the per-window Python of the reference (Szuberla & Olson 2004 as used at
lts_classes.py:715-759 for LTS and :895-968 for OLS, with ``rthEllipse`` :527-640 and the
quartic/cubic/quadratic root finders :347-524) by ONE vectorised host pass over all windows.
The geometry is solved directly instead of through quartic roots:

* the p-confidence ellipse of the slowness estimate is centred on ``s = (sx, sy)`` with semi-axes
  ``sqrt(chi2) * sigma_tau / sqrt(lambda_k)`` along the eigenvectors of ``X^T X`` (the co-array
  restricted to the pairs with unit weight for LTS);
* ``conf_int_vel = |1/d_min - 1/d_max| / 2`` with d the extremal distances from the origin to the
  ellipse — stationary points of ``|p(t)|^2`` on the parametrised ellipse, bracketed on a grid and
  polished by safeguarded Newton steps;
* ``conf_int_baz`` = half the angle between the two tangent rays from the origin — a homogeneous
  quadratic in the ray direction, solved in closed form.

Where the origin lies outside the ellipse (every window with a usable estimate) this and the reference's
closed-form machinery give the same numbers to ~1e-9.  If the origin lies INSIDE the ellipse no tangent exists;
the reference then still returns numbers — the tangency quadratic of ``rthEllipse`` (lts_classes.py:617-633) has
complex roots and it keeps their real parts, and the extremal distances (:588-615) stay real — so for identical
results ``_tangent_half_angle_inside`` evaluates that same quadratic in complex arithmetic and keeps the real
parts, and ``conf_int_vel`` uses the (finite) extremal distances.  NBLS itself discards both outputs
(narrow_band_least_squares.py:91).
"""
from __future__ import annotations

import numpy as np

_NGRID = 512


def _sym2_eigh(m00, m01, m11):
    """Ascending eigenvalues and the unit eigenvector of the SMALLER one for 2x2 symmetric matrices."""
    half_tr = 0.5 * (m00 + m11)
    d = 0.5 * (m00 - m11)
    r = np.hypot(d, m01)
    l0, l1 = half_tr - r, half_tr + r
    # eigenvector of l0: (m01, l0 - m00) or (l0 - m11, m01), whichever is better conditioned
    vx = np.where(d > 0, m01, l0 - m11)
    vy = np.where(d > 0, l0 - m00, m01)
    nrm = np.hypot(vx, vy)
    iso = nrm == 0                      # isotropic co-array: any direction
    vx = np.where(iso, 1.0, vx / np.where(iso, 1.0, nrm))
    vy = np.where(iso, 0.0, vy / np.where(iso, 1.0, nrm))
    return l0, l1, vx, vy


def _extremal_distances(a, b, x0, y0):
    """min / max distance from the origin to the ellipse (x0 + a cos t, y0 + b sin t), vectorised."""
    t = np.linspace(0.0, 2.0 * np.pi, _NGRID, endpoint=False)
    ct, st = np.cos(t), np.sin(t)
    px = x0[:, None] + a[:, None] * ct
    py = y0[:, None] + b[:, None] * st
    d2 = px * px + py * py
    out = []
    for pick in (np.argmin, np.argmax):
        tk = t[pick(d2, axis=1)]
        h = 2.0 * np.pi / _NGRID
        lo, hi = tk - h, tk + h
        for _ in range(60):
            c, s = np.cos(tk), np.sin(tk)
            # g = d/dt (|p|^2 / 2), g' its derivative
            g = -a * s * (x0 + a * c) + b * c * (y0 + b * s)
            gp = -a * c * (x0 + a * c) + a * a * s * s - b * s * (y0 + b * s) + b * b * c * c
            with np.errstate(divide="ignore", invalid="ignore"):
                step = np.where(gp != 0, g / gp, 0.0)
            tn = np.clip(tk - step, lo, hi)
            if np.all(np.abs(tn - tk) <= 1e-16 * (1.0 + np.abs(tk))):
                tk = tn
                break
            tk = tn
        out.append(np.hypot(x0 + a * np.cos(tk), y0 + b * np.sin(tk)))
    return out[0], out[1]


def _tangent_half_angle(a, b, x0, y0):
    """Half the angle (degrees) between the tangent rays from the origin; 180 if the origin is inside."""
    ia2, ib2 = 1.0 / (a * a), 1.0 / (b * b)
    e = x0 * x0 * ia2 + y0 * y0 * ib2 - 1.0            # > 0: origin outside the ellipse
    # ray (c, s) is tangent iff  P c^2 + 2 Q c s + R s^2 = 0
    p = x0 * x0 * ia2 * ia2 - e * ia2
    q = x0 * y0 * ia2 * ib2
    r = y0 * y0 * ib2 * ib2 - e * ib2
    hyp = np.hypot(0.5 * (p - r), q)
    phi = np.arctan2(q, 0.5 * (p - r))
    with np.errstate(divide="ignore", invalid="ignore"):
        arg = np.clip(-(p + r) / (2.0 * hyp), -1.0, 1.0)
    dlt = np.arccos(arg)
    ang = []
    for sgn in (1.0, -1.0):
        th = 0.5 * (phi + sgn * dlt)
        c, s = np.cos(th), np.sin(th)
        reach = (c * x0 * ia2 + s * y0 * ib2)          # sign of the tangency point's ray parameter
        ang.append(np.where(reach < 0, th + np.pi, th))
    diff = np.abs(np.degrees(ang[0] - ang[1])) % 360.0
    diff = np.where(diff > 180.0, 360.0 - diff, diff)
    return np.where(e > 0, 0.5 * diff, 180.0), e


def _tangent_half_angle_inside(a, b, x0, y0):
    """conf_int_baz (degrees) the reference returns when the origin lies inside the ellipse: the tangency conditions of
    rthEllipse (lts_classes.py:561-568 constants, :617-633) solved over the complex numbers, real parts kept.

    With the ellipse axis-aligned, G = 1/a^2, H = 1/b^2, B = 2 x0/a^2, D = 2 y0/b^2, E = x0^2/a^2 + y0^2/b^2 - 1, the
    tangent points (x, y) satisfy  (D^2 + B^2 H/G) y'^2 + 4 D E y' + (4 E^2 - B^2 E/G) = 0,  y = -y',
    x = +-sqrt(E/G - (H/G) y^2); E < 0 inside, so the discriminant is negative and the roots are complex conjugates.
    """
    G, H = 1.0 / (a * a), 1.0 / (b * b)
    B, D = 2.0 * x0 * G, 2.0 * y0 * H
    E = 0.5 * (B * x0 + D * y0) - 1.0
    qa = D * D + B * B * H / G
    qb = 4.0 * D * E
    qc = 4.0 * E * E - B * B * E / G
    disc = np.sqrt((qb * qb - 4.0 * qa * qc).astype(np.complex128))
    # the numerically stable pair of roots (Numerical Recipes 5.6, as quadraticEqn :438-457): q/a and c/q
    with np.errstate(divide="ignore", invalid="ignore"):
        q = np.where(qb != 0, -0.5 * (qb + np.sign(qb) * disc), -np.sqrt((-qa * qc).astype(np.complex128)))
        r0, r1 = q / qa, qc / q
        ys = [-r0, -r1]
        xs = [np.sqrt(E / G - (H / G) * y * y) for y in ys]
    neg = x0 < 0
    xs = [np.where(neg, -x, x) for x in xs]
    ex = [np.real(x) for x in xs]
    ey = [np.real(y) for y in ys]
    flip = (x0 == 0) | (np.abs(x0) - a < 0)           # "various quadrant fixes" (:631-632): first tangent point mirrored
    ex[0] = np.where(flip, -ex[0], ex[0])
    ang = [(np.degrees(np.arctan2(y, x)) - 360.0) % 360.0 for x, y in zip(ex, ey)]
    d = np.abs(ang[1] - ang[0])
    d = np.where(d > 180.0, np.abs(d - 360.0), d)
    return 0.5 * d


def slowness_ellipse_ci(slowness, sigma_tau, xij, weights=None, p=0.90):
    """conf_int_vel, conf_int_baz for W windows.

    slowness (W, 2) final slowness vectors; sigma_tau (W,); xij (P, 2) co-array in km;
    weights (W, P) 0/1 final LTS weights or None (OLS: every pair).  NaN estimates give NaN.
    """
    slowness = np.asarray(slowness, dtype=np.float64).reshape(-1, 2)
    sigma_tau = np.asarray(sigma_tau, dtype=np.float64).reshape(-1)
    xij = np.asarray(xij, dtype=np.float64)
    w = slowness.shape[0]
    civ, cib = np.full(w, np.nan), np.full(w, np.nan)
    if w == 0:
        return civ, cib
    mom = np.stack([xij[:, 0] ** 2, xij[:, 0] * xij[:, 1], xij[:, 1] ** 2], axis=1)      # (P, 3)
    if weights is None:
        m = np.broadcast_to(mom.sum(axis=0), (w, 3))
    else:
        m = np.asarray(weights, dtype=np.float64) @ mom
    ok = np.isfinite(slowness).all(axis=1) & np.isfinite(sigma_tau) & (sigma_tau > 0)
    l0, l1, vx, vy = _sym2_eigh(m[:, 0], m[:, 1], m[:, 2])
    ok &= l0 > 0
    if not ok.any():
        return civ, cib
    i = np.where(ok)[0]
    chi = np.sqrt(-2.0 * np.log(1.0 - p))               # lts_classes.py:904, 2 degrees of freedom
    a = chi * sigma_tau[i] / np.sqrt(l0[i])
    b = chi * sigma_tau[i] / np.sqrt(l1[i])
    sx, sy = slowness[i, 0], slowness[i, 1]
    x0 = vx[i] * sx + vy[i] * sy                        # rotate so the ellipse axes are the coordinate axes
    y0 = -vy[i] * sx + vx[i] * sy
    half, e = _tangent_half_angle(a, b, x0, y0)
    dmin, dmax = _extremal_distances(a, b, x0, y0)
    with np.errstate(divide="ignore"):
        vel_ci = 0.5 * np.abs(1.0 / dmin - 1.0 / dmax)
    civ[i] = vel_ci
    inside = ~(e > 0)
    if inside.any():
        half = half.copy()
        half[inside] = _tangent_half_angle_inside(a[inside], b[inside], x0[inside], y0[inside])
    cib[i] = half
    return civ, cib

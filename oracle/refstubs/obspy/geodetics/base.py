"""Vincenty inverse on the WGS84 ellipsoid (stand-in for obspy.geodetics.base).

Restated from Vincenty (1975), "Direct and inverse solutions of geodesics on
the ellipsoid with application of nested equations", Survey Review 23(176).
Returns (distance [m], azimuth A->B [deg], back azimuth B->A [deg]) like
obspy's function of the same name, which the reference calls at
lts_data_class.py:111 and helpers.py:271.
"""
import math

WGS84_A = 6378137.0
WGS84_F = 1.0 / 298.257223563


def calc_vincenty_inverse(lat1, lon1, lat2, lon2, a=WGS84_A, f=WGS84_F):
    if abs(lat1) > 90 or abs(lat2) > 90:
        raise ValueError("latitude out of range")
    lon1 = (lon1 + 180.0) % 360.0 - 180.0
    lon2 = (lon2 + 180.0) % 360.0 - 180.0
    if abs(lat1 - lat2) < 1e-8 and abs(lon1 - lon2) < 1e-8:
        return 0.0, 0.0, 0.0
    b = a * (1.0 - f)
    phi1, phi2 = math.radians(lat1), math.radians(lat2)
    big_l = math.radians(lon2 - lon1)
    u1 = math.atan((1.0 - f) * math.tan(phi1))
    u2 = math.atan((1.0 - f) * math.tan(phi2))
    su1, cu1, su2, cu2 = math.sin(u1), math.cos(u1), math.sin(u2), math.cos(u2)
    lam = big_l
    for _ in range(1000):
        sl, cl = math.sin(lam), math.cos(lam)
        sin_sig = math.hypot(cu2 * sl, cu1 * su2 - su1 * cu2 * cl)
        if sin_sig == 0.0:
            return 0.0, 0.0, 0.0
        cos_sig = su1 * su2 + cu1 * cu2 * cl
        sig = math.atan2(sin_sig, cos_sig)
        sin_alpha = cu1 * cu2 * sl / sin_sig
        cos2_alpha = 1.0 - sin_alpha * sin_alpha
        cos_2sm = cos_sig - 2.0 * su1 * su2 / cos2_alpha if cos2_alpha != 0.0 else 0.0
        c = f / 16.0 * cos2_alpha * (4.0 + f * (4.0 - 3.0 * cos2_alpha))
        lam_new = big_l + (1.0 - c) * f * sin_alpha * (
            sig + c * sin_sig * (cos_2sm + c * cos_sig * (-1.0 + 2.0 * cos_2sm ** 2)))
        done = abs(lam_new - lam) < 1e-12
        lam = lam_new
        if done:
            break
    sl, cl = math.sin(lam), math.cos(lam)
    usq = cos2_alpha * (a * a - b * b) / (b * b)
    big_a = 1.0 + usq / 16384.0 * (4096.0 + usq * (-768.0 + usq * (320.0 - 175.0 * usq)))
    big_b = usq / 1024.0 * (256.0 + usq * (-128.0 + usq * (74.0 - 47.0 * usq)))
    dsig = big_b * sin_sig * (cos_2sm + big_b / 4.0 * (
        cos_sig * (-1.0 + 2.0 * cos_2sm ** 2)
        - big_b / 6.0 * cos_2sm * (-3.0 + 4.0 * sin_sig ** 2) * (-3.0 + 4.0 * cos_2sm ** 2)))
    dist = b * big_a * (sig - dsig)
    az12 = math.degrees(math.atan2(cu2 * sl, cu1 * su2 - su1 * cu2 * cl)) % 360.0
    az21 = (math.degrees(math.atan2(cu1 * sl, -su1 * cu2 + cu1 * su2 * cl)) + 180.0) % 360.0
    return dist, az12, az21

// stats.cuh — device helpers shared by the OLS and LTS solve kernels (numpy-compatible medians,
// the per-window MdCCM / MAD / data-spike statistics, slowness -> velocity / back-azimuth).
#pragma once
#include <math.h>
#include <stdint.h>

namespace nbls {

__device__ __forceinline__ double qnan() { return __longlong_as_double(0x7ff8000000000000LL); }

// Median of vals[0..m) (shared memory, no NaN) by rank counting, numpy semantics: middle element,
// or the mean of the two middle elements for even m.  All 32 lanes participate; result on all lanes.
__device__ __forceinline__ double warp_median(const double* vals, int m, int lane) {
    if (m <= 0) return qnan();
    const int r_hi = m >> 1;
    const int r_lo = (m & 1) ? r_hi : r_hi - 1;
    double lo = 0.0, hi = 0.0;
    bool have_lo = false, have_hi = false;
    for (int p = lane; p < m; p += 32) {
        const double v = vals[p];
        int rank = 0;
        for (int q = 0; q < m; ++q) {
            const double u = vals[q];
            rank += (u < v) || (u == v && q < p);
        }
        if (rank == r_lo) { lo = v; have_lo = true; }
        if (rank == r_hi) { hi = v; have_hi = true; }
    }
    const unsigned mlo = __ballot_sync(0xffffffffu, have_lo);
    const unsigned mhi = __ballot_sync(0xffffffffu, have_hi);
    lo = __shfl_sync(0xffffffffu, lo, __ffs(mlo) - 1);
    hi = __shfl_sync(0xffffffffu, hi, __ffs(mhi) - 1);
    return (m & 1) ? hi : (lo + hi) / 2.0;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void vel_baz(double zx, double zy, double& vel, double& baz) {
    vel = 1.0 / sqrt(zx * zx + zy * zy);                        // lts_classes.py:928 / :706
    double v = atan2(zx, zy) * 180.0 / M_PI - 360.0;            // :931 / :713, python float modulo
    double r = fmod(v, 360.0);
    if (r != 0.0 && r < 0.0) r += 360.0;
    baz = r;
}


// per-window stats shared by the OLS and LTS kernels; `buf` holds P doubles of scratch.
// Returns skip flag; outputs mdccm and mad on all lanes.
__device__ __forceinline__ bool window_stats(const int32_t* lag, const double* cmax, int P, double fs, double* buf, int lane,
                             double& mdccm, double& mad) {
    // nanmedian of cmax (lts_classes.py:873)
    int m = 0;
    for (int base = 0; base < P; base += 32) {
        const int p = base + lane;
        double v = p < P ? cmax[p] : qnan();
        const bool ok = v == v;
        const unsigned msk = __ballot_sync(0xffffffffu, ok);
        if (ok) buf[m + __popc(msk & ((1u << lane) - 1u))] = v;
        m += __popc(msk);
    }
    __syncwarp();
    mdccm = warp_median(buf, m, lane);
    __syncwarp();
    int nz = 0;
    for (int p = lane; p < P; p += 32) {
        const int l = lag[p];
        nz += l != 0;
        buf[p] = fabs((double)(-l) / fs);      // |tau|, tau = (N - delay)/fs = -lag/fs (:876)
    }
    __syncwarp();
    mad = 1.4826 * warp_median(buf, P, lane);  // :877-878
    __syncwarp();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nz += __shfl_xor_sync(0xffffffffu, nz, o);
    return mad == 0.0 || nz < P - 2;           // data-spike test (:911, :65, :663)
}


}  // namespace nbls

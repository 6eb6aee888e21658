// solve.cu — K3: per-window statistics (MdCCM, MAD of the delays, data-spike test) and the slowness
// solve for every band-window in one batched launch: ordinary least squares against the fixed
// co-array geometry, or FAST-LTS with the reference's seeded subset stream for alpha < 1.
//
// Replaces, of the reference's lts_array/lts_array/classes/lts_classes.py:
//   :873-878   mdccm = nanmedian(max cij), tau = -lag/fs, time_delay_mad = 1.4826*median|tau|
//   :891-936   OLSEstimator.solve   (uncertainty ellipse :938-968 is not on the NBLS path)
//   :58-150    fast_LTS (numba)     500 two-point starts x <=4 C-steps, 10 best x <=100 C-steps
//   :643-764   post_process         reweighting, final weights, vel/baz/sigma_tau
// One warp owns one window (OLS) / one CTA owns one window (LTS).
#include <float.h>
#include <math.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "stats.cuh"

namespace nbls {

struct SolveArgs {
    int64_t nwin;
    int C, P;
    double fs;
    const int32_t* lag;     // [E][P]
    const double* cmax;     // [E][P]
    const double* xij;      // [P][2]
    const double* gain;     // [2][P]
    double madx, mady;      // xij_mad
    double *mdccm, *mad, *vel, *baz, *sig, *slow;
};

constexpr int OLS_WARPS = 4;

__global__ void __launch_bounds__(OLS_WARPS * 32) stats_ols_kernel(SolveArgs a) {
    extern __shared__ double sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* buf = sm + (size_t)warp * a.P;
    for (int64_t e = (int64_t)blockIdx.x * OLS_WARPS + warp; e < a.nwin; e += (int64_t)gridDim.x * OLS_WARPS) {
        const int32_t* lag = a.lag + e * a.P;
        double mdccm, mad;
        const bool skip = window_stats(lag, a.cmax + e * a.P, a.P, a.fs, buf, lane, mdccm, mad);
        double vel = qnan(), baz = qnan(), sig = qnan(), zx = qnan(), zy = qnan();
        if (!skip) {
            // z = pinv(X_std) (tau/mad), then de-standardise (:915-921)
            double sx = 0.0, sy = 0.0;
            for (int p = lane; p < a.P; p += 32) {
                const double y = ((double)(-lag[p]) / a.fs) / mad;
                sx = fma(a.gain[p], y, sx);
                sy = fma(a.gain[a.P + p], y, sy);
            }
            sx = warp_sum(sx);
            sy = warp_sum(sy);
            zx = sx * (mad / a.madx);
            zy = sy * (mad / a.mady);
            vel_baz(zx, zy, vel, baz);
            // sigma_tau = sqrt(tau^T (tau - xij z) / (P-2))   (:934-936)
            double s = 0.0;
            for (int p = lane; p < a.P; p += 32) {
                const double tau = (double)(-lag[p]) / a.fs;
                const double res = tau - (a.xij[2 * p] * zx + a.xij[2 * p + 1] * zy);
                s = fma(tau, res, s);
            }
            s = warp_sum(s);
            sig = sqrt(s / (double)(a.P - 2));
        }
        if (lane == 0) {
            if (a.mdccm) a.mdccm[e] = mdccm;
            if (a.mad) a.mad[e] = mad;
            if (a.vel) a.vel[e] = vel;
            if (a.baz) a.baz[e] = baz;
            if (a.sig) a.sig[e] = sig;
            if (a.slow) { a.slow[2 * e] = zx; a.slow[2 * e + 1] = zy; }
        }
        __syncwarp();
    }
}

int lts_impl(nbls_ctx* ctx, const nbls_solve_desc* s, int64_t nwin, const int32_t* lag, const double* cmax,
             const nbls_outputs* out, const double* dgeom);

int solve_impl(nbls_ctx* ctx, const nbls_solve_desc* s, int64_t nwin, const int32_t* lag, const double* cmax,
               const nbls_outputs* out) {
    NBLS_REQUIRE(s && lag && cmax && out, "nbls_solve: null argument");
    const int P = s->npairs, C = s->nchan;
    NBLS_REQUIRE(C >= 3 && P == C * (C - 1) / 2, "nbls_solve: need >= 3 channels and P = C(C-1)/2 (C=%d P=%d)", C, P);
    NBLS_REQUIRE(s->xij && s->xij_mad && s->ols_gain, "nbls_solve: geometry arrays missing");
    NBLS_REQUIRE(s->fs > 0, "nbls_solve: fs must be positive");
    if (nwin <= 0) return NBLS_OK;
    // geometry block on the device: xij [P][2], gain [2][P]
    NBLS_CHECK(ensure(ctx, ctx->geom, (size_t)P * 4 * sizeof(double)));
    double* dgeom = (double*)ctx->geom.p;
    NBLS_CUDA(cudaMemcpyAsync(dgeom, s->xij, (size_t)P * 2 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    NBLS_CUDA(cudaMemcpyAsync(dgeom + 2 * P, s->ols_gain, (size_t)P * 2 * sizeof(double), cudaMemcpyHostToDevice,
                              ctx->stream));
    NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
    if (s->alpha != 1.0) return lts_impl(ctx, s, nwin, lag, cmax, out, dgeom);

    SolveArgs a;
    memset(&a, 0, sizeof(a));
    a.nwin = nwin; a.C = C; a.P = P; a.fs = s->fs; a.lag = lag; a.cmax = cmax;
    a.xij = dgeom; a.gain = dgeom + 2 * P; a.madx = s->xij_mad[0]; a.mady = s->xij_mad[1];
    a.mdccm = out->mdccm; a.mad = out->mad; a.vel = out->vel; a.baz = out->baz; a.sig = out->sigma_tau;
    a.slow = out->slowness;
    const size_t smem = (size_t)OLS_WARPS * P * sizeof(double);
    NBLS_REQUIRE(smem <= ctx->smem_optin, "nbls_solve: too many pairs (%d)", P);
    NBLS_CUDA(allow_max_dynamic_smem(stats_ols_kernel, ctx->smem_optin));
    const int64_t want = (nwin + OLS_WARPS - 1) / OLS_WARPS;
    const int grid = (int)std::min<int64_t>(want, (int64_t)ctx->sm_count * 16);
    {
        KernelGroupScope ks(ctx, KG_STATS_OLS, (double)nwin * 10.0 * P);
        stats_ols_kernel<<<grid, OLS_WARPS * 32, smem, ctx->stream>>>(a);
    }
    NBLS_CUDA(cudaGetLastError());
    ctx->launches += 1;
    return NBLS_OK;
}

}  // namespace nbls

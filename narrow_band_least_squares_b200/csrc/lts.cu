// lts.cu — K3b: FAST-LTS slowness solve with the reference's seeded subset stream, one CTA per
// band-window, followed by the reweighting step that yields the final outlier flags.
//
// Follows lts_array/lts_array/classes/lts_classes.py of the reference:
//   fast_LTS :58-150 — per window, on standardised data (y = tau/mad, X = xij/xij_mad):
//     500 starts: exact fit through the 2 pairs drawn by random_set (:16-36; the LCG is restarted
//     at seed 0 for every window, :76, so the draws are a host-precomputed table), then up to 4
//     C-steps (rank |r|, keep the h smallest, refit, objective = sum of the h smallest r^2, stop
//     when the objective repeats, :89-103); the 10 smallest objectives are kept — with duplicates,
//     because check_array's de-dup never fires (:39-55) — and iterated for up to
//     100 C-steps (:111-142); the strict minimum wins.
//   post_process :643-764 (without the uncertainty ellipse): raw scale s0 from the h smallest
//     squared residuals, first weights |r/s0| <= 2.2414..., refit on them, reweighted scale,
//     final weights |r/scale| <= 2.5, vel/baz from the refit, sigma_tau over the final weights.
// A warp runs one start at a time with its lanes spread over the P pairs; ranking is by counting
// (rank_p = #{q : |r_q| < |r_p| or tie with q < p}), which yields the h-subset and the objective in
// one pass and makes every fit a deterministic function of the subset.
#include <cuda_fp16.h>
#include <float.h>
#include <math.h>
#include <math_constants.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "stats.cuh"

namespace nbls {


constexpr int LTS_WARPS = 8;

struct LtsArgs {
    int64_t nwin;
    int P, h, n_starts, n_keep, csteps1, csteps2;
    double fs, madx, mady, raw_factor, rew1;
    const int32_t* lag;
    const double* cmax;
    const double* xij;       // [P][2] km
    const int32_t* subsets;  // [(n_starts+n_keep)][2]
    const double* rew_cons;  // [P+1]
    double *mdccm, *mad, *vel, *baz, *sig, *slow;
    uint8_t* weights;        // [E][P]
};

struct Sums {
    double sxx, sxy, syy, sxt, syt, obj;
};

// The six sums of a ranking pass reduced over the warp with 8 exchanges instead of 30: at the xor-16 step a lane keeps
// three of the six values and hands the other three to its partner, then two of (three + a zero), then one; the last two
// steps are plain butterflies.  Every partial sum adds the same two operands as warp_sum()'s butterfly does (a + b = b + a
// exactly), so the totals - read back from lanes 0, 4, 8, 16, 20, 24 - are bit-identical to six warp_sum() calls.
__device__ __forceinline__ void warp_sum6(Sums& s, int lane) {
    const bool u16 = (lane & 16) != 0, u8 = (lane & 8) != 0, u4 = (lane & 4) != 0;
    double k0 = u16 ? s.sxt : s.sxx, k1 = u16 ? s.syt : s.sxy, k2 = u16 ? s.obj : s.syy;      // kept: v0 v1 v2 | v3 v4 v5
    const double g0 = u16 ? s.sxx : s.sxt, g1 = u16 ? s.sxy : s.syt, g2 = u16 ? s.syy : s.obj;
    k0 += __shfl_xor_sync(0xffffffffu, g0, 16);
    k1 += __shfl_xor_sync(0xffffffffu, g1, 16);
    k2 += __shfl_xor_sync(0xffffffffu, g2, 16);
    // four slots (k0 k1 | k2 0): lanes with bit 3 clear keep the first two
    double m0 = u8 ? k2 : k0, m1 = u8 ? 0.0 : k1;
    const double h0 = u8 ? k0 : k2, h1 = u8 ? k1 : 0.0;
    m0 += __shfl_xor_sync(0xffffffffu, h0, 8);
    m1 += __shfl_xor_sync(0xffffffffu, h1, 8);
    double t = u4 ? m1 : m0;
    const double h = u4 ? m0 : m1;
    t += __shfl_xor_sync(0xffffffffu, h, 4);
    t += __shfl_xor_sync(0xffffffffu, t, 2);
    t += __shfl_xor_sync(0xffffffffu, t, 1);
    // value index -> lanes: v0 0-3, v1 4-7, v2 8-11, v3 16-19, v4 20-23, v5 24-27
    s.sxx = __shfl_sync(0xffffffffu, t, 0);  s.sxy = __shfl_sync(0xffffffffu, t, 4);  s.syy = __shfl_sync(0xffffffffu, t, 8);
    s.sxt = __shfl_sync(0xffffffffu, t, 16); s.syt = __shfl_sync(0xffffffffu, t, 20); s.obj = __shfl_sync(0xffffffffu, t, 24);
}

// 2x2 normal equations -> coefficients; singular systems give NaN (never selected)
__device__ __forceinline__ void solve2(const Sums& s, double& zx, double& zy) {
    const double det = s.sxx * s.syy - s.sxy * s.sxy;
    zx = (s.syy * s.sxt - s.sxy * s.syt) / det;
    zy = (s.sxx * s.syt - s.sxy * s.sxt) / det;
}

// One ranking pass for coefficients (zx, zy): residuals, ranks, then over the h smallest |r| the
// objective and the normal-equation sums of the next fit.  x0/x1/y: [P] shared; absr: warp scratch.
__device__ __forceinline__ Sums rank_pass(const double* x0, const double* x1, const double* y, double* absr, int P,
                                          int h, double zx, double zy, int lane) {
    for (int p = lane; p < P; p += 32) absr[p] = fabs(y[p] - (x0[p] * zx + x1[p] * zy));
    __syncwarp();
    Sums s = {0, 0, 0, 0, 0, 0};
    for (int p = lane; p < P; p += 32) {
        const double v = absr[p];
        int rank = 0;
        for (int q = 0; q < P; ++q) {
            const double u = absr[q];
            rank += (u < v) || (u == v && q < p);
        }
        if (rank < h) {
            const double a = x0[p], b = x1[p], t = y[p];
            s.sxx = fma(a, a, s.sxx); s.sxy = fma(a, b, s.sxy); s.syy = fma(b, b, s.syy);
            s.sxt = fma(a, t, s.sxt); s.syt = fma(b, t, s.syt);
            s.obj = fma(v, v, s.obj);
        }
    }
    __syncwarp();
    warp_sum6(s, lane);
    return s;
}

// The same pass with the h-subset found through FP16 keys (NI = items per lane, P <= 32 NI <= 2016).  Rounding
// |r| to FP16 is monotone, so with le(p) = #{q : key_q <= key_p} (counted two keys per instruction, exactly, in
// half2 arithmetic: counts stay below 2048)
//   le(p) <= h                  =>  rank_p < h            (p is in the subset),
//   key_p > K*, K* = smallest key with le > h  =>  rank_p >= h   (p is not),
// and only if the FP16 tie group at K* straddles the boundary (a few per cent of the passes) are its members
// ranked with the exact FP64 rule.  The subset, and hence the sums, are identical to rank_pass's.
template <int NI>
__device__ __forceinline__ Sums rank_pass16(const double* x0, const double* x1, const double* y, double* absr,
                                            __half2* key2, int P, int h, double zx, double zy, int lane,
                                            unsigned (&member)[NI > 0 ? NI : 1]) {
    __half* key = reinterpret_cast<__half*>(key2);
    const unsigned short* keyb = reinterpret_cast<const unsigned short*>(key2);
    const __half hnan = __ushort_as_half((unsigned short)0x7E00);
    double v[NI];
    __half2 kp[NI], cnt[NI];
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int p = lane + 32 * i;
        v[i] = 0.0;
        kp[i] = __half2half2(hnan);
        cnt[i] = __float2half2_rn(0.f);
        if (p < P) {
            v[i] = fabs(y[p] - (x0[p] * zx + x1[p] * zy));
            absr[p] = v[i];
            const __half k = __double2half(v[i]);
            key[p] = k;
            kp[i] = __half2half2(k);
        }
    }
    if (lane == 0 && (P & 1)) key[P] = hnan;      // pad to a whole half2: NaN compares false
    __syncwarp();
    bool in[NI];
    if constexpr (NI >= 4) {
        // P > 96: instead of counting every key against every key (P^2 / 64 half2 compares per lane), SELECT the h-th
        // smallest key K* = min{v : #{key <= v} >= h} bit by bit over the 15-bit pattern of a non-negative FP16 (patterns
        // order like values): 15 warp-wide counts of NI ballots each.  Keys below K* are in the subset, keys above are
        // not, and the tie group at K* is in as a whole if it ends exactly at rank h - otherwise it straddles the
        // boundary and its members are ranked with the exact FP64 rule, as below.  Same subset as the counting form.
        // (NaN residuals - a NaN fit - get pattern 0: they are in the subset either way and make the sums NaN.)
        unsigned ku[NI];
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const int p = lane + 32 * i;
            const unsigned b = (unsigned)__half_as_ushort(__low2half(kp[i]));
            ku[i] = p < P ? (b > 0x7C00u ? 0u : b) : 0xFFFFu;
        }
        unsigned kstar = 0;
#pragma unroll
        for (int bit = 14; bit >= 0; --bit) {
            const unsigned t = kstar | ((1u << bit) - 1u);
            int c = 0;
#pragma unroll
            for (int i = 0; i < NI; ++i) c += __popc(__ballot_sync(0xffffffffu, ku[i] <= t));
            if (c < h) kstar |= 1u << bit;
        }
        int c_lt = 0, c_le = 0;
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            c_lt += __popc(__ballot_sync(0xffffffffu, ku[i] < kstar));
            c_le += __popc(__ballot_sync(0xffffffffu, ku[i] <= kstar));
        }
        const bool whole = c_le == h;           // the tie group at K* ends exactly at rank h
#pragma unroll
        for (int i = 0; i < NI; ++i) in[i] = ku[i] < kstar || (whole && ku[i] == kstar);
        if (!whole) {                           // warp-uniform: the tie group straddles the boundary
            const int need = h - c_lt;
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                const int p = lane + 32 * i;
                if (ku[i] == kstar) {
                    int r = 0;
                    for (int q = 0; q < P; ++q) {
                        const unsigned bq = (unsigned)keyb[q];
                        if ((bq > 0x7C00u ? 0u : bq) == kstar) {
                            const double u = absr[q];
                            r += (u < v[i]) || (u == v[i] && q < p);
                        }
                    }
                    in[i] = r < need;
                }
            }
        }
    } else {
    const int P2 = (P + 1) >> 1;
    for (int q2 = 0; q2 < P2; ++q2) {
        const __half2 kq = key2[q2];
#pragma unroll
        for (int i = 0; i < NI; ++i) cnt[i] = __hadd2(cnt[i], __hle2(kq, kp[i]));
    }
    int n_in = 0;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int p = lane + 32 * i;
        const int le = (int)(__low2float(cnt[i]) + __high2float(cnt[i]));
        in[i] = (p < P) && (le <= h);
        n_in += __popc(__ballot_sync(0xffffffffu, in[i]));
    }
    if (n_in < h) {        // warp-uniform: the tie group at K* straddles the boundary
        const int need = h - n_in;
        unsigned kmin = 0xFFFFu;
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const int p = lane + 32 * i;
            if (p < P && !in[i]) kmin = min(kmin, (unsigned)__half_as_ushort(__low2half(kp[i])));
        }
        kmin = __reduce_min_sync(0xffffffffu, kmin);
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const int p = lane + 32 * i;
            if (p < P && !in[i] && (unsigned)__half_as_ushort(__low2half(kp[i])) == kmin) {
                int r = 0;
                for (int q = 0; q < P; ++q) {
                    if ((unsigned)keyb[q] == kmin) {
                        const double u = absr[q];
                        r += (u < v[i]) || (u == v[i] && q < p);
                    }
                }
                in[i] = r < need;
            }
        }
    }
    }
    Sums s = {0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const int p = lane + 32 * i;
        member[i] = __ballot_sync(0xffffffffu, in[i]);      // the h-subset as NI lane masks (warp-uniform)
        if (in[i]) {
            const double a = x0[p], b = x1[p], t = y[p];
            s.sxx = fma(a, a, s.sxx); s.sxy = fma(a, b, s.sxy); s.syy = fma(b, b, s.syy);
            s.sxt = fma(a, t, s.sxt); s.syt = fma(b, t, s.syt);
            s.obj = fma(v[i], v[i], s.obj);
        }
    }
    __syncwarp();
    warp_sum6(s, lane);
    return s;
}

template <int NI>
__device__ __forceinline__ Sums rank_any(const double* x0, const double* x1, const double* y, double* absr, __half2* key2,
                                         int P, int h, double zx, double zy, int lane, unsigned (&member)[NI > 0 ? NI : 1]) {
    if constexpr (NI == 0) return rank_pass(x0, x1, y, absr, P, h, zx, zy, lane);
    else return rank_pass16<NI>(x0, x1, y, absr, key2, P, h, zx, zy, lane, member);
}

// C-steps from (zx, zy) until the objective repeats or `limit` steps (:89-103 / :125-139)
template <int NI>
__device__ __forceinline__ void csteps(const double* x0, const double* x1, const double* y, double* absr, __half2* key2,
                                       int P, int h, int limit, double& zx, double& zy, double& obj, int lane) {
    unsigned was[NI > 0 ? NI : 1], now[NI > 0 ? NI : 1];
    Sums s = rank_any<NI>(x0, x1, y, absr, key2, P, h, zx, zy, lane, was);
    double prev = 0.0;
    obj = s.obj;
    for (int kk = 0; kk < limit; ++kk) {
        solve2(s, zx, zy);
        s = rank_any<NI>(x0, x1, y, absr, key2, P, h, zx, zy, lane, now);
        obj = s.obj;
        if (kk >= 1 && obj == prev) break;
        if constexpr (NI > 0) {
            // The h-subset did not change: the next step would refit the same rows - the same sums, hence (zx, zy) and
            // the objective bit for bit - and then stop on the repeated objective (or on the step limit).  Stop here with
            // exactly those values and save the ranking pass.
            bool same = true;
#pragma unroll
            for (int i = 0; i < NI; ++i) { same = same && (now[i] == was[i]); was[i] = now[i]; }
            if (same) break;
        }
        prev = obj;
    }
}

// exact fit through two pairs (:84-86); a singular pair of rows gets the minimum-norm solution
// like LAPACK's lstsq
__device__ __forceinline__ void fit2(const double* x0, const double* x1, const double* y, int i0, int i1, double& zx,
                                     double& zy) {
    const double a = x0[i0], b = x1[i0], c = x0[i1], d = x1[i1], u = y[i0], v = y[i1];
    const double det = a * d - b * c;
    if (fabs(det) > 1e-14 * (fabs(a * d) + fabs(b * c))) {
        zx = (u * d - b * v) / det;
        zy = (a * v - u * c) / det;
    } else {
        const double f = a * a + b * b + c * c + d * d;
        zx = (a * u + c * v) / f;
        zy = (b * u + d * v) / f;
    }
}

template <int NI>
__global__ void __launch_bounds__(LTS_WARPS * 32) lts_kernel(LtsArgs a) {
    extern __shared__ double sm[];
    const int P = a.P, NS = a.n_starts, NK = a.n_keep;
    double* ys = sm;                   // [P] standardised delays
    double* x0 = ys + P;               // [P]
    double* x1 = x0 + P;               // [P]
    double* tau = x1 + P;              // [P] seconds
    const int KW = (P + 5) / 4;        // doubles holding P+1 FP16 keys
    double* scratch = tau + P;         // [LTS_WARPS][P + KW]: |r| and its FP16 keys, per warp
    double* sobj = scratch + (size_t)LTS_WARPS * (P + KW);   // [NS]
    double* szx = sobj + NS;           // [NS]
    double* szy = szx + NS;            // [NS]
    double* cobj = szy + NS;           // [NK]
    double* czx = cobj + NK;           // [NK]
    double* czy = czx + NK;            // [NK]
    __shared__ double sh_mdccm, sh_mad;
    __shared__ int sh_skip, sh_ncand;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* absr = scratch + (size_t)warp * (P + KW);
    __half2* key2 = reinterpret_cast<__half2*>(absr + P);

    for (int64_t e = blockIdx.x; e < a.nwin; e += gridDim.x) {
        const int32_t* lag = a.lag + e * P;
        __syncthreads();
        if (warp == 0) {
            double mdccm, mad;
            const bool skip = window_stats(lag, a.cmax + e * P, P, a.fs, absr, lane, mdccm, mad);
            if (lane == 0) { sh_mdccm = mdccm; sh_mad = mad; sh_skip = skip ? 1 : 0; sh_ncand = 0; }
        }
        __syncthreads();
        const double mad = sh_mad;
        const bool skip = sh_skip != 0;
        if (tid == 0) {
            if (a.mdccm) a.mdccm[e] = sh_mdccm;
            if (a.mad) a.mad[e] = mad;
        }
        if (skip) {   // data spike (:65-67, :663-665): NaN estimates, all-zero weights
            if (tid == 0) {
                if (a.vel) a.vel[e] = qnan();
                if (a.baz) a.baz[e] = qnan();
                if (a.sig) a.sig[e] = qnan();
                if (a.slow) { a.slow[2 * e] = qnan(); a.slow[2 * e + 1] = qnan(); }
            }
            if (a.weights) for (int p = tid; p < P; p += blockDim.x) a.weights[e * P + p] = 0;
            continue;
        }
        for (int p = tid; p < P; p += blockDim.x) {
            const double t = (double)(-lag[p]) / a.fs;
            tau[p] = t;
            ys[p] = t / mad;                       // :70
            x0[p] = a.xij[2 * p] / a.madx;         // xij_standardized (:849-850)
            x1[p] = a.xij[2 * p + 1] / a.mady;
        }
        __syncthreads();

        // ---- 500 random starts (:80-108) ----
        for (int s = warp; s < NS; s += LTS_WARPS) {
            double zx, zy, obj;
            fit2(x0, x1, ys, a.subsets[2 * s], a.subsets[2 * s + 1], zx, zy);
            csteps<NI>(x0, x1, ys, absr, key2, P, a.h, a.csteps1, zx, zy, obj, lane);
            if (lane == 0) { sobj[s] = obj; szx[s] = zx; szy[s] = zy; }
        }
        __syncthreads();
        // keep the n_keep smallest finite objectives, arrival order on ties
        for (int s = tid; s < NS; s += blockDim.x) {
            const double v = sobj[s];
            if (isfinite(v)) {
                int rank = 0;
                for (int q = 0; q < NS; ++q) {
                    const double u = sobj[q];
                    rank += isfinite(u) && ((u < v) || (u == v && q < s));
                }
                if (rank < NK) { cobj[rank] = v; czx[rank] = szx[s]; czy[rank] = szy[s]; atomicAdd(&sh_ncand, 1); }
            }
        }
        __syncthreads();
        const int ncand = min(sh_ncand, NK);
        // ---- final C-steps on the candidates (:111-142) ----
        for (int c = warp; c < NK; c += LTS_WARPS) {
            double zx, zy, obj = qnan();
            bool ok = true;
            if (c < ncand) { zx = czx[c]; zy = czy[c]; }
            else {   // fewer than n_keep finite candidates: fresh random pairs (:117-121)
                const int s = NS + (c - ncand);
                fit2(x0, x1, ys, a.subsets[2 * s], a.subsets[2 * s + 1], zx, zy);
            }
            ok = isfinite(zx);
            if (ok) csteps<NI>(x0, x1, ys, absr, key2, P, a.h, a.csteps2, zx, zy, obj, lane);
            __syncwarp();
            if (lane == 0) { cobj[c] = ok ? obj : qnan(); czx[c] = zx; czy[c] = zy; }
        }
        __syncthreads();

        // ---- strict minimum, de-standardise (:140-148), post_process (:643-764) : warp 0 ----
        if (warp == 0) {
            double best = CUDART_INF, bzx = qnan(), bzy = qnan();
            for (int c = 0; c < NK; ++c) {
                const double o = cobj[c];
                if (o < best) { best = o; bzx = czx[c]; bzy = czy[c]; }
            }
            const double cx = bzx * (mad / a.madx), cy = bzy * (mad / a.mady);
            // raw scale from the h smallest squared residuals (:668-673)
            for (int p = lane; p < P; p += 32) absr[p] = fabs(tau[p] - (a.xij[2 * p] * cx + a.xij[2 * p + 1] * cy));
            __syncwarp();
            double ssq = 0.0;
            for (int p = lane; p < P; p += 32) {
                const double v = absr[p];
                int rank = 0;
                for (int q = 0; q < P; ++q) {
                    const double u = absr[q];
                    rank += (u < v) || (u == v && q < p);
                }
                if (rank < a.h) ssq = fma(v, v, ssq);
            }
            ssq = warp_sum(ssq);
            const double s0 = sqrt(ssq / (double)a.h) * a.raw_factor;
            double zx = cx, zy = cy;
            // first weights (:679) -> refit (:687-689) -> reweighted scale (:693-700) -> final weights (:701)
            const bool exact = fabs(s0) < 1e-7;
            Sums s = {0, 0, 0, 0, 0, 0};
            int nw = 0;
            for (int p = lane, i = 0; p < P; p += 32, ++i) {
                const bool w = exact ? (absr[p] < 1e-7) : (fabs((tau[p] - (a.xij[2 * p] * cx + a.xij[2 * p + 1] * cy)) / s0) <= 2.2414027276049473);
                if (w) {
                    const double xa = a.xij[2 * p], xb = a.xij[2 * p + 1], t = tau[p];
                    s.sxx = fma(xa, xa, s.sxx); s.sxy = fma(xa, xb, s.sxy); s.syy = fma(xb, xb, s.syy);
                    s.sxt = fma(xa, t, s.sxt); s.syt = fma(xb, t, s.syt);
                    ++nw;
                }
                absr[p] = w ? 1.0 : 0.0;
            }
            __syncwarp();
            nw = __reduce_add_sync(0xffffffffu, nw);
            if (!exact) {
                s.sxx = warp_sum(s.sxx); s.sxy = warp_sum(s.sxy); s.syy = warp_sum(s.syy); s.sxt = warp_sum(s.sxt); s.syt = warp_sum(s.syt);
                solve2(s, zx, zy);
                double rs = 0.0;
                for (int p = lane; p < P; p += 32) {
                    if (absr[p] != 0.0) {
                        const double r = tau[p] - (a.xij[2 * p] * zx + a.xij[2 * p + 1] * zy);
                        rs = fma(r, r, rs);
                    }
                }
                rs = warp_sum(rs);
                double scale = sqrt(rs / (double)(nw - 1)) * a.rew1;
                if (nw != P) scale *= a.rew_cons[nw];
                __syncwarp();
                for (int p = lane; p < P; p += 32) {
                    const double r = tau[p] - (a.xij[2 * p] * zx + a.xij[2 * p + 1] * zy);
                    absr[p] = (fabs(r / scale) <= 2.5) ? 1.0 : 0.0;
                }
                __syncwarp();
            }
            // sigma_tau over the final weights (:723-730)
            double num = 0.0;
            int mw = 0;
            for (int p = lane; p < P; p += 32) {
                const bool w = absr[p] != 0.0;
                if (w) {
                    const double r = tau[p] - (a.xij[2 * p] * zx + a.xij[2 * p + 1] * zy);
                    num = fma(tau[p], r, num);
                    ++mw;
                }
                if (a.weights) a.weights[e * P + p] = w ? 1 : 0;
            }
            num = warp_sum(num);
            mw = __reduce_add_sync(0xffffffffu, mw);
            if (lane == 0) {
                double vel, baz;
                vel_baz(zx, zy, vel, baz);
                const double val = num / (double)(mw - 2);
                const double sig = (val != val || val < 0.0) ? qnan() : sqrt(val);
                if (a.vel) a.vel[e] = vel;
                if (a.baz) a.baz[e] = baz;
                if (a.sig) a.sig[e] = sig;
                if (a.slow) { a.slow[2 * e] = zx; a.slow[2 * e + 1] = zy; }
            }
        }
    }
}

int lts_impl(nbls_ctx* ctx, const nbls_solve_desc* s, int64_t nwin, const int32_t* lag, const double* cmax,
             const nbls_outputs* out, const double* dgeom) {
    const int P = s->npairs;
    NBLS_REQUIRE(s->alpha >= 0.5 && s->alpha < 1.0, "nbls_solve: alpha must lie in [0.5, 1] (got %g)", s->alpha);
    NBLS_REQUIRE(P >= 4, "nbls_solve: the co-array must have at least 4 elements for least trimmed squares");
    NBLS_REQUIRE(s->subsets && s->rew_cons, "nbls_solve: LTS tables missing");
    NBLS_REQUIRE(s->h >= 2 && s->h <= P, "nbls_solve: h=%d out of range", s->h);
    NBLS_REQUIRE(s->n_starts >= 1 && s->n_starts <= 4096 && s->n_keep >= 1 && s->n_keep <= 64,
                 "nbls_solve: n_starts/n_keep out of range");
    const size_t nsub = (size_t)(s->n_starts + s->n_keep) * 2;
    const size_t tab_bytes = round_up(nsub * sizeof(int32_t), 16) + (size_t)(P + 1) * sizeof(double);
    NBLS_CHECK(ensure(ctx, ctx->lts_tab, tab_bytes));
    int32_t* dsub = (int32_t*)ctx->lts_tab.p;
    double* drew = (double*)((char*)ctx->lts_tab.p + round_up(nsub * sizeof(int32_t), 16));
    NBLS_CUDA(cudaMemcpyAsync(dsub, s->subsets, nsub * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    NBLS_CUDA(cudaMemcpyAsync(drew, s->rew_cons, (size_t)(P + 1) * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    NBLS_CUDA(cudaStreamSynchronize(ctx->stream));

    LtsArgs a;
    memset(&a, 0, sizeof(a));
    a.nwin = nwin; a.P = P; a.h = s->h; a.n_starts = s->n_starts; a.n_keep = s->n_keep;
    a.csteps1 = s->csteps1; a.csteps2 = s->csteps2; a.fs = s->fs;
    a.madx = s->xij_mad[0]; a.mady = s->xij_mad[1]; a.raw_factor = s->raw_factor; a.rew1 = s->rew_factor1;
    a.lag = lag; a.cmax = cmax; a.xij = dgeom; a.subsets = dsub; a.rew_cons = drew;
    a.mdccm = out->mdccm; a.mad = out->mad; a.vel = out->vel; a.baz = out->baz; a.sig = out->sigma_tau;
    a.slow = out->slowness; a.weights = out->weights;
    const int KW = (P + 5) / 4;
    const size_t smem = ((size_t)4 * P + (size_t)LTS_WARPS * (P + KW) + 3 * (size_t)s->n_starts + 3 * (size_t)s->n_keep) *
                        sizeof(double);
    NBLS_REQUIRE(smem <= ctx->smem_optin, "nbls_solve: LTS working set of %zu B exceeds shared memory", smem);
    // items per lane of the FP16-key ranking; 0 = generic FP64 counting (P > 320)
    typedef void (*lts_fn)(LtsArgs);
    static const lts_fn table[11] = {lts_kernel<0>, lts_kernel<1>, lts_kernel<2>, lts_kernel<3>, lts_kernel<4>, lts_kernel<5>,
                                     lts_kernel<6>, lts_kernel<7>, lts_kernel<8>, lts_kernel<9>, lts_kernel<10>};
    const int ni = (P + 31) / 32;
    const lts_fn fn = table[(ni <= 10 && !ctx->opt[NBLS_OPT_LTS_FP64_RANK]) ? ni : 0];
    NBLS_CUDA(allow_max_dynamic_smem(fn, ctx->smem_optin));
    const int grid = (int)std::min<int64_t>(nwin, (int64_t)ctx->sm_count * 32);
    {
        KernelGroupScope ks(ctx, KG_LTS, 0.0);
        fn<<<grid, LTS_WARPS * 32, smem, ctx->stream>>>(a);
    }
    NBLS_CUDA(cudaGetLastError());
    ctx->launches += 1;
    return NBLS_OK;
}

}  // namespace nbls

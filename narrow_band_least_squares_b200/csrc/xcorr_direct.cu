// xcorr_direct.cu — K2 (time-domain): windowed pairwise cross-correlation over the co-array with an
// on-chip arg-max, FP64 throughout.
//
// Replaces LsBeam.correlate (lts_array/lts_array/classes/lts_classes.py:852-880 of the reference):
// for every window and every pair (i<j), cij = np.correlate(x_i, x_j, 'full') / sqrt(Ei*Ej),
// delay = argmax(cij) (first maximum wins), cmax = max(cij).  With c[k] = sum_n a[n+k] b[n]
// (a = x_i, b = x_j, k = lag) the 2N-1 lags of a length-N window are the two halves
// of a CIRCULAR correlation of the window zero-padded to M (multiple of 8, M >= N):
//     slot k in [0,M):  pos[k] = sum over n with n+k <  M   -> lag  k      (valid if k <= N-1)
//                       neg[k] = sum over n with n+k >= M   -> lag  k - M  (valid if k >= M-N+1)
// so every slot costs exactly M multiply-adds and the work is M^2 per pair with no triangle waste.
//
// One CTA per (window, pair).  Both windows are staged in shared memory once (a linear with 8
// wrap-around values appended, b in a padded layout).  A lane owns 8 consecutive slots
// (k0 = 8*tile) and walks the common phase phi = (n0 + k0) mod M in steps of 8, so the 16-value
// sliding window of `a` is a warp-wide broadcast and only `b` is read per lane (conflict-free
// thanks to the 80-byte row padding); each step is an 8x8 register outer product = 64 DFMA per
// lane against 8 LDS.128.  The wrap (pos -> neg) falls on the same step for all lanes.  The
// arg-max is reduced with warp shuffles, ties resolved towards the lowest lag like np.argmax.
#include <float.h>
#include <limits.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace nbls {

constexpr int XC_MAX_WARPS = 16;

struct XcArgs {
    const double* y;      // band base
    int64_t ldy;
    int C, P;
    int N, M;             // window length, padded circular length (multiple of 8)
    int hop;
    int64_t first;
    int W;
    int lag_max;          // effective |lag| bound, <= N-1
    int nitems;           // warp-tiles to process
    int split;            // items [0,split) are warp-tiles 0..split-1, the rest wt_hi0 + (it - split)
    int wt_hi0;
    int32_t* lag;         // [W][P]
    double* cmax;         // [W][P] or nullptr
    const int* list;      // list mode: items w*P+p to evaluate (device), else nullptr = every (w, p)
    const int* count;     // list mode: number of items (device)
};

__device__ __forceinline__ int pad_b(int e) { return e + ((e >> 3) << 1); }

struct Best {
    double v;
    int lag;
};
__device__ __forceinline__ void best_update(Best& b, double v, int lag) {
    if (v > b.v || (v == b.v && lag < b.lag)) { b.v = v; b.lag = lag; }
}

__global__ void __launch_bounds__(XC_MAX_WARPS * 32, 1) xcorr_direct_kernel(XcArgs g) {
    extern __shared__ __align__(16) double smem[];
    double* sa = smem;                         // [M + 8] : a padded to M, then a[0..8) again
    double* sb = smem + g.M + 8;               // padded layout of b, M*10/8 doubles
    __shared__ double red_e[2][XC_MAX_WARPS];
    __shared__ double red_v[XC_MAX_WARPS];
    __shared__ int red_l[XC_MAX_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    // one (window, pair) per CTA; in list mode a CTA strides over the items listed on the device
    const int64_t n_items = g.list ? (int64_t)*g.count : (int64_t)gridDim.x;
    for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int64_t wp = g.list ? (int64_t)g.list[item] : item;
    const int w = (int)(wp / g.P), p = (int)(wp - (int64_t)w * g.P);
    // pair index -> (i, j), i<j row-major (lts_classes.py:840)
    int ci = 0, rem = p;
    while (rem >= g.C - 1 - ci) { rem -= g.C - 1 - ci; ++ci; }
    const int cj = ci + 1 + rem;
    const int64_t t0 = g.first + (int64_t)w * g.hop;
    const double* xa = g.y + (int64_t)ci * g.ldy + t0;
    const double* xb = g.y + (int64_t)cj * g.ldy + t0;

    double ea = 0.0, eb = 0.0;
    for (int e = tid; e < g.M; e += blockDim.x) {
        const double va = e < g.N ? xa[e] : 0.0;
        const double vb = e < g.N ? xb[e] : 0.0;
        sa[e] = va;
        if (e < 8) sa[g.M + e] = va;
        sb[pad_b(e)] = vb;
        ea = fma(va, va, ea);
        eb = fma(vb, vb, eb);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ea += __shfl_xor_sync(0xffffffffu, ea, o);
        eb += __shfl_xor_sync(0xffffffffu, eb, o);
    }
    if (lane == 0) { red_e[0][warp] = ea; red_e[1][warp] = eb; }
    __syncthreads();

    const int M = g.M;
    Best best;
    best.v = -DBL_MAX;
    best.lag = INT_MAX;
    const int neg_min_slot = M - g.lag_max;   // slot k holds a wanted negative lag iff k >= neg_min_slot

    for (int it = warp; it < g.nitems; it += nwarps) {
        const int wt = it < g.split ? it : g.wt_hi0 + (it - g.split);
        const int k0 = (wt * 32 + lane) * 8;
        if (k0 < M) {
            double acc[8], accn[8], aw[16], bv[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) { acc[r] = 0.0; accn[r] = 0.0; }
            {
                const double2* pa = reinterpret_cast<const double2*>(sa);
#pragma unroll
                for (int q = 0; q < 4; ++q) { double2 d = pa[q]; aw[2 * q] = d.x; aw[2 * q + 1] = d.y; }
            }
            int n0 = k0 == 0 ? 0 : M - k0;
            for (int phi = 0; phi < M - 8; phi += 8) {
                if (phi == k0) {   // this lane's slots stop wrapping: bank the negative-lag sums
#pragma unroll
                    for (int r = 0; r < 8; ++r) { accn[r] = acc[r]; acc[r] = 0.0; }
                }
                const double2* pa = reinterpret_cast<const double2*>(sa + phi + 8);
                const double2* pb = reinterpret_cast<const double2*>(sb + (n0 >> 3) * 10);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    double2 d = pa[q]; aw[8 + 2 * q] = d.x; aw[9 + 2 * q] = d.y;
                    double2 e = pb[q]; bv[2 * q] = e.x; bv[2 * q + 1] = e.y;
                }
#pragma unroll
                for (int q = 0; q < 8; ++q)
#pragma unroll
                    for (int r = 0; r < 8; ++r) acc[r] = fma(aw[r + q], bv[q], acc[r]);
#pragma unroll
                for (int q = 0; q < 8; ++q) aw[q] = aw[8 + q];
                n0 += 8;
                if (n0 == M) n0 = 0;
            }
            {   // last step, phi = M-8: elements with r+q <= 7 stay below M, the rest wrap
                if (M - 8 == k0) {
#pragma unroll
                    for (int r = 0; r < 8; ++r) { accn[r] = acc[r]; acc[r] = 0.0; }
                }
                const double2* pa = reinterpret_cast<const double2*>(sa + M);
                const double2* pb = reinterpret_cast<const double2*>(sb + (n0 >> 3) * 10);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    double2 d = pa[q]; aw[8 + 2 * q] = d.x; aw[9 + 2 * q] = d.y;
                    double2 e = pb[q]; bv[2 * q] = e.x; bv[2 * q + 1] = e.y;
                }
#pragma unroll
                for (int q = 0; q < 8; ++q)
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        if (r + q <= 7) acc[r] = fma(aw[r + q], bv[q], acc[r]);
                        else accn[r] = fma(aw[r + q], bv[q], accn[r]);
                    }
            }
            // candidates in increasing-lag order inside the lane: negative lags first
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int k = k0 + r;
                if (k < M && k >= neg_min_slot && k - M >= -(g.N - 1)) best_update(best, accn[r], k - M);
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int k = k0 + r;
                if (k <= g.lag_max) best_update(best, acc[r], k);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Best other;
        other.v = __shfl_xor_sync(0xffffffffu, best.v, o);
        other.lag = __shfl_xor_sync(0xffffffffu, best.lag, o);
        best_update(best, other.v, other.lag);
    }
    if (lane == 0) { red_v[warp] = best.v; red_l[warp] = best.lag; }
    __syncthreads();
    if (tid == 0) {
        double Ea = 0.0, Eb = 0.0;
        for (int i = 0; i < nwarps; ++i) { Ea += red_e[0][i]; Eb += red_e[1][i]; }
        for (int i = 1; i < nwarps; ++i) best_update(best, red_v[i], red_l[i]);
        const double den = sqrt(Ea * Eb);
        const int64_t o = (int64_t)w * g.P + p;
        // zero-energy pair: cij is all-NaN, np.argmax gives index 0 = the lowest lag searched
        // (lts_classes.py:871-875)
        const bool dead = !(den > 0.0) || best.lag == INT_MAX;
        g.lag[o] = dead ? -g.lag_max : best.lag;
        if (g.cmax) g.cmax[o] = dead ? __longlong_as_double(0x7ff8000000000000LL) : best.v / den;
    }
    __syncthreads();   // shared buffers are reused by the next item
    }
}

// list == nullptr: every (window, pair) of the band.  Otherwise only the items w*P+p in list[0..*count)
// (both on the device; used by the FFT path to settle near-tied arg-maxima), with at most `grid_cap` CTAs.
int xcorr_direct_band(nbls_ctx* ctx, const double* yband, int64_t ldy, int C, int N, int hop, int64_t first, int W,
                      int max_lag, int32_t* lag, double* cmax, const int* list, const int* count, int grid_cap) {
    if (W <= 0) return NBLS_OK;
    const int P = C * (C - 1) / 2;
    XcArgs a;
    memset(&a, 0, sizeof(a));
    a.y = yband; a.ldy = ldy; a.C = C; a.P = P; a.N = N; a.M = (int)round_up(N, 8);
    a.hop = hop; a.first = first; a.W = W; a.lag = lag; a.cmax = cmax;
    a.list = list; a.count = count;
    a.lag_max = (max_lag < 0 || max_lag > N - 1) ? N - 1 : max_lag;
    const int WT = (a.M + 255) / 256;                       // warp-tiles of 256 slots
    const int lo_last = a.lag_max / 256;                    // warp-tiles holding lags 0..lag_max
    int hi_first = (a.M - a.lag_max) / 256;                 // warp-tiles holding lags -lag_max..-1
    if (a.lag_max == 0) hi_first = WT;
    if (hi_first <= lo_last + 1) { a.split = WT; a.wt_hi0 = WT; a.nitems = WT; }
    else { a.split = lo_last + 1; a.wt_hi0 = hi_first; a.nitems = a.split + (WT - hi_first); }
    const int rounds = (a.nitems + XC_MAX_WARPS - 1) / XC_MAX_WARPS;
    const int nwarps = (a.nitems + rounds - 1) / rounds;
    const size_t smem = ((size_t)a.M + 8 + (size_t)a.M / 8 * 10) * sizeof(double);
    if (smem > ctx->smem_optin - 1024) {
        set_error("nbls_xcorr(direct): window of %d samples needs %zu B of shared memory (> %zu)", N, smem,
                  ctx->smem_optin - 1024);
        return NBLS_ERR_UNSUPPORTED;
    }
    NBLS_CUDA(allow_max_dynamic_smem(xcorr_direct_kernel, ctx->smem_optin));
    const int64_t nblocks = list ? std::min<int64_t>((int64_t)W * P, std::max(1, grid_cap)) : (int64_t)W * P;
    NBLS_REQUIRE(nblocks < (1LL << 31), "nbls_xcorr: too many (window, pair) items in one band: %lld", (long long)nblocks);
    {
        // executed work is known on the host only for whole bands (M^2 multiply-adds per pair-window); list mode adds none
        KernelGroupScope ks(ctx, KG_DIRECT, list ? 0.0 : (double)W * P * 2.0 * (double)a.M * (double)a.M);
        xcorr_direct_kernel<<<(unsigned)nblocks, nwarps * 32, smem, ctx->stream>>>(a);
    }
    NBLS_CUDA(cudaGetLastError());
    ctx->launches += 1;
    return NBLS_OK;
}

}  // namespace nbls

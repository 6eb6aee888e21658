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
#include <mutex>
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


// ---- long windows ---------------------------------------------------------------------------
// Windows too long for the kernel above (its shared memory grows with N: ~18 N bytes) are cut into tiles: a CTA owns
// one (window, pair) and XL_LAGS consecutive lags, walks x_j in tiles of XL_TILE samples and keeps the matching
// XL_TILE + XL_LAGS samples of x_i beside it (zero outside the window), so shared memory is constant in N.  A thread
// owns 8 consecutive lags and slides a 16-value register window of x_i over the tile: 64 DFMA per 8 + 8 shared loads.
// Per-CTA maxima go to a [item][lag block] table; xcorr_long_finish_kernel reduces them in rising-lag order (strict >,
// so the first maximum wins like np.argmax) and normalises by the window energies.  Any N is served.
constexpr int XL_THREADS = 256, XL_LAGS = XL_THREADS * 8, XL_TILE = 1024;

struct XlArgs {
    const double* y;
    int64_t ldy;
    int C, P, N, hop;
    int64_t first;
    int lag_max, nblk;        // lags -lag_max .. lag_max in nblk blocks of XL_LAGS
    double* part_v;           // [W*P][nblk]
    int* part_l;
    int32_t* lag;
    double* cmax;
};

__global__ void __launch_bounds__(XL_THREADS) xcorr_long_kernel(XlArgs g) {
    __shared__ __align__(16) double sb[XL_TILE];
    __shared__ __align__(16) double sa[(XL_TILE + XL_LAGS + 8) / 8 * 10];      // rows of 8 padded to 10 (pad_b)
    __shared__ double red_v[XL_THREADS / 32];
    __shared__ int red_l[XL_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t item = blockIdx.x / g.nblk;
    const int blk = blockIdx.x - (int)(item * g.nblk);
    const int w = (int)(item / g.P), p = (int)(item - (int64_t)w * g.P);
    int ci = 0, rem = p;
    while (rem >= g.C - 1 - ci) { rem -= g.C - 1 - ci; ++ci; }
    const int cj = ci + 1 + rem;
    const int64_t t0 = g.first + (int64_t)w * g.hop;
    const double* xa = g.y + (int64_t)ci * g.ldy + t0;
    const double* xb = g.y + (int64_t)cj * g.ldy + t0;
    const int N = g.N;
    const int k_lo = -g.lag_max + blk * XL_LAGS;          // first lag of the block
    const int k0 = k_lo + tid * 8;                        // this thread's lags k0 .. k0+7
    double acc[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc[r] = 0.0;
    // c[k] = sum_n a[n+k] b[n]: only tiles of b whose shifted range meets [0, N) for some lag of the block matter
    const int n_first = max(0, -(k_lo + XL_LAGS - 1)), n_last = min(N, N - k_lo);      // n in [n_first, n_last)
    for (int nt = n_first / XL_TILE * XL_TILE; nt < n_last; nt += XL_TILE) {
        __syncthreads();
        for (int e = tid; e < XL_TILE; e += XL_THREADS) sb[e] = (nt + e < N) ? xb[nt + e] : 0.0;
        for (int e = tid; e < XL_TILE + XL_LAGS + 8; e += XL_THREADS) {
            const int64_t ia = (int64_t)nt + k_lo + e;
            sa[pad_b(e)] = (ia >= 0 && ia < N) ? xa[ia] : 0.0;
        }
        __syncthreads();
        const double* pa = sa + tid * 10;                 // a[nt + k0 + 8 m + q] = pa[10 m + q]: conflict-free LDS.128
        double aw[16];
        {
            const double2* q2 = reinterpret_cast<const double2*>(pa);
#pragma unroll
            for (int q = 0; q < 4; ++q) { const double2 d = q2[q]; aw[2 * q] = d.x; aw[2 * q + 1] = d.y; }
        }
        for (int n = 0; n < XL_TILE; n += 8) {
            double bv[8];
            const double2* a2 = reinterpret_cast<const double2*>(pa + (n / 8 + 1) * 10);
            const double2* b2 = reinterpret_cast<const double2*>(sb + n);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 d = a2[q]; aw[8 + 2 * q] = d.x; aw[9 + 2 * q] = d.y;
                const double2 e = b2[q]; bv[2 * q] = e.x; bv[2 * q + 1] = e.y;
            }
#pragma unroll
            for (int q = 0; q < 8; ++q)
#pragma unroll
                for (int r = 0; r < 8; ++r) acc[r] = fma(aw[r + q], bv[q], acc[r]);
#pragma unroll
            for (int q = 0; q < 8; ++q) aw[q] = aw[8 + q];
        }
    }
    Best best;
    best.v = -DBL_MAX;
    best.lag = INT_MAX;
#pragma unroll
    for (int r = 0; r < 8; ++r)
        if (k0 + r <= g.lag_max) best_update(best, acc[r], k0 + r);
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
        for (int i = 1; i < XL_THREADS / 32; ++i) best_update(best, red_v[i], red_l[i]);
        g.part_v[item * g.nblk + blk] = best.v;
        g.part_l[item * g.nblk + blk] = best.lag;
    }
}

__global__ void __launch_bounds__(256) xcorr_long_finish_kernel(XlArgs g) {
    __shared__ double red_e[2][8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t item = blockIdx.x;
    const int w = (int)(item / g.P), p = (int)(item - (int64_t)w * g.P);
    int ci = 0, rem = p;
    while (rem >= g.C - 1 - ci) { rem -= g.C - 1 - ci; ++ci; }
    const int cj = ci + 1 + rem;
    const int64_t t0 = g.first + (int64_t)w * g.hop;
    const double* xa = g.y + (int64_t)ci * g.ldy + t0;
    const double* xb = g.y + (int64_t)cj * g.ldy + t0;
    double ea = 0.0, eb = 0.0;
    for (int e = tid; e < g.N; e += 256) {
        ea = fma(xa[e], xa[e], ea);
        eb = fma(xb[e], xb[e], eb);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ea += __shfl_xor_sync(0xffffffffu, ea, o);
        eb += __shfl_xor_sync(0xffffffffu, eb, o);
    }
    if (lane == 0) { red_e[0][warp] = ea; red_e[1][warp] = eb; }
    __syncthreads();
    if (tid == 0) {
        double Ea = 0.0, Eb = 0.0;
        for (int i = 0; i < 8; ++i) { Ea += red_e[0][i]; Eb += red_e[1][i]; }
        Best best;
        best.v = -DBL_MAX;
        best.lag = INT_MAX;
        for (int b = 0; b < g.nblk; ++b) best_update(best, g.part_v[item * g.nblk + b], g.part_l[item * g.nblk + b]);
        const double den = sqrt(Ea * Eb);
        // zero-energy pair: cij is all-NaN, np.argmax gives index 0 = the lowest lag searched (lts_classes.py:871-875)
        const bool dead = !(den > 0.0) || best.lag == INT_MAX;
        g.lag[item] = dead ? -g.lag_max : best.lag;
        if (g.cmax) g.cmax[item] = dead ? __longlong_as_double(0x7ff8000000000000LL) : best.v / den;
    }
}

static int xcorr_long_band(nbls_ctx* ctx, const double* yband, int64_t ldy, int C, int N, int hop, int64_t first, int W,
                           int lag_max, int32_t* lag, double* cmax) {
    const int P = C * (C - 1) / 2;
    XlArgs a;
    memset(&a, 0, sizeof(a));
    a.y = yband; a.ldy = ldy; a.C = C; a.P = P; a.N = N; a.hop = hop; a.first = first; a.lag_max = lag_max;
    a.nblk = (2 * lag_max + 1 + XL_LAGS - 1) / XL_LAGS;
    // windows per launch bounded so that the partial table and the grid stay small
    const int64_t per_win = (int64_t)P * a.nblk;
    const int wb = (int)std::max<int64_t>(1, std::min<int64_t>(W, (1LL << 26) / per_win));
    NBLS_CHECK(ensure(ctx, ctx->scratch, (size_t)wb * per_win * (sizeof(double) + sizeof(int)) + 256));
    a.part_v = (double*)ctx->scratch.p;
    a.part_l = (int*)(a.part_v + (size_t)wb * per_win);
    for (int w0 = 0; w0 < W; w0 += wb) {
        const int wn = std::min(wb, W - w0);
        a.first = first + (int64_t)w0 * hop;
        a.lag = lag + (int64_t)w0 * P;
        a.cmax = cmax ? cmax + (int64_t)w0 * P : nullptr;
        KernelGroupScope ks(ctx, KG_DIRECT, (double)wn * P * 2.0 * (double)N * (double)N);
        xcorr_long_kernel<<<(unsigned)(wn * per_win), XL_THREADS, 0, ctx->stream>>>(a);
        NBLS_CUDA(cudaGetLastError());
        xcorr_long_finish_kernel<<<(unsigned)(wn * P), 256, 0, ctx->stream>>>(a);
        NBLS_CUDA(cudaGetLastError());
        ctx->launches += 2;
    }
    return NBLS_OK;
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
        // the window does not fit in one SM's shared memory: tiled kernel (any N)
        NBLS_REQUIRE(!list, "nbls_xcorr(direct): list mode is not available for windows of %d samples", N);
        return xcorr_long_band(ctx, yband, ldy, C, N, hop, first, W, a.lag_max, lag, cmax);
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

// ---- allocation-free variant (nbls_xcorr_static) -------------------------------------------
// The tiled kernels above with their [item][lag block] partial table in a file-scope device array instead of the
// context's growable scratch buffer: for callers that may not allocate (CUDA graph capture, serving loops).  Every GPU
// that loads the library holds the array (3 MiB) until exit.  In host code the array's NAME is only a stand-in for the
// symbol; the device address comes from cudaGetSymbolAddress, looked up once per device with that device current.
// All calls share the array: two calls that overlap on the GPU (different streams / contexts) race.  Nothing in it has
// to start at zero: xcorr_long_kernel writes every (item, block) entry before xcorr_long_finish_kernel reads it.
constexpr size_t XS_BYTES = (size_t)NBLS_XCORR_STATIC_MAX_ENTRIES * (sizeof(double) + sizeof(int));
static __device__ __align__(256) unsigned char xs_scratch[XS_BYTES];

constexpr int XS_MAX_DEVICES = 64;
static void* xs_addr[XS_MAX_DEVICES];       // device address of xs_scratch on each device, nullptr until looked up
static std::mutex xs_mutex;

// the caller has made `device` current
static int xs_scratch_address(int device, void** out) {
    NBLS_REQUIRE(device >= 0 && device < XS_MAX_DEVICES, "nbls_xcorr_static: device %d out of range", device);
    std::lock_guard<std::mutex> lock(xs_mutex);
    if (!xs_addr[device]) {
        void* p = nullptr;
        NBLS_CUDA(cudaGetSymbolAddress(&p, xs_scratch));
        NBLS_REQUIRE(p, "nbls_xcorr_static: no device address for the static scratch array on device %d", device);
        xs_addr[device] = p;
    }
    *out = xs_addr[device];
    return NBLS_OK;
}

static int xcorr_static_impl(nbls_ctx* ctx, const double* y_dev, int C, int64_t ldy, int N, int hop, int64_t first, int W,
                             int max_lag, int32_t* lag_dev, double* cmax_dev) {
    NBLS_REQUIRE(ctx, "nbls_xcorr_static: null context");
    NBLS_REQUIRE(y_dev && lag_dev, "nbls_xcorr_static: null input or output pointer");
    NBLS_REQUIRE(C >= 2 && N >= 1 && hop >= 1 && W >= 0 && first >= 0, "nbls_xcorr_static: bad sizes (nchan %d, win_len %d, "
                 "win_hop %d, win_count %d, win_first %lld)", C, N, hop, W, (long long)first);
    NBLS_REQUIRE(first + (int64_t)(W > 0 ? W - 1 : 0) * hop + N <= ldy, "nbls_xcorr_static: the last window ends after ldy = %lld",
                 (long long)ldy);
    if (W == 0) return NBLS_OK;
    NBLS_CUDA(cudaSetDevice(ctx->device));
    const int P = C * (C - 1) / 2;
    XlArgs a;
    memset(&a, 0, sizeof(a));
    a.y = y_dev; a.ldy = ldy; a.C = C; a.P = P; a.N = N; a.hop = hop; a.first = first;
    a.lag_max = (max_lag < 0 || max_lag > N - 1) ? N - 1 : max_lag;
    a.nblk = (2 * a.lag_max + 1 + XL_LAGS - 1) / XL_LAGS;
    const int64_t entries = (int64_t)W * P * a.nblk;
    NBLS_REQUIRE(entries <= NBLS_XCORR_STATIC_MAX_ENTRIES,
                 "nbls_xcorr_static: %d windows x %d pairs x %d lag blocks = %lld partial results exceed the static scratch "
                 "array (NBLS_XCORR_STATIC_MAX_ENTRIES = %d); split the windows over several calls or use nbls_xcorr",
                 W, P, a.nblk, (long long)entries, NBLS_XCORR_STATIC_MAX_ENTRIES);
    void* scratch = nullptr;
    NBLS_CHECK(xs_scratch_address(ctx->device, &scratch));
    a.part_v = static_cast<double*>(scratch);
    a.part_l = reinterpret_cast<int*>(a.part_v + NBLS_XCORR_STATIC_MAX_ENTRIES);
    a.lag = lag_dev;
    a.cmax = cmax_dev;
    xcorr_long_kernel<<<(unsigned)entries, XL_THREADS, 0, ctx->stream>>>(a);
    NBLS_CUDA(cudaGetLastError());
    xcorr_long_finish_kernel<<<(unsigned)((int64_t)W * P), 256, 0, ctx->stream>>>(a);
    NBLS_CUDA(cudaGetLastError());
    ctx->launches += 2;
    return NBLS_OK;
}

}  // namespace nbls

extern "C" int nbls_xcorr_static(nbls_ctx* ctx, const double* y_dev, int32_t nchan, int64_t ldy, int32_t win_len,
                                 int32_t win_hop, int64_t win_first, int32_t win_count, int32_t max_lag, int32_t* lag_dev,
                                 double* cmax_dev) {
    return nbls::xcorr_static_impl(ctx, y_dev, nchan, ldy, win_len, win_hop, win_first, win_count, max_lag, lag_dev,
                                   cmax_dev);
}

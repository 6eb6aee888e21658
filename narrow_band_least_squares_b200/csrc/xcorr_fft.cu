// xcorr_fft.cu — K2 (frequency domain): windowed pairwise cross-correlation over the co-array through
// shared-memory FFTs, with an on-chip arg-max and an exactness guard.  Two flavours share the code:
//   NBLS_XCORR_FFT32 (default)  FP32 FFTs SCREEN the 2N-1 lags: every lag whose FP32 value lies within
//                    XF_TAU32 of the FP32 maximum (>= 50x the FP32 pipeline's worst observed error) is a
//                    candidate - one lag for ~95 % of the pair-windows - and xf_verify_kernel evaluates the
//                    candidates EXACTLY (FP64 time-domain sums over the staged windows), so the integer
//                    lag and cmax come from FP64 arithmetic on the original samples.
//   NBLS_XCORR_FFT   FP64 FFTs decide directly; near-ties (gap < XF_TAU) go to the time-domain kernel.
//
// Replaces LsBeam.correlate (lts_array/lts_array/classes/lts_classes.py:852-880 of the reference):
// cij = np.correlate(x_i, x_j, 'full') / sqrt(Ei*Ej); delay = argmax(cij); cmax = max(cij).
// np.correlate costs N^2 multiply-adds per pair-window; here a window of N samples is zero-padded to
// L >= N + lag_max (L = R1*R2*R3 with factors 2 and 3) and
//   spectra kernel : one CTA per (window, channel pair): both channels, scaled to unit energy, ride in
//                    one complex FFT (z = x_a + i x_b); the two half-spectra are unpacked and stored.
//   pair kernel    : one CTA per (window, two co-array pairs): Z = X_i1 conj(X_j1) + i X_i2 conj(X_j2)
//                    is built on the fly while loading, one inverse complex FFT returns the two
//                    correlation functions in its real and imaginary parts, and the arg-max over the
//                    wanted lags is taken from registers — the correlation function never exists in
//                    memory.
// FFT structure (both kernels): three register-radix passes R1, R2, R3 with two exchanges through a
// padded shared-memory array; pass 1 of the pair kernel handles the butterflies f0 and M1-f0
// together because the Hermitian halves Z[f], Z[L-f] come from the same spectrum entries.
//
// Exactness guard: the FFT value of a lag differs from the exact sum by far less than TAU (in units
// of the normalised correlation).  Each thread tracks the best and the second-best value; if the
// two best lags of a pair are closer than TAU the arg-max is not decided here — the pair-window is
// appended to a list and re-evaluated by the time-domain FP64 kernel (xcorr_direct.cu, list mode),
// whose first-maximum-wins semantics match np.argmax.  For every other pair-window the FFT arg-max
// provably equals the time-domain one.
#include <cuda_pipeline.h>
#include <float.h>
#include <limits.h>
#include <math.h>
#include <math_constants.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <vector>

#include "common.cuh"
#include "stats.cuh"
#include "fft_radix.cuh"

namespace nbls {

using namespace fft;

constexpr double XF_TAU = 1e-10;   // FP64 FFT: near-tie threshold, normalised correlation units
constexpr float XF_TAU32 = 2e-5f;  // FP32 screen: candidate band below the FP32 maximum (error seen: <= 4e-7)
constexpr double XF_EPS32 = 0.5 * 2e-5;   // = XF_TAU32 / 2: the error of one FP32 screen value the candidate band allows for
// FP32 pair kernel: true = pass 1 multiplies the spectra as it loads them (Hermitian partner tasks, as the FP64 kernel);
// false = a staging pass writes the cross-spectra to shared memory in natural order first.  Measured on the c3 step:
// 244.8 ms against 254.0 ms (fewer shared-memory passes and one barrier less outweigh the shorter load bursts)
constexpr bool XF_FUSED_PASS1 = true;
constexpr int XF_KC = 4;           // candidate lags kept per pair-window (more -> time-domain kernel)

struct XfArgs {
    const double* y;      // band base, [C] rows of stride ldy
    int64_t ldy;
    int C, P;
    int N, hop;
    int64_t first;        // sample index of window 0 of the band
    int w0, wn;           // windows [w0, w0+wn) of the band are processed by this launch
    int lag_max;
    int SP;               // spectrum row length L/2+1
    const void* tw;       // [L] e^{+2 pi i k/L}, complex of the transform's precision
    void* spec;           // [wn][C][SP] complex of the transform's precision
    double* energy;       // [wn][C]
    int32_t* lag;         // [W][P] band level
    double* cmax;         // [W][P] or nullptr
    int* flag_list;       // [>= wn*P] items w*P+p (band level) to re-evaluate in the time domain
    int* flag_count;
    // spec, energy, part, cand and sval rows are indexed by wl = (band-level window index) - w0, whatever list the
    // launch walks, so a later launch over a subset of the batch finds the rows of an earlier one
    int32_t* cand;        // FP32 screen: [wb*P][XF_KC] candidate lags of batch item wl*P+p, INT_MIN = unused
    float* sval;          // FP32 screen: [wb*P] the FP32 maximum (normalised)
    unsigned long long* stat;   // [0] candidates verified, [1] items sent to the time-domain kernel,
                                // [2] max |FP32 max - exact max| as float bits
    // two-stage lag search (FP32 screen only).  Stage A searches |lag| <= k0 with a shorter transform and proves per
    // pair-window, by Cauchy-Schwarz over partial energies, that no lag beyond k0 can reach the maximum found; the
    // pair-windows where the proof fails are listed (pf_list) and redone over all lags (stage B: spectra of their windows
    // in window-list mode, pair kernel and xf_verify_items_kernel in item mode).
    int k0;               // stage A: lag bound of the search (>= 0); otherwise -1
    int lag_dead;         // lag written for a zero-energy pair: -(N-1) or -max_lag of the request
    double* part;         // stage A: [wn][C][2] = sum of x^2 over n < N-k0-1 (head) and over n > k0 (tail)
    int* pf_list;         // stage B in: batch items wl*P+p whose proof failed, ascending (xf_compact_* kernels)
    int* pf_count;        // [0] number of such items, [1] number of windows holding at least one
    unsigned char* pf_flag;   // stage A out: [wn*P] 1 = the proof failed for batch item wl*P+p
    const int* wlist;     // window-list mode: band-level window index of launch slot s; nullptr = w0 + s
    const int* items;     // item mode (pair kernel, xf_verify_items_kernel): sorted batch items wl*P+p, two per pair CTA
    int n_items;
    // audit (NBLS_OPT_SCREEN_AUDIT): the pair kernel also stores the whole circular correlation function of every
    // pair-window with wl < dump_wn - [dump_wn * P][L] reals of the transform's precision, unnormalised (value * L)
    void* dump;
    int dump_wn;
    // lazy verification (NBLS_OPT_LAZY_VERIFY, FP32 screen, per-pair cmax not requested): a pair-window with ONE candidate
    // keeps the screen's lag and value (apx = 1); only the others (need_list) are evaluated in FP64, plus, per window, the
    // pairs whose value decides the median (xf_median_select_kernel).  The screen's values are taken to be within XF_EPS32
    // of the exact ones - the same premise the candidate band rests on.
    int lazy;
    unsigned char* apx;   // [wb*P] 1 = cmax of batch item wl*P+p is the FP32 screen value
    int* need_list;       // [wb*P] batch items to evaluate exactly (any order: items are independent)
    int* need_count;
    const int* items_count;   // item mode of xf_verify_items_kernel: number of items on the device (else n_items)
};

__device__ __forceinline__ int win_index(const XfArgs& g, int s) { return g.wlist ? g.wlist[s] : g.w0 + s; }

template <int R1_, int R2_, int R3_>
struct Shape {
    static constexpr int R1 = R1_, R2 = R2_, R3 = R3_;
    static constexpr int L = R1 * R2 * R3;
    static constexpr int M1 = R2 * R3;              // length of the sub-transforms after pass 1
    static constexpr int M2 = R3;                   // ... after pass 2
    static constexpr int ROW = R3 + 1;              // padded block stride: one spare slot per R3
    static constexpr int SMEM_ELEMS = ROW * R1 * R2;
    // position of natural index i (pass-1 output order n1*M1 + f) in the padded array
    static __device__ __forceinline__ int pad(int i) { return i + i / R3; }
    // position of output index n after the in-place pass 3
    static __device__ __forceinline__ int pos(int n) {
        return ROW * ((n % R1) * R2 + (n / R1) % R2) + n / (R1 * R2);
    }
};

__device__ __forceinline__ double2 ldg2(const double2* p) { return __ldg(p); }
__device__ __forceinline__ float2 ldg2(const float2* p) { return __ldg(p); }

// pass 1 tail: radix-R1 butterfly of the inputs f0 + M1*k, twiddle from the table, store
template <class S, class V>
__device__ __forceinline__ void pass1_table(V (&z)[S::R1], int f0, V* sm, const V* tw) {
    Dft<S::R1>::run(z);
#pragma unroll
    for (int n1 = 0; n1 < S::R1; ++n1) {
        const V v = n1 == 0 ? z[0] : cmul(z[n1], ldg2(tw + f0 * n1));
        sm[S::pad(n1 * S::M1 + f0)] = v;
    }
}

// pass 2: radix-R2 butterflies inside every sub-transform, in place
template <class S, int T, class V>
__device__ __forceinline__ void pass2(V* sm, const V* tw, int tid) {
    typedef real_t<V> F;
    for (int u = tid; u < S::L / S::R2; u += T) {
        const int n1 = u / S::M2, f1 = u - n1 * S::M2;
        V* base = sm + S::ROW * (n1 * S::R2) + f1;
        V z[S::R2];
#pragma unroll
        for (int k = 0; k < S::R2; ++k) z[k] = base[S::ROW * k];
        // twiddles w_L^{R1 f1 n2} = W1^n2: one table entry, the powers by a product tree
        V Wp[S::R2];
        Wp[0] = mk2((F)1, (F)0);
        Wp[1] = ldg2(tw + S::R1 * f1);
        Dft<S::R2>::run(z);
        powers<S::R2>(Wp);
#pragma unroll
        for (int n2 = 0; n2 < S::R2; ++n2) base[S::ROW * n2] = n2 == 0 ? z[0] : cmul(z[n2], Wp[n2]);
    }
}

// ---- spectra kernel -------------------------------------------------------------------------
// grid = wn * ceil(C/2); CTA (w, cp) transforms channels a = 2cp and b = 2cp+1 of window w.
// Forward DFT through the inverse machinery: Y = IDFT(conj z)  =>  Z = conj(Y).
template <class S, int T, int MINB, class V>
__global__ void __launch_bounds__(T, MINB) xf_spectra_kernel(XfArgs g) {
    typedef real_t<V> F;
    extern __shared__ __align__(16) unsigned char xf_smem_raw[];
    V* sm = reinterpret_cast<V*>(xf_smem_raw);
    const V* tw = static_cast<const V*>(g.tw);
    __shared__ double red[2][T / 32];
    __shared__ double rs_sh[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ncp = (g.C + 1) >> 1;
    const int wslot = blockIdx.x / ncp, cp = blockIdx.x - wslot * ncp;
    const int wband = win_index(g, wslot), wl = wband - g.w0;
    const int ca = 2 * cp, cb = 2 * cp + 1;
    const bool has_b = cb < g.C;
    const int64_t t0 = g.first + (int64_t)wband * g.hop;
    const double* xa = g.y + (int64_t)ca * g.ldy + t0;
    const double* xb = g.y + (int64_t)(has_b ? cb : ca) * g.ldy + t0;

    // ONE pass over the window: energies (same accumulation as the time-domain kernel: strided fma, then tree) and the
    // samples, rounded to the transform's precision but not yet scaled, into the transform array in natural order.  (A
    // first pass for the energies and a second, strided one for the transform inputs left the CTA waiting twice for
    // the same DRAM-resident samples.)  The unit-energy scale is applied when pass 1 picks its inputs up.
    double ea = 0.0, eb = 0.0;
    if (g.k0 >= 0) {
        // stage A of the two-stage search: also the head (n < N-k0-1) and tail (n > k0) energies of both channels
        __shared__ double redp[4][T / 32];
        const int nh = g.N - g.k0 - 1, nt = g.k0;
        double ha = 0.0, ta = 0.0, hb = 0.0, tb = 0.0;
        for (int n = tid; n < g.N; n += T) {
            const double va = xa[n], vb = xb[n];
            const double qa = va * va, qb = vb * vb;
            ea = fma(va, va, ea);           // the window energies keep the accumulation of every other path
            eb = fma(vb, vb, eb);
            if (n < nh) { ha += qa; hb += qb; }
            if (n > nt) { ta += qa; tb += qb; }
            sm[S::pad(n)] = mk2((F)va, (F)vb);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            ha += __shfl_xor_sync(0xffffffffu, ha, o); ta += __shfl_xor_sync(0xffffffffu, ta, o);
            hb += __shfl_xor_sync(0xffffffffu, hb, o); tb += __shfl_xor_sync(0xffffffffu, tb, o);
        }
        if (lane == 0) { redp[0][warp] = ha; redp[1][warp] = ta; redp[2][warp] = hb; redp[3][warp] = tb; }
        __syncthreads();
        if (tid < 4 && (tid < 2 || has_b)) {
            double v = 0.0;
            for (int i = 0; i < T / 32; ++i) v += redp[tid][i];
            g.part[((int64_t)wl * g.C + (tid < 2 ? ca : cb)) * 2 + (tid & 1)] = v;
        }
    } else {
        for (int n = tid; n < g.N; n += T) {
            const double va = xa[n], vb = xb[n];
            ea = fma(va, va, ea);
            eb = fma(vb, vb, eb);
            sm[S::pad(n)] = mk2((F)va, (F)vb);
        }
    }
    for (int n = g.N + tid; n < S::L; n += T) sm[S::pad(n)] = mk2((F)0, (F)0);      // zero padding of the window
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ea += __shfl_xor_sync(0xffffffffu, ea, o);
        eb += __shfl_xor_sync(0xffffffffu, eb, o);
    }
    if (lane == 0) { red[0][warp] = ea; red[1][warp] = eb; }
    __syncthreads();
    if (tid == 0) {
        double Ea = 0.0, Eb = 0.0;
        for (int i = 0; i < T / 32; ++i) { Ea += red[0][i]; Eb += red[1][i]; }
        if (!has_b) Eb = 0.0;
        g.energy[(int64_t)wl * g.C + ca] = Ea;
        if (has_b) g.energy[(int64_t)wl * g.C + cb] = Eb;
        // unit-energy scaling keeps both riders of a packed transform at the same magnitude
        rs_sh[0] = (Ea > 0.0 && isfinite(Ea)) ? 1.0 / sqrt(Ea) : 0.0;
        rs_sh[1] = (Eb > 0.0 && isfinite(Eb)) ? 1.0 / sqrt(Eb) : 0.0;
    }
    __syncthreads();
    const F rsa = (F)rs_sh[0], rsb = (F)rs_sh[1];

    // pass 1 in place: inputs conj(z)[n] = (x_a[n], -x_b[n]) * scale from the array (pad(f0 + M1 k) = pad(f0) + k (M1 + R2));
    // a thread reads and writes the same R1 slots
    for (int f0 = tid; f0 < S::M1; f0 += T) {
        V z[S::R1];
        const V* base = sm + S::pad(f0);
#pragma unroll
        for (int k = 0; k < S::R1; ++k) {
            // a non-finite or zero-energy rider has scale 0: its samples must not turn 0 * inf into NaN for the other rider
            V v = cmul_elem(base[k * (S::M1 + S::R2)], mk2(rsa, -rsb));
            if (rsa == (F)0) v.x = (F)0;
            if (rsb == (F)0) v.y = (F)0;
            z[k] = v;
        }
        pass1_table<S>(z, f0, sm, tw);
    }
    __syncthreads();
    pass2<S, T>(sm, tw, tid);
    __syncthreads();
    // pass 3 in place
    for (int u = tid; u < S::L / S::R3; u += T) {
        V* base = sm + S::ROW * u;
        V z[S::R3];
#pragma unroll
        for (int k = 0; k < S::R3; ++k) z[k] = base[k];
        Dft<S::R3>::run(z);
#pragma unroll
        for (int k = 0; k < S::R3; ++k) base[k] = z[k];
    }
    __syncthreads();
    // unpack: X_a[f] = (conj(Y[f]) + Y[L-f]) / 2,  X_b[f] = (conj(Y[f]) - Y[L-f]) / (2i)
    V* Xa = static_cast<V*>(g.spec) + ((int64_t)wl * g.C + ca) * g.SP;
    V* Xb = static_cast<V*>(g.spec) + ((int64_t)wl * g.C + cb) * g.SP;
    constexpr F half = (F)0.5;
    // Consecutive f sit R2 rows apart in the in-place layout (pos(): the row is led by f % R1), which is a multiple of 32
    // banks.  A warp therefore takes A consecutive f (one full 32-byte sector of each spectrum) times 32/A values of
    // (f / R1) % R2, which are one row apart: A-way instead of R1-way bank conflicts, sectors still written whole.
    constexpr int A = S::R1 % 4 == 0 ? 4 : (S::R1 % 2 == 0 ? 2 : 1), B = 32 / A;
    constexpr int NJ = (S::L / 2 + 1 + S::R1 * B - 1) / (S::R1 * B);
    const int la = lane % A, lb = lane / A;
    for (int tile = warp; tile < (S::R1 / A) * NJ; tile += T / 32) {
        const int i = tile % (S::R1 / A), j = tile / (S::R1 / A);
        const int f = A * i + la + S::R1 * (B * j + lb);
        if (f > S::L / 2) continue;
        const V yf = sm[S::pos(f)];
        const V ym = sm[S::pos(f == 0 ? 0 : S::L - f)];
        Xa[f] = cscale(half, cadd(mk2(yf.x, -yf.y), ym));
        if (has_b) Xb[f] = cscale(-half, cadd(mk2(yf.y, yf.x), mk2(ym.y, -ym.x)));
    }
}

// ---- pair kernel ----------------------------------------------------------------------------
struct Top {       // best and second-best value seen, lag of the best
    double b1, b2;
    int l1;
};
__device__ __forceinline__ void top_push(Top& t, double v, int lag, bool ok) {
    const bool p1 = ok & (v > t.b1), p2 = ok & (v > t.b2);     // bitwise: no short-circuit branches
    t.b2 = p1 ? t.b1 : (p2 ? v : t.b2);
    t.b1 = p1 ? v : t.b1;
    t.l1 = p1 ? lag : t.l1;
}
__device__ __forceinline__ void top_merge(Top& a, double b1, double b2, int l1) {
    if (b1 > a.b1) {
        a.b2 = fmax(a.b1, b2);
        a.b1 = b1;
        a.l1 = l1;
    } else {
        a.b2 = fmax(a.b2, b1);     // an exact tie lands here: gap 0 -> decided in the time domain
    }
}
__device__ __forceinline__ void pair_channels(int p, int C, int& ci, int& cj) {
    // i<j row-major (lts_classes.py:840): row i starts at s(i) = i (2C - i - 1) / 2.  Closed form through the root of
    // s(i) = p in FP32 (exact integers below 2^24), then at most one step either way
    const float b = (float)(2 * C - 1);
    int i = (int)((b - sqrtf(fmaxf(b * b - 8.0f * (float)p, 0.0f))) * 0.5f);
    i = max(0, min(i, C - 2));
    if (i * (2 * C - i - 1) / 2 > p) --i;
    else if (i + 1 <= C - 2 && (i + 1) * (2 * C - i - 2) / 2 <= p) ++i;
    ci = i;
    cj = i + 1 + p - i * (2 * C - i - 1) / 2;
}
__device__ __forceinline__ int pair_index(int ci, int cj, int C) { return ci * (2 * C - ci - 1) / 2 + (cj - ci - 1); }

// pass-1 stores of the partner butterflies f0 = t (A) and M1 - t (B); W[n] = w_L^{t n}, so B's twiddle
// is w_L^{(M1-t) n} = e^{2 pi i n/R1} conj(W[n]); the constant factor e^{2 pi i n/R1} was applied for free
// by feeding B's inputs to the butterfly rotated by one position (DFT shift theorem)
// pA = sm + pad(t), pB = sm + pad(M1 - t); pad(n1 M1 + x) = pad(x) + n1 (M1 + R2): constant offsets from the two bases
template <class S, int N1, class V>
__device__ __forceinline__ void store_pair(const V (&zA)[S::R1], const V (&zB)[S::R1], const V (&W)[S::R1], V* pA, V* pB) {
    if constexpr (N1 < S::R1) {
        if constexpr (N1 == 0) {
            pA[0] = zA[0];
            pB[0] = zB[0];
        } else {
            pA[N1 * (S::M1 + S::R2)] = cmul(zA[N1], W[N1]);
            pB[N1 * (S::M1 + S::R2)] = cmulc(zB[N1], W[N1]);
        }
        store_pair<S, N1 + 1>(zA, zB, W, pA, pB);
    }
}

// grid = wn * ceil(P/2); CTA (w, q) correlates pairs 2q (real part) and 2q+1 (imaginary part)
template <class S, int T, int MINB, class V>
__global__ void __launch_bounds__(T, MINB) xf_pair_kernel(XfArgs g) {
    typedef real_t<V> F;
    constexpr bool SCREEN = sizeof(F) == 4;
    extern __shared__ __align__(16) unsigned char xf_smem_raw[];
    V* sm = reinterpret_cast<V*>(xf_smem_raw);
    const V* tw = static_cast<const V*>(g.tw);
    constexpr int R1 = S::R1, L = S::L, M1 = S::M1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int npp = (g.P + 1) >> 1;
    // the two riders of this CTA: (data row wl0, pair p0) on the real part, (wl1, p1) on the imaginary part
    int wl0, wl1, p0, p1;
    bool has1;
    if (g.items) {
        const int e0 = 2 * (int)blockIdx.x, e1 = e0 + 1;
        has1 = e1 < g.n_items;
        const int it0 = g.items[e0], it1 = g.items[has1 ? e1 : e0];
        wl0 = it0 / g.P; p0 = it0 - wl0 * g.P;
        wl1 = it1 / g.P; p1 = it1 - wl1 * g.P;
    } else {
        const int wslot = blockIdx.x / npp, q = blockIdx.x - wslot * npp;
        wl0 = wl1 = win_index(g, wslot) - g.w0;
        p0 = 2 * q; p1 = 2 * q + 1;
        has1 = p1 < g.P;
        if (!has1) p1 = p0;
    }
    int i0, j0, i1, j1;
    pair_channels(p0, g.C, i0, j0);
    if (g.items) pair_channels(p1, g.C, i1, j1);
    else if (!has1) { i1 = i0; j1 = j0; }
    else if (j0 + 1 < g.C) { i1 = i0; j1 = j0 + 1; }      // p1 = p0 + 1: the next pair of the same row, or the first of the next
    else { i1 = i0 + 1; j1 = i0 + 2; }
    const V* sbase0 = static_cast<const V*>(g.spec) + (int64_t)wl0 * g.C * g.SP;
    const V* sbase1 = static_cast<const V*>(g.spec) + (int64_t)wl1 * g.C * g.SP;
    const V *Xi0 = sbase0 + (int64_t)i0 * g.SP, *Xj0 = sbase0 + (int64_t)j0 * g.SP;
    const V *Xi1 = sbase1 + (int64_t)i1 * g.SP, *Xj1 = sbase1 + (int64_t)j1 * g.SP;
    const bool same_i = i0 == i1 && wl0 == wl1;

    // cross-spectra of entry gidx: a = X_i0 conj(X_j0), b = X_i1 conj(X_j1)
    auto cross = [&](int gidx, V& a, V& b) {
        const V xi = ldg2(Xi0 + gidx);
        a = cmulc(xi, ldg2(Xj0 + gidx));
        if (has1) b = cmulc(same_i ? xi : ldg2(Xi1 + gidx), ldg2(Xj1 + gidx));
        else b = mk2((F)0, (F)0);
    };
    // Z[f] for any f in [0, L): Z = a + i b on the stored half, conj(a) + i conj(b) on the mirrored one
    auto fetch = [&](int f) -> V {
        V a, b;
        if (f <= L / 2) { cross(f, a, b); return cadd_i(a, b); }
        cross(L - f, a, b);
        return cadd_i_conj(a, b);
    };

    if constexpr (SCREEN && !XF_FUSED_PASS1) {
        // ---- FP32: cross-spectra straight into shared memory (natural order; every load independent, so the L2
        // latency is hidden by memory-level parallelism), then a plain in-place radix-R1 pass.  Register use stays
        // low enough for 24 warps per SM. ----
        // A CTA without a second rider carries its first one twice (p1 = p0 above), so the loop has no rider test; f and
        // L - f advance by T per step, hence pad() of both by the constant PSTEP: two base addresses, immediate offsets.
        // Thread 0's first mirror store (f = 0, "Z[L]") lands in the spare slot behind the array; f = L/2 has no mirror
        // and is done by one thread after the loop.
        static_assert((L / 2) % T == 0 && T % S::R3 == 0, "load loop: whole steps, constant pad() stride");
        constexpr int NK = L / 2 / T, PSTEP = T + T / S::R3;
        V* pf = sm + S::pad(tid);
        V* pm = sm + S::pad(L - tid);
        const V *gi0 = Xi0 + tid, *gj0 = Xj0 + tid, *gi1 = Xi1 + tid, *gj1 = Xj1 + tid;
        if (same_i) {
#pragma unroll 4
            for (int k = 0; k < NK; ++k) {
                const V xi = ldg2(gi0 + k * T), xj = ldg2(gj0 + k * T), yj = ldg2(gj1 + k * T);
                const V a = cmulc(xi, xj), b = cmulc(xi, yj);
                pf[k * PSTEP] = cadd_i(a, b);                    // Z[f]   = a + i b
                pm[-k * PSTEP] = cadd_i_conj(a, b);              // Z[L-f] = conj(a) + i conj(b)
            }
        } else {
#pragma unroll 4
            for (int k = 0; k < NK; ++k) {
                const V xi = ldg2(gi0 + k * T), xj = ldg2(gj0 + k * T), yi = ldg2(gi1 + k * T), yj = ldg2(gj1 + k * T);
                const V a = cmulc(xi, xj), b = cmulc(yi, yj);
                pf[k * PSTEP] = cadd_i(a, b);
                pm[-k * PSTEP] = cadd_i_conj(a, b);
            }
        }
        if (tid == 0) {
            const V a = cmulc(ldg2(Xi0 + L / 2), ldg2(Xj0 + L / 2)), b = cmulc(ldg2(Xi1 + L / 2), ldg2(Xj1 + L / 2));
            sm[S::pad(L / 2)] = cadd_i(a, b);
        }
        __syncthreads();
        for (int f0 = tid; f0 < M1; f0 += T) {
            V z[R1], W[R1];
            V* base = sm + S::pad(f0);           // pad(f0 + M1 k) = pad(f0) + k (M1 + R2)
#pragma unroll
            for (int k = 0; k < R1; ++k) z[k] = base[k * (M1 + S::R2)];
            W[0] = mk2((F)1, (F)0);
            if constexpr (R1 > 1) W[1] = ldg2(tw + f0);
            Dft<R1>::run(z);
            powers<R1>(W);
#pragma unroll
            for (int n1 = 0; n1 < R1; ++n1) base[n1 * (M1 + S::R2)] = n1 == 0 ? z[0] : cmul(z[n1], W[n1]);
        }
    } else {
    // ---- pass 1: task t owns the butterflies f0 = t and M1 - t (Hermitian partners) ----
    if (warp == T / 32 - 1) {
        // task 0: the two self-mirrored butterflies f0 = 0 and f0 = M1/2.  Their 2 R1 inputs are fetched by 2 R1 lanes
        // at once (one L2 round trip instead of 2 R1 on the CTA's critical path) and handed to lanes 0 and 1; the
        // last warp has the fewest regular tasks
        static_assert(2 * R1 <= 32, "special task needs 2 R1 lanes");
        V zf = mk2((F)0, (F)0);
        if (lane < 2 * R1) zf = fetch((lane / R1) * (M1 / 2) + M1 * (lane % R1));
        V z[R1];
#pragma unroll
        for (int k = 0; k < R1; ++k) {
            z[k].x = __shfl_sync(0xffffffffu, zf.x, k + R1 * (lane & 1));
            z[k].y = __shfl_sync(0xffffffffu, zf.y, k + R1 * (lane & 1));
        }
        if (lane < 2) pass1_table<S>(z, lane * (M1 / 2), sm, tw);
    }
    {
        // f = t + M1 k lies in the stored half iff k < R1/2 (R1 even) or k <= (R1-1)/2 (R1 odd); otherwise its
        // mirror L - f is loaded: two base offsets per task, the rest are immediate offsets
        constexpr int HALF = (R1 % 2 == 0) ? R1 / 2 : (R1 + 1) / 2;
        V vi0[R1], vj0[R1], vi1[R1], vj1[R1], w1;
        // a CTA without a second rider carries its first one twice (p1 = p0), so no rider test here.  Row bases are uniform;
        // per task two pointers per row (direct: + t, mirrored: + L - t), every load an immediate offset from one of them
        auto load_task = [&](int t) {
            const V *di0 = Xi0 + t, *dj0 = Xj0 + t, *dj1 = Xj1 + t, *di1 = Xi1 + t;
            const V *mi0 = Xi0 + (L - t), *mj0 = Xj0 + (L - t), *mj1 = Xj1 + (L - t), *mi1 = Xi1 + (L - t);
#pragma unroll
            for (int k = 0; k < R1; ++k) {
                if (k < HALF) {
                    vi0[k] = ldg2(di0 + M1 * k);
                    vj0[k] = ldg2(dj0 + M1 * k);
                    vj1[k] = ldg2(dj1 + M1 * k);
                    vi1[k] = same_i ? vi0[k] : ldg2(di1 + M1 * k);
                } else {
                    vi0[k] = ldg2(mi0 - M1 * k);
                    vj0[k] = ldg2(mj0 - M1 * k);
                    vj1[k] = ldg2(mj1 - M1 * k);
                    vi1[k] = same_i ? vi0[k] : ldg2(mi1 - M1 * k);
                }
            }
            w1 = ldg2(tw + t);
        };
        // software pipeline: the loads of the task after this one are in flight while this one's butterflies run.
        // The prefetch buffer is 4 R1 complex registers: affordable up to R1 = 8 (FP64) / 12 (FP32); beyond, load then use.
        // (the prefetch buffer on top of the FP32 kernel spills under its 80-register cap)
        constexpr bool PIPE = !SCREEN && R1 <= 8;
        int t = tid == 0 ? T : tid;      // task 0 is the special one above
        if (PIPE && t < M1 / 2) load_task(t);
        while (t < M1 / 2) {
            if (!PIPE) load_task(t);
            V zA[R1], zB[R1], W[R1];
#pragma unroll
            for (int k = 0; k < R1; ++k) {
                const bool direct = k < HALF;
                const V a = cmulc(vi0[k], vj0[k]), b = cmulc(vi1[k], vj1[k]);
                const V zg = cadd_i(a, b);     // Z[g]
                const V zm = cadd_i_conj(a, b);     // Z[L-g]
                zA[k] = direct ? zg : zm;
                zB[(R1 - k) % R1] = direct ? zm : zg;      // slot R1-1-k, rotated by one (see store_pair)
            }
            W[0] = mk2((F)1, (F)0);
            if constexpr (R1 > 1) W[1] = w1;
            const int tn = t + T;
            if (PIPE && tn < M1 / 2) load_task(tn);
            Dft<R1>::run(zA);
            Dft<R1>::run(zB);
            powers<R1>(W);
            store_pair<S, 0>(zA, zB, W, sm + S::pad(t), sm + S::pad(M1 - t));
            t = tn;
        }
    }
    }
    __syncthreads();
    pass2<S, T>(sm, tw, tid);
    // with R2 = R3 = 32 and one warp per n1, a warp's pass-3 rows are exactly the rows it wrote in pass 2
    if constexpr (S::R2 == 32 && S::R3 == 32 && T == 32 * R1) __syncwarp();
    else __syncthreads();

    const int lag_max = g.lag_max;
    constexpr int STEP = R1 * S::R2;
    if constexpr (SCREEN) {
        // ---- pass 3 + screening from registers: the FP32 maximum over the wanted lags, then every lag within
        // XF_TAU32 of it becomes a candidate for the exact FP64 evaluation (xf_verify_kernel) ----
        static_assert(T * S::R3 >= L, "at most one pass-3 row per thread");
        __shared__ float smax[2][T / 32];
        __shared__ int scnt[2];
        __shared__ int scl[2][XF_KC];
        const bool act = tid < L / S::R3;     // T is a multiple of 32, the row count need not be
        const int u = act ? tid : 0;
        V z[S::R3];
        {
            const V* base = sm + S::ROW * u;
#pragma unroll
            for (int k = 0; k < S::R3; ++k) z[k] = base[k];
        }
        Dft<S::R3>::run(z);
        const int n1 = u / S::R2, n2 = u - n1 * S::R2;
        const int nb = n1 + R1 * n2;
        if (g.dump && act && !g.items) {
            F* d = static_cast<F*>(g.dump);
#pragma unroll
            for (int n3 = 0; n3 < S::R3; ++n3) {
                if (wl0 < g.dump_wn) d[((int64_t)wl0 * g.P + p0) * L + nb + STEP * n3] = z[n3].x;
                if (has1 && wl1 < g.dump_wn) d[((int64_t)wl1 * g.P + p1) * L + nb + STEP * n3] = z[n3].y;
            }
        }
        // circular index n = nb + R1 R2 n3 stands for lag n (n <= lag_max) or n - L (n >= L - lag_max): the wanted
        // n3 are [0, n3a] and [n3b, R3)
        const int n3a = !act ? -1 : (lag_max >= nb ? (lag_max - nb) / STEP : -1);
        const int n3b = !act ? S::R3 : (max(L - lag_max - nb, 0) + STEP - 1) / STEP;
        const float NEG = -CUDART_INF_F;
        float mr = NEG, mi = NEG;
#pragma unroll
        for (int n3 = 0; n3 < S::R3; ++n3) {
            const bool ok = (n3 <= n3a) | (n3 >= n3b);
            mr = fmaxf(mr, ok ? z[n3].x : NEG);
            mi = fmaxf(mi, ok ? z[n3].y : NEG);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mr = fmaxf(mr, __shfl_xor_sync(0xffffffffu, mr, o));
            mi = fmaxf(mi, __shfl_xor_sync(0xffffffffu, mi, o));
        }
        if (lane == 0) { smax[0][warp] = mr; smax[1][warp] = mi; }
        if (tid < 2) scnt[tid] = 0;
        // the two threads that finish the CTA need their rider's energies (and the stage-A partial energies): asynchronous
        // 8-byte copies into shared memory start here, fly under the two barriers and cost no registers, instead of a
        // chain of dependent global loads holding the CTA's 68 KB of shared memory at the very end
        __shared__ double s_en[2][6];
        if (tid < 2) {
            const int wl_ = tid == 0 ? wl0 : wl1, ci_ = tid == 0 ? i0 : i1, cj_ = tid == 0 ? j0 : j1;
            __pipeline_memcpy_async(&s_en[tid][0], g.energy + (int64_t)wl_ * g.C + ci_, 8);
            __pipeline_memcpy_async(&s_en[tid][1], g.energy + (int64_t)wl_ * g.C + cj_, 8);
            if (g.lazy && g.k0 >= 0) {
                const double* pi = g.part + ((int64_t)wl_ * g.C + ci_) * 2;
                const double* pj = g.part + ((int64_t)wl_ * g.C + cj_) * 2;
                __pipeline_memcpy_async(&s_en[tid][2], pi, 8);
                __pipeline_memcpy_async(&s_en[tid][3], pi + 1, 8);
                __pipeline_memcpy_async(&s_en[tid][4], pj, 8);
                __pipeline_memcpy_async(&s_en[tid][5], pj + 1, 8);
            }
            __pipeline_commit();
        }
        __syncthreads();
        const float lmr = mr, lmi = mi;       // warp-level maxima
        mr = smax[0][0]; mi = smax[1][0];
#pragma unroll
        for (int i = 1; i < T / 32; ++i) { mr = fmaxf(mr, smax[0][i]); mi = fmaxf(mi, smax[1][i]); }
        const float thr_r = mr - XF_TAU32 * (float)L, thr_i = mi - XF_TAU32 * (float)L;
        if ((lmr >= thr_r) | (lmi >= thr_i)) {      // warp-uniform: only warps holding a value inside the band rescan
            // The unrolled part only builds two hit masks (z[] lives in registers, so n3 must be a constant); the rare hits
            // are then handled by a rolled loop.  One warp of every CTA comes through here with a cold instruction cache
            // while the other warps wait at the barrier below, so this code has to stay small.
            static_assert(S::R3 <= 32, "one mask bit per pass-3 output");
            unsigned hbr = 0u, hbi = 0u;
#pragma unroll
            for (int n3 = 0; n3 < S::R3; ++n3) {
                hbr |= (z[n3].x >= thr_r ? 1u : 0u) << n3;
                hbi |= (z[n3].y >= thr_i ? 1u : 0u) << n3;
            }
            unsigned okm = 0u;                         // wanted n3: [0, n3a] and [n3b, R3)
            if (n3a >= 0) okm |= n3a >= 31 ? 0xffffffffu : ((2u << n3a) - 1u);
            if (n3b < S::R3) okm |= n3b <= 0 ? 0xffffffffu : ~((1u << n3b) - 1u);
            if (S::R3 < 32) okm &= (1u << (S::R3 & 31)) - 1u;
            hbr &= okm;
            hbi &= okm;
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                unsigned hb = c == 0 ? hbr : hbi;
                while (hb) {
                    const int n3 = __ffs(hb) - 1;
                    hb &= hb - 1u;
                    const int n = nb + STEP * n3;
                    const int lg = n <= lag_max ? n : n - L;
                    const int s_ = atomicAdd(&scnt[c], 1);
                    if (s_ < XF_KC) scl[c][s_] = lg;
                }
            }
        }
        __syncthreads();
        if (tid < 2 && (tid == 0 || has1)) {
            const int p = tid == 0 ? p0 : p1, wl = tid == 0 ? wl0 : wl1;
            const int ci = tid == 0 ? i0 : i1, cj = tid == 0 ? j0 : j1;
            __pipeline_wait_prior(0);
            const double den = sqrt(s_en[tid][0] * s_en[tid][1]);
            const int64_t item = (int64_t)(g.w0 + wl) * g.P + p;
            const int64_t loc = (int64_t)wl * g.P + p;
            const int cnt = scnt[tid];
            int32_t* cd = g.cand + loc * XF_KC;
            bool give = false;
            if (!(den > 0.0)) {
                // zero-energy pair: cij is all-NaN, np.argmax gives index 0 = the lowest lag searched
                g.lag[item] = g.lag_dead;
                if (g.cmax) g.cmax[item] = __longlong_as_double(0x7ff8000000000000LL);
            } else if (!isfinite(den) || cnt < 1 || cnt > XF_KC) {
                // non-finite data, or too many near-tied lags: the time-domain kernel evaluates every lag
                const int slot = atomicAdd(g.flag_count, 1);
                g.flag_list[slot] = (int)item;
                atomicAdd(g.stat + 1, 1ULL);
            } else {
                give = true;
            }
            // the candidates arrive in the order the atomics ran in; sorted by lag, slot s always holds the same lag, so the
            // exact sums (slot 0: slice tree, slots 1..: one warp-strided pass) do not depend on the arrival order
            int cl[XF_KC];
#pragma unroll
            for (int s_ = 0; s_ < XF_KC; ++s_) cl[s_] = (give && s_ < cnt) ? scl[tid][s_] : INT_MAX;
#pragma unroll
            for (int i_ = 1; i_ < XF_KC; ++i_)
#pragma unroll
                for (int j_ = XF_KC - 1; j_ >= i_; --j_)
                    if (cl[j_] < cl[j_ - 1]) { const int t_ = cl[j_]; cl[j_] = cl[j_ - 1]; cl[j_ - 1] = t_; }
#pragma unroll
            for (int s_ = 0; s_ < XF_KC; ++s_) cd[s_] = cl[s_] == INT_MAX ? INT_MIN : cl[s_];
            const float sv = (tid == 0 ? mr : mi) / (float)L;
            g.sval[loc] = sv;
            if (g.lazy) {
                bool approx = false;
                // one pair-window in 256 (a fixed hash of its index) is evaluated exactly although it has a single candidate:
                // the consistency guard and the max_screen_err statistic of nbls_xcorr_stats keep watching the screen
                const bool sampled = ((unsigned)loc * 2654435761u) >> 24 == 0u;
                if (give && cnt == 1 && !sampled) {
                    // one lag inside the candidate band: it is the arg-max; its value is known to XF_EPS32
                    approx = true;
                    g.lag[item] = cl[0];
                    if (g.cmax) g.cmax[item] = (double)sv;
                    if (g.k0 >= 0) {
                        // stage-A proof (see xf_verify_kernel) on the screen value less its error, in normalised units
                        const double b2 = fmax(s_en[tid][3] * s_en[tid][4], s_en[tid][2] * s_en[tid][5]);
                        if (!(((double)sv - XF_EPS32) * den > sqrt(b2) * (1.0 + 1e-12))) g.pf_flag[loc] = 1;
                    }
                } else if (give) {
                    g.need_list[atomicAdd(g.need_count, 1)] = (int)loc;
                }
                g.apx[loc] = approx ? 1 : 0;
            }
        }
    } else {
    // ---- pass 3 + arg-max straight from registers ----
    __shared__ double rb1[2][T / 32], rb2[2][T / 32];
    __shared__ int rl1[2][T / 32];
    constexpr int NCH = 4;     // independent arg-max chains per component (the update is a serial dependency)
    Top accr = {-CUDART_INF, -CUDART_INF, INT_MAX}, acci = {-CUDART_INF, -CUDART_INF, INT_MAX};
    for (int u = tid; u < L / S::R3; u += T) {
        Top tr[NCH], ti[NCH];
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            tr[c].b1 = tr[c].b2 = ti[c].b1 = ti[c].b2 = -CUDART_INF;
            tr[c].l1 = ti[c].l1 = INT_MAX;
        }
        const V* base = sm + S::ROW * u;
        V z[S::R3];
#pragma unroll
        for (int k = 0; k < S::R3; ++k) z[k] = base[k];
        Dft<S::R3>::run(z);
        const int n1 = u / S::R2, n2 = u - n1 * S::R2;
        const int nb = n1 + R1 * n2;
        if (g.dump && !g.items) {
            F* d = static_cast<F*>(g.dump);
#pragma unroll
            for (int n3 = 0; n3 < S::R3; ++n3) {
                if (wl0 < g.dump_wn) d[((int64_t)wl0 * g.P + p0) * L + nb + STEP * n3] = z[n3].x;
                if (has1 && wl1 < g.dump_wn) d[((int64_t)wl1 * g.P + p1) * L + nb + STEP * n3] = z[n3].y;
            }
        }
        // circular index n = nb + R1 R2 n3 stands for lag n (n <= lag_max) or n - L (n >= L - lag_max): the wanted
        // n3 are [0, n3a] and [n3b, R3); the chains record n3, the lag is rebuilt afterwards
        const int n3a = lag_max >= nb ? (lag_max - nb) / STEP : -1;
        const int n3b = (max(L - lag_max - nb, 0) + STEP - 1) / STEP;
#pragma unroll
        for (int n3 = 0; n3 < S::R3; ++n3) {
            const bool ok = (n3 <= n3a) | (n3 >= n3b);
            top_push(tr[n3 % NCH], z[n3].x, n3, ok);
            top_push(ti[n3 % NCH], z[n3].y, n3, ok);
        }
#pragma unroll
        for (int c = 0; c < NCH; ++c) {      // n3 -> lag, then fold the chain into the thread's result
            const int nr = nb + STEP * tr[c].l1, ni = nb + STEP * ti[c].l1;
            const int lr = tr[c].l1 == INT_MAX ? INT_MAX : (nr <= lag_max ? nr : nr - L);
            const int li = ti[c].l1 == INT_MAX ? INT_MAX : (ni <= lag_max ? ni : ni - L);
            top_merge(accr, tr[c].b1, tr[c].b2, lr);
            top_merge(acci, ti[c].b1, ti[c].b2, li);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        top_merge(accr, __shfl_xor_sync(0xffffffffu, accr.b1, o), __shfl_xor_sync(0xffffffffu, accr.b2, o),
                  __shfl_xor_sync(0xffffffffu, accr.l1, o));
        top_merge(acci, __shfl_xor_sync(0xffffffffu, acci.b1, o), __shfl_xor_sync(0xffffffffu, acci.b2, o),
                  __shfl_xor_sync(0xffffffffu, acci.l1, o));
    }
    if (lane == 0) {
        rb1[0][warp] = accr.b1; rb2[0][warp] = accr.b2; rl1[0][warp] = accr.l1;
        rb1[1][warp] = acci.b1; rb2[1][warp] = acci.b2; rl1[1][warp] = acci.l1;
    }
    __syncthreads();
    if (tid < 2 && (tid == 0 || has1)) {
        Top b = {rb1[tid][0], rb2[tid][0], rl1[tid][0]};
        for (int i = 1; i < T / 32; ++i) top_merge(b, rb1[tid][i], rb2[tid][i], rl1[tid][i]);
        const int p = tid == 0 ? p0 : p1, wl = tid == 0 ? wl0 : wl1;
        const int ci = tid == 0 ? i0 : i1, cj = tid == 0 ? j0 : j1;
        const double Ei = g.energy[(int64_t)wl * g.C + ci], Ej = g.energy[(int64_t)wl * g.C + cj];
        const double den = sqrt(Ei * Ej);
        const int64_t item = (int64_t)(g.w0 + wl) * g.P + p;
        if (!(den > 0.0)) {
            // zero-energy pair: cij is all-NaN, np.argmax gives index 0 = the lowest lag searched
            g.lag[item] = g.lag_dead;
            if (g.cmax) g.cmax[item] = __longlong_as_double(0x7ff8000000000000LL);
        } else {
            g.lag[item] = b.l1 == INT_MAX ? -lag_max : b.l1;
            if (g.cmax) g.cmax[item] = b.b1 / (double)L;
            // undecided (near-tie, or non-finite data): the time-domain kernel settles it
            if (!(b.b1 - b.b2 > XF_TAU * (double)L) || !isfinite(den)) {
                const int slot = atomicAdd(g.flag_count, 1);
                g.flag_list[slot] = (int)item;
                atomicAdd(g.stat + 1, 1ULL);
            }
        }
    }
    }
}

// ---- verification kernel (FP32 screen only) -------------------------------------------------
// Exact FP64 value of every candidate lag: c[k] = sum_n x_i[n+k] x_j[n] over the original filtered samples.
// grid = wn * ceil((C-1)/IB); CTA (w, ib) stages the windows of channels i = IB ib .. IB ib + IB-1 in shared
// memory.  Warp (slice, group) owns one of XV_SL time slices of the window and every XV_JG-th channel j > i: it
// streams its slice of x_j once from L2 (8 loads in flight per lane) and accumulates the first candidate of the
// (up to IB) pairs (i, j) together; the part of a slice whose shifted indices all stay inside the window runs
// without bounds checks.  Slice sums are combined in a fixed order (deterministic); further candidates (near-ties,
// ~5 % of the pairs) are evaluated warp by warp afterwards.  The largest exact value wins, the lowest lag on an
// exact tie (np.argmax); cmax = value / sqrt(Ei Ej).
constexpr int XV_SL = 8;      // time slices of a window; a CTA has 32 warps (one per SM) or 16 (two per SM)
static_assert(XF_KC <= 4, "work-list entries pack the candidate slot in two bits");

static size_t verify_smem(int IB, int N, int C) {
    return (size_t)IB * N * sizeof(double) + (size_t)(C - 1) * IB * (XV_SL + XF_KC - 1) * sizeof(double) +
           (size_t)(C - 1) * IB * (XF_KC + 1) * sizeof(int) + 16;
}

// lanes walk [a, b) with stride 32; CHECK: shifted indices may fall outside [0, N) and then contribute 0
// (rows are addressed as xs[off + n] with INTEGER offsets off = ii N + k: a pointer xs + off formed for a negative
// lag would lie before the array - undefined behaviour the compiler exploits to "prove" k >= 0)
template <int IB, bool CHECK>
__device__ __forceinline__ void verify_slice(const double* xj, const double* xs, const int (&off)[IB], const int (&k)[IB],
                                             int a, int b, int N, int lane, double (&acc)[IB]) {
    constexpr int U = 8;
    int n = a + lane;
    for (; n + 32 * (U - 1) < b; n += 32 * U) {
        double xv[U];
#pragma unroll
        for (int q = 0; q < U; ++q) xv[q] = __ldg(xj + n + 32 * q);
#pragma unroll
        for (int q = 0; q < U; ++q) {
#pragma unroll
            for (int ii = 0; ii < IB; ++ii) {
                double v;
                if constexpr (CHECK) v = (unsigned)(n + 32 * q + k[ii]) < (unsigned)N ? xs[off[ii] + n + 32 * q] : 0.0;
                else v = xs[off[ii] + n + 32 * q];
                acc[ii] = fma(v, xv[q], acc[ii]);
            }
        }
    }
    for (; n < b; n += 32) {
        const double xv = __ldg(xj + n);
#pragma unroll
        for (int ii = 0; ii < IB; ++ii) {
            const double v = (!CHECK || (unsigned)(n + k[ii]) < (unsigned)N) ? xs[off[ii] + n] : 0.0;
            acc[ii] = fma(v, xv, acc[ii]);
        }
    }
}

template <int IB, int NW>
__global__ void __launch_bounds__(NW * 32, 32 / NW) xf_verify_kernel(XfArgs g) {
    constexpr int XV_WARPS = NW, XV_JG = NW / XV_SL;
    extern __shared__ __align__(16) unsigned char xf_smem_raw[];
    const int N = g.N, NJ = g.C - 1;
    double* xs = reinterpret_cast<double*>(xf_smem_raw);     // [IB][N]
    double* sp = xs + (size_t)IB * N;                        // [NJ][IB][XV_SL] slice sums of the first candidate
    double* se = sp + (size_t)NJ * IB * XV_SL;               // [NJ][IB][XF_KC-1] sums of the further candidates
    int* elist = reinterpret_cast<int*>(se + (size_t)NJ * IB * (XF_KC - 1));   // [NJ*IB*(XF_KC-1)] (jj, ii, s) packed
    int* k0s = elist + (size_t)NJ * IB * (XF_KC - 1);        // [NJ][IB] first candidate k of pair (i_base+ii, i_base+1+jj),
    int* nk0s = k0s + (size_t)NJ * IB;                       // 0 if none; and -k (see the note on VIMNMX3 below)
    __shared__ int ecount;
    __shared__ float werr[XV_WARPS];
    __shared__ unsigned int wcnt[XV_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nib = (g.C - 1 + IB - 1) / IB;
    const int wslot = blockIdx.x / nib, ib = blockIdx.x - wslot * nib;
    const int wband = win_index(g, wslot), wl = wband - g.w0;
    const int i_base = ib * IB;
    const int64_t t0 = g.first + (int64_t)wband * g.hop;
    if (tid == 0) ecount = 0;
    const int njj = g.C - 1 - i_base;                    // channels j = i_base+1 .. C-1
    for (int e = tid; e < njj * IB; e += XV_WARPS * 32) {
        const int jj = e / IB, ii = e - jj * IB;
        const int ci = i_base + ii, j = i_base + 1 + jj;
        int kk = ci < j ? g.cand[((int64_t)wl * g.P + pair_index(ci, j, g.C)) * XF_KC] : INT_MIN;
        if (kk == INT_MIN) kk = 0;
        k0s[e] = kk;
        nk0s[e] = -kk;
    }
#pragma unroll
    for (int ii = 0; ii < IB; ++ii) {
        const int ci = i_base + ii;
        if (ci < g.C - 1) {
            const double* src = g.y + (int64_t)ci * g.ldy + t0;
#pragma unroll 8
            for (int n = tid; n < N; n += XV_WARPS * 32) xs[ii * N + n] = __ldg(src + n);
        }
    }
    __syncthreads();
    {
        const int sl = warp % XV_SL, jg = warp / XV_SL;
        const int SLN = ((N + XV_SL - 1) / XV_SL + 31) / 32 * 32;
        const int a = min(N, sl * SLN), b = min(N, a + SLN);
        for (int j = i_base + 1 + jg; j < g.C; j += XV_JG) {
            const double* xj = g.y + (int64_t)j * g.ldy + t0;
            int k[IB], off[IB];
            int lo = a, hi = b;          // [lo, hi): the part of the slice where every pair's shifted index is inside
#pragma unroll
            for (int ii = 0; ii < IB; ++ii) {
                // -k comes from its own shared-memory array: written as max(max(lo, -k0), -k1), ptxas 12.9 (sm_100a) folds
                // the expression into VIMNMX3 and drops the negation of one operand (seen in SASS and at run time)
                const int kk = k0s[(j - i_base - 1) * IB + ii], nk = nk0s[(j - i_base - 1) * IB + ii];
                k[ii] = kk;
                off[ii] = ii * N + kk;
                lo = max(lo, nk);
                hi = min(hi, N + nk);
            }
            // whole blocks of 256 samples for the unchecked part, so that none of the three calls leaves a ragged tail
            // inside the slice (the checked path costs more per sample but the edges are short)
            lo = a + (lo - a + 255) / 256 * 256;
            lo = min(lo, b);
            hi = hi < b ? lo + max(hi - lo, 0) / 256 * 256 : b;
            hi = max(hi, lo);
            double acc[IB];
#pragma unroll
            for (int ii = 0; ii < IB; ++ii) acc[ii] = 0.0;
            if (a < lo) verify_slice<IB, true>(xj, xs, off, k, a, lo, N, lane, acc);
            verify_slice<IB, false>(xj, xs, off, k, lo, hi, N, lane, acc);
            if (hi < b) verify_slice<IB, true>(xj, xs, off, k, hi, b, N, lane, acc);
            double* dst = sp + (size_t)(j - i_base - 1) * IB * XV_SL + sl;
            if constexpr (IB == 4) {
                // packed butterfly: after the exchange over 16 a lane carries two of the four sums, after 8 one; the
                // totals end in lanes 0, 8, 16, 24 (ii = 0, 1, 2, 3) in a fixed order: 12 shuffles instead of 40
                const bool up16 = (lane & 16) != 0, up8 = (lane & 8) != 0;
                double s0 = up16 ? acc[0] : acc[2], s1 = up16 ? acc[1] : acc[3];     // what this lane gives away
                double k0 = up16 ? acc[2] : acc[0], k1 = up16 ? acc[3] : acc[1];     // what it keeps
                k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
                k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
                const double g = up8 ? k0 : k1;
                double t = up8 ? k1 : k0;
                t += __shfl_xor_sync(0xffffffffu, g, 8);
                t += __shfl_xor_sync(0xffffffffu, t, 4);
                t += __shfl_xor_sync(0xffffffffu, t, 2);
                t += __shfl_xor_sync(0xffffffffu, t, 1);
                if ((lane & 7) == 0) dst[((lane >> 4) * 2 + ((lane >> 3) & 1)) * XV_SL] = t;
            } else {
#pragma unroll
                for (int ii = 0; ii < IB; ++ii) {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) acc[ii] += __shfl_xor_sync(0xffffffffu, acc[ii], o);
                }
                if (lane == 0) {
#pragma unroll
                    for (int ii = 0; ii < IB; ++ii) dst[ii * XV_SL] = acc[ii];
                }
            }
        }
    }
    // further candidates -> work list
    for (int e = tid; e < njj * IB; e += XV_WARPS * 32) {
        const int jj = e / IB, ii = e - jj * IB;
        const int ci = i_base + ii, j = i_base + 1 + jj;
        if (ci < j) {
            const int32_t* cd = g.cand + ((int64_t)wl * g.P + pair_index(ci, j, g.C)) * XF_KC;
            if (cd[0] != INT_MIN)
                for (int s_ = 1; s_ < XF_KC && cd[s_] != INT_MIN; ++s_) elist[atomicAdd(&ecount, 1)] = (e << 2) | s_;
        }
    }
    __syncthreads();
    const int ne = ecount;
    for (int q = warp; q < ne; q += XV_WARPS) {
        const int e = elist[q] >> 2, s_ = elist[q] & 3;
        const int jj = e / IB, ii = e - jj * IB;
        const int ci = i_base + ii, j = i_base + 1 + jj;
        const int ks = g.cand[((int64_t)wl * g.P + pair_index(ci, j, g.C)) * XF_KC + s_];
        const double* xj = g.y + (int64_t)j * g.ldy + t0;
        const int offs = ii * N + ks;
        const int nlo = max(0, -ks), nhi = min(N, N - ks);
        double d = 0.0;
#pragma unroll 4
        for (int n = nlo + lane; n < nhi; n += 32) d = fma(xs[offs + n], __ldg(xj + n), d);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (lane == 0) se[e * (XF_KC - 1) + s_ - 1] = d;
    }
    __syncthreads();
    // decide: one thread per pair (i, j) of the block
    float err = 0.f;
    unsigned int ncand = 0;
    for (int e = tid; e < njj * IB; e += XV_WARPS * 32) {
        const int jj = e / IB, ii = e - jj * IB;
        const int ci = i_base + ii, j = i_base + 1 + jj;
        if (ci >= j) continue;
        const int p = pair_index(ci, j, g.C);
        const int32_t* cd = g.cand + ((int64_t)wl * g.P + p) * XF_KC;
        if (cd[0] == INT_MIN) continue;
        double best = 0.0;
#pragma unroll
        for (int sl = 0; sl < XV_SL; ++sl) best += sp[e * XV_SL + sl];
        int bl = cd[0];
        ++ncand;
        for (int s_ = 1; s_ < XF_KC && cd[s_] != INT_MIN; ++s_) {
            const double d = se[e * (XF_KC - 1) + s_ - 1];
            const int ks = cd[s_];
            if (d > best || (d == best && ks < bl)) { best = d; bl = ks; }
            ++ncand;
        }
        const double Ei = g.energy[(int64_t)wl * g.C + ci], Ej = g.energy[(int64_t)wl * g.C + j];
        const double c = best / sqrt(Ei * Ej);
        const int64_t item = (int64_t)wband * g.P + p;
        g.lag[item] = bl;
        if (g.cmax) g.cmax[item] = c;
        if (g.k0 >= 0) {
            // |c[k]| <= sqrt(sum x_i^2 * sum x_j^2) over the overlap of lag k: for k > k0 that is <= sqrt(T_i H_j), for
            // k < -k0 <= sqrt(H_i T_j).  Unless the maximum found beats both bounds the window is redone over all lags.
            const double* pi = g.part + ((int64_t)wl * g.C + ci) * 2;
            const double* pj = g.part + ((int64_t)wl * g.C + j) * 2;
            const double b2 = fmax(pi[1] * pj[0], pi[0] * pj[1]);
            if (!(best > sqrt(b2) * (1.0 + 1e-12))) g.pf_flag[(int64_t)wl * g.P + p] = 1;
        }
        // consistency guard: the exact maximum must sit where the screen saw it
        const float ee = fabsf((float)(c - (double)g.sval[(int64_t)wl * g.P + p]));
        if (!(ee <= 0.25f * XF_TAU32)) {
            const int slot = atomicAdd(g.flag_count, 1);
            g.flag_list[slot] = (int)item;
            atomicAdd(g.stat + 1, 1ULL);
        } else {
            err = fmaxf(err, ee);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        err = fmaxf(err, __shfl_xor_sync(0xffffffffu, err, o));
        ncand += __shfl_xor_sync(0xffffffffu, ncand, o);
    }
    if (lane == 0) { werr[warp] = err; wcnt[warp] = ncand; }
    __syncthreads();
    if (tid == 0) {
        unsigned long long n = 0;
        for (int i = 0; i < XV_WARPS; ++i) { err = fmaxf(err, werr[i]); n += wcnt[i]; }
        atomicAdd(g.stat, n);
        atomicMax(g.stat + 2, (unsigned long long)__float_as_uint(err));
    }
}

// Item mode of the verification (stage B of the two-stage search): one CTA of XV_SL warps per listed pair-window.  The sums
// are formed exactly like xf_verify_kernel forms them (same time slices, same per-lane order, same trees), so a
// pair-window gets the same bits whichever kernel evaluates its candidates.
__global__ void __launch_bounds__(XV_SL * 32) xf_verify_items_kernel(XfArgs g) {
    extern __shared__ __align__(16) unsigned char xv_smem_raw[];
    double* xs = reinterpret_cast<double*>(xv_smem_raw);     // [N] window of channel i
    __shared__ double sp[XV_SL], se[XF_KC - 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = g.N;
    // the list length is a launch argument (grid = items) or sits on the device (any grid; CTAs stride over the list)
    const int n_list = g.items_count ? *g.items_count : g.n_items;
    for (int e = blockIdx.x; e < n_list; e += gridDim.x) {
        const int it = g.items[e];
        const int wl = it / g.P, p = it - wl * g.P;
        const int32_t* cd = g.cand + (int64_t)it * XF_KC;
        if (cd[0] == INT_MIN) continue;          // dead pair or handed to the time-domain kernel by the pair kernel
        int ci, j;
        pair_channels(p, g.C, ci, j);
        const int64_t t0 = g.first + (int64_t)(g.w0 + wl) * g.hop;
        {
            const double* src = g.y + (int64_t)ci * g.ldy + t0;
#pragma unroll 8
            for (int n = tid; n < N; n += XV_SL * 32) xs[n] = __ldg(src + n);
        }
        __syncthreads();
        const double* xj = g.y + (int64_t)j * g.ldy + t0;
        {
            const int SLN = ((N + XV_SL - 1) / XV_SL + 31) / 32 * 32;
            const int a = min(N, warp * SLN), b = min(N, a + SLN);
            const int k[1] = {cd[0]}, off[1] = {cd[0]};
            double acc[1] = {0.0};
            verify_slice<1, true>(xj, xs, off, k, a, b, N, lane, acc);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], o);
            if (lane == 0) sp[warp] = acc[0];
        }
        if (warp + 1 < XF_KC && cd[warp + 1] != INT_MIN) {
            const int ks = cd[warp + 1];
            const int nlo = max(0, -ks), nhi = min(N, N - ks);
            double d = 0.0;
#pragma unroll 4
            for (int n = nlo + lane; n < nhi; n += 32) d = fma(xs[ks + n], __ldg(xj + n), d);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
            if (lane == 0) se[warp] = d;
        }
        __syncthreads();
        if (tid == 0) {
            double best = 0.0;
#pragma unroll
            for (int sl = 0; sl < XV_SL; ++sl) best += sp[sl];
            int bl = cd[0];
            unsigned long long ncand = 1;
            for (int s_ = 1; s_ < XF_KC && cd[s_] != INT_MIN; ++s_) {
                const double d = se[s_ - 1];
                const int ks = cd[s_];
                if (d > best || (d == best && ks < bl)) { best = d; bl = ks; }
                ++ncand;
            }
            const double Ei = g.energy[(int64_t)wl * g.C + ci], Ej = g.energy[(int64_t)wl * g.C + j];
            const double c = best / sqrt(Ei * Ej);
            const int64_t item = (int64_t)(g.w0 + wl) * g.P + p;
            g.lag[item] = bl;
            if (g.cmax) g.cmax[item] = c;
            if (g.apx) g.apx[it] = 0;
            if (g.k0 >= 0) {      // stage-A proof, as in xf_verify_kernel
                const double* pi = g.part + ((int64_t)wl * g.C + ci) * 2;
                const double* pj = g.part + ((int64_t)wl * g.C + j) * 2;
                const double b2 = fmax(pi[1] * pj[0], pi[0] * pj[1]);
                if (!(best > sqrt(b2) * (1.0 + 1e-12))) g.pf_flag[it] = 1;
            }
            atomicAdd(g.stat, ncand);
            const float ee = fabsf((float)(c - (double)g.sval[it]));
            if (!(ee <= 0.25f * XF_TAU32)) {
                g.flag_list[atomicAdd(g.flag_count, 1)] = (int)item;
                atomicAdd(g.stat + 1, 1ULL);
            } else {
                atomicMax(g.stat + 2, (unsigned long long)__float_as_uint(ee));
            }
        }
        __syncthreads();      // xs, sp and se are reused by the next item
    }
}

// Lazy verification, last step: per window, which pairs still carry a screen value (apx) that could be one of the middle
// order statistics of cmax?  v_p = stored values, each within XF_EPS32 of the exact e_p; m = a middle order statistic of v.
// The same order statistic m* of e lies within XF_EPS32 of m, and a pair with |v_p - m| > 2 XF_EPS32 is on the same side of
// m* before and after it is replaced by its exact value.  So once every apx pair within 2 XF_EPS32 of either middle
// statistic is evaluated exactly, the median of the stored values IS the exact median (numpy nanmedian semantics as in
// stats.cuh).  One warp per window; the listed pairs go to xf_verify_items_kernel.
constexpr int XM_WARPS = 4;
__global__ void __launch_bounds__(XM_WARPS * 32) xf_median_select_kernel(const double* cmax, const unsigned char* apx, int P,
                                                                        int wn, int* need_list, int* need_count) {
    extern __shared__ double xm_sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* buf = xm_sm + (size_t)warp * P;
    for (int w = blockIdx.x * XM_WARPS + warp; w < wn; w += gridDim.x * XM_WARPS) {
        const double* cm = cmax + (int64_t)w * P;
        const unsigned char* ap = apx + (int64_t)w * P;
        int m = 0;
        for (int base = 0; base < P; base += 32) {
            const int p = base + lane;
            const double v = p < P ? cm[p] : qnan();
            const bool ok = v == v;
            const unsigned msk = __ballot_sync(0xffffffffu, ok);
            if (ok) buf[m + __popc(msk & ((1u << lane) - 1u))] = v;
            m += __popc(msk);
        }
        __syncwarp();
        if (m > 0) {
            const int r_hi = m >> 1, r_lo = (m & 1) ? r_hi : r_hi - 1;
            double lo = 0.0, hi = 0.0;
            bool have_lo = false, have_hi = false;
            for (int p = lane; p < m; p += 32) {
                const double v = buf[p];
                int rank = 0;
                for (int q = 0; q < m; ++q) {
                    const double u = buf[q];
                    rank += (u < v) || (u == v && q < p);
                }
                if (rank == r_lo) { lo = v; have_lo = true; }
                if (rank == r_hi) { hi = v; have_hi = true; }
            }
            const unsigned mlo = __ballot_sync(0xffffffffu, have_lo), mhi = __ballot_sync(0xffffffffu, have_hi);
            lo = __shfl_sync(0xffffffffu, lo, __ffs(mlo) - 1);
            hi = __shfl_sync(0xffffffffu, hi, __ffs(mhi) - 1);
            const double band = 2.0 * XF_EPS32 * (1.0 + 1e-9);
            for (int p = lane; p < P; p += 32) {
                if (!ap[p]) continue;
                const double v = cm[p];
                if (fabs(v - lo) <= band || fabs(v - hi) <= band) need_list[atomicAdd(need_count, 1)] = w * P + p;
            }
        }
        __syncwarp();
    }
}

// ---- stage A -> stage B work lists, built on the device -----------------------------------------
// The failing pair-windows of a batch (pf_flag) become an ASCENDING item list and the list of windows that own one,
// without atomics (the pairing of items in the packed transforms of stage B, and with it every FP32 value, must not
// depend on scheduling) and without a host pass: count per window (one warp each), one-block scan, ordered write.
__global__ void __launch_bounds__(256) xf_compact_count_kernel(const unsigned char* flag, int P, int wn, int* cnt) {
    const int w = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= wn) return;
    const unsigned char* f = flag + (int64_t)w * P;
    int c = 0;
    for (int p = lane; p < P; p += 32) c += f[p] ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) cnt[w] = c;
}
// cnt[w] -> off[w] (exclusive prefix), counts[0] = items, counts[1] = windows, wlist[] = band-level indices of those windows
__global__ void __launch_bounds__(1024) xf_compact_scan_kernel(const int* cnt, int wn, int w0, int* off, int* wlist, int* counts) {
    __shared__ int wsum[32], wnum[32];
    __shared__ int base_items, base_wins;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { base_items = 0; base_wins = 0; }
    __syncthreads();
    for (int w_ = 0; w_ < wn; w_ += 1024) {
        const int w = w_ + tid;
        const int c = w < wn ? cnt[w] : 0, has = c > 0 ? 1 : 0;
        int s = c, n = has;                      // inclusive scans inside the warp
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int a = __shfl_up_sync(0xffffffffu, s, o), b = __shfl_up_sync(0xffffffffu, n, o);
            if (lane >= o) { s += a; n += b; }
        }
        if (lane == 31) { wsum[warp] = s; wnum[warp] = n; }
        __syncthreads();
        if (warp == 0) {
            int a = wsum[lane], b = wnum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(0xffffffffu, a, o), y = __shfl_up_sync(0xffffffffu, b, o);
                if (lane >= o) { a += x; b += y; }
            }
            wsum[lane] = a; wnum[lane] = b;      // inclusive over warps
        }
        __syncthreads();
        const int pre_s = base_items + (warp ? wsum[warp - 1] : 0) + s - c;
        const int pre_n = base_wins + (warp ? wnum[warp - 1] : 0) + n - has;
        if (w < wn) {
            off[w] = pre_s;
            if (has) wlist[pre_n] = w0 + w;
        }
        __syncthreads();
        if (tid == 1023) { base_items += wsum[31]; base_wins += wnum[31]; }
        __syncthreads();
    }
    if (tid == 0) { counts[0] = base_items; counts[1] = base_wins; }
}
__global__ void __launch_bounds__(256) xf_compact_write_kernel(const unsigned char* flag, int P, int wn, const int* off, int* items) {
    const int w = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= wn) return;
    const unsigned char* f = flag + (int64_t)w * P;
    int pos = off[w];
    for (int p0 = 0; p0 < P; p0 += 32) {
        const int p = p0 + lane;
        const bool hit = p < P && f[p];
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (hit) items[pos + __popc(m & ((1u << lane) - 1u))] = w * P + p;
        pos += __popc(m);
    }
}

// ---- audit (NBLS_OPT_SCREEN_AUDIT) ------------------------------------------------------------
// Largest |FP32 screen value - FP64 FFT value| over EVERY searched lag of the dumped pair-windows, in normalised
// correlation units, into stat[3] (float bits); stat[4] counts the lags compared.  One CTA per pair-window.
__global__ void __launch_bounds__(256) xf_audit_kernel(const float* d32, const double* d64, int L, int lag_max,
                                                       unsigned long long* stat) {
    __shared__ float wmax[8];
    const float* a = d32 + (int64_t)blockIdx.x * L;
    const double* b = d64 + (int64_t)blockIdx.x * L;
    float e = 0.f;
    unsigned int n_ok = 0;
    for (int n = threadIdx.x; n < L; n += 256) {
        if (n <= lag_max || n >= L - lag_max) {
            const double d = fabs((double)a[n] - b[n]) / (double)L;
            if (d == d) { e = fmaxf(e, (float)d); ++n_ok; }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        e = fmaxf(e, __shfl_xor_sync(0xffffffffu, e, o));
        n_ok += __shfl_xor_sync(0xffffffffu, n_ok, o);
    }
    if ((threadIdx.x & 31) == 0) {
        wmax[threadIdx.x >> 5] = e;
        atomicAdd(stat + 4, (unsigned long long)n_ok);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < 8; ++i) e = fmaxf(e, wmax[i]);
        atomicMax(stat + 3, (unsigned long long)__float_as_uint(e));
    }
}

// ---- host side ------------------------------------------------------------------------------
typedef void (*xf_kernel)(XfArgs);
struct XfPlan {
    int L, threads;
    size_t smem;
    xf_kernel spectra, pair;
};

template <class V, int R1, int R2, int R3, int T, int MINB>
static XfPlan make_plan() {
    using S = Shape<R1, R2, R3>;
    XfPlan p;
    p.L = S::L;
    p.threads = T;
    p.smem = (size_t)(S::SMEM_ELEMS + 2) * sizeof(V);     // + the spare slot of the pair kernel's load loop (kept 16-byte sized)
    p.spectra = xf_spectra_kernel<S, T, MINB, V>;
    p.pair = xf_pair_kernel<S, T, MINB, V>;
    return p;
}

// transform lengths 2^a, 3*2^a and 5*2^a from 512 to 12288 (192 KB of complex FP64 is what one SM holds)
static const std::vector<XfPlan>& plans(bool fp32) {
    static const std::vector<XfPlan> v64 = {
        make_plan<double2, 2, 16, 16, 32, 4>(),   make_plan<double2, 3, 16, 16, 64, 4>(),
        make_plan<double2, 4, 16, 16, 64, 4>(),   make_plan<double2, 6, 16, 16, 96, 4>(),
        make_plan<double2, 8, 16, 16, 128, 3>(),  make_plan<double2, 10, 16, 16, 160, 2>(),
        make_plan<double2, 12, 16, 16, 192, 2>(), make_plan<double2, 16, 16, 16, 256, 2>(),
        make_plan<double2, 10, 16, 32, 160, 2>(), make_plan<double2, 6, 32, 32, 192, 2>(),
        make_plan<double2, 8, 32, 32, 256, 1>(),  make_plan<double2, 10, 32, 32, 320, 1>(),
        make_plan<double2, 12, 32, 32, 384, 1>(),
    };
    static const std::vector<XfPlan> v32 = {
        make_plan<float2, 2, 16, 16, 32, 8>(),    make_plan<float2, 3, 16, 16, 64, 8>(),
        make_plan<float2, 4, 16, 16, 64, 8>(),    make_plan<float2, 6, 16, 16, 96, 8>(),
        make_plan<float2, 8, 16, 16, 128, 6>(),   make_plan<float2, 10, 16, 16, 160, 4>(),
        make_plan<float2, 12, 16, 16, 192, 4>(),  make_plan<float2, 16, 16, 16, 256, 3>(),
        make_plan<float2, 10, 16, 32, 160, 4>(),  make_plan<float2, 6, 32, 32, 192, 4>(),
        make_plan<float2, 8, 32, 32, 256, 3>(),   make_plan<float2, 10, 32, 32, 320, 2>(),
        make_plan<float2, 12, 32, 32, 384, 2>(),
    };
    return fp32 ? v32 : v64;
}

static const XfPlan* pick_plan(int N, int max_lag, bool fp32) {
    if (N < 1) return nullptr;
    const int lag_max = (max_lag < 0 || max_lag > N - 1) ? N - 1 : max_lag;
    const int need = N + lag_max;          // circular length with no wrap-around inside |lag| <= lag_max
    for (const XfPlan& p : plans(fp32))
        if (p.L >= need) return &p;
    return nullptr;
}

int xcorr_fft_length(int N, int max_lag) {
    const XfPlan* p = pick_plan(N, max_lag, false);
    return p ? p->L : 0;
}

template <class V>
static int twiddles(nbls_ctx* ctx, int L, const void** out) {
    const int key = sizeof(V) == sizeof(float2) ? -L : L;
    auto it = ctx->fft_tw.find(key);
    if (it != ctx->fft_tw.end()) { *out = it->second; return NBLS_OK; }
    std::vector<V> h(L);
    const long double two_pi = 6.283185307179586476925286766559005768L;
    for (int k = 0; k < L; ++k) {
        const long double a = two_pi * (long double)k / (long double)L;
        h[k].x = (real_t<V>)cosl(a);
        h[k].y = (real_t<V>)sinl(a);
    }
    void* d = nullptr;
    cudaError_t e = cudaMalloc(&d, (size_t)L * sizeof(V));
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("nbls_xcorr(fft): twiddle table allocation failed: %s", cudaGetErrorString(e));
        return NBLS_ERR_NOMEM;
    }
    NBLS_CUDA(cudaMemcpyAsync(d, h.data(), (size_t)L * sizeof(V), cudaMemcpyHostToDevice, ctx->stream));
    NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->fft_tw[key] = d;
    *out = d;
    return NBLS_OK;
}

typedef void (*xv_kernel)(XfArgs);
static xv_kernel pick_verify(int N, int C, size_t smem_optin, int* ib, int* nw) {
    const size_t sm_total = 228 * 1024, half = sm_total / 2 - 2048;     // two CTAs per SM must both fit
    // two 16-warp CTAs per SM (one stages while the other computes) beat one 32-warp CTA even at half the channel block:
    // c3 2 h 8.8 -> 7.5 ms
    *nw = 32;
    if (verify_smem(4, N, C) <= half) { *ib = 4; *nw = 16; return xf_verify_kernel<4, 16>; }
    if (verify_smem(2, N, C) <= half) { *ib = 2; *nw = 16; return xf_verify_kernel<2, 16>; }
    if (verify_smem(4, N, C) <= smem_optin - 1024) { *ib = 4; return xf_verify_kernel<4, 32>; }
    if (verify_smem(2, N, C) <= smem_optin - 1024) { *ib = 2; return xf_verify_kernel<2, 32>; }
    if (verify_smem(1, N, C) <= smem_optin - 1024) { *ib = 1; return xf_verify_kernel<1, 32>; }
    return nullptr;
}

// Stage-A plan of the two-stage lag search: the longest transform that is at most 0.7 (else 0.8) of the full one and
// still covers |lag| <= N/4; nullptr if there is none.
static const XfPlan* pick_stage_a(int N, const XfPlan* full) {
    const XfPlan* best = nullptr;
    for (double frac : {0.7, 0.8}) {
        for (const XfPlan& p : plans(true))
            if (p.L < full->L && p.L <= frac * full->L && p.L >= N + N / 4) best = &p;
        if (best) return best;
    }
    return nullptr;
}

int xcorr_fft_band(nbls_ctx* ctx, const double* yband, int64_t ldy, int C, int N, int hop, int64_t first, int W,
                   int max_lag, int32_t* lag, double* cmax, bool fp32, bool two_stage) {
    if (W <= 0) return NBLS_OK;
    const XfPlan* pl = pick_plan(N, max_lag, fp32);
    if (!pl) {
        set_error("nbls_xcorr(fft): no transform length for a window of %d samples", N);
        return NBLS_ERR_UNSUPPORTED;
    }
    NBLS_REQUIRE(pl->smem <= ctx->smem_optin, "nbls_xcorr(fft): transform of %d points needs %zu B of shared memory", pl->L,
                 pl->smem);
    const int P = C * (C - 1) / 2;
    NBLS_REQUIRE((int64_t)W * P < (1LL << 31), "nbls_xcorr: too many (window, pair) items in one band: %lld",
                 (long long)W * P);
    const int lag_full = (max_lag < 0 || max_lag > N - 1) ? N - 1 : max_lag;
    // two-stage search: only for the full-lag request of the FP32 screen, and not while recent bands kept failing
    const XfPlan* pa = nullptr;
    if (fp32 && two_stage && lag_full == N - 1 && ctx->opt[NBLS_OPT_TWO_STAGE]) {
        if (ctx->ts_skip > 0) --ctx->ts_skip;
        else pa = pick_stage_a(N, pl);
    }
    const void *tw = nullptr, *twa = nullptr;
    if (fp32) NBLS_CHECK(twiddles<float2>(ctx, pl->L, &tw));
    else NBLS_CHECK(twiddles<double2>(ctx, pl->L, &tw));
    if (pa) NBLS_CHECK(twiddles<float2>(ctx, pa->L, &twa));
    int ib = 0, vnw = 32;
    xv_kernel verify = nullptr;
    size_t vsmem = 0;
    if (fp32) {
        verify = pick_verify(N, C, ctx->smem_optin, &ib, &vnw);
        NBLS_REQUIRE(verify, "nbls_xcorr(fft32): a window of %d samples does not fit in shared memory", N);
        vsmem = verify_smem(ib, N, C);
        NBLS_CUDA(allow_max_dynamic_smem(verify, ctx->smem_optin));
    }
    // windows per batch: spectra workspace bounded (they are re-read from L2 window by window)
    const int SP = pl->L / 2 + 1;
    const size_t per_win = (size_t)C * SP * (fp32 ? sizeof(float2) : sizeof(double2));
    const size_t budget = (size_t)4 << 30;
    const int ncp = (C + 1) / 2, npp = (P + 1) / 2;
    int wb = (int)std::min<int64_t>(W, std::max<int64_t>(1, (int64_t)(budget / per_win)));
    wb = (int)std::min<int64_t>(wb, ((1LL << 31) - 1) / std::max(npp, ncp));
    NBLS_CHECK(ensure(ctx, ctx->spec, (size_t)wb * per_win));
    NBLS_CHECK(ensure(ctx, ctx->energy, (size_t)wb * C * 3 * sizeof(double)));
    NBLS_CHECK(ensure(ctx, ctx->flags, ((size_t)wb * P + 4) * sizeof(int)));
    if (!ctx->xstat.p) {
        NBLS_CHECK(ensure(ctx, ctx->xstat, 8 * sizeof(unsigned long long)));
        NBLS_CUDA(cudaMemsetAsync(ctx->xstat.p, 0, 8 * sizeof(unsigned long long), ctx->stream));
    }
    if (fp32) {
        NBLS_CHECK(ensure(ctx, ctx->cand, (size_t)wb * P * XF_KC * sizeof(int32_t)));
        NBLS_CHECK(ensure(ctx, ctx->sval, (size_t)wb * P * sizeof(float)));
    }
    // stage A -> B bookkeeping: [4] counters, [wb*P] failing-pair items (ascending), [wb] window list, [wb] per-window
    // counts, [wb] per-window offsets, then [wb*P] bytes of failure flags
    if (pa) {
        NBLS_CHECK(ensure(ctx, ctx->wfail, ((size_t)wb * P + 3 * (size_t)wb + 8) * sizeof(int) + (size_t)wb * P + 64));
        NBLS_CUDA(allow_max_dynamic_smem(xf_verify_items_kernel, ctx->smem_optin));
    }
    for (const XfPlan* q : {pl, pa}) {
        if (!q) continue;
        NBLS_CUDA(cudaFuncSetAttribute(q->spectra, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)q->smem));
        NBLS_CUDA(cudaFuncSetAttribute(q->pair, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)q->smem));
    }

    // lazy verification: only where the caller wants no per-pair cmax (the audit keeps the dense evaluation)
    const bool lazy = fp32 && ctx->opt[NBLS_OPT_LAZY_VERIFY] != 0 && ctx->opt[NBLS_OPT_SCREEN_AUDIT] == 0 &&
                      (ctx->cmax_median_only || cmax == nullptr);
    if (lazy) {
        NBLS_CHECK(ensure(ctx, ctx->lazy, round_up((size_t)wb * P, 256) + ((size_t)wb * P + 8) * sizeof(int)));
        NBLS_CUDA(allow_max_dynamic_smem(xf_verify_items_kernel, ctx->smem_optin));
        NBLS_CUDA(allow_max_dynamic_smem(xf_median_select_kernel, ctx->smem_optin));
    }
    XfArgs a;
    memset(&a, 0, sizeof(a));
    a.y = yband; a.ldy = ldy; a.C = C; a.P = P; a.N = N; a.hop = hop; a.first = first;
    a.lazy = lazy ? 1 : 0;
    if (lazy) {
        a.apx = (unsigned char*)ctx->lazy.p;
        a.need_count = (int*)((char*)ctx->lazy.p + round_up((size_t)wb * P, 256));
        a.need_list = a.need_count + 8;
    }
    a.spec = ctx->spec.p;
    a.energy = (double*)ctx->energy.p;
    a.part = a.energy + (size_t)wb * C;
    a.lag = lag; a.cmax = cmax;
    a.flag_count = (int*)ctx->flags.p;
    a.flag_list = a.flag_count + 4;
    a.cand = (int32_t*)ctx->cand.p;
    a.sval = (float*)ctx->sval.p;
    a.stat = (unsigned long long*)ctx->xstat.p;
    a.lag_dead = -lag_full;
    a.pf_count = (int*)ctx->wfail.p;
    a.pf_list = a.pf_count ? a.pf_count + 4 : nullptr;
    int* dwlist = a.pf_list ? a.pf_list + (size_t)wb * P : nullptr;
    int* dcnt = dwlist ? dwlist + wb : nullptr;
    int* doff = dcnt ? dcnt + wb : nullptr;
    a.pf_flag = doff ? (unsigned char*)(doff + wb + 4) : nullptr;

    // undecided arg-maxima are settled by the time-domain kernel (normally an empty list)
    auto settle = [&]() -> int {
        return xcorr_direct_band(ctx, yband, ldy, C, N, hop, first, W, max_lag, lag, cmax, a.flag_list, a.flag_count,
                                 ctx->sm_count * 2);
    };
    // one pass of spectra -> pair -> (verify) -> time-domain list over the windows described by `a`
    // executed algorithmic flop of one complex transform of q->L points: 5 L log2 L; the pair kernel adds 16 L (two
    // cross-spectra, packing, two screening compares per point); one exact candidate costs 2 N FP64 flop
    auto fft_flop = [](const XfPlan* q) { return 5.0 * q->L * log2((double)q->L); };
    // audit: the first `wa` windows of the pass also go through the FP64 transform of the same length and every
    // searched lag is compared (xf_audit_kernel)
    const bool audit = fp32 && ctx->opt[NBLS_OPT_SCREEN_AUDIT] != 0;
    auto audit_pass = [&](const XfPlan* q, int wa) -> int {
        const XfPlan* q64 = nullptr;
        for (const XfPlan& p64 : plans(false))
            if (p64.L == q->L) q64 = &p64;
        if (!q64 || q64->smem > ctx->smem_optin) return NBLS_OK;
        const void* tw64 = nullptr;
        NBLS_CHECK(twiddles<double2>(ctx, q->L, &tw64));
        const size_t n_it = (size_t)wa * P, SPq = (size_t)q->L / 2 + 1;
        const size_t o_d64 = round_up(n_it * q->L * sizeof(float), 256);
        const size_t o_spec = o_d64 + round_up(n_it * q->L * sizeof(double), 256);
        const size_t o_int = o_spec + round_up((size_t)wa * C * SPq * sizeof(double2), 256);
        char* base = (char*)ctx->audit.p;
        XfArgs b = a;
        b.first = a.first + (int64_t)a.w0 * a.hop; b.w0 = 0; b.wn = wa; b.k0 = -1; b.wlist = nullptr; b.items = nullptr;
        b.tw = tw64; b.spec = base + o_spec;
        b.lag = (int32_t*)(base + o_int); b.cmax = nullptr;
        b.flag_count = b.lag + n_it; b.flag_list = b.flag_count + 4;
        b.stat = (unsigned long long*)(base + o_int + round_up((2 * n_it + 8) * sizeof(int), 256));
        b.dump = base + o_d64; b.dump_wn = wa;
        NBLS_CUDA(cudaMemsetAsync(b.flag_count, 0, sizeof(int), ctx->stream));
        NBLS_CUDA(cudaFuncSetAttribute(q64->spectra, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)q64->smem));
        NBLS_CUDA(cudaFuncSetAttribute(q64->pair, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)q64->smem));
        q64->spectra<<<(unsigned)(wa * ncp), q64->threads, q64->smem, ctx->stream>>>(b);
        NBLS_CUDA(cudaGetLastError());
        q64->pair<<<(unsigned)(wa * npp), q64->threads, q64->smem, ctx->stream>>>(b);
        NBLS_CUDA(cudaGetLastError());
        xf_audit_kernel<<<(unsigned)n_it, 256, 0, ctx->stream>>>((const float*)base, (const double*)(base + o_d64), q->L,
                                                               a.lag_max, a.stat);
        NBLS_CUDA(cudaGetLastError());
        ctx->launches += 3;
        return NBLS_OK;
    };
    // lazy mode: exact evaluation of the pair-windows on the need list (length on the device: a fixed grid strides over it)
    auto verify_needed = [&]() -> int {
        XfArgs v = a;
        v.items = a.need_list; v.items_count = a.need_count; v.n_items = 0; v.wlist = nullptr;
        KernelGroupScope ks(ctx, KG_VERIFY_ITEMS, 0.0);
        xf_verify_items_kernel<<<(unsigned)(ctx->sm_count * 8), XV_SL * 32, (size_t)N * sizeof(double), ctx->stream>>>(v);
        NBLS_CUDA(cudaGetLastError());
        ctx->launches += 1;
        return NBLS_OK;
    };
    auto pass = [&](const XfPlan* q) -> int {
        NBLS_CUDA(cudaMemsetAsync(a.flag_count, 0, sizeof(int), ctx->stream));
        if (lazy) NBLS_CUDA(cudaMemsetAsync(a.need_count, 0, sizeof(int), ctx->stream));
        int wa = 0;
        a.dump = nullptr; a.dump_wn = 0;
        if (audit) {
            const size_t per_w = (size_t)P * q->L * 12 + (size_t)C * (q->L / 2 + 1) * 16 + (size_t)P * 8 + 1024;
            wa = (int)std::min<size_t>((size_t)a.wn, std::max<size_t>(1, ((size_t)1 << 30) / per_w));
            NBLS_CHECK(ensure(ctx, ctx->audit, (size_t)wa * per_w + 8192));
            a.dump = ctx->audit.p; a.dump_wn = wa;
        }
        {
            KernelGroupScope ks(ctx, KG_SPECTRA, (double)a.wn * ncp * fft_flop(q));
            q->spectra<<<(unsigned)(a.wn * ncp), q->threads, q->smem, ctx->stream>>>(a);
        }
        NBLS_CUDA(cudaGetLastError());
        {
            KernelGroupScope ks(ctx, KG_PAIR, (double)a.wn * npp * (fft_flop(q) + 16.0 * q->L));
            q->pair<<<(unsigned)(a.wn * npp), q->threads, q->smem, ctx->stream>>>(a);
        }
        NBLS_CUDA(cudaGetLastError());
        ctx->launches += 2;
        if (lazy) {
            NBLS_CHECK(verify_needed());
        } else if (fp32) {
            const int nib = (C - 1 + ib - 1) / ib;
            KernelGroupScope ks(ctx, KG_VERIFY, (double)a.wn * P * 2.0 * N);
            verify<<<(unsigned)(a.wn * nib), vnw * 32, vsmem, ctx->stream>>>(a);
            NBLS_CUDA(cudaGetLastError());
            ctx->launches += 1;
        }
        if (wa > 0) NBLS_CHECK(audit_pass(q, wa));
        a.dump = nullptr; a.dump_wn = 0;
        return settle();
    };
    // lazy mode, end of a batch: the pairs that decide the median of a window get their exact value
    auto median_step = [&](int w0, int wn) -> int {
        if (!lazy || !cmax) return NBLS_OK;
        NBLS_CUDA(cudaMemsetAsync(a.flag_count, 0, sizeof(int), ctx->stream));
        NBLS_CUDA(cudaMemsetAsync(a.need_count, 0, sizeof(int), ctx->stream));
        a.w0 = w0; a.wn = wn; a.wlist = nullptr; a.items = nullptr; a.n_items = 0; a.k0 = -1; a.lag_max = lag_full;
        {
            KernelGroupScope ks(ctx, KG_VERIFY_ITEMS, 0.0);
            const int grid = std::min((wn + XM_WARPS - 1) / XM_WARPS, ctx->sm_count * 16);
            xf_median_select_kernel<<<grid, XM_WARPS * 32, (size_t)XM_WARPS * P * sizeof(double), ctx->stream>>>(
                cmax + (int64_t)w0 * P, a.apx, P, wn, a.need_list, a.need_count);
            NBLS_CUDA(cudaGetLastError());
            ctx->launches += 1;
        }
        NBLS_CHECK(verify_needed());
        return settle();
    };
    int64_t n_redo = 0, n_redo_pairs = 0;
    for (int w0 = 0; w0 < W; w0 += wb) {
        const int wn = std::min(wb, W - w0);
        a.w0 = w0; a.wn = wn; a.wlist = nullptr; a.items = nullptr; a.n_items = 0;
        if (!pa) {
            a.k0 = -1; a.lag_max = lag_full;
            a.SP = SP; a.tw = tw;
            NBLS_CHECK(pass(pl));
            NBLS_CHECK(median_step(w0, wn));
            continue;
        }
        // stage A: |lag| <= k0 with the shorter transform, plus the proof per pair-window that nothing beyond k0 can win
        a.k0 = std::min(pa->L - N, N - 1); a.lag_max = a.k0;
        a.SP = pa->L / 2 + 1; a.tw = twa;
        NBLS_CUDA(cudaMemsetAsync(a.pf_flag, 0, (size_t)wn * P, ctx->stream));
        NBLS_CHECK(pass(pa));
        // stage B work lists on the device; the host only needs the two counts for the launch grids
        xf_compact_count_kernel<<<(wn + 7) / 8, 256, 0, ctx->stream>>>(a.pf_flag, P, wn, dcnt);
        NBLS_CUDA(cudaGetLastError());
        xf_compact_scan_kernel<<<1, 1024, 0, ctx->stream>>>(dcnt, wn, w0, doff, dwlist, a.pf_count);
        NBLS_CUDA(cudaGetLastError());
        xf_compact_write_kernel<<<(wn + 7) / 8, 256, 0, ctx->stream>>>(a.pf_flag, P, wn, doff, a.pf_list);
        NBLS_CUDA(cudaGetLastError());
        ctx->launches += 3;
        int hcount[2] = {0, 0};
        NBLS_CUDA(cudaMemcpyAsync(hcount, a.pf_count, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
        const int nfail = hcount[0], nwin_fail = hcount[1];
        if (nfail == 0) {
            NBLS_CHECK(median_step(w0, wn));
            continue;
        }
        n_redo += nwin_fail;
        n_redo_pairs += nfail;
        a.k0 = -1; a.lag_max = lag_full;
        a.SP = SP; a.tw = tw;
        NBLS_CUDA(cudaMemsetAsync(a.flag_count, 0, sizeof(int), ctx->stream));
        a.wn = nwin_fail; a.wlist = dwlist;
        {
            KernelGroupScope ks(ctx, KG_SPECTRA, (double)a.wn * ncp * fft_flop(pl));
            pl->spectra<<<(unsigned)(a.wn * ncp), pl->threads, pl->smem, ctx->stream>>>(a);
        }
        NBLS_CUDA(cudaGetLastError());
        a.wlist = nullptr; a.items = a.pf_list; a.n_items = nfail;
        if (lazy) NBLS_CUDA(cudaMemsetAsync(a.need_count, 0, sizeof(int), ctx->stream));
        {
            KernelGroupScope ks(ctx, KG_PAIR, (double)((nfail + 1) / 2) * (fft_flop(pl) + 16.0 * pl->L));
            pl->pair<<<(unsigned)((nfail + 1) / 2), pl->threads, pl->smem, ctx->stream>>>(a);
        }
        NBLS_CUDA(cudaGetLastError());
        if (lazy) {
            a.items = nullptr; a.n_items = 0;
            NBLS_CHECK(verify_needed());
        } else {
            KernelGroupScope ks(ctx, KG_VERIFY_ITEMS, (double)nfail * 2.0 * N);
            xf_verify_items_kernel<<<(unsigned)nfail, XV_SL * 32, (size_t)N * sizeof(double), ctx->stream>>>(a);
            NBLS_CUDA(cudaGetLastError());
            ctx->launches += 1;
        }
        ctx->launches += 2;
        NBLS_CHECK(settle());
        NBLS_CHECK(median_step(w0, wn));
    }
    if (pa) {
        ctx->ts_windows += W;
        ctx->ts_redone += n_redo;
        ctx->ts_redone_pairs += n_redo_pairs;
        // a band where most pair-windows had to be redone: leave the next three bands to the single-stage search
        if (n_redo_pairs * 2 > (int64_t)W * P) ctx->ts_skip = 3;
    }
    return NBLS_OK;
}

}  // namespace nbls

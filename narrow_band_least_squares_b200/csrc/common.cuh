// common.cuh — context, error plumbing and small device helpers shared by the libnbls kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <map>
#include <string>
#include <vector>

#include "nbls.h"

namespace nbls {

void set_error(const char* fmt, ...);

#define NBLS_CUDA(expr)                                                                      \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            ::nbls::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,             \
                              cudaGetErrorString(e__));                                      \
            return NBLS_ERR_CUDA;                                                            \
        }                                                                                    \
    } while (0)

#define NBLS_CHECK(expr)            \
    do {                            \
        int rc__ = (expr);          \
        if (rc__ != NBLS_OK) return rc__; \
    } while (0)

#define NBLS_REQUIRE(cond, ...)             \
    do {                                    \
        if (!(cond)) {                      \
            ::nbls::set_error(__VA_ARGS__); \
            return NBLS_ERR_ARG;            \
        }                                   \
    } while (0)

// A grow-only device buffer cached in the context (bench steps pay no cudaMalloc).
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

enum { EV_START = 0, EV_H2D, EV_FILTER, EV_XCORR, EV_SOLVE, EV_D2H, EV_COUNT };

// kernel groups of the per-group profile (nbls_profile) — keep nbls_kernel_group_name() in api.cu in step
enum {
    KG_FILTER_P1F = 0,   // forward phase 1 (chunk end states)
    KG_FILTER_P3F,       // forward phase 3 (output)
    KG_FILTER_P1B,       // backward phase 1
    KG_FILTER_P3B,       // backward phase 3 (output + taper)
    KG_FILTER_CARRY,     // both carry kernels
    KG_SPECTRA,          // xf_spectra_kernel
    KG_PAIR,             // xf_pair_kernel
    KG_VERIFY,           // xf_verify_kernel
    KG_VERIFY_ITEMS,     // xf_verify_items_kernel (stage B)
    KG_DIRECT,           // xcorr_direct_kernel (whole bands or the near-tie list)
    KG_STATS_OLS,        // stats_ols_kernel
    KG_LTS,              // lts_kernel
    KG_COUNT
};

}  // namespace nbls

namespace nbls {
// Lets a kernel use the device's whole opt-in shared memory.  Always the same value for a given kernel, so contexts working
// on different jobs from different host threads cannot lower each other's limit between this call and the launch (the
// attribute is per function, not per context).
template <class K>
inline cudaError_t allow_max_dynamic_smem(K fn, size_t smem_optin) {
    cudaFuncAttributes at;
    cudaError_t e = cudaFuncGetAttributes(&at, fn);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem_optin - at.sharedSizeBytes));
}
}  // namespace nbls

struct nbls_ctx {
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;         // second stream of the filter (band groups side by side), created on first use
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    size_t ws_limit = 0;
    int64_t launches = 0;
    // cached buffers
    nbls::DevBuf coef, state, carry, taper, filt, geom, lts_tab, xin, res, lagbuf, cmaxbuf, wbuf, scratch;
    nbls::DevBuf spec, energy, flags;       // FFT cross-correlation workspace
    nbls::DevBuf cand, sval, xstat;         // FP32 screen: candidate lags, screen maxima; xcorr statistics
    nbls::DevBuf audit;                     // NBLS_OPT_SCREEN_AUDIT workspace
    nbls::DevBuf wfail;                     // two-stage lag search: failing pair-window items and the window list of stage B
    nbls::DevBuf lazy;                      // NBLS_OPT_LAZY_VERIFY: approximate-cmax flags and the list of pair-windows to evaluate
    bool cmax_median_only = false;          // set by nbls_run_* when the caller wants no per-pair cmax
    int ts_skip = 0;                        // bands left to the single-stage search after a band that mostly failed
    int64_t ts_windows = 0, ts_redone = 0;  // two-stage statistics: windows searched / redone over all lags
    std::map<int, void*> fft_tw;            // twiddle tables by transform length (device)
    cudaEvent_t ev[nbls::EV_COUNT] = {};
    bool ev_valid = false;
    float stage_ms[6] = {};
    // options (nbls_set_option) — per context, never read from the environment
    int64_t opt[NBLS_OPT_COUNT] = {};
    // per-kernel-group profile (NBLS_OPT_PROFILE): device ms, launches, executed algorithmic work (flop, or bytes for the
    // filter groups) since the last reset
    cudaEvent_t pev[2] = {};
    cudaEvent_t mark[8] = {};               // nbls_mark slots, created on first use
    double prof_ms[nbls::KG_COUNT] = {};
    int64_t prof_n[nbls::KG_COUNT] = {};
    double prof_work[nbls::KG_COUNT] = {};
    double audit_err = 0.0;                 // NBLS_OPT_SCREEN_AUDIT: largest |FP32 - FP64| over every searched lag
    int64_t audit_lags = 0;
    std::vector<unsigned char> filter_key;  // host image of the filter tables now on the device (filter.cu)
    const void *filter_key_coef = nullptr, *filter_key_taper = nullptr;
    int64_t last_E = -1, last_P = 0;        // rows / pairs of the lags kept by the last nbls_run_host (nbls_resolve_host)
    int64_t ts_redone_pairs = 0;            // two-stage: pair-windows redone over all lags
};

namespace nbls {
// Brackets the launches of one kernel group: an NVTX range always (visible in nsys / ncu --nvtx), and with
// NBLS_OPT_PROFILE a CUDA-event pair on the context's stream whose elapsed time is added to the group (this
// synchronises after every group, so a profiled run is slower as a whole; the per-group times are what it is for).
struct KernelGroupScope {
    nbls_ctx* c;
    int g;
    KernelGroupScope(nbls_ctx* ctx, int group, double work);
    ~KernelGroupScope();
};
}  // namespace nbls

namespace nbls {

int ensure(nbls_ctx* ctx, DevBuf& b, size_t bytes);

// stage implementations (each in its own .cu)
int filter_impl(nbls_ctx* ctx, const nbls_filter_desc* d, const double* x, int C, int64_t T,
                int64_t ldx, double* y, int64_t ldy, int64_t ystride_band);
int xcorr_impl(nbls_ctx* ctx, const nbls_window_desc* w, const double* y, int C, int64_t ldy,
               int64_t ystride_band, int32_t* lag, double* cmax);
int solve_impl(nbls_ctx* ctx, const nbls_solve_desc* s, int64_t nwin, const int32_t* lag,
               const double* cmax, const nbls_outputs* out);
int experiment_dft_pass_impl(nbls_ctx* ctx, int mode, const float* x_dev, int64_t rows, int reps, int iters, float* y_dev,
                             float* ms_out);
int fp64_peak_impl(nbls_ctx* ctx, int iters, double* tflops);
int fp32_peak_impl(nbls_ctx* ctx, int iters, double* tflops);
int xcorr_direct_band(nbls_ctx* ctx, const double* yband, int64_t ldy, int C, int N, int hop, int64_t first, int W,
                      int max_lag, int32_t* lag, double* cmax, const int* list, const int* count, int grid_cap);
int xcorr_fft_band(nbls_ctx* ctx, const double* yband, int64_t ldy, int C, int N, int hop, int64_t first, int W,
                   int max_lag, int32_t* lag, double* cmax, bool fp32, bool two_stage);
// transform length the FFT path would use for a window of N samples and lag bound, 0 if unsupported
int xcorr_fft_length(int N, int max_lag);

static inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

}  // namespace nbls

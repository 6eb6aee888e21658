/* nbls.h — C ABI of libnbls, the B200 (sm_100a) narrow-band least-squares hot path.
 *
 * This is the drop-in boundary.  The reference
 * (amiezzi/narrow_band_least_squares, pure Python) has no FFI of its own; each entry
 * point below replaces one Python-level call on its hot path and is what a ctypes stub
 * inside the reference would bind.  Citations are relative to the
 * reference tree.
 *
 *   nbls_filter      <- helpers.py:108-141 filter_data (obspy zero-phase Butterworth
 *                       band-pass / causal cheby1 + 1 % Hann taper), all bands at once
 *   nbls_xcorr       <- lts_array/lts_array/classes/lts_classes.py:852-880
 *                       LsBeam.correlate (np.correlate 'full' + argmax), all bands/windows
 *   nbls_solve       <- lts_classes.py:873-878 (mdccm, MAD), :891-936 OLSEstimator.solve,
 *                       :58-150 fast_LTS, :643-764 post_process (no uncertainty ellipse)
 *   nbls_run_device  <- narrow_band_least_squares.py:67-124 band loop body, device buffers
 *   nbls_run_host    <- the same with HOST buffers (H2D + path + D2H)
 *
 * Conventions: plain pointers and sizes, no C++ or torch types.  `*_dev` pointers are
 * device memory owned by the caller; every other pointer is host memory and is read
 * before the call returns.  All calls are ordered on the context's stream and
 * return after enqueueing unless stated; nbls_synchronize() waits.  Return value 0 = OK,
 * otherwise an NBLS_ERR_* code and nbls_last_error() describes it (per host thread).  A
 * context owns one stream and its workspaces and is driven by one host thread at a time;
 * several contexts on one GPU are independent and may run concurrently from different
 * threads (e.g. two records in flight).  There is no CPU fallback: without a CUDA device
 * nbls_create fails.
 */
#ifndef NBLS_H_
#define NBLS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NBLS_VERSION 100

#define NBLS_OK 0
#define NBLS_ERR_ARG 1
#define NBLS_ERR_CUDA 2
#define NBLS_ERR_NOMEM 3
#define NBLS_ERR_UNSUPPORTED 4

#define NBLS_MAX_SECTIONS 8
#define NBLS_MAX_BANDS 256

#define NBLS_XCORR_AUTO 0   /* FFT32 with the two-stage lag search where it applies (see below), else DIRECT */
#define NBLS_XCORR_DIRECT 1 /* time-domain FP64 contraction over the circularly padded window */
#define NBLS_XCORR_FFT 2    /* shared-memory FP64 FFT + exact re-evaluation of near-tied lags */
#define NBLS_XCORR_FFT32 3  /* FP32 FFT screens the lags, every candidate within 2e-5 of the FP32 maximum is
                               evaluated exactly (FP64 time-domain sums); lags and cmax are FP64 results.
                               NBLS_XCORR_AUTO picks this wherever the FFT path applies. */

typedef struct nbls_ctx nbls_ctx;

/* ---- K1: per-band IIR band-pass (cascaded biquads) + taper ------------------------- */
typedef struct nbls_filter_desc {
    int32_t nbands;        /* B */
    int32_t nsections;     /* S second-order sections per band (<= NBLS_MAX_SECTIONS) */
    const double* sos;     /* [B][S][6] scipy layout b0 b1 b2 a0 a1 a2, a0 == 1 */
    int32_t zerophase;     /* 1: forward pass then pass over the reversed signal, zero state
                              twice (obspy zerophase=True, helpers.py:127); 0: one causal
                              pass (cheby1 branch, helpers.py:131-137) */
    int32_t exact_scan;    /* 1: chunk end-states from whole chunks (exact linear scan);
                              0: from the last K_b samples only, K_b = decay length of band b
                              to 2^-64 (contribution below FP64 rounding) */
    int32_t chunk_len;     /* scan chunk length in samples (multiple of 16); 0 = auto */
    int32_t reserved0;
    int64_t taper_len;     /* wlen of obspy Trace.taper(0.01) (helpers.py:139); 0 = none */
    const double* taper_up;   /* [taper_len] multipliers of global samples 0 .. wlen-1 */
    const double* taper_down; /* [taper_len] multipliers of global samples G-wlen .. G-1 */
    int64_t global_t0;     /* global index of local sample 0 (time-sharded segments) */
    int64_t global_npts;   /* G: length of the whole record the taper refers to */
} nbls_filter_desc;

/* x_dev: [nchan] rows of npts doubles, row stride ldx.  y_dev: [B][nchan] rows, row stride
 * ldy, band stride ystride_band (doubles).  Rows must be 16-byte aligned: pointers aligned
 * to 16 bytes and strides even.  y may not alias x. */
int nbls_filter(nbls_ctx* ctx, const nbls_filter_desc* desc, const double* x_dev, int32_t nchan,
                int64_t npts, int64_t ldx, double* y_dev, int64_t ldy, int64_t ystride_band);

/* ---- K2: windowed pairwise cross-correlation --------------------------------------- */
typedef struct nbls_window_desc {
    int32_t nbands;
    int32_t max_lag;          /* < 0: all 2N-1 lags (reference behaviour); else |lag| <= max_lag */
    int32_t method;           /* NBLS_XCORR_* */
    int32_t reserved0;
    const int32_t* win_len;   /* [B] N_b  = int(window_length*fs)       lts_data_class.py:58 */
    const int32_t* win_hop;   /* [B] H_b  = int(round((1-ov)*N_b))      lts_data_class.py:60 */
    const int64_t* win_first; /* [B] local sample index of window 0 of the band */
    const int32_t* win_count; /* [B] W_b windows in the band            lts_data_class.py:63 */
} nbls_window_desc;

/* Outputs are window-major: row e = (sum of W_b' for b' < b) + w holds the P = C(C-1)/2
 * pairs in the reference's idx_pair order (i<j, row-major; lts_classes.py:840).
 * lag = argmax_m cij[m] - (N-1)  (first maximum wins; zero-energy pair -> -(N-1));
 * tau = -lag/fs (lts_classes.py:875-876).  cmax = max_m cij[m] (normalised), NaN for a
 * zero-energy pair.  cmax_dev may be NULL. */
int nbls_xcorr(nbls_ctx* ctx, const nbls_window_desc* win, const double* y_dev, int32_t nchan,
               int64_t ldy, int64_t ystride_band, int32_t* lag_dev, double* cmax_dev);

/* Transform length the FFT path uses for windows of `win_len` samples and lag bound `max_lag` (as in
 * nbls_window_desc); 0 if the FFT path does not cover that window length (NBLS_XCORR_AUTO then runs the
 * time-domain kernel).  Pure host arithmetic, needs no context. */
int nbls_xcorr_fft_length(int32_t win_len, int32_t max_lag);

/* Diagnostics of the cross-correlation calls since the context was created (or last reset): candidate lags
 * evaluated exactly by the FP32-screen path, (window, pair) items handed to the time-domain kernel (near-ties,
 * non-finite data, screen/exact disagreement), and the largest |FP32 screen maximum - exact maximum| seen
 * (normalised correlation units; the screen's candidate band is 2e-5).  Synchronises.  Any pointer may be NULL. */
int nbls_xcorr_stats(nbls_ctx* ctx, int64_t* n_verified, int64_t* n_time_domain, double* max_screen_err,
                     int32_t reset);

/* NBLS_OPT_SCREEN_AUDIT results: the largest |FP32 screen value - FP64 FFT value| over EVERY searched lag of the audited
 * pair-windows (normalised correlation units; the candidate band assumes this stays below 1e-5) and the number of lags
 * compared, since creation / the last reset.  Synchronises. */
int nbls_xcorr_audit(nbls_ctx* ctx, double* max_err_all_lags, int64_t* n_lags, int32_t reset);

/* Two-stage lag search of NBLS_XCORR_AUTO (all 2N-1 lags requested, FP32 screen): stage A searches |lag| <= k0 with a
 * shorter transform and proves for every pair-window - Cauchy-Schwarz over the partial energies of the two windows - that no
 * lag beyond k0 can reach the maximum found; pair-windows without a proof are redone over all lags, so the result is the
 * arg-max over all 2N-1 lags either way.  A band where most pair-windows needed the second stage switches the next three
 * bands to the single-stage search.  Returns the windows searched in two stages and those that held at least one redone
 * pair-window, since creation / the last reset. */
int nbls_xcorr_two_stage_stats(nbls_ctx* ctx, int64_t* n_windows, int64_t* n_redone, int32_t reset);
/* Pair-windows the second stage redid over all lags since creation / the last reset (the unit stage B works in). */
int nbls_xcorr_two_stage_pairs(nbls_ctx* ctx, int64_t* n_redone_pairs, int32_t reset);

/* ---- K3: per-window statistics and slowness solve ---------------------------------- */
typedef struct nbls_solve_desc {
    int32_t nchan;            /* C */
    int32_t npairs;           /* P */
    double fs;
    const double* xij;        /* [P][2] co-array, km          lts_classes.py:842-843 */
    const double* xij_mad;    /* [2]                          lts_classes.py:846-850 */
    const double* ols_gain;   /* [2][P] pseudo-inverse of xij_standardized (OLS, :889,:917) */
    double alpha;             /* 1.0 = OLS; [0.5,1) = FAST-LTS */
    /* FAST-LTS only (alpha < 1): */
    int32_t h;                /* subset size                  lts_classes.py:1013 */
    int32_t n_starts;         /* 500                          lts_classes.py:991  */
    int32_t n_keep;           /* 10                           lts_classes.py:993  */
    int32_t csteps1;          /* 4                            lts_classes.py:995  */
    int32_t csteps2;          /* 100                          lts_classes.py:997  */
    int32_t reserved0;
    const int32_t* subsets;   /* [n_starts + n_keep][2] the LCG stream of random_set(P,2,seed)
                                 restarted at seed 0 (lts_classes.py:16-36,76) */
    double raw_factor;        /* raw_corfactor*raw_consfactor lts_classes.py:646-648 */
    double rew_factor1;       /* rew_corfactor                lts_classes.py:650 */
    const double* rew_cons;   /* [P+1] rew_consfactor by number of unit weights (:326-344) */
} nbls_solve_desc;

/* Device output block of one run, all window-major over E = sum_b W_b rows.
 * NULL members are not written / not copied. */
typedef struct nbls_outputs {
    double* mdccm;      /* [E] nanmedian of cmax over pairs          lts_classes.py:873 */
    double* mad;        /* [E] 1.4826*median|tau|                    lts_classes.py:877-878 */
    double* vel;        /* [E] trace velocity km/s, NaN if skipped   lts_classes.py:928,:706 */
    double* baz;        /* [E] back-azimuth deg                      lts_classes.py:931,:713 */
    double* sigma_tau;  /* [E]                                       lts_classes.py:934-936,:723-730 */
    double* slowness;   /* [E][2] final slowness vector (s/km) */
    int32_t* lag;       /* [E][P] */
    double* cmax;       /* [E][P] */
    uint8_t* weights;   /* [E][P] final LTS weights (element_weights, :701,:762); alpha<1 only */
} nbls_outputs;

int nbls_solve(nbls_ctx* ctx, const nbls_solve_desc* sol, int64_t nwin, const int32_t* lag_dev,
               const double* cmax_dev, const nbls_outputs* out_dev);

/* ---- whole path --------------------------------------------------------------------- */
typedef struct nbls_job {
    nbls_filter_desc filter;
    nbls_window_desc window;
    nbls_solve_desc solve;
} nbls_job;

/* Filter -> xcorr -> solve for every band and window of the job, data resident on the
 * device.  x_dev as in nbls_filter.  `out_dev` members are device pointers (lag must be
 * non-NULL; cmax may be NULL and is then kept in workspace). */
int nbls_run_device(nbls_ctx* ctx, const nbls_job* job, const double* x_dev, int32_t nchan,
                    int64_t npts, int64_t ldx, const nbls_outputs* out_dev);

/* The same with host buffers: copies x (nchan rows, stride ldx_host) to the device, runs
 * the path, copies every non-NULL member of `out_host` back, and synchronises.  Use
 * nbls_malloc_host / nbls_host_register for pinned buffers. */
int nbls_run_host(nbls_ctx* ctx, const nbls_job* job, const double* x_host, int32_t nchan,
                  int64_t npts, int64_t ldx_host, const nbls_outputs* out_host);

/* Solve again on the integer lags and cmax the context kept from its last nbls_run_host call (same record, same window
 * grid) with another solver description - e.g. the FAST-LTS pass of an OLS + LTS sweep over one record
 * (the cross-correlation, > 90 % of a run, is shared; the reference would call ltsva twice, ltsva.py:45-54) - and
 * copy the non-NULL estimate members of `out_host` back (lag and cmax members are ignored).  Synchronises. */
int nbls_resolve_host(nbls_ctx* ctx, const nbls_solve_desc* sol, const nbls_outputs* out_host);

/* ---- context, memory, diagnostics --------------------------------------------------- */
int nbls_version(void);
const char* nbls_last_error(void);
int nbls_device_count(int32_t* count);
int nbls_create(nbls_ctx** ctx, int32_t device);
int nbls_destroy(nbls_ctx* ctx);
int nbls_synchronize(nbls_ctx* ctx);
/* Upper bound of device workspace the context may hold (bytes); 0 = 80 % of free memory. */
int nbls_set_workspace_limit(nbls_ctx* ctx, size_t bytes);

int nbls_malloc_device(nbls_ctx* ctx, void** ptr, size_t bytes);
int nbls_free_device(nbls_ctx* ctx, void* ptr);
int nbls_malloc_host(void** ptr, size_t bytes); /* pinned */
int nbls_free_host(void* ptr);
int nbls_host_register(void* ptr, size_t bytes);
int nbls_host_unregister(void* ptr);
int nbls_memcpy_h2d(nbls_ctx* ctx, void* dst_dev, const void* src_host, size_t bytes);
int nbls_memcpy_d2h(nbls_ctx* ctx, void* dst_host, const void* src_dev, size_t bytes);

/* ---- options: per context, set explicitly; libnbls reads nothing from the environment -------------------- */
#define NBLS_OPT_PROFILE 0           /* 1: time every kernel group with CUDA events (nbls_profile); the context then
                                        synchronises after each group, so whole-run times are not representative */
#define NBLS_OPT_TWO_STAGE 1         /* 1 (default): two-stage lag search under NBLS_XCORR_AUTO; 0: single stage */
#define NBLS_OPT_FILTER_SPLIT 2      /* blocks per (32 chunks, channel) tile of the shared forward phase 1; default 1 */
#define NBLS_OPT_FILTER_SHARED_P1 3  /* 1 (default): forward phase 1 reads the raw record once for all bands */
#define NBLS_OPT_FILTER_LANES 4      /* target number of time chunks in flight (auto chunk length); default 131072 */
#define NBLS_OPT_LTS_FP64_RANK 5     /* 1: h-subsets by FP64 rank counting instead of FP16 keys (same subsets) */
#define NBLS_OPT_SCREEN_AUDIT 6      /* 1: the FP32-screen path also runs the FP64 transform of every pair-window and
                                        records the largest |FP32 - FP64| over ALL searched lags (nbls_xcorr_stats'
                                        max_screen_err then covers every lag, not only the maximum); slow, for tests */
#define NBLS_OPT_FILTER_P1_WARPS 7   /* warps per block of the shared forward phase 1: 4 or 8; 0 (default) = automatic */
#define NBLS_OPT_FILTER_P1_SUB 8     /* sub-chunks per chunk of the forward state pass: 1 (default, also 0), 2, 4 or 8 */
#define NBLS_OPT_FILTER_OVERLAP 9    /* 1 (default): the filter runs two band groups side by side on two streams (the state
                                        passes of the low bands hide under the output passes of the high bands); 0: one stream */
#define NBLS_OPT_LAZY_VERIFY 10      /* 1 (default): when the caller asks for no per-pair cmax (nbls_outputs.cmax = NULL), the FP32-screen
                                        path evaluates in FP64 only the pair-windows whose result needs it: those with more
                                        than one candidate lag, and per window the pairs whose screen value lies within the
                                        screen's error band of the median (MdCCM is the only consumer of cmax then and stays
                                        exact).  0: every pair-window is evaluated in FP64 */
#define NBLS_OPT_COUNT 16
int nbls_set_option(nbls_ctx* ctx, int32_t key, int64_t value);
int nbls_get_option(nbls_ctx* ctx, int32_t key, int64_t* value);

/* Per-kernel-group profile, filled while NBLS_OPT_PROFILE is 1: device milliseconds (CUDA events on the context's
 * stream around the group's launches), launches, and the ALGORITHMIC work the launches executed - flop for the
 * cross-correlation and solve groups, bytes for the filter groups (DESIGN.md section 5 gives the formulas).
 * Groups are numbered 0 .. nbls_kernel_group_count()-1; any output pointer may be NULL; reset != 0 clears the group. */
int nbls_kernel_group_count(void);
const char* nbls_kernel_group_name(int32_t group);
int nbls_profile(nbls_ctx* ctx, int32_t group, double* ms, int64_t* launches, double* work, int32_t reset);

/* Device time of the stages of the last nbls_run_* call, milliseconds, measured with CUDA
 * events on the context's stream: [0] H2D, [1] filter, [2] xcorr, [3] solve, [4] D2H,
 * [5] total.  Synchronises.  Returns the number of values written (<= n). */
int nbls_last_timings(nbls_ctx* ctx, float* ms, int32_t n);
/* Caller-placed time stamps on the context's stream (CUDA events; slots 0..7): nbls_mark records one, nbls_mark_elapsed
 * waits for slot `b` and returns the device milliseconds between slots `a` and `b`.  For timing sequences of calls
 * (e.g. a batch of arrays) on the stream the kernels actually run on. */
int nbls_mark(nbls_ctx* ctx, int32_t slot);
int nbls_mark_elapsed(nbls_ctx* ctx, int32_t a, int32_t b, float* ms);
/* Kernel launches issued by this context since creation (for bench.py's gpu_launches). */
int64_t nbls_launch_count(nbls_ctx* ctx);
/* Register-resident DFMA throughput microbenchmark: returns FP64 TFLOP/s (2 flop per FMA)
 * measured over `iters` dependent-chain iterations on every SM. */
int nbls_measure_fp64_peak(nbls_ctx* ctx, int32_t iters, double* tflops);
/* The same with FP32 FMAs (roofline denominator of the FP32 screening FFTs). */
int nbls_measure_fp32_peak(nbls_ctx* ctx, int32_t iters, double* tflops);

/* ---- experiment: tensor-core DFT pass ("tensor-pipe verdict") ---------------------------
 * Runs `reps` radix-32 passes (32-point inverse DFT of every row + inter-pass twiddles) over x_dev [rows][32] complex
 * FP32 -> y_dev, tile by tile in shared memory, `iters` times, and returns the best launch time.
 *   mode 0: the production way - register butterflies on the CUDA cores (fft_radix.cuh)
 *   mode 1: tcgen05.mma DFT-matrix product, FP16 operands, one product  (error ~2^-11)
 *   mode 2: the same with split FP16 operands, three products           (error ~2^-21, FP32-grade)
 * Not used by any production path; kept so that the verdict can be re-measured. */
int nbls_experiment_dft_pass(nbls_ctx* ctx, int32_t mode, const float* x_dev, int64_t rows, int32_t reps, int32_t iters,
                             float* y_dev, float* ms);

/* ---- allocation-free cross-correlation ------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Time-domain FP64 cross-correlation of ONE band (the tiled kernel pair nbls_xcorr uses for long windows: same lags, same
 * tie rule, same cmax), whose scratch - one partial maximum per (window, pair, block of 2048 lags) - lives in a
 * file-scope `static __device__` array of the library instead of the context's growable workspace.  The call performs no
 * cudaMalloc / cudaFree and calls no allocating routine; it enqueues two kernels on the context's stream and returns.
 *   y_dev [nchan][ldy] filtered band (device); windows w = 0..win_count-1 start at win_first + w*win_hop, win_len samples
 *   max_lag as in nbls_window_desc; lag_dev [win_count][P] int32, cmax_dev [win_count][P] double or NULL (device)
 * Limit: win_count * P * ceil((2*L+1)/2048) <= NBLS_XCORR_STATIC_MAX_ENTRIES (L = the lag bound searched, win_len-1 for
 * max_lag < 0); larger requests fail with NBLS_ERR_* and a message, nothing is launched - split the windows over
 * several calls or use nbls_xcorr.  The array (3 MiB) is held by every GPU that loads the library until exit and is
 * SHARED by all calls and contexts on a GPU: calls that overlap on the GPU (different contexts/streams) race; calls on
 * one context are ordered by its stream.  The array keeps the last call's contents; nothing in it needs clearing (every
 * entry is written before it is read).  The first call on a device looks the array's address up (cudaGetSymbolAddress,
 * no allocation); make one call per device at set-up if later calls run under stream capture. */
#define NBLS_XCORR_STATIC_MAX_ENTRIES 262144
int nbls_xcorr_static(nbls_ctx* ctx, const double* y_dev, int32_t nchan, int64_t ldy, int32_t win_len, int32_t win_hop,
                      int64_t win_first, int32_t win_count, int32_t max_lag, int32_t* lag_dev, double* cmax_dev);

#ifdef __cplusplus
}
#endif
#endif /* NBLS_H_ */

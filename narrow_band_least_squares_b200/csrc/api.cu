// api.cu — the extern "C" surface of libnbls (include/nbls.h): context, memory helpers, stage entry
// points and the whole-path drivers.  No kernels here except the FP64 peak microbenchmark.
#include <stdarg.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "common.cuh"

namespace nbls {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int ensure(nbls_ctx* ctx, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return NBLS_OK;
    if (b.p) {
        NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
        NBLS_CUDA(cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    const size_t want = (size_t)round_up((int64_t)bytes, 256);
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        b.p = nullptr;
        cudaGetLastError();
        set_error("device allocation of %zu bytes failed: %s", want, cudaGetErrorString(e));
        return NBLS_ERR_NOMEM;
    }
    b.cap = want;
    return NBLS_OK;
}


static const char* const kGroupNames[KG_COUNT] = {
    "filter_fwd_phase1", "filter_fwd_phase3", "filter_bwd_phase1", "filter_bwd_phase3", "filter_carry",
    "xf_spectra", "xf_pair", "xf_verify", "xf_verify_items", "xcorr_direct", "stats_ols", "lts"};

KernelGroupScope::KernelGroupScope(nbls_ctx* ctx, int group, double work) : c(ctx), g(group) {
    nvtxRangePushA(kGroupNames[group]);
    c->prof_work[g] += work;
    if (c->opt[NBLS_OPT_PROFILE]) cudaEventRecord(c->pev[0], c->stream);
}
KernelGroupScope::~KernelGroupScope() {
    if (c->opt[NBLS_OPT_PROFILE]) {
        float ms = 0.f;
        if (cudaEventRecord(c->pev[1], c->stream) == cudaSuccess && cudaEventSynchronize(c->pev[1]) == cudaSuccess &&
            cudaEventElapsedTime(&ms, c->pev[0], c->pev[1]) == cudaSuccess) {
            c->prof_ms[g] += ms;
            c->prof_n[g] += 1;
        }
    }
    nvtxRangePop();
}

static int validate_windows(const nbls_window_desc* w, int C, int64_t ldy, int64_t* total) {
    NBLS_REQUIRE(w && w->win_len && w->win_hop && w->win_first && w->win_count, "nbls_xcorr: null window arrays");
    NBLS_REQUIRE(w->nbands >= 1 && w->nbands <= NBLS_MAX_BANDS, "nbls_xcorr: nbands %d out of range", w->nbands);
    NBLS_REQUIRE(C >= 2 && C <= 2048, "nbls_xcorr: channel count %d out of range", C);
    int64_t E = 0;
    for (int b = 0; b < w->nbands; ++b) {
        const int N = w->win_len[b], H = w->win_hop[b], W = w->win_count[b];
        NBLS_REQUIRE(N >= 1 && W >= 0 && (H >= 1 || W <= 1), "nbls_xcorr: band %d: bad window grid N=%d hop=%d W=%d", b, N, H, W);
        if (W > 0)
            NBLS_REQUIRE(w->win_first[b] >= 0 && w->win_first[b] + (int64_t)(W - 1) * H + N <= ldy,
                         "nbls_xcorr: band %d: windows run past the row (first=%lld hop=%d W=%d N=%d ld=%lld)", b,
                         (long long)w->win_first[b], H, W, N, (long long)ldy);
        E += W;
    }
    *total = E;
    return NBLS_OK;
}

int xcorr_impl(nbls_ctx* ctx, const nbls_window_desc* w, const double* y, int C, int64_t ldy, int64_t ystride_band,
               int32_t* lag, double* cmax) {
    NBLS_REQUIRE(y && lag, "nbls_xcorr: null argument");
    int64_t E = 0;
    NBLS_CHECK(validate_windows(w, C, ldy, &E));
    const int P = C * (C - 1) / 2;
    int64_t row = 0;
    ctx->ts_skip = 0;      // the two-stage search adapts within a call only: identical calls give identical bits
    for (int b = 0; b < w->nbands; ++b) {
        const int N = w->win_len[b], W = w->win_count[b];
        int method = w->method;
        // AUTO: the FFT path wherever it applies and the window is long enough for N log N to win
        if (method == NBLS_XCORR_AUTO)
            method = (N >= 128 && xcorr_fft_length(N, w->max_lag) > 0) ? NBLS_XCORR_FFT32 : NBLS_XCORR_DIRECT;
        NBLS_REQUIRE(method == NBLS_XCORR_DIRECT || method == NBLS_XCORR_FFT || method == NBLS_XCORR_FFT32,
                     "nbls_xcorr: unknown method %d", method);
        if (method != NBLS_XCORR_DIRECT && xcorr_fft_length(N, w->max_lag) == 0) {
            set_error("nbls_xcorr: FFT path does not support windows of %d samples", N);
            return NBLS_ERR_UNSUPPORTED;
        }
        int32_t* lb = lag + row * P;
        double* cb = cmax ? cmax + row * P : nullptr;
        const double* yb = y + (int64_t)b * ystride_band;
        if (method != NBLS_XCORR_DIRECT)
            NBLS_CHECK(xcorr_fft_band(ctx, yb, ldy, C, N, w->win_hop[b], w->win_first[b], W, w->max_lag, lb, cb,
                                      method == NBLS_XCORR_FFT32, w->method == NBLS_XCORR_AUTO));
        else
            NBLS_CHECK(xcorr_direct_band(ctx, yb, ldy, C, N, w->win_hop[b], w->win_first[b], W, w->max_lag, lb, cb,
                                         nullptr, nullptr, 0));
        row += W;
    }
    return NBLS_OK;
}

// ---- FMA peak microbenchmarks (FP64 and FP32) ---------------------------------------------
template <class F>
__global__ void __launch_bounds__(256) fma_peak_kernel(F* out, int iters, F seed) {
    F a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6,
      a7 = a0 + 7;
    const F m = (F)0.999999, c = (F)1e-9;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
            a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

template <class F>
static int fma_peak(nbls_ctx* ctx, int iters, double* tflops) {
    NBLS_REQUIRE(tflops && iters > 0, "nbls_measure_fma_peak: bad argument");
    const int blocks = ctx->sm_count * 8, threads = 256;
    NBLS_CHECK(ensure(ctx, ctx->scratch, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t e0, e1;
    NBLS_CUDA(cudaEventCreate(&e0));
    NBLS_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        NBLS_CUDA(cudaEventRecord(e0, ctx->stream));
        fma_peak_kernel<F><<<blocks, threads, 0, ctx->stream>>>((F*)ctx->scratch.p, iters, (F)(1.0 + rep));
        NBLS_CUDA(cudaEventRecord(e1, ctx->stream));
        NBLS_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        NBLS_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 64.0 * (double)iters * blocks * threads;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
        ctx->launches += 1;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = best;
    return NBLS_OK;
}

int fp64_peak_impl(nbls_ctx* ctx, int iters, double* tflops) { return fma_peak<double>(ctx, iters, tflops); }
int fp32_peak_impl(nbls_ctx* ctx, int iters, double* tflops) { return fma_peak<float>(ctx, iters, tflops); }

// ---- whole path ---------------------------------------------------------------------------
static int record(nbls_ctx* ctx, int which) {
    NBLS_CUDA(cudaEventRecord(ctx->ev[which], ctx->stream));
    return NBLS_OK;
}

static int run_device_inner(nbls_ctx* ctx, const nbls_job* job, const double* x, int C, int64_t T, int64_t ldx,
                            const nbls_outputs* out, int64_t E) {
    const nbls_filter_desc& f = job->filter;
    const nbls_window_desc& w = job->window;
    const int B = f.nbands, P = C * (C - 1) / 2;
    NBLS_REQUIRE(w.nbands == B, "nbls_run: filter has %d bands, window grid has %d", B, w.nbands);
    NBLS_REQUIRE(job->solve.nchan == C && job->solve.npairs == P, "nbls_run: solve geometry does not match %d channels", C);
    NBLS_REQUIRE(out && out->lag, "nbls_run: outputs.lag is required");
    const int64_t ldy = round_up(T, 16);
    // band groups bounded by the workspace limit
    const size_t per_band = (size_t)C * ldy * sizeof(double);
    int G = B;
    // the free-memory query costs ~1 ms of idle GPU: skipped when the workspace of an earlier call already holds every band
    if ((size_t)B * per_band > ctx->filt.cap || (ctx->ws_limit && (size_t)B * per_band > ctx->ws_limit)) {
        size_t freeb = 0, totalb = 0;
        NBLS_CUDA(cudaMemGetInfo(&freeb, &totalb));
        size_t limit = ctx->ws_limit ? ctx->ws_limit : (size_t)((double)(freeb + ctx->filt.cap) * 0.8);
        G = (int)std::min<size_t>((size_t)B, std::max<size_t>(1, limit / per_band));
        if ((size_t)G * per_band > freeb + ctx->filt.cap) {
            set_error("nbls_run: one band of filtered data (%zu B) does not fit in device memory", per_band);
            return NBLS_ERR_NOMEM;
        }
    }
    NBLS_CHECK(ensure(ctx, ctx->filt, (size_t)G * per_band));
    double* yf = (double*)ctx->filt.p;
    double* cmax = out->cmax;
    if (!cmax) {
        NBLS_CHECK(ensure(ctx, ctx->cmaxbuf, (size_t)std::max<int64_t>(E, 1) * P * sizeof(double)));
        cmax = (double*)ctx->cmaxbuf.p;
    }
    float ms_f = 0, ms_x = 0;
    // band-grouped runs time every group with the profile event pair + EV_FILTER (EV_START must keep the run's start)
    cudaEvent_t ea = ctx->pev[0], eb = ctx->ev[EV_FILTER], ec = ctx->ev[EV_XCORR];
    int64_t row = 0;
    const bool grouped = G < B;
    for (int b0 = 0; b0 < B; b0 += G) {
        const int nb = std::min(G, B - b0);
        nbls_filter_desc fg = f;
        fg.nbands = nb;
        fg.sos = f.sos + (size_t)b0 * f.nsections * 6;
        nbls_window_desc wg = w;
        wg.nbands = nb;
        wg.win_len = w.win_len + b0; wg.win_hop = w.win_hop + b0; wg.win_first = w.win_first + b0;
        wg.win_count = w.win_count + b0;
        if (grouped) NBLS_CUDA(cudaEventRecord(ea, ctx->stream));
        NBLS_CHECK(filter_impl(ctx, &fg, x, C, T, ldx, yf, ldy, (int64_t)C * ldy));
        NBLS_CUDA(cudaEventRecord(eb, ctx->stream));
        NBLS_CHECK(xcorr_impl(ctx, &wg, yf, C, ldy, (int64_t)C * ldy, out->lag + row * P, cmax + row * P));
        NBLS_CUDA(cudaEventRecord(ec, ctx->stream));
        if (grouped) {   // accumulate per-group stage times (events are reused)
            NBLS_CUDA(cudaEventSynchronize(ec));
            float a = 0, c = 0;
            NBLS_CUDA(cudaEventElapsedTime(&a, ea, eb));
            NBLS_CUDA(cudaEventElapsedTime(&c, eb, ec));
            ms_f += a; ms_x += c;
        }
        for (int b = 0; b < nb; ++b) row += wg.win_count[b];
    }
    ctx->stage_ms[1] = grouped ? ms_f : -1.f;
    ctx->stage_ms[2] = grouped ? ms_x : -1.f;
    NBLS_CHECK(solve_impl(ctx, &job->solve, E, out->lag, cmax, out));
    NBLS_CHECK(record(ctx, EV_SOLVE));
    return NBLS_OK;
}

static int count_windows(const nbls_job* job, int C, int64_t ldy, int64_t* E) {
    return validate_windows(&job->window, C, ldy, E);
}

}  // namespace nbls

using namespace nbls;

extern "C" {

int nbls_version(void) { return NBLS_VERSION; }
const char* nbls_last_error(void) { return g_err; }

int nbls_device_count(int32_t* count) {
    NBLS_REQUIRE(count, "nbls_device_count: null argument");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    *count = n;
    return NBLS_OK;
}

int nbls_create(nbls_ctx** out, int32_t device) {
    NBLS_REQUIRE(out, "nbls_create: null argument");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        set_error("nbls_create: no CUDA device available (%s); libnbls has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return NBLS_ERR_CUDA;
    }
    NBLS_REQUIRE(device >= 0 && device < n, "nbls_create: device %d out of range [0,%d)", device, n);
    NBLS_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    NBLS_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        set_error("nbls_create: device %d is sm_%d%d; libnbls is built for sm_100a (B200) only", device, prop.major,
                  prop.minor);
        return NBLS_ERR_UNSUPPORTED;
    }
    nbls_ctx* ctx = new nbls_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    NBLS_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (int i = 0; i < EV_COUNT; ++i) NBLS_CUDA(cudaEventCreate(&ctx->ev[i]));
    for (int i = 0; i < 2; ++i) NBLS_CUDA(cudaEventCreate(&ctx->pev[i]));
    ctx->opt[NBLS_OPT_TWO_STAGE] = 1;
    ctx->opt[NBLS_OPT_FILTER_SPLIT] = 1;
    ctx->opt[NBLS_OPT_FILTER_SHARED_P1] = 1;
    ctx->opt[NBLS_OPT_FILTER_LANES] = 131072;
    ctx->opt[NBLS_OPT_FILTER_OVERLAP] = 1;
    ctx->opt[NBLS_OPT_LAZY_VERIFY] = 1;
    *out = ctx;
    return NBLS_OK;
}

int nbls_destroy(nbls_ctx* ctx) {
    if (!ctx) return NBLS_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf* bufs[] = {&ctx->coef, &ctx->state, &ctx->carry, &ctx->taper, &ctx->filt, &ctx->geom, &ctx->lts_tab,
                      &ctx->xin, &ctx->res, &ctx->lagbuf, &ctx->cmaxbuf, &ctx->wbuf, &ctx->scratch, &ctx->spec,
                      &ctx->energy, &ctx->flags, &ctx->cand, &ctx->sval, &ctx->xstat, &ctx->wfail, &ctx->audit, &ctx->lazy};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    for (auto& kv : ctx->fft_tw)
        if (kv.second) cudaFree(kv.second);
    for (int i = 0; i < EV_COUNT; ++i)
        if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    for (int i = 0; i < 2; ++i)
        if (ctx->pev[i]) cudaEventDestroy(ctx->pev[i]);
    for (int i = 0; i < 8; ++i)
        if (ctx->mark[i]) cudaEventDestroy(ctx->mark[i]);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return NBLS_OK;
}

int nbls_synchronize(nbls_ctx* ctx) {
    NBLS_REQUIRE(ctx, "nbls_synchronize: null context");
    NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
    return NBLS_OK;
}

int nbls_set_workspace_limit(nbls_ctx* ctx, size_t bytes) {
    NBLS_REQUIRE(ctx, "nbls_set_workspace_limit: null context");
    ctx->ws_limit = bytes;
    return NBLS_OK;
}

int nbls_set_option(nbls_ctx* ctx, int32_t key, int64_t value) {
    NBLS_REQUIRE(ctx, "nbls_set_option: null context");
    NBLS_REQUIRE(key >= 0 && key < NBLS_OPT_COUNT, "nbls_set_option: unknown option %d", key);
    if (key == NBLS_OPT_FILTER_SPLIT) NBLS_REQUIRE(value >= 1 && value <= 64, "nbls_set_option: filter split %lld out of range", (long long)value);
    if (key == NBLS_OPT_FILTER_LANES) NBLS_REQUIRE(value >= 32, "nbls_set_option: filter lanes %lld out of range", (long long)value);
    ctx->opt[key] = value;
    return NBLS_OK;
}
int nbls_get_option(nbls_ctx* ctx, int32_t key, int64_t* value) {
    NBLS_REQUIRE(ctx && value, "nbls_get_option: null argument");
    NBLS_REQUIRE(key >= 0 && key < NBLS_OPT_COUNT, "nbls_get_option: unknown option %d", key);
    *value = ctx->opt[key];
    return NBLS_OK;
}
int nbls_kernel_group_count(void) { return KG_COUNT; }
const char* nbls_kernel_group_name(int32_t group) { return group >= 0 && group < KG_COUNT ? kGroupNames[group] : ""; }
int nbls_profile(nbls_ctx* ctx, int32_t group, double* ms, int64_t* launches, double* work, int32_t reset) {
    NBLS_REQUIRE(ctx, "nbls_profile: null context");
    NBLS_REQUIRE(group >= 0 && group < KG_COUNT, "nbls_profile: group %d out of range", group);
    if (ms) *ms = ctx->prof_ms[group];
    if (launches) *launches = ctx->prof_n[group];
    if (work) *work = ctx->prof_work[group];
    if (reset) { ctx->prof_ms[group] = 0.0; ctx->prof_n[group] = 0; ctx->prof_work[group] = 0.0; }
    return NBLS_OK;
}

int nbls_malloc_device(nbls_ctx* ctx, void** ptr, size_t bytes) {
    NBLS_REQUIRE(ctx && ptr, "nbls_malloc_device: null argument");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    cudaError_t e = cudaMalloc(ptr, bytes ? bytes : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *ptr = nullptr;
        set_error("nbls_malloc_device: %zu bytes: %s", bytes, cudaGetErrorString(e));
        return NBLS_ERR_NOMEM;
    }
    return NBLS_OK;
}
int nbls_free_device(nbls_ctx* ctx, void* ptr) {
    NBLS_REQUIRE(ctx, "nbls_free_device: null context");
    if (ptr) {
        NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
        NBLS_CUDA(cudaFree(ptr));
    }
    return NBLS_OK;
}
int nbls_malloc_host(void** ptr, size_t bytes) {
    NBLS_REQUIRE(ptr, "nbls_malloc_host: null argument");
    cudaError_t e = cudaMallocHost(ptr, bytes ? bytes : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *ptr = nullptr;
        set_error("nbls_malloc_host: %zu bytes: %s", bytes, cudaGetErrorString(e));
        return NBLS_ERR_NOMEM;
    }
    return NBLS_OK;
}
int nbls_free_host(void* ptr) {
    if (ptr) NBLS_CUDA(cudaFreeHost(ptr));
    return NBLS_OK;
}
int nbls_host_register(void* ptr, size_t bytes) {
    NBLS_REQUIRE(ptr, "nbls_host_register: null argument");
    NBLS_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
    return NBLS_OK;
}
int nbls_host_unregister(void* ptr) {
    NBLS_REQUIRE(ptr, "nbls_host_unregister: null argument");
    NBLS_CUDA(cudaHostUnregister(ptr));
    return NBLS_OK;
}
int nbls_memcpy_h2d(nbls_ctx* ctx, void* dst, const void* src, size_t bytes) {
    NBLS_REQUIRE(ctx && (bytes == 0 || (dst && src)), "nbls_memcpy_h2d: null argument");
    if (bytes) {
        NBLS_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return NBLS_OK;
}
int nbls_memcpy_d2h(nbls_ctx* ctx, void* dst, const void* src, size_t bytes) {
    NBLS_REQUIRE(ctx && (bytes == 0 || (dst && src)), "nbls_memcpy_d2h: null argument");
    if (bytes) {
        NBLS_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return NBLS_OK;
}

int nbls_filter(nbls_ctx* ctx, const nbls_filter_desc* desc, const double* x_dev, int32_t nchan, int64_t npts,
                int64_t ldx, double* y_dev, int64_t ldy, int64_t ystride_band) {
    NBLS_REQUIRE(ctx, "nbls_filter: null context");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    return filter_impl(ctx, desc, x_dev, nchan, npts, ldx, y_dev, ldy, ystride_band);
}

int nbls_xcorr(nbls_ctx* ctx, const nbls_window_desc* win, const double* y_dev, int32_t nchan, int64_t ldy,
               int64_t ystride_band, int32_t* lag_dev, double* cmax_dev) {
    NBLS_REQUIRE(ctx, "nbls_xcorr: null context");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    return xcorr_impl(ctx, win, y_dev, nchan, ldy, ystride_band, lag_dev, cmax_dev);
}

int nbls_xcorr_fft_length(int32_t win_len, int32_t max_lag) { return xcorr_fft_length(win_len, max_lag); }

int nbls_xcorr_stats(nbls_ctx* ctx, int64_t* n_verified, int64_t* n_time_domain, double* max_screen_err, int32_t reset) {
    NBLS_REQUIRE(ctx, "nbls_xcorr_stats: null context");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    unsigned long long h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ctx->xstat.p) {
        NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
        NBLS_CUDA(cudaMemcpy(h, ctx->xstat.p, sizeof(h), cudaMemcpyDeviceToHost));
        if (reset) NBLS_CUDA(cudaMemset(ctx->xstat.p, 0, sizeof(h)));
    }
    if (n_verified) *n_verified = (int64_t)h[0];
    if (n_time_domain) *n_time_domain = (int64_t)h[1];
    if (max_screen_err) {
        const unsigned int bits = (unsigned int)h[2];
        float f;
        memcpy(&f, &bits, sizeof(f));
        *max_screen_err = (double)f;
    }
    if (ctx->xstat.p && (h[3] || h[4])) {
        // audit counters are folded into the host copies and cleared on the device every time they are read
        const unsigned int bits = (unsigned int)h[3];
        float f;
        memcpy(&f, &bits, sizeof(f));
        ctx->audit_err = std::max(ctx->audit_err, (double)f);
        ctx->audit_lags += (int64_t)h[4];
        NBLS_CUDA(cudaMemset((char*)ctx->xstat.p + 3 * sizeof(unsigned long long), 0, 2 * sizeof(unsigned long long)));
    }
    return NBLS_OK;
}

int nbls_xcorr_audit(nbls_ctx* ctx, double* max_err_all_lags, int64_t* n_lags, int32_t reset) {
    NBLS_REQUIRE(ctx, "nbls_xcorr_audit: null context");
    NBLS_CHECK(nbls_xcorr_stats(ctx, nullptr, nullptr, nullptr, 0));     // folds the device counters into the host copies
    if (max_err_all_lags) *max_err_all_lags = ctx->audit_err;
    if (n_lags) *n_lags = ctx->audit_lags;
    if (reset) { ctx->audit_err = 0.0; ctx->audit_lags = 0; }
    return NBLS_OK;
}

int nbls_xcorr_two_stage_stats(nbls_ctx* ctx, int64_t* n_windows, int64_t* n_redone, int32_t reset) {
    NBLS_REQUIRE(ctx, "nbls_xcorr_two_stage_stats: null context");
    if (n_windows) *n_windows = ctx->ts_windows;
    if (n_redone) *n_redone = ctx->ts_redone;
    if (reset) { ctx->ts_windows = 0; ctx->ts_redone = 0; ctx->ts_skip = 0; }
    return NBLS_OK;
}

int nbls_xcorr_two_stage_pairs(nbls_ctx* ctx, int64_t* n_redone_pairs, int32_t reset) {
    NBLS_REQUIRE(ctx, "nbls_xcorr_two_stage_pairs: null context");
    if (n_redone_pairs) *n_redone_pairs = ctx->ts_redone_pairs;
    if (reset) ctx->ts_redone_pairs = 0;
    return NBLS_OK;
}

int nbls_solve(nbls_ctx* ctx, const nbls_solve_desc* sol, int64_t nwin, const int32_t* lag_dev, const double* cmax_dev,
               const nbls_outputs* out_dev) {
    NBLS_REQUIRE(ctx, "nbls_solve: null context");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    return solve_impl(ctx, sol, nwin, lag_dev, cmax_dev, out_dev);
}

int nbls_run_device(nbls_ctx* ctx, const nbls_job* job, const double* x_dev, int32_t nchan, int64_t npts, int64_t ldx,
                    const nbls_outputs* out_dev) {
    NBLS_REQUIRE(ctx && job && x_dev && out_dev, "nbls_run_device: null argument");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    int64_t E = 0;
    NBLS_CHECK(count_windows(job, nchan, round_up(npts, 16), &E));
    NBLS_CHECK(record(ctx, EV_START));
    NBLS_CHECK(record(ctx, EV_H2D));
    ctx->cmax_median_only = out_dev->cmax == nullptr;
    const int rc_run = run_device_inner(ctx, job, x_dev, nchan, npts, ldx, out_dev, E);
    ctx->cmax_median_only = false;
    NBLS_CHECK(rc_run);
    NBLS_CHECK(record(ctx, EV_D2H));
    ctx->ev_valid = true;
    return NBLS_OK;
}

int nbls_run_host(nbls_ctx* ctx, const nbls_job* job, const double* x_host, int32_t nchan, int64_t npts,
                  int64_t ldx_host, const nbls_outputs* oh) {
    NBLS_REQUIRE(ctx && job && x_host && oh, "nbls_run_host: null argument");
    NBLS_REQUIRE(ldx_host >= npts && nchan >= 1, "nbls_run_host: bad input shape");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    const int C = nchan, P = C * (C - 1) / 2;
    int64_t E = 0;
    const int64_t ld = round_up(npts, 16);
    NBLS_CHECK(count_windows(job, C, ld, &E));
    const int64_t En = std::max<int64_t>(E, 1);
    NBLS_CHECK(ensure(ctx, ctx->xin, (size_t)C * ld * sizeof(double)));
    NBLS_CHECK(ensure(ctx, ctx->res, (size_t)En * 7 * sizeof(double)));
    NBLS_CHECK(ensure(ctx, ctx->lagbuf, (size_t)En * P * sizeof(int32_t)));
    NBLS_CHECK(ensure(ctx, ctx->cmaxbuf, (size_t)En * P * sizeof(double)));
    const bool lts = job->solve.alpha != 1.0;
    if (lts) NBLS_CHECK(ensure(ctx, ctx->wbuf, (size_t)En * P));
    double* r = (double*)ctx->res.p;
    nbls_outputs od;
    memset(&od, 0, sizeof(od));
    od.mdccm = r; od.mad = r + En; od.vel = r + 2 * En; od.baz = r + 3 * En; od.sigma_tau = r + 4 * En;
    od.slowness = r + 5 * En;
    od.lag = (int32_t*)ctx->lagbuf.p;
    od.cmax = (double*)ctx->cmaxbuf.p;
    od.weights = lts ? (uint8_t*)ctx->wbuf.p : nullptr;

    NBLS_CHECK(record(ctx, EV_START));
    NBLS_CUDA(cudaMemcpy2DAsync(ctx->xin.p, (size_t)ld * sizeof(double), x_host, (size_t)ldx_host * sizeof(double),
                                (size_t)npts * sizeof(double), C, cudaMemcpyHostToDevice, ctx->stream));
    NBLS_CHECK(record(ctx, EV_H2D));
    ctx->cmax_median_only = oh->cmax == nullptr;
    const int rc_run = run_device_inner(ctx, job, (const double*)ctx->xin.p, C, npts, ld, &od, E);
    ctx->cmax_median_only = false;
    NBLS_CHECK(rc_run);
    if (E > 0) {
        auto back = [&](void* dst, const void* src, size_t bytes) -> int {
            if (dst) NBLS_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
            return NBLS_OK;
        };
        NBLS_CHECK(back(oh->mdccm, od.mdccm, E * sizeof(double)));
        NBLS_CHECK(back(oh->mad, od.mad, E * sizeof(double)));
        NBLS_CHECK(back(oh->vel, od.vel, E * sizeof(double)));
        NBLS_CHECK(back(oh->baz, od.baz, E * sizeof(double)));
        NBLS_CHECK(back(oh->sigma_tau, od.sigma_tau, E * sizeof(double)));
        NBLS_CHECK(back(oh->slowness, od.slowness, 2 * E * sizeof(double)));
        NBLS_CHECK(back(oh->lag, od.lag, (size_t)E * P * sizeof(int32_t)));
        NBLS_CHECK(back(oh->cmax, od.cmax, (size_t)E * P * sizeof(double)));
        if (lts) NBLS_CHECK(back(oh->weights, od.weights, (size_t)E * P));
    }
    NBLS_CHECK(record(ctx, EV_D2H));
    NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->ev_valid = true;
    ctx->last_E = E;
    ctx->last_P = P;
    return NBLS_OK;
}

int nbls_resolve_host(nbls_ctx* ctx, const nbls_solve_desc* sol, const nbls_outputs* oh) {
    NBLS_REQUIRE(ctx && sol && oh, "nbls_resolve_host: null argument");
    NBLS_REQUIRE(ctx->last_E >= 0, "nbls_resolve_host: no nbls_run_host call to re-solve");
    NBLS_REQUIRE(sol->npairs == ctx->last_P, "nbls_resolve_host: %d pairs, the retained lags have %d", sol->npairs,
                 (int)ctx->last_P);
    NBLS_CUDA(cudaSetDevice(ctx->device));
    const int64_t E = ctx->last_E, En = std::max<int64_t>(E, 1);
    const int P = sol->npairs;
    const bool lts = sol->alpha != 1.0;
    if (lts) NBLS_CHECK(ensure(ctx, ctx->wbuf, (size_t)En * P));
    double* r = (double*)ctx->res.p;
    nbls_outputs od;
    memset(&od, 0, sizeof(od));
    od.mdccm = r; od.mad = r + En; od.vel = r + 2 * En; od.baz = r + 3 * En; od.sigma_tau = r + 4 * En;
    od.slowness = r + 5 * En;
    od.weights = lts ? (uint8_t*)ctx->wbuf.p : nullptr;
    if (E > 0) {
        NBLS_CHECK(solve_impl(ctx, sol, E, (const int32_t*)ctx->lagbuf.p, (const double*)ctx->cmaxbuf.p, &od));
        auto back = [&](void* dst, const void* src, size_t bytes) -> int {
            if (dst) NBLS_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
            return NBLS_OK;
        };
        NBLS_CHECK(back(oh->mdccm, od.mdccm, E * sizeof(double)));
        NBLS_CHECK(back(oh->mad, od.mad, E * sizeof(double)));
        NBLS_CHECK(back(oh->vel, od.vel, E * sizeof(double)));
        NBLS_CHECK(back(oh->baz, od.baz, E * sizeof(double)));
        NBLS_CHECK(back(oh->sigma_tau, od.sigma_tau, E * sizeof(double)));
        NBLS_CHECK(back(oh->slowness, od.slowness, 2 * E * sizeof(double)));
        if (lts) NBLS_CHECK(back(oh->weights, od.weights, (size_t)E * P));
    }
    NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
    return NBLS_OK;
}

int nbls_last_timings(nbls_ctx* ctx, float* ms, int32_t n) {
    NBLS_REQUIRE(ctx && ms && n >= 1, "nbls_last_timings: bad argument");
    NBLS_REQUIRE(ctx->ev_valid, "nbls_last_timings: no run recorded yet");
    NBLS_CUDA(cudaEventSynchronize(ctx->ev[EV_D2H]));
    float v[6] = {0, 0, 0, 0, 0, 0};
    NBLS_CUDA(cudaEventElapsedTime(&v[0], ctx->ev[EV_START], ctx->ev[EV_H2D]));
    if (ctx->stage_ms[1] >= 0.f) {   // band-grouped run: stage times were accumulated per group
        v[1] = ctx->stage_ms[1];
        v[2] = ctx->stage_ms[2];
    } else {
        NBLS_CUDA(cudaEventElapsedTime(&v[1], ctx->ev[EV_H2D], ctx->ev[EV_FILTER]));
        NBLS_CUDA(cudaEventElapsedTime(&v[2], ctx->ev[EV_FILTER], ctx->ev[EV_XCORR]));
    }
    NBLS_CUDA(cudaEventElapsedTime(&v[3], ctx->ev[EV_XCORR], ctx->ev[EV_SOLVE]));
    NBLS_CUDA(cudaEventElapsedTime(&v[4], ctx->ev[EV_SOLVE], ctx->ev[EV_D2H]));
    NBLS_CUDA(cudaEventElapsedTime(&v[5], ctx->ev[EV_START], ctx->ev[EV_D2H]));
    const int m = std::min<int>(n, 6);
    for (int i = 0; i < m; ++i) ms[i] = v[i];
    return m;
}

int nbls_mark(nbls_ctx* ctx, int32_t slot) {
    NBLS_REQUIRE(ctx && slot >= 0 && slot < 8, "nbls_mark: bad argument");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    if (!ctx->mark[slot]) NBLS_CUDA(cudaEventCreate(&ctx->mark[slot]));
    NBLS_CUDA(cudaEventRecord(ctx->mark[slot], ctx->stream));
    return NBLS_OK;
}
int nbls_mark_elapsed(nbls_ctx* ctx, int32_t a, int32_t b, float* ms) {
    NBLS_REQUIRE(ctx && ms && a >= 0 && a < 8 && b >= 0 && b < 8 && ctx->mark[a] && ctx->mark[b],
                 "nbls_mark_elapsed: bad argument or unrecorded slot");
    NBLS_CUDA(cudaEventSynchronize(ctx->mark[b]));
    NBLS_CUDA(cudaEventElapsedTime(ms, ctx->mark[a], ctx->mark[b]));
    return NBLS_OK;
}

int64_t nbls_launch_count(nbls_ctx* ctx) { return ctx ? ctx->launches : -1; }

int nbls_measure_fp64_peak(nbls_ctx* ctx, int32_t iters, double* tflops) {
    NBLS_REQUIRE(ctx, "nbls_measure_fp64_peak: null context");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    return fp64_peak_impl(ctx, iters, tflops);
}

int nbls_experiment_dft_pass(nbls_ctx* ctx, int32_t mode, const float* x_dev, int64_t rows, int32_t reps, int32_t iters,
                             float* y_dev, float* ms) {
    NBLS_REQUIRE(ctx, "nbls_experiment_dft_pass: null context");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    return experiment_dft_pass_impl(ctx, mode, x_dev, rows, reps, iters, y_dev, ms);
}

int nbls_measure_fp32_peak(nbls_ctx* ctx, int32_t iters, double* tflops) {
    NBLS_REQUIRE(ctx, "nbls_measure_fp32_peak: null context");
    NBLS_CUDA(cudaSetDevice(ctx->device));
    return fp32_peak_impl(ctx, iters, tflops);
}

}  // extern "C"

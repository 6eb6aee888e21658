// fp32x2_microbench.cu — issue rate of the sm_100 packed-FP32 instructions (FADD2 / FMUL2 / FFMA2: two FP32 lanes per
// register pair and per issue slot) against their scalar forms, alone and mixed with ALU-pipe work, at the occupancy of
// the pair kernel (24 warps per SM).  Decides whether the FFT butterflies of xcorr_fft.cu are worth packing.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32x2_microbench scripts/fp32x2_microbench.cu && ./fp32x2_microbench
#include <cuda_runtime.h>
#include <cstdio>

constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(256, 3) bench(float* out, float seed, int iters) {
    // 16 independent FP32 chains per thread = 8 packed chains
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(seed + i + threadIdx.x, seed - i);
    const float2 m = make_float2(seed * 0.999f, seed * 1.001f), c = make_float2(seed * 1e-3f, -seed * 1e-3f);
    unsigned x = threadIdx.x, y = blockIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) { a[i].x = fmaf(a[i].x, m.x, c.x); a[i].y = fmaf(a[i].y, m.y, c.y); }          // 16 FFMA
            if (MODE == 1) a[i] = __ffma2_rn(a[i], m, c);                                                // 8 FFMA2
            if (MODE == 2) { a[i].x = a[i].x + c.x; a[i].y = a[i].y + c.y; }                              // 16 FADD
            if (MODE == 3) a[i] = __fadd2_rn(a[i], c);                                                   // 8 FADD2
            if (MODE == 4) { a[i].x = fmaf(a[i].x, m.x, c.x); a[i].y = fmaf(a[i].y, m.y, c.y); x = (x ^ (x >> 3)) + 0x9e37u; y = (y ^ (y >> 5)) + 0x79b9u; }
            if (MODE == 5) { a[i] = __ffma2_rn(a[i], m, c); x = (x ^ (x >> 3)) + 0x9e37u; y = (y ^ (y >> 5)) + 0x79b9u; }
            if (MODE == 6) { a[i].x = a[i].x * m.x; a[i].y = a[i].y * m.y; }                              // 16 FMUL
            if (MODE == 7) a[i] = __fmul2_rn(a[i], m);                                                   // 8 FMUL2
        }
    }
    float s = (float)(x + y);
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    if (s == 12345.678f) out[0] = s;
}

template <int MODE>
static void run(const char* name, float* d, int sms) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sms * 3;
    bench<MODE><<<grid, 256>>>(d, 1.0f, 64);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    bench<MODE><<<grid, 256>>>(d, 1.0f, ITERS);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double lane_ops = (double)grid * 256 * ITERS * 16;      // FP32 lane operations
    printf("{\"mode\": \"%s\", \"ms\": %.4f, \"fp32_lane_ops_per_s\": %.4e, \"per_sm_per_clk_at_1965MHz\": %.1f}\n", name, ms,
           lane_ops / (ms * 1e-3), lane_ops / (ms * 1e-3) / sms / 1.965e9);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* d;
    cudaMalloc(&d, 4);
    run<0>("FFMA  x16", d, sms);
    run<1>("FFMA2 x8", d, sms);
    run<2>("FADD  x16", d, sms);
    run<3>("FADD2 x8", d, sms);
    run<6>("FMUL  x16", d, sms);
    run<7>("FMUL2 x8", d, sms);
    run<4>("FFMA  x16 + 16 int updates", d, sms);
    run<5>("FFMA2 x8 + 16 int updates", d, sms);
    return cudaGetLastError() != cudaSuccess;
}

// fft_radix.cuh — register-resident DFT butterflies (radix 2..32, factors 2 and 3) used by the
// shared-memory FFTs of xcorr_fft.cu, written once for complex FP64 (double2) and complex FP32 (float2).  All transforms have the INVERSE sign, e^{+2 pi i kn/R}, and
// are unnormalised; every index is a compile-time constant after unrolling so the operands stay in
// registers and trivial twiddles (1, i, -1, -i) cost nothing.
#pragma once
#include <cuda_runtime.h>

namespace nbls {
namespace fft {

// cos(2 pi k / 96), k = 0..24, correctly rounded (generated in 80-bit arithmetic).  96 = lcm(32, 24)
// covers every root of unity the radix-2/3 butterflies below need.
constexpr double Q96[25] = {
    1.0,
    0.9978589232386035,
    0.9914448613738104,
    0.9807852804032304,
    0.9659258262890683,
    0.9469301294951057,
    0.9238795325112867,
    0.8968727415326883,
    0.8660254037844386,
    0.8314696123025452,
    0.7933533402912352,
    0.7518398074789774,
    0.7071067811865476,
    0.6593458151000688,
    0.6087614290087207,
    0.5555702330196022,
    0.5,
    0.4422886902190013,
    0.3826834323650898,
    0.3214394653031616,
    0.25881904510252074,
    0.19509032201612828,
    0.1305261922200516,
    0.06540312923014306,
    0.0};

constexpr int mod96(int k) { return ((k % 96) + 96) % 96; }
constexpr double cos96(int k) {
    k = mod96(k);
    if (k > 48) k = 96 - k;                      // cos is even about 180 deg
    return k <= 24 ? Q96[k] : -Q96[48 - k];      // and odd about 90 deg
}
constexpr double sin96(int k) { return cos96(k - 24); }

// complex helpers for both precisions: V is double2 or float2, real_t<V> its component type
template <class V> struct real_of;
template <> struct real_of<double2> { typedef double type; };
template <> struct real_of<float2> { typedef float type; };
template <class V> using real_t = typename real_of<V>::type;
template <class F> struct cplx_of;
template <> struct cplx_of<double> { typedef double2 type; };
template <> struct cplx_of<float> { typedef float2 type; };
template <class F> using cplx_t = typename cplx_of<F>::type;

__device__ __forceinline__ double2 mk2(double x, double y) { return make_double2(x, y); }
__device__ __forceinline__ float2 mk2(float x, float y) { return make_float2(x, y); }
__device__ __forceinline__ double fmad(double a, double b, double c) { return fma(a, b, c); }
__device__ __forceinline__ float fmad(float a, float b, float c) { return fmaf(a, b, c); }

template <class V> __device__ __forceinline__ V cadd(V a, V b) { return mk2(a.x + b.x, a.y + b.y); }
template <class V> __device__ __forceinline__ V csub(V a, V b) { return mk2(a.x - b.x, a.y - b.y); }
template <class V> __device__ __forceinline__ V cconj(V a) { return mk2(a.x, -a.y); }
template <class V> __device__ __forceinline__ V cmul(V a, V b) {
    return mk2(fmad(a.x, b.x, -(a.y * b.y)), fmad(a.x, b.y, a.y * b.x));
}
// a * conj(b)
template <class V> __device__ __forceinline__ V cmulc(V a, V b) {
    return mk2(fmad(a.x, b.x, a.y * b.y), fmad(a.y, b.x, -(a.x * b.y)));
}

// v * e^{+2 pi i NUM/DEN}, DEN | 96
template <int NUM, int DEN, class V>
__device__ __forceinline__ V mul_root(V v) {
    typedef real_t<V> F;
    static_assert(96 % DEN == 0, "root order must divide 96");
    constexpr int k = mod96(NUM * (96 / DEN));
    if constexpr (k == 0) return v;
    else if constexpr (k == 24) return mk2(-v.y, v.x);
    else if constexpr (k == 48) return mk2(-v.x, -v.y);
    else if constexpr (k == 72) return mk2(v.y, -v.x);
    else {
        constexpr F c = (F)cos96(k), s = (F)sin96(k);
        return mk2(fmad(c, v.x, -(s * v.y)), fmad(c, v.y, s * v.x));
    }
}

template <int R>
struct Dft {
    static_assert(R % 2 == 0, "radix must be 2^a or 3*2^a");
    // decimation in time: even/odd halves, then one layer of radix-2 butterflies
    template <class V>
    static __device__ __forceinline__ void run(V (&v)[R]) {
        constexpr int H = R / 2;
        V e[H], o[H];
#pragma unroll
        for (int k = 0; k < H; ++k) { e[k] = v[2 * k]; o[k] = v[2 * k + 1]; }
        Dft<H>::run(e);
        Dft<H>::run(o);
        combine<0>(v, e, o);
    }
    template <int N, class V>
    static __device__ __forceinline__ void combine(V (&v)[R], const V (&e)[R / 2], const V (&o)[R / 2]) {
        if constexpr (N < R / 2) {
            const V t = mul_root<N, R>(o[N]);
            v[N] = cadd(e[N], t);
            v[N + R / 2] = csub(e[N], t);
            combine<N + 1>(v, e, o);
        }
    }
};
template <>
struct Dft<1> {
    template <class V>
    static __device__ __forceinline__ void run(V (&)[1]) {}
};
template <>
struct Dft<2> {
    template <class V>
    static __device__ __forceinline__ void run(V (&v)[2]) {
        const V a = v[0], b = v[1];
        v[0] = cadd(a, b);
        v[1] = csub(a, b);
    }
};
template <>
struct Dft<3> {
    template <class V>
    static __device__ __forceinline__ void run(V (&v)[3]) {
        typedef real_t<V> F;
        const V t = cadd(v[1], v[2]), d = csub(v[1], v[2]);
        const V m = mk2(fmad((F)-0.5, t.x, v[0].x), fmad((F)-0.5, t.y, v[0].y));
        constexpr F s = (F)Q96[8];   // sin 120 deg
        const V ie = mk2(-s * d.y, s * d.x);          // i * s * d
        v[0] = cadd(v[0], t);
        v[1] = cadd(m, ie);
        v[2] = csub(m, ie);
    }
};
template <>
struct Dft<4> {
    template <class V>
    static __device__ __forceinline__ void run(V (&v)[4]) {
        const V a = cadd(v[0], v[2]), b = csub(v[0], v[2]), c = cadd(v[1], v[3]), d = csub(v[1], v[3]);
        const V id = mk2(-d.y, d.x);
        v[0] = cadd(a, c);
        v[2] = csub(a, c);
        v[1] = cadd(b, id);
        v[3] = csub(b, id);
    }
};

// W[n] = W[1]^n for n = 2..R-1 by a balanced product tree (depth <= log2 R roundings)
template <int R, int N = 2, class V>
__device__ __forceinline__ void powers(V (&W)[R]) {
    if constexpr (N < R) {
        W[N] = cmul(W[N / 2], W[N - N / 2]);
        powers<R, N + 1>(W);
    }
}

}  // namespace fft
}  // namespace nbls

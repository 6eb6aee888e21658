// fft_radix.cuh — register-resident DFT butterflies (radix 2..32, factors 2 and 3) used by the
// shared-memory FFTs of xcorr_fft.cu, written once for complex FP64 (double2) and complex FP32 (float2).  All transforms have the INVERSE sign, e^{+2 pi i kn/R}, and
// are unnormalised; every index is a compile-time constant after unrolling so the operands stay in
// registers and trivial twiddles (1, i, -1, -i) cost nothing.
#pragma once
#include <cuda_runtime.h>

namespace nbls {
namespace fft {

// cos(2 pi k / 480), k = 0..120, correctly rounded (generated with 50-digit arithmetic).  480 = lcm(32, 24, 20)
// covers every root of unity the radix-2/3/5 butterflies below need.
constexpr int QN = 480;
constexpr double QCOS[QN / 4 + 1] = {
    1.0, 0.999914327574007, 0.9996573249755573, 0.9992290362407229,
    0.9986295347545738, 0.9978589232386035, 0.996917333733128, 0.9958049275746618,
    0.9945218953682733, 0.9930684569549263, 0.9914448613738104, 0.9896513868196702,
    0.9876883405951378, 0.9855560590580777, 0.9832549075639546, 0.9807852804032304,
    0.9781476007338057, 0.9753423205085127, 0.9723699203976766, 0.9692309097067544,
    0.9659258262890683, 0.9624552364536473, 0.958819734868193, 0.9550199444571866,
    0.9510565162951535, 0.9469301294951057, 0.9426414910921784, 0.9381913359224842,
    0.9335804264972017, 0.9288095528719242, 0.9238795325112867, 0.9187912101488983,
    0.9135454576426009, 0.9081431738250813, 0.9025852843498606, 0.8968727415326883,
    0.8910065241883679, 0.8849876374630419, 0.8788171126619654, 0.8724960070727971,
    0.8660254037844386, 0.8594064115014527, 0.8526401643540922, 0.8457278217039732,
    0.838670567945424, 0.8314696123025452, 0.8241261886220157, 0.8166415551616789,
    0.8090169943749475, 0.8012538126910607, 0.7933533402912352, 0.785316930880745,
    0.7771459614569709, 0.7688418320734596, 0.7604059656000309, 0.7518398074789774,
    0.7431448254773942, 0.7343225094356856, 0.7253743710122876, 0.7163019434246543,
    0.7071067811865476, 0.6977904598416801, 0.688354575693754, 0.6788007455329418,
    0.6691306063588582, 0.6593458151000688, 0.6494480483301837, 0.6394390019805848,
    0.6293203910498375, 0.619093949309834, 0.6087614290087207, 0.5983246005706591,
    0.5877852522924731, 0.5771451900372336, 0.5664062369248328, 0.5555702330196022,
    0.5446390350150271, 0.5336145159156115, 0.5224985647159489, 0.511293086077052,
    0.5, 0.48862124149695496, 0.4771587602596084, 0.4656145203251114,
    0.4539904997395468, 0.4422886902190013, 0.43051109680829514, 0.4186597375374281,
    0.4067366430758002, 0.39474385638426723, 0.3826834323650898, 0.3705574375098362,
    0.35836794954530027, 0.34611705707749296, 0.3338068592337709, 0.3214394653031616,
    0.30901699437494745, 0.29654157497557093, 0.2840153447039226, 0.27144044986507426,
    0.25881904510252074, 0.24615329302899303, 0.23344536385590542, 0.2206974350215011,
    0.20791169081775934, 0.19509032201612828, 0.18223552549214744, 0.1693495038490246,
    0.15643446504023087, 0.14349262199117932, 0.1305261922200516, 0.11753739745783764,
    0.10452846326765347, 0.09150161866340238, 0.07845909572784494, 0.06540312923014306,
    0.052335956242943835, 0.03925981575906861, 0.026176948307873153, 0.01308959557134444,
    0.0};

constexpr int modq(int k) { return ((k % QN) + QN) % QN; }
constexpr double cosq(int k) {
    k = modq(k);
    if (k > QN / 2) k = QN - k;                          // cos is even about 180 deg
    return k <= QN / 4 ? QCOS[k] : -QCOS[QN / 2 - k];    // and odd about 90 deg
}
constexpr double sinq(int k) { return cosq(k - QN / 4); }

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
// i * a
template <class V> __device__ __forceinline__ V cmul_i(V a) { return mk2(-a.y, a.x); }
// a + i b  and  conj(a) + i conj(b)
template <class V> __device__ __forceinline__ V cadd_i(V a, V b) { return mk2(a.x - b.y, a.y + b.x); }
template <class V> __device__ __forceinline__ V cadd_i_conj(V a, V b) { return mk2(a.x + b.y, b.x - a.y); }
// half-by-half product
template <class V> __device__ __forceinline__ V cmul_elem(V a, V b) { return mk2(a.x * b.x, a.y * b.y); }
// c * t  and  a + c * t  with a real c
template <class V> __device__ __forceinline__ V cscale(real_t<V> c, V t) { return mk2(c * t.x, c * t.y); }
template <class V> __device__ __forceinline__ V caxpy(real_t<V> c, V t, V a) {
    return mk2(fmad(c, t.x, a.x), fmad(c, t.y, a.y));
}

// ---- FP32: the same operations on the packed instructions of sm_100 (FADD2 / FMUL2 / FFMA2) ---------------------
// A complex FP32 number is a register pair (re, im) and the packed instructions work on both halves in ONE issue slot;
// their operands take a half swap (.LO_HI), a per-half negation and a scalar broadcast for free, which is all complex
// arithmetic needs (times i = swap + negate one half).  Every overload below performs, half by half, exactly the
// operations of its generic twin above (same products, same fused multiply-adds, IEEE round-to-nearest), so the FP32
// transforms are bit-identical to the scalar formulation at half the floating-point instruction count.
__device__ __forceinline__ float2 bc2(float s) { return make_float2(s, s); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    const float2 t = __fmul2_rn(bc2(a.y), make_float2(b.y, b.x));                  // (ay by, ay bx)
    return __ffma2_rn(bc2(a.x), b, make_float2(-t.x, t.y));
}
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {
    const float2 t = __fmul2_rn(make_float2(a.y, a.x), bc2(b.y));                  // (ay by, ax by)
    return __ffma2_rn(a, bc2(b.x), make_float2(t.x, -t.y));
}
__device__ __forceinline__ float2 cadd_i(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.y, b.x)); }
__device__ __forceinline__ float2 cadd_i_conj(float2 a, float2 b) {
    return __fadd2_rn(make_float2(a.x, -a.y), make_float2(b.y, b.x));
}
__device__ __forceinline__ float2 cmul_elem(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 cscale(float c, float2 t) { return __fmul2_rn(bc2(c), t); }
__device__ __forceinline__ float2 caxpy(float c, float2 t, float2 a) { return __ffma2_rn(bc2(c), t, a); }

// v * e^{+2 pi i NUM/DEN}, DEN | 480
template <int NUM, int DEN, class V>
__device__ __forceinline__ V mul_root(V v) {
    typedef real_t<V> F;
    static_assert(QN % DEN == 0, "root order must divide 480");
    constexpr int k = modq(NUM * (QN / DEN));
    if constexpr (k == 0) return v;
    else if constexpr (k == QN / 4) return mk2(-v.y, v.x);
    else if constexpr (k == QN / 2) return mk2(-v.x, -v.y);
    else if constexpr (k == 3 * QN / 4) return mk2(v.y, -v.x);
    else {
        constexpr F c = (F)cosq(k), s = (F)sinq(k);
        if constexpr (sizeof(F) == 4) {
            const float2 t = __fmul2_rn(bc2(s), make_float2(v.y, v.x));            // (s vy, s vx)
            return __ffma2_rn(bc2(c), v, make_float2(-t.x, t.y));
        } else {
            return mk2(fmad(c, v.x, -(s * v.y)), fmad(c, v.y, s * v.x));
        }
    }
}

template <int R>
struct Dft {
    static_assert(R % 2 == 0, "radix must be 2^a, 3*2^a or 5*2^a");
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
        const V m = caxpy((F)-0.5, t, v[0]);
        constexpr F s = (F)sinq(QN / 3);   // sin 120 deg
        const V ie = cscale(s, cmul_i(d));            // i * s * d
        v[0] = cadd(v[0], t);
        v[1] = cadd(m, ie);
        v[2] = csub(m, ie);
    }
};
template <>
struct Dft<5> {
    template <class V>
    static __device__ __forceinline__ void run(V (&v)[5]) {
        typedef real_t<V> F;
        constexpr F c1 = (F)cosq(QN / 5), c2 = (F)cosq(2 * QN / 5), s1 = (F)sinq(QN / 5), s2 = (F)sinq(2 * QN / 5);
        const V t1 = cadd(v[1], v[4]), t2 = cadd(v[2], v[3]), t3 = csub(v[1], v[4]), t4 = csub(v[2], v[3]);
        const V m1 = caxpy(c2, t2, caxpy(c1, t1, v[0]));
        const V m2 = caxpy(c1, t2, caxpy(c2, t1, v[0]));
        const V a1 = caxpy(s2, t4, cscale(s1, t3));        // s1 t3 + s2 t4
        const V a2 = caxpy(-s1, t4, cscale(s2, t3));       // s2 t3 - s1 t4
        const V n1 = cmul_i(a1), n2 = cmul_i(a2);          // times i
        v[0] = cadd(v[0], cadd(t1, t2));
        v[1] = cadd(m1, n1);
        v[4] = csub(m1, n1);
        v[2] = cadd(m2, n2);
        v[3] = csub(m2, n2);
    }
};
template <>
struct Dft<4> {
    template <class V>
    static __device__ __forceinline__ void run(V (&v)[4]) {
        const V a = cadd(v[0], v[2]), b = csub(v[0], v[2]), c = cadd(v[1], v[3]), d = csub(v[1], v[3]);
        const V id = cmul_i(d);
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

// Register-resident DFT codelets for the shared-memory mixed-radix FFT of the acquisition pipeline.
//
// Dft<R, INV>::run(v) transforms v[0..R) in place, natural order in and out:
//     INV = false:  X[k] = sum_n v[n] exp(-2*pi*i*n*k/R)      (FFTW forward)
//     INV = true :  X[k] = sum_n v[n] exp(+2*pi*i*n*k/R)      (FFTW backward, unnormalised)
// This is what gr::fft::fft_complex_fwd / _rev compute for the reference
// (reference: src/algorithms/libs/gnss_sdr_fft.h:26-63; pcps_acquisition.cc:662,668).
//
// Primitive codelets: 2, 4 and any odd prime (symmetric form, compile-time cos/sin). Composite radices are
// Cooley-Tukey products of those, fully unrolled, with every internal twiddle a compile-time constant (and
// multiples of pi/4 strength-reduced). Everything stays in registers.
#pragma once
#include <cuda_runtime.h>
#include <type_traits>
#include <utility>

namespace b200
{
namespace fft
{
// ---------------- compile-time trigonometry (double, then rounded once to float) ----------------
constexpr double kPi = 3.14159265358979323846264338327950288;

constexpr double cx_sin_small(double x)  // |x| <= pi/4
{
    const double x2 = x * x;
    double term = x, sum = x;
    for (int i = 1; i < 14; i++)
        {
            term *= -x2 / ((2.0 * i) * (2.0 * i + 1.0));
            sum += term;
        }
    return sum;
}
constexpr double cx_cos_small(double x)
{
    const double x2 = x * x;
    double term = 1.0, sum = 1.0;
    for (int i = 1; i < 14; i++)
        {
            term *= -x2 / ((2.0 * i - 1.0) * (2.0 * i));
            sum += term;
        }
    return sum;
}
// cos / sin of 2*pi*k/n with octant reduction on the integer fraction (exact argument reduction)
constexpr double cx_cos_frac(long long k, long long n)
{
    k %= n;
    if (k < 0) k += n;
    if (2 * k > n) k = n - k;                                             // cos(2pi - t) = cos t
    if (4 * k > n) return -cx_cos_frac(n - 2 * k, 2 * n);                 // cos t = -cos(pi - t): (n/2 - k)/n
    if (8 * k > n) return cx_sin_small(2.0 * kPi * (double)(n - 4 * k) / (4.0 * (double)n));  // cos t = sin(pi/2 - t)
    return cx_cos_small(2.0 * kPi * (double)k / (double)n);
}
constexpr double cx_sin_frac(long long k, long long n)
{
    // sin t = cos(t - pi/2) = cos(2pi*(4k - n)/(4n))
    return cx_cos_frac(4 * k - n, 4 * n);
}

template <int N>
struct IC : std::integral_constant<int, N>
{
};

template <class F, int... I>
__device__ __forceinline__ void static_for_impl(F&& f, std::integer_sequence<int, I...>)
{
    (f(IC<I>{}), ...);
}
template <int N, class F>
__device__ __forceinline__ void static_for(F&& f)
{
    static_for_impl(static_cast<F&&>(f), std::make_integer_sequence<int, N>{});
}

// ---------------- value types ----------------
// The codelets are written once over a scalar type T. The kernels use T = float (complex value = float2). A packed
// FP32x2 scalar (two transforms per thread in the lanes of FADD2 / FFMA2) was measured with tools/scratch/dft_pack_bench.cu:
// the packed instructions issue at half rate, so points/s of the codelets do not change (radix 16: 3.37 vs 3.39 T points/s
// on B200) and the kernels, which are not issue-bound, keep the scalar form.
using ::fmaf;
template <class T>
struct cx  // complex value over T; cx<float> has the layout of float2
{
    T x, y;
};
template <class T>
__device__ __forceinline__ cx<T> make_cx(T x, T y)
{
    return cx<T>{x, y};
}
template <class T>
__device__ __forceinline__ T zero_of();
template <>
__device__ __forceinline__ float zero_of<float>()
{
    return 0.f;
}

template <class T>
__device__ __forceinline__ cx<T> cmul(const cx<T> a, const cx<T> b)
{
    return make_cx<T>(fmaf(a.x, b.x, -(a.y * b.y)), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul(const float2 a, const float2 b)
{
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul_conj(const float2 a, const float2 b)  // a * conj(b)
{
    return make_float2(fmaf(a.x, b.x, a.y * b.y), fmaf(a.y, b.x, -a.x * b.y));
}

// x *= exp(-/+ 2*pi*i*E/R) with E, R compile-time (sign '-' forward, '+' inverse)
template <int E_, int R, bool INV, class T>
__device__ __forceinline__ void mul_root(cx<T>& x)
{
    constexpr int E = ((E_ % R) + R) % R;
    if constexpr (E == 0)
        {
        }
    else if constexpr (2 * E == R)
        {
            x = make_cx<T>(-x.x, -x.y);
        }
    else if constexpr (4 * E == R)  // -i (fwd) / +i (inv)
        {
            x = INV ? make_cx<T>(-x.y, x.x) : make_cx<T>(x.y, -x.x);
        }
    else if constexpr (4 * E == 3 * R)  // +i (fwd) / -i (inv)
        {
            x = INV ? make_cx<T>(x.y, -x.x) : make_cx<T>(-x.y, x.x);
        }
    else if constexpr ((8 * E) % R == 0)  // odd multiples of pi/4
        {
            constexpr float h = 0.70710678118654752440f;
            constexpr int o = (8 * E) / R;  // 1,3,5,7
            // forward root: (c - i s); c = +-h, s = +-h
            constexpr float c = (o == 1 || o == 7) ? h : -h;
            constexpr float s0 = (o == 1 || o == 3) ? h : -h;
            constexpr float s = INV ? -s0 : s0;  // forward multiplies by (c - i s0)
            // (x.x + i x.y)(c - i s) = (x.x c + x.y s) + i (x.y c - x.x s)
            const T re = x.x * c + x.y * s;
            const T im = x.y * c - x.x * s;
            x = make_cx<T>(re, im);
        }
    else
        {
            constexpr float c = (float)cx_cos_frac(E, R);
            constexpr float s0 = (float)cx_sin_frac(E, R);
            constexpr float s = INV ? -s0 : s0;
            const T re = fmaf(x.x, c, x.y * s);
            const T im = fmaf(x.y, c, x.x * (-s));  // the same bits as -(x.x * s)
            x = make_cx<T>(re, im);
        }
}

template <int R, bool INV, class Enable = void>
struct Dft;

template <bool INV>
struct Dft<1, INV>
{
    template <class T>
    static __device__ __forceinline__ void run(cx<T> (&)[1])
    {
    }
    static __device__ __forceinline__ void run(float2 (&)[1]) {}
};

template <bool INV>
struct Dft<2, INV>
{
    static __device__ __forceinline__ void run(float2 (&v)[2]) { run(reinterpret_cast<cx<float>(&)[2]>(v)); }
    template <class T>
    static __device__ __forceinline__ void run(cx<T> (&v)[2])
    {
        const cx<T> a = v[0], b = v[1];
        v[0] = make_cx<T>(a.x + b.x, a.y + b.y);
        v[1] = make_cx<T>(a.x - b.x, a.y - b.y);
    }
};

template <bool INV>
struct Dft<4, INV>
{
    static __device__ __forceinline__ void run(float2 (&v)[4]) { run(reinterpret_cast<cx<float>(&)[4]>(v)); }
    template <class T>
    static __device__ __forceinline__ void run(cx<T> (&v)[4])
    {
        const cx<T> s02 = make_cx<T>(v[0].x + v[2].x, v[0].y + v[2].y);
        const cx<T> d02 = make_cx<T>(v[0].x - v[2].x, v[0].y - v[2].y);
        const cx<T> s13 = make_cx<T>(v[1].x + v[3].x, v[1].y + v[3].y);
        const cx<T> d13 = make_cx<T>(v[1].x - v[3].x, v[1].y - v[3].y);
        // forward: X1 = d02 - i d13, X3 = d02 + i d13
        const cx<T> r = INV ? make_cx<T>(-d13.y, d13.x) : make_cx<T>(d13.y, -d13.x);
        v[0] = make_cx<T>(s02.x + s13.x, s02.y + s13.y);
        v[2] = make_cx<T>(s02.x - s13.x, s02.y - s13.y);
        v[1] = make_cx<T>(d02.x + r.x, d02.y + r.y);
        v[3] = make_cx<T>(d02.x - r.x, d02.y - r.y);
    }
};

constexpr bool is_odd_prime(int r)
{
    if (r < 3 || r % 2 == 0) return false;
    for (int d = 3; d * d <= r; d += 2)
        if (r % d == 0) return false;
    return true;
}

// odd prime P: pair v[k], v[P-k]; X[m], X[P-m] share the cosine sum and differ in the sign of the sine sum
template <int P, bool INV>
struct Dft<P, INV, std::enable_if_t<is_odd_prime(P)>>
{
    static constexpr int H = (P - 1) / 2;
    static __device__ __forceinline__ void run(float2 (&v)[P]) { run(reinterpret_cast<cx<float>(&)[P]>(v)); }
    template <class T>
    static __device__ __forceinline__ void run(cx<T> (&v)[P])
    {
        cx<T> tp[H], tm[H];
        static_for<H>([&](auto kc) {
            constexpr int k = decltype(kc)::value + 1;
            tp[k - 1] = make_cx<T>(v[k].x + v[P - k].x, v[k].y + v[P - k].y);
            tm[k - 1] = make_cx<T>(v[k].x - v[P - k].x, v[k].y - v[P - k].y);
        });
        const cx<T> v0 = v[0];
        cx<T> dc = v0;
        static_for<H>([&](auto kc) {
            constexpr int k = decltype(kc)::value;
            dc.x = dc.x + tp[k].x;
            dc.y = dc.y + tp[k].y;
        });
        v[0] = dc;
        static_for<H>([&](auto mc) {
            constexpr int m = decltype(mc)::value + 1;
            cx<T> a = v0;                        // cosine part
            cx<T> b = make_cx<T>(zero_of<T>(), zero_of<T>());     // sine part: sum_k sin(2 pi m k / P) * tm[k]
            static_for<H>([&](auto kc) {
                constexpr int k = decltype(kc)::value + 1;
                constexpr float c = (float)cx_cos_frac((long long)m * k, P);
                constexpr float s = (float)cx_sin_frac((long long)m * k, P);
                a.x = fmaf(c, tp[k - 1].x, a.x);
                a.y = fmaf(c, tp[k - 1].y, a.y);
                b.x = fmaf(s, tm[k - 1].x, b.x);
                b.y = fmaf(s, tm[k - 1].y, b.y);
            });
            // forward: X[m] = a - i b = (a.x + b.y, a.y - b.x); X[P-m] = a + i b
            if (INV)
                {
                    v[m] = make_cx<T>(a.x - b.y, a.y + b.x);
                    v[P - m] = make_cx<T>(a.x + b.y, a.y - b.x);
                }
            else
                {
                    v[m] = make_cx<T>(a.x + b.y, a.y - b.x);
                    v[P - m] = make_cx<T>(a.x - b.y, a.y + b.x);
                }
        });
    }
};

// Cooley-Tukey composite R = R1*R2: input n = R2*n1 + n2, output k = k1 + R1*k2
template <int R1, int R2, bool INV>
struct DftCT
{
    static constexpr int R = R1 * R2;
    template <class T>
    static __device__ __forceinline__ void run(cx<T> (&v)[R])
    {
        static_for<R2>([&](auto n2c) {
            constexpr int n2 = decltype(n2c)::value;
            cx<T> t[R1];
            static_for<R1>([&](auto n1c) {
                constexpr int n1 = decltype(n1c)::value;
                t[n1] = v[R2 * n1 + n2];
            });
            Dft<R1, INV>::run(t);
            static_for<R1>([&](auto k1c) {
                constexpr int k1 = decltype(k1c)::value;
                mul_root<k1 * n2, R, INV, T>(t[k1]);
                v[R2 * k1 + n2] = t[k1];
            });
        });
        cx<T> o[R];
        static_for<R1>([&](auto k1c) {
            constexpr int k1 = decltype(k1c)::value;
            cx<T> t[R2];
            static_for<R2>([&](auto n2c) {
                constexpr int n2 = decltype(n2c)::value;
                t[n2] = v[R2 * k1 + n2];
            });
            Dft<R2, INV>::run(t);
            static_for<R2>([&](auto k2c) {
                constexpr int k2 = decltype(k2c)::value;
                o[k1 + R1 * k2] = t[k2];
            });
        });
        static_for<R>([&](auto ic) {
            constexpr int i = decltype(ic)::value;
            v[i] = o[i];
        });
    }
};

#define B200_DFT_COMPOSITE(R, R1, R2)                                                   \
    template <bool INV>                                                                 \
    struct Dft<R, INV>                                                                  \
    {                                                                                   \
        template <class T>                                                              \
        static __device__ __forceinline__ void run(cx<T> (&v)[R])                        \
        {                                                                               \
            DftCT<R1, R2, INV>::run(v);                                                 \
        }                                                                               \
        static __device__ __forceinline__ void run(float2 (&v)[R])                       \
        {                                                                               \
            DftCT<R1, R2, INV>::run(reinterpret_cast<cx<float>(&)[R]>(v));              \
        }                                                                               \
    };
B200_DFT_COMPOSITE(6, 2, 3)
B200_DFT_COMPOSITE(8, 4, 2)
B200_DFT_COMPOSITE(9, 3, 3)
B200_DFT_COMPOSITE(10, 2, 5)
B200_DFT_COMPOSITE(12, 4, 3)
B200_DFT_COMPOSITE(15, 3, 5)
B200_DFT_COMPOSITE(16, 4, 4)
B200_DFT_COMPOSITE(20, 4, 5)
B200_DFT_COMPOSITE(25, 5, 5)
#undef B200_DFT_COMPOSITE

}  // namespace fft
}  // namespace b200

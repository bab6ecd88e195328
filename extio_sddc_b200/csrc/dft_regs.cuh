// dft_regs.cuh -- compile-time unrolled, register-resident small DFTs for sm_100a.
//
// The r2iq kernels decompose every transform on the path (the 1024-point complex FFT behind the
// 2048-point real forward FFT of Core/fft_mt_r2iq_impl.hpp:88 and the mfft-point backward FFT of
// impl.hpp:98) into passes of radix-R DFTs (R = 2..32) that live entirely in one thread's
// registers.  All twiddles inside such a pass are powers of exp(-+2*pi*i/R) known at compile
// time, so they are generated here as constexpr double values (Taylor series after quadrant
// reduction; no runtime sincos, no fast-math) and folded into FFMA immediates by ptxas.
//
// Butterfly forms are chosen for instruction count on the FP32 pipe, the co-roof of this path:
//   * trivial twiddles (1, -+i) cost 4 FADD;
//   * (+-1 +- i)/sqrt(2) twiddles cost 2 FADD + 4 FFMA;
//   * general twiddles cost 6 FFMA: a' = a + w*b as two chained FFMA per component, and
//     b' = 2a - a' as one FFMA per component instead of recomputing a - w*b.
// A radix-32 pass is 388 FP32 instructions per thread (vs 5*N*log2(N)/... = 800 flops nominal).
#pragma once
#include <cuda_runtime.h>
#include <utility>

namespace r2iq {

// ---------------------------------------------------------------------------------------------
// constexpr trigonometry (double), exact at multiples of pi/2 by construction
// ---------------------------------------------------------------------------------------------
constexpr double kPi = 3.14159265358979323846264338327950288;

__host__ __device__ constexpr double cx_sin_small(double x) {   // |x| <= pi/4
    double x2 = x * x;
    double term = x, sum = x;
    for (int i = 1; i <= 11; i++) {
        term *= -x2 / ((2 * i) * (2 * i + 1));
        sum += term;
    }
    return sum;
}
__host__ __device__ constexpr double cx_cos_small(double x) {   // |x| <= pi/4
    double x2 = x * x;
    double term = 1.0, sum = 1.0;
    for (int i = 1; i <= 11; i++) {
        term *= -x2 / ((2 * i - 1) * (2 * i));
        sum += term;
    }
    return sum;
}
// cos / sin of 2*pi*k/n for integers, n a power of two (any k)
__host__ __device__ constexpr double cx_cos2pi(long k, long n) {
    k %= n;
    if (k < 0) k += n;
    // octant reduction on integers: angle = 2*pi*k/n; work with e = 8*k/n in eighths of a turn
    // quadrant q and remainder r in (-n/8, n/8] measured in units of 2*pi/n
    long q = (4 * k + n / 2) / n;         // nearest quarter turn
    long r = k * 4 - q * n;               // 4*(k - q*n/4), i.e. remainder * 4
    double x = 2.0 * kPi * (double)r / (4.0 * (double)n);
    q &= 3;
    return q == 0 ? cx_cos_small(x) : q == 1 ? -cx_sin_small(x) : q == 2 ? -cx_cos_small(x) : cx_sin_small(x);
}
__host__ __device__ constexpr double cx_sin2pi(long k, long n) {
    k %= n;
    if (k < 0) k += n;
    long q = (4 * k + n / 2) / n;
    long r = k * 4 - q * n;
    double x = 2.0 * kPi * (double)r / (4.0 * (double)n);
    q &= 3;
    return q == 0 ? cx_sin_small(x) : q == 1 ? cx_cos_small(x) : q == 2 ? -cx_sin_small(x) : -cx_cos_small(x);
}

__host__ __device__ constexpr int bitrev(int v, int bits) {
    int r = 0;
    for (int b = 0; b < bits; b++)
        if (v & (1 << b)) r |= 1 << (bits - 1 - b);
    return r;
}
__host__ __device__ constexpr int ilog2(int n) {
    int l = 0;
    while ((1 << l) < n) l++;
    return l;
}

// ---------------------------------------------------------------------------------------------
// complex helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul_conj(float2 a, float2 b) {   // a * conj(b)
    return make_float2(fmaf(a.x, b.x, a.y * b.y), fmaf(a.y, b.x, -a.x * b.y));
}

// ---------------------------------------------------------------------------------------------
// radix-2 DIT butterfly with compile-time twiddle w = exp(-+2*pi*i*J/LEN)  (INV: + sign)
//   a' = a + w*b ; b' = a - w*b
// ---------------------------------------------------------------------------------------------
template <int J, int LEN, bool INV>
__device__ __forceinline__ void bfly(float2& a, float2& b) {
    constexpr float kInvSqrt2 = 0.70710678118654752440f;
    if constexpr (J == 0) {
        float2 t = b;
        b = make_float2(a.x - t.x, a.y - t.y);
        a = make_float2(a.x + t.x, a.y + t.y);
    } else if constexpr (4 * J == LEN) {
        // w = -i (forward) / +i (inverse):  w*b = (b.y, -b.x) / (-b.y, b.x)
        float2 t = INV ? make_float2(-b.y, b.x) : make_float2(b.y, -b.x);
        b = make_float2(a.x - t.x, a.y - t.y);
        a = make_float2(a.x + t.x, a.y + t.y);
    } else if constexpr (8 * J == LEN) {
        // w = (1 -+ i)/sqrt2 : forward w*b = c*(b.x+b.y, b.y-b.x) ; inverse c*(b.x-b.y, b.x+b.y)
        float tx = INV ? (b.x - b.y) : (b.x + b.y);
        float ty = INV ? (b.x + b.y) : (b.y - b.x);
        b = make_float2(fmaf(-kInvSqrt2, tx, a.x), fmaf(-kInvSqrt2, ty, a.y));
        a = make_float2(fmaf(kInvSqrt2, tx, a.x), fmaf(kInvSqrt2, ty, a.y));
    } else if constexpr (8 * J == 3 * LEN) {
        // w = (-1 -+ i)/sqrt2 : forward w*b = c*(b.y-b.x, -(b.x+b.y)) ; inverse c*(-(b.x+b.y), b.x-b.y)
        float tx = INV ? (b.x + b.y) : (b.y - b.x);      // inverse carries an extra minus, applied below
        float ty = INV ? (b.x - b.y) : (b.x + b.y);      // forward carries an extra minus, applied below
        if constexpr (INV) {
            b = make_float2(fmaf(kInvSqrt2, tx, a.x), fmaf(-kInvSqrt2, ty, a.y));
            a = make_float2(fmaf(-kInvSqrt2, tx, a.x), fmaf(kInvSqrt2, ty, a.y));
        } else {
            b = make_float2(fmaf(-kInvSqrt2, tx, a.x), fmaf(kInvSqrt2, ty, a.y));
            a = make_float2(fmaf(kInvSqrt2, tx, a.x), fmaf(-kInvSqrt2, ty, a.y));
        }
    } else {
        constexpr float wr = (float)cx_cos2pi(J, LEN);
        constexpr float wi = (float)(INV ? cx_sin2pi(J, LEN) : -cx_sin2pi(J, LEN));
        // a' = a + w*b, two chained FFMA per component
        float ax = fmaf(wr, b.x, fmaf(-wi, b.y, a.x));
        float ay = fmaf(wr, b.y, fmaf(wi, b.x, a.y));
        // b' = 2a - a'
        b = make_float2(fmaf(2.0f, a.x, -ax), fmaf(2.0f, a.y, -ay));
        a = make_float2(ax, ay);
    }
}

template <int N, int LEN, bool INV, int... T>
__device__ __forceinline__ void dit_stage(float2 (&v)[N], std::integer_sequence<int, T...>) {
    constexpr int half = LEN / 2;
    (bfly<(T % half), LEN, INV>(v[(T / half) * LEN + (T % half)], v[(T / half) * LEN + (T % half) + half]), ...);
}

// In-place radix-2 decimation-in-time DFT of size N over a register array.
// INPUT must be placed in bit-reversed order (v[bitrev(n)] = x[n]); OUTPUT is in natural order.
// INV selects exp(+2*pi*i*nk/N); no normalisation in either direction (FFTW convention).
template <int N, bool INV, int LEN = 2>
__device__ __forceinline__ void dit(float2 (&v)[N]) {
    if constexpr (N >= 2) {
        dit_stage<N, LEN, INV>(v, std::make_integer_sequence<int, N / 2>{});
        if constexpr (LEN < N) dit<N, INV, LEN * 2>(v);
    }
}

}  // namespace r2iq

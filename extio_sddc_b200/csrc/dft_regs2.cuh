// dft_regs2.cuh -- the register-resident DFTs of dft_regs.cuh on PACKED fp32x2 lanes (sm_100a).
//
// Blackwell adds packed single-precision instructions (PTX add/sub/mul/fma.rn.f32x2, SASS
// FADD2/FMUL2/FFMA2).  Measured on B200 (tools/ubench/fp32_pipes.cu): a warp-wide FFMA2 retires
// 64 FMAs and occupies the FMA pipe for two cycles, so the peak FLOP rate equals scalar FFMA but
// it costs half the ISSUE slots -- the other slot is free for the loads, stores, shuffles and
// integer work that otherwise compete with the butterflies.  The r2iq kernels therefore run TWO
// independent frames per warp side by side: every value is a pair (frame A, frame B) in one
// 64-bit register, every butterfly one packed instruction, and all twiddles (compile-time
// immediates or per-lane scalars broadcast by the instruction's .F32 operand form) are shared.
//
// Negation is written as unpack/negate/repack; ptxas folds it into the operand's negate modifier
// (verified in SASS: FFMA2 R6, R4.F32x2.HI_LO, 2, -R6.F32x2.HI_LO), so it is free.
#pragma once
#include <cuda_runtime.h>
#include <utility>

#include "dft_regs.cuh"

namespace r2iq {

typedef unsigned long long u64;

__device__ __forceinline__ u64 pk(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 bc(float s) { return pk(s, s); }   // broadcast a scalar to both frames
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
    u64 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 sub2(u64 a, u64 b) {
    u64 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
    u64 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {   // a*b + c
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 neg2(u64 a) {
    float lo, hi;
    upk(a, lo, hi);
    return pk(-lo, -hi);
}

// complex value of two frames: re = (A.re, B.re), im = (A.im, B.im)
struct c2 {
    u64 re, im;
};

// v * w for a per-lane scalar complex w shared by both frames (4 packed ops)
__device__ __forceinline__ c2 cmul2s(c2 v, float2 w) {
    c2 r;
    r.re = fma2(bc(w.x), v.re, mul2(bc(-w.y), v.im));
    r.im = fma2(bc(w.x), v.im, mul2(bc(w.y), v.re));
    return r;
}

// v[k] *= w^k, k = 1..31, for a per-lane complex w shared by both frames -- the inter-pass twiddles of a
// 32 x 32 decomposition.  `tab` points at this lane's column of a [32][32] table whose row k holds w^k; only
// rows 1, 2, 16 and 17 are loaded (4 instead of 31 loads: the L1 data pipe, not the FMA pipe, is the busier
// unit, and the table rows no longer evict the filter tables from L1).  The other powers come from the packed
// recurrence (w^(2j+2), w^(2j+3)) = (w^(2j), w^(2j+1)) * w^2, seven steps from each loaded pair, so no power is
// more than 7 roundings away from a table value.
__device__ __forceinline__ void twiddle_powers32(c2 (&v)[32], const float2* tab) {
    const float2 w1 = __ldg(tab + 32 * 1);
    const float2 w2 = __ldg(tab + 32 * 2);
    const float2 w16 = __ldg(tab + 32 * 16);
    const float2 w17 = __ldg(tab + 32 * 17);
    const u64 cr = bc(w2.x), ci = bc(w2.y), cni = bc(-w2.y);
#pragma unroll
    for (int half = 0; half < 2; half++) {
        u64 pr = half ? pk(w16.x, w17.x) : pk(1.0f, w1.x);   // (re w^(2j), re w^(2j+1))
        u64 pi = half ? pk(w16.y, w17.y) : pk(0.0f, w1.y);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int k = 16 * half + 2 * j;
            float r0, r1, i0, i1;
            upk(pr, r0, r1);
            upk(pi, i0, i1);
            if (k > 0) v[k] = cmul2s(v[k], make_float2(r0, i0));
            v[k + 1] = cmul2s(v[k + 1], make_float2(r1, i1));
            if (j < 7) {
                const u64 nr = fma2(cr, pr, mul2(cni, pi));
                pi = fma2(cr, pi, mul2(ci, pr));
                pr = nr;
            }
        }
    }
}

// radix-2 DIT butterfly, compile-time twiddle exp(-+2*pi*i*J/LEN), packed version of bfly<>
template <int J, int LEN, bool INV>
__device__ __forceinline__ void bfly2(c2& a, c2& b) {
    constexpr float kInvSqrt2 = 0.70710678118654752440f;
    if constexpr (J == 0) {
        c2 t = b;
        b.re = sub2(a.re, t.re);
        b.im = sub2(a.im, t.im);
        a.re = add2(a.re, t.re);
        a.im = add2(a.im, t.im);
    } else if constexpr (4 * J == LEN) {
        // forward w = -i: w*b = (b.im, -b.re); inverse w = +i: w*b = (-b.im, b.re)
        c2 t = b;
        if constexpr (INV) {
            b.re = add2(a.re, t.im);
            b.im = sub2(a.im, t.re);
            a.re = sub2(a.re, t.im);
            a.im = add2(a.im, t.re);
        } else {
            b.re = sub2(a.re, t.im);
            b.im = add2(a.im, t.re);
            a.re = add2(a.re, t.im);
            a.im = sub2(a.im, t.re);
        }
    } else if constexpr (8 * J == LEN) {
        u64 tx = INV ? sub2(b.re, b.im) : add2(b.re, b.im);
        u64 ty = INV ? add2(b.re, b.im) : sub2(b.im, b.re);
        b.re = fma2(bc(-kInvSqrt2), tx, a.re);
        b.im = fma2(bc(-kInvSqrt2), ty, a.im);
        a.re = fma2(bc(kInvSqrt2), tx, a.re);
        a.im = fma2(bc(kInvSqrt2), ty, a.im);
    } else if constexpr (8 * J == 3 * LEN) {
        u64 tx = INV ? add2(b.re, b.im) : sub2(b.im, b.re);
        u64 ty = INV ? sub2(b.re, b.im) : add2(b.re, b.im);
        if constexpr (INV) {
            b.re = fma2(bc(kInvSqrt2), tx, a.re);
            b.im = fma2(bc(-kInvSqrt2), ty, a.im);
            a.re = fma2(bc(-kInvSqrt2), tx, a.re);
            a.im = fma2(bc(kInvSqrt2), ty, a.im);
        } else {
            b.re = fma2(bc(-kInvSqrt2), tx, a.re);
            b.im = fma2(bc(kInvSqrt2), ty, a.im);
            a.re = fma2(bc(kInvSqrt2), tx, a.re);
            a.im = fma2(bc(-kInvSqrt2), ty, a.im);
        }
    } else {
        constexpr float wr = (float)cx_cos2pi(J, LEN);
        constexpr float wi = (float)(INV ? cx_sin2pi(J, LEN) : -cx_sin2pi(J, LEN));
        u64 ax = fma2(bc(wr), b.re, fma2(bc(-wi), b.im, a.re));
        u64 ay = fma2(bc(wr), b.im, fma2(bc(wi), b.re, a.im));
        b.re = fma2(bc(2.0f), a.re, neg2(ax));     // b' = 2a - a'
        b.im = fma2(bc(2.0f), a.im, neg2(ay));
        a.re = ax;
        a.im = ay;
    }
}

template <int N, int LEN, bool INV, int... T>
__device__ __forceinline__ void dit2_stage(c2 (&v)[N], std::integer_sequence<int, T...>) {
    constexpr int half = LEN / 2;
    (bfly2<(T % half), LEN, INV>(v[(T / half) * LEN + (T % half)], v[(T / half) * LEN + (T % half) + half]), ...);
}

// packed counterpart of dit<N, INV>: bit-reversed input order, natural output order, unnormalised
template <int N, bool INV, int LEN = 2>
__device__ __forceinline__ void dit2(c2 (&v)[N]) {
    if constexpr (N >= 2) {
        dit2_stage<N, LEN, INV>(v, std::make_integer_sequence<int, N / 2>{});
        if constexpr (LEN < N) dit2<N, INV, LEN * 2>(v);
    }
}

// dit2 on inputs that still lack their per-register twiddles: v[i] stands for v[i] * w(i), w(0) = 1.  The twiddles
// are folded into the first stage (its butterflies are the trivial a +- b):  a' = w_a a + w_b b  (4 + 4 packed
// instructions; 4 when w_a = 1),  b' = 2 w_a a - a'  (2)  -- 10 per butterfly instead of 4 + 4 for the two products
// and 4 for the sums: a 32-point pass with its inter-pass twiddles costs 480 instead of 512 packed instructions.
// `w(i)` is any callable returning the float2 twiddle of register i (a table load, shared by both frames).
// kFirstOne = false: w(0) is a real twiddle too (the rotated backward transform of d = 0: 4 more instructions).
template <int N, bool INV, bool kFirstOne = true, class Tw>
__device__ __forceinline__ void dit2_tw(c2 (&v)[N], Tw w) {
    static_assert(N >= 4, "first stage folded into a radix-2 pass of at least two stages");
#pragma unroll
    for (int i = 0; i < N / 2; i++) {
        c2& a = v[2 * i];
        c2& b = v[2 * i + 1];
        c2 ta = a;
        if (i || !kFirstOne) ta = cmul2s(a, w(2 * i));
        const float2 wb = w(2 * i + 1);
        const u64 ax = fma2(bc(wb.x), b.re, fma2(bc(-wb.y), b.im, ta.re));
        const u64 ay = fma2(bc(wb.x), b.im, fma2(bc(wb.y), b.re, ta.im));
        b.re = fma2(bc(2.0f), ta.re, neg2(ax));
        b.im = fma2(bc(2.0f), ta.im, neg2(ay));
        a.re = ax;
        a.im = ay;
    }
    dit2<N, INV, 4>(v);
}

// dit2 whose first stage (the trivial a +- b butterflies) is written as fma(+-one, b, a) with a run-time packed 1.0:
// the same bits as the additions (1.0 * b is exact) on the same pipe, but every value of the pass now depends on
// `one`.  The FMA-token kernels (r2iq_tok.cu) load `one` from shared memory right after they acquire the token, which
// keeps ptxas from hoisting the whole register pass above the barrier (it does: only memory operations are ordered
// against BAR.SYNC).
template <int N, bool INV>
__device__ __forceinline__ void dit2_pinned(c2 (&v)[N], u64 one) {
    if constexpr (N >= 2) {
        const u64 mone = neg2(one);
#pragma unroll
        for (int i = 0; i < N / 2; i++) {
            c2& a = v[2 * i];
            c2& b = v[2 * i + 1];
            const c2 t = b;
            b.re = fma2(mone, t.re, a.re);
            b.im = fma2(mone, t.im, a.im);
            a.re = fma2(one, t.re, a.re);
            a.im = fma2(one, t.im, a.im);
        }
        if constexpr (N > 2) dit2<N, INV, 4>(v);
    }
}

}  // namespace r2iq

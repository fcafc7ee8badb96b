// 400-point real DFT of one Hann-windowed frame, split 16 x 25 so that every butterfly is
// register-resident and every twiddle inside a sub-transform is a compile-time constant.
//
// Replaces the ATen `stft` + `abs().pow(2)` call sites of reference src/audio.rs:236-247.
//
//   n = 16 m + t   (t = 0..15, m = 0..24)        k = q + 25 s   (q = 0..24, s = 0..15)
//   W400^{nk} = W25^{mq} * W400^{tq} * W16^{ts}
//
//   stage A (item t):  Y_t[q]  = sum_m xw[16m+t] W25^{mq}          real 25-point DFT, q = 0..12 kept
//                      Yh_t[q] = Y_t[q] * W400^{tq}
//   stage B (item q):  X[q+25s] = sum_t Yh_t[q] W16^{ts}           complex 16-point DFT
//
// For q = 1..12, s = 0..7 gives bins q+25s (<= 187) and s = 8..15 gives the conjugates of bins
// 400-(q+25s) = (25-q)+25(15-s) (13..199); q = 0 gives bins 25s, s = 0..8.  12*16 + 9 = 201 bins.
//
// Every routine is a template over the real type T.  T = float is one frame per thread (also what the host
// test, tests/test_fft_host.py, compiles with g++ and checks against a float64 DFT).  T = f32x2 (device only)
// is TWO frames per thread on Blackwell's packed FP32 pipe: add/mul/fma.rn.f32x2 -> FADD2 / FMUL2 / FFMA2,
// one issue slot for both frames, compile-time twiddles folded into the instruction as broadcast immediates.
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define WRS_HD __host__ __device__ __forceinline__
#define WRS_D __device__ __forceinline__
#else
#define WRS_HD inline
#endif

namespace wrs {

template <class T> struct cx { T re, im; };
typedef cx<float> cf;

// ---- the arithmetic the butterflies are written in: cmul_(c, a) = c*a, cfma_(c, a, b) = c*a + b with a
// compile-time constant c; fma_(a, b, c) = a*b + c with run-time operands
WRS_HD float add_(float a, float b) { return a + b; }
WRS_HD float sub_(float a, float b) { return a - b; }
WRS_HD float mul_(float a, float b) { return a * b; }
WRS_HD float fma_(float a, float b, float c) { return fmaf(a, b, c); }
WRS_HD float cmul_(float c, float a) { return c * a; }
WRS_HD float cfma_(float c, float a, float b) { return fmaf(c, a, b); }

#if defined(__CUDACC__)
// two frames side by side in one 64-bit register pair
struct f32x2 { unsigned long long v; };
WRS_D f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi)); return r; }
WRS_D void unpack2(f32x2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v)); }
WRS_D f32x2 add_(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
WRS_D f32x2 sub_(f32x2 a, f32x2 b) { f32x2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
WRS_D f32x2 mul_(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
WRS_D f32x2 fma_(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r;
}
WRS_D f32x2 cmul_(float c, f32x2 a) { return mul_(a, pack2(c, c)); }
WRS_D f32x2 cfma_(float c, f32x2 a, f32x2 b) { return fma_(a, pack2(c, c), b); }
#endif

// (a.re + i a.im)(b.re + i b.im), both run-time
template <class T> WRS_HD cx<T> cmul(cx<T> a, cx<T> b) {
    return cx<T>{sub_(mul_(a.re, b.re), mul_(a.im, b.im)), fma_(a.re, b.im, mul_(a.im, b.re))};
}
// conj(a) * (b.re + i b.im), both run-time
template <class T> WRS_HD cx<T> cmul_conj(cx<T> a, cx<T> b) {
    return cx<T>{fma_(a.im, b.im, mul_(a.re, b.re)), sub_(mul_(a.re, b.im), mul_(a.im, b.re))};
}
// a * (wr + i wi) with a compile-time twiddle
template <class T> WRS_HD cx<T> cmulc(cx<T> a, float wr, float wi) {
    return cx<T>{cfma_(-wi, a.im, cmul_(wr, a.re)), cfma_(wi, a.re, cmul_(wr, a.im))};
}

// cos/sin(2 pi j / 5)
#define WRS_C1 0.30901699437494745f
#define WRS_C2 (-0.80901699437494745f)
#define WRS_S1 0.95105651629515353f
#define WRS_S2 0.58778525229247314f

// real 5-point DFT: X0 (real), X1, X2 (X3 = conj X2, X4 = conj X1)
template <class T> WRS_HD void rdft5(T a0, T a1, T a2, T a3, T a4, T& x0, cx<T>& x1, cx<T>& x2) {
    const T sa = add_(a1, a4), da = sub_(a1, a4), sb = add_(a2, a3), db = sub_(a2, a3);
    x0 = add_(a0, add_(sa, sb));
    x1.re = cfma_(WRS_C2, sb, cfma_(WRS_C1, sa, a0));
    x2.re = cfma_(WRS_C1, sb, cfma_(WRS_C2, sa, a0));
    x1.im = cfma_(-WRS_S2, db, cmul_(-WRS_S1, da));      // -(S1 da + S2 db)
    x2.im = cfma_(WRS_S1, db, cmul_(-WRS_S2, da));       // -(S2 da - S1 db)
}

// complex 5-point DFT, forward (W = exp(-2 pi i / 5))
template <class T>
WRS_HD void cdft5(cx<T> a0, cx<T> a1, cx<T> a2, cx<T> a3, cx<T> a4, cx<T>& x0, cx<T>& x1, cx<T>& x2, cx<T>& x3, cx<T>& x4) {
    const cx<T> sa{add_(a1.re, a4.re), add_(a1.im, a4.im)}, da{sub_(a1.re, a4.re), sub_(a1.im, a4.im)};
    const cx<T> sb{add_(a2.re, a3.re), add_(a2.im, a3.im)}, db{sub_(a2.re, a3.re), sub_(a2.im, a3.im)};
    x0 = cx<T>{add_(a0.re, add_(sa.re, sb.re)), add_(a0.im, add_(sa.im, sb.im))};
    const cx<T> t1{cfma_(WRS_C2, sb.re, cfma_(WRS_C1, sa.re, a0.re)), cfma_(WRS_C2, sb.im, cfma_(WRS_C1, sa.im, a0.im))};
    const cx<T> t2{cfma_(WRS_C1, sb.re, cfma_(WRS_C2, sa.re, a0.re)), cfma_(WRS_C1, sb.im, cfma_(WRS_C2, sa.im, a0.im))};
    const cx<T> u1{cfma_(WRS_S2, db.re, cmul_(WRS_S1, da.re)), cfma_(WRS_S2, db.im, cmul_(WRS_S1, da.im))};
    const cx<T> u2{cfma_(-WRS_S1, db.re, cmul_(WRS_S2, da.re)), cfma_(-WRS_S1, db.im, cmul_(WRS_S2, da.im))};
    // X1 = t1 - i u1, X4 = t1 + i u1, X2 = t2 - i u2, X3 = t2 + i u2
    x1 = cx<T>{add_(t1.re, u1.im), sub_(t1.im, u1.re)};
    x4 = cx<T>{sub_(t1.re, u1.im), add_(t1.im, u1.re)};
    x2 = cx<T>{add_(t2.re, u2.im), sub_(t2.im, u2.re)};
    x3 = cx<T>{sub_(t2.re, u2.im), add_(t2.im, u2.re)};
}

WRS_HD float neg_(float a) { return -a; }
WRS_HD void zero_(float& a) { a = 0.0f; }
#if defined(__CUDACC__)
WRS_D f32x2 neg_(f32x2 a) { return cmul_(-1.0f, a); }
WRS_D void zero_(f32x2& a) { a.v = 0ull; }
#endif

// W25^p = exp(-2 pi i p / 25), p = 1..8: (cos, -sin)
#define WRS_W25_1R 0.96858316112863108f
#define WRS_W25_1I (-0.24868988716485479f)
#define WRS_W25_2R 0.87630668004386358f
#define WRS_W25_2I (-0.48175367410171532f)
#define WRS_W25_3R 0.72896862742141155f
#define WRS_W25_3I (-0.68454710592868873f)
#define WRS_W25_4R 0.53582679497899666f
#define WRS_W25_4I (-0.84432792550201508f)
#define WRS_W25_6R 0.06279051952931337f
#define WRS_W25_6I (-0.99802672842827156f)
#define WRS_W25_8R (-0.42577929156507272f)
#define WRS_W25_8I (-0.90482705246601958f)

// Real 25-point DFT of v[0..24]; returns Y[0..12] (Y[0] real, its .im = 0).
// m = 5a + b: first a real 5-point DFT over a for each residue b, then (after W25^{b q0}) a 5-point
// DFT over b for q0 = 0 (real), 1, 2; q0 = 3, 4 are the conjugates of outputs of q0 = 2, 1.
// RAW_CONJ: Y[3], Y[4], Y[8], Y[9] are left conjugated (rdft25_conj_slot(q) tells which); the caller folds the
// conjugation into the twiddle multiplication that follows (cmul_conj), which saves the four negations.
WRS_HD bool rdft25_conj_slot(int q) { return q == 3 || q == 4 || q == 8 || q == 9; }
template <bool RAW_CONJ = false, class T> WRS_HD void rdft25(const T (&v)[25], cx<T> (&y)[13]) {
    T u0[5];
    cx<T> u1[5], u2[5];
#pragma unroll
    for (int b = 0; b < 5; ++b) rdft5(v[b], v[5 + b], v[10 + b], v[15 + b], v[20 + b], u0[b], u1[b], u2[b]);
    // q0 = 0
    {
        T y0; cx<T> y5, y10;
        rdft5(u0[0], u0[1], u0[2], u0[3], u0[4], y0, y5, y10);
        y[0].re = y0; zero_(y[0].im); y[5] = y5; y[10] = y10;
    }
    // q0 = 1 : outputs Y[1], Y[6], Y[11], Y[16] = conj Y[9], Y[21] = conj Y[4]
    {
        cx<T> o0, o1, o2, o3, o4;
        cdft5(u1[0], cmulc(u1[1], WRS_W25_1R, WRS_W25_1I), cmulc(u1[2], WRS_W25_2R, WRS_W25_2I),
              cmulc(u1[3], WRS_W25_3R, WRS_W25_3I), cmulc(u1[4], WRS_W25_4R, WRS_W25_4I), o0, o1, o2, o3, o4);
        y[1] = o0; y[6] = o1; y[11] = o2;
        if (RAW_CONJ) { y[9] = o3; y[4] = o4; } else { y[9] = cx<T>{o3.re, neg_(o3.im)}; y[4] = cx<T>{o4.re, neg_(o4.im)}; }
    }
    // q0 = 2 : outputs Y[2], Y[7], Y[12], Y[17] = conj Y[8], Y[22] = conj Y[3]
    {
        cx<T> o0, o1, o2, o3, o4;
        cdft5(u2[0], cmulc(u2[1], WRS_W25_2R, WRS_W25_2I), cmulc(u2[2], WRS_W25_4R, WRS_W25_4I),
              cmulc(u2[3], WRS_W25_6R, WRS_W25_6I), cmulc(u2[4], WRS_W25_8R, WRS_W25_8I), o0, o1, o2, o3, o4);
        y[2] = o0; y[7] = o1; y[12] = o2;
        if (RAW_CONJ) { y[8] = o3; y[3] = o4; } else { y[8] = cx<T>{o3.re, neg_(o3.im)}; y[3] = cx<T>{o4.re, neg_(o4.im)}; }
    }
}

// radix-4 forward butterfly (W4 = -i)
template <class T> WRS_HD void dft4(cx<T> a, cx<T> b, cx<T> c, cx<T> d, cx<T>& x0, cx<T>& x1, cx<T>& x2, cx<T>& x3) {
    const cx<T> s0{add_(a.re, c.re), add_(a.im, c.im)}, d0{sub_(a.re, c.re), sub_(a.im, c.im)};
    const cx<T> s1{add_(b.re, d.re), add_(b.im, d.im)}, d1{sub_(b.re, d.re), sub_(b.im, d.im)};
    x0 = cx<T>{add_(s0.re, s1.re), add_(s0.im, s1.im)};
    x2 = cx<T>{sub_(s0.re, s1.re), sub_(s0.im, s1.im)};
    x1 = cx<T>{add_(d0.re, d1.im), sub_(d0.im, d1.re)};   // d0 - i d1
    x3 = cx<T>{sub_(d0.re, d1.im), add_(d0.im, d1.re)};   // d0 + i d1
}

// W16^p = exp(-2 pi i p / 16)
#define WRS_COS8 0.92387953251128674f   // cos(pi/8)
#define WRS_SIN8 0.38268343236508977f   // sin(pi/8)
#define WRS_R2 0.70710678118654752f

template <class T> WRS_HD cx<T> mul_w16(cx<T> a, int p) {   // p is a compile-time constant after unrolling
    switch (p) {
        case 1: return cmulc(a, WRS_COS8, -WRS_SIN8);
        case 2: return cx<T>{cmul_(WRS_R2, add_(a.re, a.im)), cmul_(WRS_R2, sub_(a.im, a.re))};
        case 3: return cmulc(a, WRS_SIN8, -WRS_COS8);
        case 4: return cx<T>{a.im, neg_(a.re)};
        case 6: return cx<T>{cmul_(WRS_R2, sub_(a.im, a.re)), cmul_(-WRS_R2, add_(a.re, a.im))};
        case 9: return cmulc(a, -WRS_COS8, WRS_SIN8);
        default: return a;
    }
}

// Complex 16-point forward DFT, in place: x[s] = sum_t z[t] W16^{ts}.
// t = t0 + 4 t1, s = s1 + 4 s0:  W16^{ts} = W4^{t1 s1} * W16^{t0 s1} * W4^{t0 s0}
template <class T> WRS_HD void cdft16(cx<T> (&z)[16]) {
    cx<T> a[4][4];   // a[t0][s1]
#pragma unroll
    for (int t0 = 0; t0 < 4; ++t0) {
        dft4(z[t0], z[t0 + 4], z[t0 + 8], z[t0 + 12], a[t0][0], a[t0][1], a[t0][2], a[t0][3]);
#pragma unroll
        for (int s1 = 1; s1 < 4; ++s1) a[t0][s1] = mul_w16(a[t0][s1], t0 * s1);
    }
#pragma unroll
    for (int s1 = 0; s1 < 4; ++s1)
        dft4(a[0][s1], a[1][s1], a[2][s1], a[3][s1], z[s1], z[s1 + 4], z[s1 + 8], z[s1 + 12]);
}

// |z|^2
template <class T> WRS_HD T power(cx<T> z) { return fma_(z.re, z.re, mul_(z.im, z.im)); }

// bin index produced by stage-B item q at output slot s (power is conjugation-invariant)
WRS_HD int stageb_bin(int q, int s) {
    const int k = q + 25 * s;
    return k <= 200 ? k : 400 - k;
}

}  // namespace wrs

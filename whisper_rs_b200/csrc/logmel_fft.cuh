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
// The same source compiles for the device (nvcc) and for the host (g++, tests/test_fft_host.py) so
// the index algebra is checked on CPU against a float64 DFT before it ever reaches a GPU.
#pragma once

#if defined(__CUDACC__)
#define WRS_HD __host__ __device__ __forceinline__
#else
#define WRS_HD inline
#endif

namespace wrs {

struct cf { float re, im; };

WRS_HD cf cmul(cf a, cf b) { return cf{a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }

// cos/sin(2 pi j / 5)
#define WRS_C1 0.30901699437494745f
#define WRS_C2 (-0.80901699437494745f)
#define WRS_S1 0.95105651629515353f
#define WRS_S2 0.58778525229247314f

// real 5-point DFT: X0 (real), X1, X2 (X3 = conj X2, X4 = conj X1)
WRS_HD void rdft5(float a0, float a1, float a2, float a3, float a4, float& x0, cf& x1, cf& x2) {
    const float sa = a1 + a4, da = a1 - a4, sb = a2 + a3, db = a2 - a3;
    x0 = a0 + (sa + sb);
    x1.re = a0 + WRS_C1 * sa + WRS_C2 * sb;
    x2.re = a0 + WRS_C2 * sa + WRS_C1 * sb;
    x1.im = -(WRS_S1 * da + WRS_S2 * db);
    x2.im = -(WRS_S2 * da - WRS_S1 * db);
}

// complex 5-point DFT, forward (W = exp(-2 pi i / 5))
WRS_HD void cdft5(cf a0, cf a1, cf a2, cf a3, cf a4, cf& x0, cf& x1, cf& x2, cf& x3, cf& x4) {
    const cf sa{a1.re + a4.re, a1.im + a4.im}, da{a1.re - a4.re, a1.im - a4.im};
    const cf sb{a2.re + a3.re, a2.im + a3.im}, db{a2.re - a3.re, a2.im - a3.im};
    x0 = cf{a0.re + (sa.re + sb.re), a0.im + (sa.im + sb.im)};
    const cf t1{a0.re + WRS_C1 * sa.re + WRS_C2 * sb.re, a0.im + WRS_C1 * sa.im + WRS_C2 * sb.im};
    const cf t2{a0.re + WRS_C2 * sa.re + WRS_C1 * sb.re, a0.im + WRS_C2 * sa.im + WRS_C1 * sb.im};
    const cf u1{WRS_S1 * da.re + WRS_S2 * db.re, WRS_S1 * da.im + WRS_S2 * db.im};
    const cf u2{WRS_S2 * da.re - WRS_S1 * db.re, WRS_S2 * da.im - WRS_S1 * db.im};
    // X1 = t1 - i u1, X4 = t1 + i u1, X2 = t2 - i u2, X3 = t2 + i u2
    x1 = cf{t1.re + u1.im, t1.im - u1.re};
    x4 = cf{t1.re - u1.im, t1.im + u1.re};
    x2 = cf{t2.re + u2.im, t2.im - u2.re};
    x3 = cf{t2.re - u2.im, t2.im + u2.re};
}

// W25^p = exp(-2 pi i p / 25), p = 1..8 (literal constants: cos, -sin)
#define WRS_W25_1 cf{0.96858316112863108f, -0.24868988716485479f}
#define WRS_W25_2 cf{0.87630668004386358f, -0.48175367410171532f}
#define WRS_W25_3 cf{0.72896862742141155f, -0.68454710592868873f}
#define WRS_W25_4 cf{0.53582679497899666f, -0.84432792550201508f}
#define WRS_W25_6 cf{0.06279051952931337f, -0.99802672842827156f}
#define WRS_W25_8 cf{-0.42577929156507272f, -0.90482705246601958f}

// Real 25-point DFT of v[0..24]; returns Y[0..12] (Y[0] real, its .im = 0).
// m = 5a + b: first a real 5-point DFT over a for each residue b, then (after W25^{b q0}) a 5-point
// DFT over b for q0 = 0 (real), 1, 2; q0 = 3, 4 are the conjugates of outputs of q0 = 2, 1.
WRS_HD void rdft25(const float (&v)[25], cf (&y)[13]) {
    float u0[5];
    cf u1[5], u2[5];
#pragma unroll
    for (int b = 0; b < 5; ++b) rdft5(v[b], v[5 + b], v[10 + b], v[15 + b], v[20 + b], u0[b], u1[b], u2[b]);
    // q0 = 0
    {
        float y0; cf y5, y10;
        rdft5(u0[0], u0[1], u0[2], u0[3], u0[4], y0, y5, y10);
        y[0] = cf{y0, 0.0f}; y[5] = y5; y[10] = y10;
    }
    // q0 = 1 : outputs Y[1], Y[6], Y[11], Y[16] = conj Y[9], Y[21] = conj Y[4]
    {
        cf o0, o1, o2, o3, o4;
        cdft5(u1[0], cmul(u1[1], WRS_W25_1), cmul(u1[2], WRS_W25_2), cmul(u1[3], WRS_W25_3), cmul(u1[4], WRS_W25_4),
              o0, o1, o2, o3, o4);
        y[1] = o0; y[6] = o1; y[11] = o2; y[9] = cf{o3.re, -o3.im}; y[4] = cf{o4.re, -o4.im};
    }
    // q0 = 2 : outputs Y[2], Y[7], Y[12], Y[17] = conj Y[8], Y[22] = conj Y[3]
    {
        cf o0, o1, o2, o3, o4;
        cdft5(u2[0], cmul(u2[1], WRS_W25_2), cmul(u2[2], WRS_W25_4), cmul(u2[3], WRS_W25_6), cmul(u2[4], WRS_W25_8),
              o0, o1, o2, o3, o4);
        y[2] = o0; y[7] = o1; y[12] = o2; y[8] = cf{o3.re, -o3.im}; y[3] = cf{o4.re, -o4.im};
    }
}

// radix-4 forward butterfly (W4 = -i)
WRS_HD void dft4(cf a, cf b, cf c, cf d, cf& x0, cf& x1, cf& x2, cf& x3) {
    const cf s0{a.re + c.re, a.im + c.im}, d0{a.re - c.re, a.im - c.im};
    const cf s1{b.re + d.re, b.im + d.im}, d1{b.re - d.re, b.im - d.im};
    x0 = cf{s0.re + s1.re, s0.im + s1.im};
    x2 = cf{s0.re - s1.re, s0.im - s1.im};
    x1 = cf{d0.re + d1.im, d0.im - d1.re};   // d0 - i d1
    x3 = cf{d0.re - d1.im, d0.im + d1.re};   // d0 + i d1
}

// W16^p = exp(-2 pi i p / 16)
#define WRS_COS8 0.92387953251128674f   // cos(pi/8)
#define WRS_SIN8 0.38268343236508977f   // sin(pi/8)
#define WRS_R2 0.70710678118654752f

WRS_HD cf mul_w16(cf a, int p) {   // p is a compile-time constant after unrolling
    switch (p) {
        case 0: return a;
        case 1: return cmul(a, cf{WRS_COS8, -WRS_SIN8});
        case 2: return cf{WRS_R2 * (a.re + a.im), WRS_R2 * (a.im - a.re)};
        case 3: return cmul(a, cf{WRS_SIN8, -WRS_COS8});
        case 4: return cf{a.im, -a.re};
        case 6: return cf{WRS_R2 * (a.im - a.re), -WRS_R2 * (a.re + a.im)};
        case 9: return cmul(a, cf{-WRS_COS8, WRS_SIN8});
        default: return a;
    }
}

// Complex 16-point forward DFT, in place: x[s] = sum_t z[t] W16^{ts}.
// t = t0 + 4 t1, s = s1 + 4 s0:  W16^{ts} = W4^{t1 s1} * W16^{t0 s1} * W4^{t0 s0}
WRS_HD void cdft16(cf (&z)[16]) {
    cf a[4][4];   // a[t0][s1]
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

// bin index produced by stage-B item q at output slot s (power is conjugation-invariant)
WRS_HD int stageb_bin(int q, int s) {
    const int k = q + 25 * s;
    return k <= 200 ? k : 400 - k;
}

}  // namespace wrs

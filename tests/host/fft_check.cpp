// Host-side check of whisper_rs_b200/csrc/logmel_fft.cuh (the exact code the CUDA kernel runs per thread)
// against a float64 direct DFT.  Built and run by tests/test_fft_host.py with g++; no GPU involved.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../whisper_rs_b200/csrc/logmel_fft.cuh"

int main() {
    const double PI = 3.14159265358979323846;
    double worst_rel_to_max = 0, worst_rel = 0;
    srand(1234);
    for (int trial = 0; trial < 64; ++trial) {
        float xw[400];
        for (int n = 0; n < 400; ++n) {
            double w = 0.5 - 0.5 * cos(2 * PI * n / 400.0);
            double x;
            switch (trial % 4) {
                case 0: x = (rand() / (double)RAND_MAX - 0.5); break;
                case 1: x = sin(2 * PI * 440.0 * n / 16000.0 + trial); break;
                case 2: x = sin(2 * PI * (37.3 + trial) * n / 400.0) + 1e-3 * (rand() / (double)RAND_MAX - 0.5); break;
                default: x = (n == trial) ? 1.0 : 0.0; break;
            }
            xw[n] = (float)(x * w);
        }
        // stage A
        wrs::cf yh[16][13];
        for (int t = 0; t < 16; ++t) {
            float v[25];
            for (int m = 0; m < 25; ++m) v[m] = xw[16 * m + t];
            wrs::cf y[13];
            // even trials: plain outputs; odd trials: the form the kernel uses (four outputs left conjugated, the
            // conjugation folded into the twiddle multiplication)
            if (trial & 1) wrs::rdft25<true>(v, y); else wrs::rdft25(v, y);
            for (int q = 0; q < 13; ++q) {
                double a = -2 * PI * (double)(t * q) / 400.0;
                wrs::cf tw{(float)cos(a), (float)sin(a)};
                if ((trial & 1) && wrs::rdft25_conj_slot(q)) yh[t][q] = wrs::cmul_conj(y[q], tw);
                else yh[t][q] = (t == 0 || q == 0) ? y[q] : wrs::cmul(y[q], tw);
            }
        }
        // stage B
        std::vector<double> p(201, -1.0);
        for (int q = 0; q < 13; ++q) {
            wrs::cf z[16];
            for (int t = 0; t < 16; ++t) z[t] = yh[t][q];
            wrs::cdft16(z);
            const int ns = (q == 0) ? 9 : 16;
            for (int s = 0; s < ns; ++s) {
                int k = wrs::stageb_bin(q, s);
                if (p[k] >= 0) { printf("bin %d produced twice\n", k); return 2; }
                p[k] = (double)(z[s].re * z[s].re + z[s].im * z[s].im);
            }
        }
        // reference
        double pmax = 0;
        std::vector<double> pr(201);
        for (int k = 0; k <= 200; ++k) {
            double re = 0, im = 0;
            for (int n = 0; n < 400; ++n) {
                double a = -2 * PI * (double)((n * k) % 400) / 400.0;
                re += xw[n] * cos(a); im += xw[n] * sin(a);
            }
            pr[k] = re * re + im * im;
            if (pr[k] > pmax) pmax = pr[k];
        }
        for (int k = 0; k <= 200; ++k) {
            if (p[k] < 0) { printf("bin %d missing\n", k); return 2; }
            double e = fabs(p[k] - pr[k]);
            if (e / pmax > worst_rel_to_max) worst_rel_to_max = e / pmax;
            if (pr[k] > 1e-8 * pmax && e / pr[k] > worst_rel) worst_rel = e / pr[k];
        }
    }
    printf("worst |dp|/pmax = %.3e   worst |dp|/p (bins above max-80dB) = %.3e\n", worst_rel_to_max, worst_rel);
    return (worst_rel_to_max < 2e-6 && worst_rel < 5e-3) ? 0 : 1;
}

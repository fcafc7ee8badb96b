// Microbenchmark: FFMA vs packed FFMA2 (fma.rn.f32x2) issue/throughput on sm_100a, alone and mixed with LDS.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void fma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
template <int MODE>
__global__ void k(float* out, int iters, float s) {
    __shared__ float sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = s * i;
    __syncthreads();
    float a[8]; unsigned long long p[8];
    for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x + i; p[i] = ((unsigned long long)__float_as_uint(a[i]) << 32) | __float_as_uint(a[i] + 1); }
    const float m = s; const unsigned long long mm = ((unsigned long long)__float_as_uint(s) << 32) | __float_as_uint(s);
    float acc = 0;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], m, 0.5f);
        } else if (MODE == 1) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 8; ++i) fma2(p[i], mm, mm);
        } else if (MODE == 2) {   // 32 FFMA + 8 LDS
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], m, 0.5f);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc += sm[(threadIdx.x + 32 * i + it) & 1023];
        } else {                  // 32 FFMA2 (=64 FMA) + 8 LDS
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 8; ++i) fma2(p[i], mm, mm);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc += sm[(threadIdx.x + 32 * i + it) & 1023];
        }
    }
    float r = acc;
    for (int i = 0; i < 8; ++i) r += a[i] + __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32));
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    const int iters = 20000;
    const char* names[4] = {"FFMA", "FFMA2", "FFMA+LDS", "FFMA2+LDS"};
    for (int mode = 0; mode < 4; ++mode) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(out, iters, 0.999f);
            if (mode == 1) k<1><<<148 * 8, 256>>>(out, iters, 0.999f);
            if (mode == 2) k<2><<<148 * 8, 256>>>(out, iters, 0.999f);
            if (mode == 3) k<3><<<148 * 8, 256>>>(out, iters, 0.999f);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double fmas = (double)148 * 8 * 256 * iters * 32 * ((mode & 1) ? 2 : 1);
        printf("%-10s %8.3f ms  %7.2f TFMA/s  (%.1f FMA/clk/SM at 1.965 GHz)\n", names[mode], ms, fmas / ms / 1e9, fmas / (ms * 1e-3) / 148 / 1.965e9);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}

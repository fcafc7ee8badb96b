// Step-0 no-speech probability: reference src/decode/mod.rs:250-255
//     probs_at_sot = logits.i((.., sot_idx)).softmax(-1);  no_speech_probs = probs_at_sot[.., no_speech]
// i.e. one softmax over the UNFILTERED row per row (the filters run afterwards on the same row), reduced to a single
// probability.  HBM-bound: a row is read once (4 V bytes), one float per row is written.
//
// A CTA takes one eighth of a row into registers as 16-byte chunks that start on 16-byte boundaries of the row's memory
// (rows are only 4-byte aligned when V is odd: the chunk in front of id 0 and the one around id V - 1 are read id by id),
// every thread folds its <= 28 values into {max, sum exp(x - max)} on its own, pairs are combined by shuffles and one
// shared-memory stage (ONE CTA barrier), and the last CTA of the row (self-resetting counter) folds the eight pairs and
// writes exp(x[no_speech] - max) / sum.  avg_logprob (decode/mod.rs:389-397) is host arithmetic.
#include <algorithm>
#include <cmath>

#include <math_constants.h>

#include "wrs_internal.cuh"

struct wrs_decoder;

namespace wrs {

constexpr int kNsThreads = 256;
constexpr int kNsSplit = 8;                              // CTAs per row (4, 16, 32 and a one-wave 7 were measured: slower)
constexpr int kNsVec = 8;                                // 16-byte chunks per thread: covers V <= 8 * 256 * 8 * 4 - 3
constexpr float kNsLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ns_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// {m, s} <- {m, s} (+) {m2, s2}: max and sum of exp(x - max) of the union; exp(-inf - -inf) counts as 0
__device__ __forceinline__ void ns_combine(float& m, float& s, float m2, float s2) {
    const float M = fmaxf(m, m2);
    const float a = m == -CUDART_INF_F ? 0.0f : s * ns_ex2((m - M) * kNsLog2e);
    const float b = m2 == -CUDART_INF_F ? 0.0f : s2 * ns_ex2((m2 - M) * kNsLog2e);
    m = M; s = a + b;
}

__global__ void __launch_bounds__(kNsThreads) no_speech_kernel(const float* __restrict__ logits, int64_t row_stride, int n_rows,
                                                                int n_vocab, int no_speech, int split, float2* __restrict__ part,
                                                                int* __restrict__ arrive, float* __restrict__ out) {
    __shared__ float2 red[kNsThreads / 32];
    const int row = blockIdx.x / split, piece = blockIdx.x - row * split;
    const float* __restrict__ rp = logits + (int64_t)row * row_stride;
    const int mis = (int)((reinterpret_cast<uintptr_t>(rp) >> 2) & 3);       // ids in front of id 0 in its 16-byte chunk
    const int n_chunks = (n_vocab + mis + 3) >> 2;                            // chunk c holds ids 4c - mis .. 4c - mis + 3
    const int per = (n_chunks + split - 1) / split, c0 = piece * per, c1 = min(n_chunks, c0 + per);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float v[kNsVec][4];
#pragma unroll
    for (int k = 0; k < kNsVec; ++k) {
        const int c = c0 + threadIdx.x + k * kNsThreads, e = 4 * c - mis;
        if (c < c1 && e >= 0 && e + 3 < n_vocab) {
            const float4 t = __ldcs(reinterpret_cast<const float4*>(rp + e));
            v[k][0] = t.x; v[k][1] = t.y; v[k][2] = t.z; v[k][3] = t.w;
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) v[k][i] = (c < c1 && e + i >= 0 && e + i < n_vocab) ? __ldcs(rp + e + i) : -CUDART_INF_F;
        }
    }
    float m = -CUDART_INF_F;
#pragma unroll
    for (int k = 0; k < kNsVec; ++k) m = fmaxf(m, fmaxf(fmaxf(v[k][0], v[k][1]), fmaxf(v[k][2], v[k][3])));
    float s = 0.0f;
    if (m > -CUDART_INF_F) {
        const float nm = -m * kNsLog2e;
        float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
        for (int k = 0; k < kNsVec; ++k) {                                    // exp(-inf) = 0 for the padding; NaN stays NaN
            s0 += ns_ex2(fmaf(v[k][0], kNsLog2e, nm)) + ns_ex2(fmaf(v[k][1], kNsLog2e, nm));
            s1 += ns_ex2(fmaf(v[k][2], kNsLog2e, nm)) + ns_ex2(fmaf(v[k][3], kNsLog2e, nm));
        }
        s = s0 + s1;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ns_combine(m, s, __shfl_xor_sync(0xffffffffu, m, o), __shfl_xor_sync(0xffffffffu, s, o));
    if (lane == 0) red[warp] = make_float2(m, s);
    __syncthreads();
    if (warp != 0) return;
    {
        float2 q = lane < kNsThreads / 32 ? red[lane] : make_float2(-CUDART_INF_F, 0.0f);
        m = q.x; s = q.y;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) ns_combine(m, s, __shfl_xor_sync(0xffffffffu, m, o), __shfl_xor_sync(0xffffffffu, s, o));
    }
    int last = 0;
    if (lane == 0) {
        part[(int64_t)row * kNsSplit + piece] = make_float2(m, s);
        __threadfence();
        const int prev = atomicAdd(arrive + row, 1);
        last = prev == split - 1;
        if (last) { arrive[row] = 0; __threadfence(); }                      // self-reset for the next launch
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    // the row's last CTA: lane i folds pair i
    float2 q = lane < split ? __ldcg(part + (int64_t)row * kNsSplit + lane) : make_float2(-CUDART_INF_F, 0.0f);
    float M = q.x, S = q.y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ns_combine(M, S, __shfl_xor_sync(0xffffffffu, M, o), __shfl_xor_sync(0xffffffffu, S, o));
    if (lane == 0) out[row] = expf(rp[no_speech] - M) / S;
}

}  // namespace wrs

extern "C" {

// decoder internals this file needs (defined in decode.cu)
int wrs_decoder_scratch_(wrs_decoder* dec, wrs_ctx** ctx, int32_t* max_rows, int32_t* n_vocab, int32_t* no_speech,
                         void** pairs, int32_t** counters);

int wrs_no_speech_probs(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, float* out_probs,
                        wrs_stream_t stream) {
    using namespace wrs;
    if (!dec) return fail(nullptr, WRS_ERR_INVALID, "wrs_no_speech_probs: null decoder");
    wrs_ctx* ctx = nullptr;
    int32_t max_rows = 0, n_vocab = 0, no_speech = 0;
    void* pairs = nullptr;
    int32_t* counters = nullptr;
    wrs_decoder_scratch_(dec, &ctx, &max_rows, &n_vocab, &no_speech, &pairs, &counters);
    if (n_rows == 0) return WRS_OK;
    if (!logits || !out_probs) return fail(ctx, WRS_ERR_INVALID, "wrs_no_speech_probs: null buffer");
    if (n_rows < 0 || n_rows > max_rows || row_stride < n_vocab)
        return fail(ctx, WRS_ERR_INVALID, "wrs_no_speech_probs: need 0 <= n_rows <= max_rows and row_stride >= n_vocab");
    if (no_speech < 0 || no_speech >= n_vocab) return fail(ctx, WRS_ERR_INVALID, "wrs_no_speech_probs: cfg.no_speech is not a token id");
    if (n_vocab > kNsSplit * kNsThreads * kNsVec * 4 - 64) return fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_no_speech_probs: vocabulary above 65472");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    const int split = kNsSplit;
    no_speech_kernel<<<(unsigned)(n_rows * split), kNsThreads, 0, (cudaStream_t)stream>>>(
        logits, row_stride, (int)n_rows, n_vocab, no_speech, split, static_cast<float2*>(pairs), counters, out_probs);
    WRS_CUDA(ctx, cudaGetLastError());
    return WRS_OK;
}

int wrs_avg_logprob_host(const double* sum_logprobs_host, const int32_t* n_tokens_host, int32_t n, double* out_host) {
    if (n < 0 || (n > 0 && (!sum_logprobs_host || !n_tokens_host || !out_host)))
        return wrs::fail(nullptr, WRS_ERR_INVALID, "wrs_avg_logprob_host: bad arguments");
    for (int32_t i = 0; i < n; ++i) out_host[i] = sum_logprobs_host[i] / ((double)n_tokens_host[i] + 1.0);
    return WRS_OK;
}

}  // extern "C"

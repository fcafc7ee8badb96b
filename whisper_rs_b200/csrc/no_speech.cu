// Step-0 no-speech probability: reference src/decode/mod.rs:250-255
//     probs_at_sot = logits.i((.., sot_idx)).softmax(-1);  no_speech_probs = probs_at_sot[.., no_speech]
// i.e. one softmax over the UNFILTERED row per row (the filters run afterwards on the same row), reduced to a single
// probability.  HBM-bound: a row is read once (4 V bytes), one float per row is written.
//
// A CTA takes one eighth of a row into registers (coalesced 4-byte loads: rows are only 4-byte aligned when V is odd),
// reduces it to {max, sum exp(x - max)} and leaves that pair; the last CTA of the row (self-resetting counter) folds the
// eight pairs and writes exp(x[no_speech] - max) / sum.  avg_logprob (decode/mod.rs:389-397) is host arithmetic.
#include <algorithm>
#include <cmath>

#include <math_constants.h>

#include "wrs_internal.cuh"

struct wrs_decoder;

namespace wrs {

constexpr int kNsThreads = 256;
constexpr int kNsSplit = 8;                              // CTAs per row
constexpr int kNsPer = 32;                               // elements per thread: covers V <= 8 * 256 * 32 = 65 536

__global__ void __launch_bounds__(kNsThreads) no_speech_kernel(const float* __restrict__ logits, int64_t row_stride, int n_vocab,
                                                                int no_speech, float2* __restrict__ part, int* __restrict__ arrive,
                                                                float* __restrict__ out) {
    __shared__ float red[kNsThreads / 32];
    __shared__ float bcast;
    __shared__ int is_last;
    const int row = blockIdx.x / kNsSplit, piece = blockIdx.x % kNsSplit;
    const int per = (n_vocab + kNsSplit - 1) / kNsSplit, e0 = piece * per, e1 = min(n_vocab, e0 + per);
    const float* __restrict__ rp = logits + (int64_t)row * row_stride;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float v[kNsPer];
    float m = -CUDART_INF_F;
#pragma unroll
    for (int k = 0; k < kNsPer; ++k) {
        const int e = e0 + threadIdx.x + k * kNsThreads;
        v[k] = e < e1 ? __ldcs(rp + e) : -CUDART_INF_F;
        m = fmaxf(m, v[k]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) red[warp] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        float b = red[0];
        for (int w = 1; w < kNsThreads / 32; ++w) b = fmaxf(b, red[w]);
        bcast = b;
    }
    __syncthreads();
    const float bm = bcast;
    float s = 0.0f;
    if (bm > -CUDART_INF_F) {
#pragma unroll
        for (int k = 0; k < kNsPer; ++k) s += __expf(v[k] - bm);            // exp(-inf) = 0 for the padding; NaN stays NaN
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __syncthreads();
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < kNsThreads / 32; ++w) t += red[w];
        part[(int64_t)row * kNsSplit + piece] = make_float2(bm, t);
        __threadfence();
        const int prev = atomicAdd(arrive + row, 1);
        is_last = prev == kNsSplit - 1;
        if (is_last) { arrive[row] = 0; __threadfence(); }                   // self-reset for the next launch
    }
    __syncthreads();
    if (!is_last || threadIdx.x != 0) return;
    float M = -CUDART_INF_F;
    float2 q[kNsSplit];
#pragma unroll
    for (int i = 0; i < kNsSplit; ++i) { q[i] = __ldcg(part + (int64_t)row * kNsSplit + i); M = fmaxf(M, q[i].x); }
    float S = 0.0f;
#pragma unroll
    for (int i = 0; i < kNsSplit; ++i) S += q[i].x == -CUDART_INF_F ? 0.0f : q[i].y * expf(q[i].x - M);
    out[row] = expf(rp[no_speech] - M) / S;
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
    if (n_vocab > kNsSplit * kNsThreads * kNsPer) return fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_no_speech_probs: vocabulary above 65536");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    no_speech_kernel<<<(unsigned)(n_rows * kNsSplit), kNsThreads, 0, (cudaStream_t)stream>>>(
        logits, row_stride, n_vocab, no_speech, static_cast<float2*>(pairs), counters, out_probs);
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

// Context, error plumbing and host-side helpers of libwrs_b200.so.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "wrs_internal.cuh"

namespace wrs {

static thread_local std::string g_thread_err;

void set_thread_error(const std::string& s) { g_thread_err = s; }

int fail(wrs_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->err = msg;
    g_thread_err = msg;
    return code;
}

int cuda_fail(wrs_ctx* ctx, cudaError_t e, const char* what) {
    return fail(ctx, WRS_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

// audio::mel_filter()  reference src/audio.rs:175-188 -> mel_filter crate 0.1.1 `mel(16000, 400, Some(80),
// None, None, htk=false, NormalizationFactor::One)`: slaney-scale triangles, area-normalised.  Computed
// in double and rounded once to f32 (the reference KAT, src/audio.rs:266-279, pins the 8x8 corner).
static double hz_to_mel(double f) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = std::log(6.4) / 27.0;
    return f >= min_log_hz ? min_log_mel + std::log(f / min_log_hz) / logstep : f / f_sp;
}
static double mel_to_hz(double m) {
    const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = std::log(6.4) / 27.0;
    return m >= min_log_mel ? min_log_hz * std::exp(logstep * (m - min_log_mel)) : f_sp * m;
}

void mel_filter_slaney(float* out) {
    const int n_mels = WRS_N_MELS, n_bins = WRS_N_BINS;
    const double sr = WRS_SAMPLE_RATE;
    double pts[WRS_N_MELS + 2];
    const double mlo = hz_to_mel(0.0), mhi = hz_to_mel(sr / 2);
    for (int i = 0; i < n_mels + 2; ++i) pts[i] = mel_to_hz(mlo + (mhi - mlo) * i / (n_mels + 1));
    for (int m = 0; m < n_mels; ++m) {
        const double enorm = 2.0 / (pts[m + 2] - pts[m]);
        for (int k = 0; k < n_bins; ++k) {
            const double f = (sr / 2) * k / (n_bins - 1);
            const double lower = (f - pts[m]) / (pts[m + 1] - pts[m]);
            const double upper = (pts[m + 2] - f) / (pts[m + 2] - pts[m + 1]);
            const double v = std::fmax(0.0, std::fmin(lower, upper));
            out[m * n_bins + k] = (float)(v * enorm);
        }
    }
}

bool build_mel_bank(const float* dense, MelBank* bank) {
    constexpr int kPRows = WRS_N_BINS + 3;                  // rows of K1's power tile: 201 bins + zero rows behind them
    int total = 0;
    for (int m = 0; m < WRS_N_MELS; ++m) {
        int lo = -1, hi = -1;
        for (int k = 0; k < WRS_N_BINS; ++k)
            if (dense[m * WRS_N_BINS + k] != 0.0f) { if (lo < 0) lo = k; hi = k; }
        if (lo < 0) { lo = 0; hi = -1; }
        const int n = hi - lo + 1;
        if (total + n > MelBank::kMaxWeights) return false;
        bank->lo[m] = lo; bank->n[m] = n; bank->off[m] = total;
        for (int j = 0; j < n; ++j) bank->w[total + j] = dense[m * WRS_N_BINS + lo + j];
        total += n;
    }
    bank->total = total;
    // classes, padded weights, and the deal: rows in descending cost, each to the warp with the least so far
    int cls[WRS_N_MELS], len[WRS_N_MELS], order[WRS_N_MELS];
    double cost[WRS_N_MELS];
    for (int m = 0; m < WRS_N_MELS; ++m) {
        const int n = bank->n[m];
        static const int kLen[MelBank::kClasses - 1] = {2, 4, 6, 8, 12, 16};
        cls[m] = MelBank::kClasses - 1; len[m] = (n + 1) & ~1;
        for (int c = 0; c < MelBank::kClasses - 1; ++c)
            if (n <= kLen[c]) { cls[m] = c; len[m] = kLen[c]; break; }
        cost[m] = 22.0 + 2.25 * len[m];                     // issue slots of one mel row, roughly
        order[m] = m;
    }
    std::stable_sort(order, order + WRS_N_MELS, [&](int a, int b2) { return cost[a] > cost[b2]; });
    double load[MelBank::kCWarps] = {0};
    int owner[WRS_N_MELS];
    for (int i = 0; i < WRS_N_MELS; ++i) {
        int best = 0;
        for (int w = 1; w < MelBank::kCWarps; ++w) if (load[w] < load[best]) best = w;
        owner[order[i]] = best; load[best] += cost[order[i]];
    }
    for (int i = 0; i < MelBank::kMaxPadded; ++i) bank->wp[i] = 0.0f;
    int nr = 0, nw = 0;
    for (int w = 0; w < MelBank::kCWarps; ++w) {
        for (int c = 0; c < MelBank::kClasses; ++c) {
            bank->cls_begin[w][c] = nr;
            for (int m = 0; m < WRS_N_MELS; ++m) {
                if (owner[m] != w || cls[m] != c) continue;
                const int L = len[m];
                if (nw + L > MelBank::kMaxPadded || L > kPRows) return false;
                const int first = std::max(0, std::min(bank->lo[m], kPRows - L));   // the window of L bins stays inside the power tile
                for (int j = 0; j < bank->n[m]; ++j) bank->wp[nw + bank->lo[m] - first + j] = bank->w[bank->off[m] + j];
                bank->rows[nr++] = make_int4(first * 32 * (int)sizeof(float2), nw * (int)sizeof(float), m, L);
                nw += (L + 3) & ~3;
            }
        }
        bank->cls_begin[w][MelBank::kClasses] = nr;
    }
    bank->rows[WRS_N_MELS] = bank->rows[WRS_N_MELS + 1] = bank->rows[0];   // (looked at, never used)
    return true;
}

}  // namespace wrs

extern "C" {

int wrs_abi_version(void) { return WRS_ABI_VERSION; }

const char* wrs_last_error(const wrs_ctx* ctx) { return ctx ? ctx->err.c_str() : wrs::g_thread_err.c_str(); }

int wrs_mel_filter_host(float* out) {
    if (!out) return wrs::fail(nullptr, WRS_ERR_INVALID, "wrs_mel_filter_host: null output");
    wrs::mel_filter_slaney(out);
    return WRS_OK;
}

int wrs_ctx_create(int device, wrs_ctx** out) {
    if (!out) return wrs::fail(nullptr, WRS_ERR_INVALID, "wrs_ctx_create: null out");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return wrs::fail(nullptr, WRS_ERR_CUDA, std::string("wrs_ctx_create: no CUDA device (") +
                                                   (e != cudaSuccess ? cudaGetErrorString(e) : "count 0") +
                                                   "); this library has no CPU fallback");
    if (device < 0 || device >= n) return wrs::fail(nullptr, WRS_ERR_INVALID, "wrs_ctx_create: bad device index");
    wrs_ctx* ctx = new wrs_ctx();
    ctx->device = device;
    // any failure from here on: the message goes to the thread-local error (the caller never sees this ctx), and the
    // half-built context is released
    auto init = [&]() -> int {
        WRS_CUDA(ctx, cudaSetDevice(device));
        cudaDeviceProp prop;
        WRS_CUDA(ctx, cudaGetDeviceProperties(&prop, device));
        ctx->sm_count = prop.multiProcessorCount;
        WRS_CUDA(ctx, cudaMalloc(&ctx->d_bank, sizeof(wrs::MelBank)));
        WRS_CUDA(ctx, cudaMalloc(&ctx->d_status, sizeof(wrs::DevStatus)));
        WRS_CUDA(ctx, cudaMemset(ctx->d_status, 0, sizeof(wrs::DevStatus)));
        WRS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
        return wrs_logmel_set_filter(ctx, nullptr);
    };
    const int rc = init();
    if (rc != WRS_OK) { const std::string m = ctx->err; wrs_ctx_destroy(ctx); return wrs::fail(nullptr, rc, m); }
    *out = ctx;
    return WRS_OK;
}

void wrs_ctx_destroy(wrs_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->d_bank);
    cudaFree(ctx->d_seg_state);
    cudaFree(ctx->d_status);
    cudaFree(ctx->d_stage_in);
    cudaFree(ctx->d_stage_out);
    cudaFree(ctx->d_stage_i16);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    if (ctx->h2d_stream) cudaStreamDestroy(ctx->h2d_stream);
    if (ctx->d2h_stream) cudaStreamDestroy(ctx->d2h_stream);
    for (int i = 0; i < wrs_ctx::kHostChunks; ++i) {
        if (ctx->ev_in[i]) cudaEventDestroy(ctx->ev_in[i]);
        if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
    }
    delete ctx;
}

int wrs_ctx_check(wrs_ctx* ctx, wrs_stream_t stream) {
    if (!ctx) return wrs::fail(nullptr, WRS_ERR_INVALID, "wrs_ctx_check: null ctx");
    wrs::DevStatus st;
    WRS_CUDA(ctx, cudaMemcpyAsync(&st, ctx->d_status, sizeof(st), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    WRS_CUDA(ctx, cudaStreamSynchronize((cudaStream_t)stream));
    if (st.timeout || st.nan_rows || st.bad_index || st.group_violation) {
        WRS_CUDA(ctx, cudaMemsetAsync(ctx->d_status, 0, sizeof(st), (cudaStream_t)stream));
        if (st.timeout) return wrs::fail(ctx, WRS_ERR_TIMEOUT, "device-side bounded wait expired");
        if (st.bad_index) return wrs::fail(ctx, WRS_ERR_INVALID, "wrs_kv_reorder: a source index was outside [0, n_rows)");
        if (st.group_violation) return wrs::fail(ctx, WRS_ERR_INVALID, "wrs_kv_reorder: a tensor flagged group_invariant has differing rows within a beam group, or a source index leaves its group");
        return wrs::fail(ctx, WRS_ERR_NUMERIC, "NaN reached the softmax (the reference panics here)");
    }
    return WRS_OK;
}

int wrs_ctx_set_option(wrs_ctx* ctx, const char* key, int64_t value) {
    if (!ctx || !key) return wrs::fail(ctx, WRS_ERR_INVALID, "wrs_ctx_set_option: null argument");
    if (!strcmp(key, "logmel_fused")) { ctx->opt_logmel_fused = (int)value; return WRS_OK; }
    if (!strcmp(key, "kv_verify_group")) { ctx->opt_kv_verify_group = (int)value; return WRS_OK; }
    if (!strcmp(key, "host_chunk_segments")) { ctx->opt_host_chunk_segments = (int)value; return WRS_OK; }
    if (!strcmp(key, "proj_cluster")) { ctx->opt_proj_cluster = (int)value; return WRS_OK; }
    return wrs::fail(ctx, WRS_ERR_INVALID, std::string("unknown option ") + key);
}

int64_t wrs_ctx_get_stat(wrs_ctx* ctx, const char* key) {
    if (!ctx || !key) return -1;
    if (!strcmp(key, "logmel_mode")) return ctx->stat_last_logmel_mode;
    if (!strcmp(key, "logmel_grid")) return ctx->stat_last_logmel_grid;
    if (!strcmp(key, "sm_count")) return ctx->sm_count;
    return -1;
}

}  // extern "C"

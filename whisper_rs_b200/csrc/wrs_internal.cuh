// Internal declarations shared by the translation units of libwrs_b200.so (not part of the ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/wrs.h"

namespace wrs {

// Banded form of an [80,201] filterbank: row m has nonzeros only in [lo, lo+n).
struct MelBank {
    static constexpr int kMaxWeights = 1024;
    int lo[WRS_N_MELS];
    int n[WRS_N_MELS];
    int off[WRS_N_MELS];
    static constexpr int kMaxPadded = 768;
    float w[kMaxWeights];
    // the same bands padded with zero weights to even counts (the mel loop takes two bins per step, a weight is the
    // scalar operand of a packed two-frame FMA).  Row m: n2[m] >= 2 weights from off2[m] (both even).
    int n2[WRS_N_MELS];
    int off2[WRS_N_MELS];
    float wp[kMaxPadded];
    // K1 stage C: mel rows [row_begin[w], row_begin[w+1]) belong to consumer warp w of 8, cut so that a warp's mel work
    // plus its share of stage B (warps 0-4 take two 16-point DFT items, warps 5-7 one) comes out even
    int row_begin[9];
    int total;
};

// Device-side status word written by kernels (bounded-wait expiry, NaN); read by wrs_ctx_check.
struct DevStatus {
    int timeout;
    int nan_rows;
    int bad_index;      // K3: a source index outside [0, n_rows)
    int pad;
};

}  // namespace wrs

struct wrs_ctx {
    int device = 0;
    int sm_count = 0;
    std::string err;
    // log-mel
    wrs::MelBank* d_bank = nullptr;        // device copy of the active filterbank
    int* d_seg_state = nullptr;            // [2][cap] : max keys, arrival counters
    int64_t seg_cap = 0;
    wrs::DevStatus* d_status = nullptr;
    int opt_logmel_fused = -1;             // -1 auto, 0 two-pass, 1 fused
    int64_t stat_last_logmel_mode = -1;
    int64_t stat_last_logmel_grid = 0;
    int logmel_max_ctas_per_sm = 0;
    // host staging for *_host entry points
    float* d_stage_in = nullptr; size_t stage_in_bytes = 0;
    float* d_stage_out = nullptr; size_t stage_out_bytes = 0;
    cudaStream_t own_stream = nullptr;
    // *_host pipeline: copy-in and copy-out streams beside own_stream, one event pair per chunk in flight
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
    static constexpr int kHostChunks = 64;
    cudaEvent_t ev_in[kHostChunks] = {}, ev_done[kHostChunks] = {};
};

namespace wrs {

void set_thread_error(const std::string& s);
int fail(wrs_ctx* ctx, int code, const std::string& msg);
int cuda_fail(wrs_ctx* ctx, cudaError_t e, const char* what);
void mel_filter_slaney(float* out_80x201);
bool build_mel_bank(const float* dense_80x201, MelBank* bank);

#define WRS_CUDA(ctx, expr)                                              \
    do {                                                                 \
        cudaError_t _e = (expr);                                         \
        if (_e != cudaSuccess) return ::wrs::cuda_fail(ctx, _e, #expr);  \
    } while (0)

}  // namespace wrs

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
    float w[kMaxWeights];
    // K1 stage C.  Every row is projected over a fixed number of bins -- classes of 2, 4, 6, 8, 12 and 16 (its band
    // zero-padded to that length: fully unrolled, all loads up front, two rows in flight) -- or, wider than 16 bins, over
    // its own even count (the last class).  rows[] lists the rows warp by warp and class by class; consumer warp w of 7
    // takes rows[cls_begin[w][c] .. cls_begin[w][c+1]).  The rows are dealt so that the warps' costs come out even.
    static constexpr int kMaxPadded = 1536, kClasses = 7, kCWarps = 7;
    float wp[kMaxPadded];                  // the padded weights, row after row (16-byte aligned starts)
    int4 rows[WRS_N_MELS + 2];             // {byte offset of the first bin's row of P, byte offset into wp, mel row, bins}; two spare entries (look-ahead)
    int cls_begin[kCWarps][kClasses + 1];
    int total;
};

// Device-side status word written by kernels (bounded-wait expiry, NaN); read by wrs_ctx_check.
struct DevStatus {
    int timeout;
    int nan_rows;
    int bad_index;      // K3: a source index outside [0, n_rows)
    int group_violation;   // K3 (option kv_verify_group): a tensor flagged group-invariant is not, or an index leaves its group
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
    int opt_host_chunk_segments = 0;       // > 0: segments per chunk of the *_host pipelines
    int opt_proj_cluster = 1;              // 0: never launch the projection as two-CTA clusters (A/B timing, debugging)
    int opt_kv_verify_group = 0;           // > 0: beam size to verify group-invariant KV tensors against (debug)
    int64_t stat_last_logmel_mode = -1;
    int64_t stat_last_logmel_grid = 0;
    int logmel_max_ctas_per_sm = 0;
    // host staging for *_host entry points
    float* d_stage_in = nullptr; size_t stage_in_bytes = 0;
    float* d_stage_out = nullptr; size_t stage_out_bytes = 0;
    int16_t* d_stage_i16 = nullptr; size_t stage_i16_bytes = 0;
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

// K2/K3: fused per-step logit processing (filters + log-softmax + top-k + beam merge) for sm_100a.
//
//   decode_scan_kernel   grid (chunks, rows): one read of the logits, masks as predicates, per-chunk
//                        (max, sum-exp, top-K) for the text and the timestamp partition
//   beam_merge_kernel    grid (n_audio): per row merge of the chunk partials, the timestamp-vs-text rule
//                        (src/decode/filter/timestamp_tokens.rs:83-96) on the partition sums, log-softmax
//                        of the K candidates; then BeamSearchTokenExtractor::update's host loop
//                        (src/decode/extractor/beam_search.rs:49-146) on device
//   greedy_finish_kernel the same row merge followed by GreedyTokenExtractor::update (greedy.rs:24-50)
//   filter_*_kernel      in-place LogitFilter::apply equivalents (src/decode/filter/*.rs)
#include <algorithm>
#include <cmath>
#include <cstring>

#include "decode_scan.cuh"
#include "wrs_internal.cuh"

struct wrs_decoder {
    wrs_ctx* ctx = nullptr;
    wrs_decode_cfg cfg{};
    int max_rows = 0, max_audio = 0, max_cand = 0, kcap = 8, chunks = 8;
    int bm_words = 0;
    uint32_t* d_bm_base = nullptr;    // suppress list (+ no_timestamps when timestamps are on)
    uint32_t* d_bm_first = nullptr;   // base + blank ids (first step, suppress_blank)
    uint32_t* d_bm_ts = nullptr;      // only no_timestamps (in-place timestamp filter)
    uint32_t* d_bm_none = nullptr;    // empty
    int32_t* d_suppress = nullptr; int n_suppress = 0;   // raw list for the in-place sparse fill
    void* d_partials = nullptr;       // [max_rows][chunks][2] Partial<kcap>
    // finished-sequence store (BeamSearchTokenExtractor::finished_sequences, beam_search.rs:13)
    int64_t* d_fin_tokens = nullptr;  // [max_audio][max_cand][max_len]
    int32_t* d_fin_len = nullptr;     // [max_audio][max_cand]
    double* d_fin_score = nullptr;    // [max_audio][max_cand]
    int32_t* d_fin_n = nullptr;       // [max_audio]
    int32_t* d_sync = nullptr;        // [2] arrive counter, full counter
};

namespace wrs {

struct ScanParams {
    const float* logits; int64_t row_stride;
    const int64_t* tokens; int64_t tok_stride;
    int cur_len, sample_begin, n_rows, chunks, chunk_len, kout;
    RuleCfg rc;
    const uint32_t* bm;        // bitmap for regular steps
    const uint32_t* bm_first;  // bitmap when cur_len == sample_begin
    void* partials;
};

template <int KCAP>
__device__ __forceinline__ void block_reduce_store(Stats<KCAP>& st, int kout, Partial<KCAP>* __restrict__ dst,
                                                   Partial<KCAP>* sm_part) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    warp_stats_merge<KCAP>(st, kout);
    if (lane == 0) {
        sm_part[warp].m = st.m; sm_part[warp].s = st.s;
#pragma unroll
        for (int i = 0; i < KCAP; ++i) { sm_part[warp].v[i] = st.v[i]; sm_part[warp].id[i] = st.id[i]; }
    }
    __syncthreads();
    if (warp == 0) {
        Stats<KCAP> a;
        a.init();
        if (lane < kScanWarps) {
            a.m = sm_part[lane].m; a.s = sm_part[lane].s;
#pragma unroll
            for (int i = 0; i < KCAP; ++i) { a.v[i] = sm_part[lane].v[i]; a.id[i] = sm_part[lane].id[i]; }
        }
        warp_stats_merge<KCAP>(a, kout);
        if (lane == 0) {
            dst->m = a.m; dst->s = a.s;
#pragma unroll
            for (int i = 0; i < KCAP; ++i) { dst->v[i] = a.v[i]; dst->id[i] = a.id[i]; }
        }
    }
    __syncthreads();
}

template <int KCAP>
__global__ void __launch_bounds__(kScanThreads) decode_scan_kernel(const ScanParams p) {
    __shared__ Partial<KCAP> sm_part[kScanWarps];
    const int row = blockIdx.y, chunk = blockIdx.x;
    const int64_t* rt = p.tokens + (int64_t)row * p.tok_stride;
    const RowRules rr = row_rules(p.rc, rt, p.cur_len, p.sample_begin);
    const uint32_t* bm = (p.cur_len == p.sample_begin) ? p.bm_first : p.bm;
    const float* rp = p.logits + (int64_t)row * p.row_stride;
    const int c_lo = chunk * p.chunk_len, c_hi = min(p.rc.n_vocab, c_lo + p.chunk_len);
    Partial<KCAP>* out = reinterpret_cast<Partial<KCAP>*>(p.partials) + ((int64_t)row * p.chunks + chunk) * 2;

    Stats<KCAP> st;
    st.init();
    scan_range<KCAP>(rp, max(c_lo, rr.t_lo), min(c_hi, rr.t_hi), bm, st);
    block_reduce_store<KCAP>(st, p.kout, out, sm_part);
    st.init();
    scan_range<KCAP>(rp, max(c_lo, rr.s_lo), min(c_hi, rr.s_hi), bm, st);
    block_reduce_store<KCAP>(st, p.kout, out + 1, sm_part);
}

// ---- row finish: merge chunk partials, apply the timestamp-vs-text rule, produce K log-probs ----
template <int KCAP>
struct RowResult {
    float lp[KCAP];
    int tok[KCAP];
    unsigned flags;
};

// executed by one full warp per row; every lane returns the same result
template <int KCAP>
__device__ __forceinline__ RowResult<KCAP> finish_row(const Partial<KCAP>* __restrict__ parts, int chunks, int kout,
                                                      unsigned rule_flags, bool timestamps) {
    const int lane = threadIdx.x & 31;
    Stats<KCAP> T, S;
    T.init(); S.init();
    if (lane < chunks) {
        const Partial<KCAP>& a = parts[lane * 2];
        const Partial<KCAP>& b = parts[lane * 2 + 1];
        T.m = a.m; T.s = a.s; S.m = b.m; S.s = b.s;
#pragma unroll
        for (int i = 0; i < KCAP; ++i) { T.v[i] = a.v[i]; T.id[i] = a.id[i]; S.v[i] = b.v[i]; S.id[i] = b.id[i]; }
    }
    warp_stats_merge<KCAP>(T, kout);
    warp_stats_merge<KCAP>(S, kout);
    RowResult<KCAP> r;
    r.flags = rule_flags;
    // logsumexp(timestamps) > max(text): both sides carry the same -logsumexp(all), which cancels
    const float lse_s = (S.s > 0.0f) ? S.m + __logf(S.s) : -CUDART_INF_F;
    const bool ts_over_text = timestamps && (lse_s > T.m);
    if (T.s != T.s || S.s != S.s) r.flags |= 128u;
    if (ts_over_text) { r.flags |= 8u; T.init(); }
    // final log-softmax over what survived: (x - M) - log(sum exp(x - M)), the order ATen uses
    const float M = fmaxf(T.m, S.m);
    const float tot = T.s * exp_diff(T.m, M) + S.s * exp_diff(S.m, M);
    const float logtot = logf(tot);
    if (M == -CUDART_INF_F) r.flags |= 128u;   // everything masked: the reference yields NaN log-probs
    // merge the two sorted lists (text ids < timestamp ids)
    int it = 0, is = 0;
#pragma unroll
    for (int i = 0; i < KCAP; ++i) {
        float tv = -CUDART_INF_F, sv = -CUDART_INF_F; int ti = INT32_MAX, si = INT32_MAX;
#pragma unroll
        for (int j = 0; j < KCAP; ++j) {
            if (j == it) { tv = T.v[j]; ti = T.id[j]; }
            if (j == is) { sv = S.v[j]; si = S.id[j]; }
        }
        const bool take_t = better(tv, ti, sv, si);
        const float x = take_t ? tv : sv;
        r.tok[i] = take_t ? ti : si;
        r.lp[i] = (x - M) - logtot;
        if (take_t) ++it; else ++is;
    }
    return r;
}

struct MergeParams {
    const void* partials; int chunks, kout, beam, n_audio;
    const int64_t* tokens; int64_t tok_stride; int cur_len, sample_begin;
    RuleCfg rc; int eot, max_cand, max_len;
    const float* sum_in; float* sum_out;
    int64_t* next_tokens; int64_t next_stride; int32_t* source_idx; int32_t* completed;
    // finished store
    int64_t* fin_tokens; int32_t* fin_len; double* fin_score; int32_t* fin_n; int32_t* sync;
    // diagnostics
    float* dbg_lp; int32_t* dbg_tok; uint8_t* dbg_flags; int32_t* dbg_fin_count; int32_t* dbg_fin_source; double* dbg_fin_score;
    DevStatus* status;
};

constexpr int kMergeThreads = 256;
constexpr int kMaxBeam = 15;
constexpr int kMaxCandAudio = kMaxBeam * (kMaxBeam + 1);   // 240

template <int KCAP>
__global__ void __launch_bounds__(kMergeThreads) beam_merge_kernel(const MergeParams p) {
    __shared__ float s_lp[kMaxBeam][KCAP];
    __shared__ int s_tok[kMaxBeam][KCAP];
    __shared__ unsigned char s_same[kMaxBeam][kMaxBeam];   // prefix j == prefix j'
    __shared__ double c_score[kMaxCandAudio];
    __shared__ short c_row[kMaxCandAudio], c_slot[kMaxCandAudio];
    __shared__ int c_tok[kMaxCandAudio];
    __shared__ short order[kMaxCandAudio];
    __shared__ int sel_src[kMaxBeam], sel_tok[kMaxBeam], fin_src[kMaxBeam];
    __shared__ double sel_score[kMaxBeam], fin_sc[kMaxBeam];
    __shared__ int n_fin, store_base, n_store, s_nan;

    const int audio = blockIdx.x, k = p.beam, K1 = p.kout;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int row0 = audio * k;
    if (tid == 0) s_nan = 0;
    for (int i = tid; i < k * k; i += kMergeThreads) s_same[i / k][i % k] = 1;
    __syncthreads();

    // 1. per-row finish (one warp per row)
    for (int j = warp; j < k; j += kMergeThreads / 32) {
        const int row = row0 + j;
        const RowRules rr = row_rules(p.rc, p.tokens + (int64_t)row * p.tok_stride, p.cur_len, p.sample_begin);
        const Partial<KCAP>* parts = reinterpret_cast<const Partial<KCAP>*>(p.partials) + (int64_t)row * p.chunks * 2;
        const RowResult<KCAP> r = finish_row<KCAP>(parts, p.chunks, K1, rr.flags, p.rc.timestamps != 0);
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < KCAP; ++i) { s_lp[j][i] = r.lp[i]; s_tok[j][i] = r.tok[i]; }
            if (r.flags & 128u) s_nan = 1;
            if (p.dbg_flags) p.dbg_flags[row] = (uint8_t)r.flags;
        }
        if (lane < K1) {
            float lpv = 0.f; int tk = 0;
#pragma unroll
            for (int i = 0; i < KCAP; ++i) if (i == lane) { lpv = r.lp[i]; tk = r.tok[i]; }
            if (p.dbg_lp) p.dbg_lp[(int64_t)row * K1 + lane] = lpv;
            if (p.dbg_tok) p.dbg_tok[(int64_t)row * K1 + lane] = tk;
        }
    }
    // 2. which beams of this audio share an identical prefix (HashMap key = whole sequence, beam_search.rs:74-77)
    for (int pr = 0; pr < k * (k - 1) / 2; ++pr) {
        int a = 0, b = 0, q = pr;
        for (a = 0; a < k; ++a) { if (q < k - 1 - a) { b = a + 1 + q; break; } q -= k - 1 - a; }
        const int64_t* ta = p.tokens + (int64_t)(row0 + a) * p.tok_stride;
        const int64_t* tb = p.tokens + (int64_t)(row0 + b) * p.tok_stride;
        bool diff = false;
        for (int i = tid; i < p.cur_len; i += kMergeThreads) diff |= (ta[i] != tb[i]);
        if (diff) { s_same[a][b] = 0; s_same[b][a] = 0; }
    }
    __syncthreads();

    // 3. candidates: score = f64(sum_logprobs[row]) + f64(logprob)   (beam_search.rs:70-71)
    const int n_c = k * K1;
    for (int c = tid; c < n_c; c += kMergeThreads) {
        const int j = c / K1, i = c - j * K1;
        const int tok = s_tok[j][i];
        // a later beam with the same prefix proposing the same token overwrites this map entry (:77)
        bool shadowed = false;
        for (int j2 = j + 1; j2 < k; ++j2)
            if (s_same[j][j2])
                for (int i2 = 0; i2 < K1; ++i2) shadowed |= (s_tok[j2][i2] == tok);
        c_score[c] = (double)p.sum_in[row0 + j] + (double)s_lp[j][i];
        c_row[c] = (short)j; c_slot[c] = (short)(shadowed ? 1 : 0); c_tok[c] = tok;
    }
    __syncthreads();
    // 4. rank sort by (score desc, source row asc, token asc); shadowed entries go last
    for (int c = tid; c < n_c; c += kMergeThreads) {
        int rank = 0;
        const double sc = c_score[c];
        for (int d = 0; d < n_c; ++d) {
            if (d == c) continue;
            bool before;
            if (c_slot[d] != c_slot[c]) before = c_slot[d] < c_slot[c];
            else if (c_score[d] != sc) before = c_score[d] > sc;
            else if (c_row[d] != c_row[c]) before = c_row[d] < c_row[c];
            else before = c_tok[d] < c_tok[c];
            rank += before ? 1 : 0;
        }
        order[rank] = (short)c;
    }
    __syncthreads();
    // 5. walk (beam_search.rs:87-104): EOT-terminated -> finished, others kept until k are saved
    if (tid == 0) {
        int saved = 0, nf = 0;
        for (int o = 0; o < n_c && saved < k; ++o) {
            const int c = order[o];
            if (c_slot[c]) break;   // only shadowed entries remain
            if (c_tok[c] == p.eot) { fin_src[nf] = c_row[c]; fin_sc[nf] = c_score[c]; ++nf; }
            else { sel_src[saved] = c_row[c]; sel_tok[saved] = c_tok[c]; sel_score[saved] = c_score[c]; ++saved; }
        }
        for (; saved < k; ++saved) { sel_src[saved] = 0; sel_tok[saved] = p.eot; sel_score[saved] = -CUDART_INF; }   // cannot happen with >= k+1 finite logits per row
        n_fin = nf;
        // finished store: append in descending score until round(patience*k) entries (:124-138)
        const int have = p.fin_n[audio];
        const int add = min(nf, max(0, p.max_cand - have));
        store_base = have; n_store = add;
        p.fin_n[audio] = have + add;
        if (p.dbg_fin_count) p.dbg_fin_count[audio] = nf;
        for (int i = 0; i < k; ++i) {
            if (p.dbg_fin_source) p.dbg_fin_source[audio * k + i] = i < nf ? row0 + fin_src[i] : -1;
            if (p.dbg_fin_score) p.dbg_fin_score[audio * k + i] = i < nf ? fin_sc[i] : 0.0;
        }
        if (s_nan) atomicAdd(&p.status->nan_rows, 1);
    }
    __syncthreads();
    // 6. outputs
    for (int i = tid; i < k; i += kMergeThreads) {
        p.source_idx[row0 + i] = row0 + sel_src[i];
        p.sum_out[row0 + i] = (float)sel_score[i];   // index_put_(.., score as f32) :91-95
    }
    for (int i = 0; i < k; ++i) {
        const int64_t* src = p.tokens + (int64_t)(row0 + sel_src[i]) * p.tok_stride;
        int64_t* dst = p.next_tokens + (int64_t)(row0 + i) * p.next_stride;
        for (int t = tid; t < p.cur_len; t += kMergeThreads) dst[t] = src[t];
        if (tid == 0) dst[p.cur_len] = sel_tok[i];
    }
    for (int i = 0; i < n_store; ++i) {
        const int slot = store_base + i;
        const int64_t* src = p.tokens + (int64_t)(row0 + fin_src[i]) * p.tok_stride;
        int64_t* dst = p.fin_tokens + ((int64_t)audio * p.max_cand + slot) * p.max_len;
        for (int t = tid; t < p.cur_len; t += kMergeThreads) dst[t] = src[t];
        if (tid == 0) {
            dst[p.cur_len] = p.eot;
            p.fin_len[audio * p.max_cand + slot] = p.cur_len + 1;
            p.fin_score[audio * p.max_cand + slot] = fin_sc[i];
        }
    }
    // 7. completed = all audios hold max_cand finished sequences (:141-146); last CTA to arrive publishes
    if (tid == 0) {
        const int full = (store_base + n_store >= p.max_cand) ? 1 : 0;
        if (full) atomicAdd(p.sync + 1, 1);
        __threadfence();
        const int arrived = atomicAdd(p.sync, 1);
        if (arrived == p.n_audio - 1) {
            __threadfence();
            const int nfull = atomicExch(p.sync + 1, 0);
            *p.completed = (nfull == p.n_audio) ? 1 : 0;
            atomicExch(p.sync, 0);
        }
    }
}

// ---- greedy: row finish + GreedyTokenExtractor::update (src/decode/extractor/greedy.rs:24-50) ----
struct GreedyParams {
    const void* partials; int chunks, n_rows;
    const int64_t* tokens; int64_t tok_stride; int cur_len, sample_begin;
    RuleCfg rc; int eot;
    const float* sum_in; float* sum_out;
    int64_t* next_tokens; int64_t next_stride; int32_t* completed; int32_t* sync;
    float* dbg_lp; int32_t* dbg_tok; uint8_t* dbg_flags;
    DevStatus* status;
};

template <int KCAP>
__global__ void __launch_bounds__(kMergeThreads) greedy_finish_kernel(const GreedyParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row = blockIdx.x * (kMergeThreads / 32) + warp;
    __shared__ int s_eot_rows, s_rows;
    if (threadIdx.x == 0) { s_eot_rows = 0; s_rows = 0; }
    __syncthreads();
    if (row < p.n_rows) {
        const int64_t* rt = p.tokens + (int64_t)row * p.tok_stride;
        const RowRules rr = row_rules(p.rc, rt, p.cur_len, p.sample_begin);
        const Partial<KCAP>* parts = reinterpret_cast<const Partial<KCAP>*>(p.partials) + (int64_t)row * p.chunks * 2;
        const RowResult<KCAP> r = finish_row<KCAP>(parts, p.chunks, 1, rr.flags, p.rc.timestamps != 0);
        int64_t* dst = p.next_tokens + (int64_t)row * p.next_stride;
        for (int t = lane; t < p.cur_len; t += 32) dst[t] = rt[t];
        if (lane == 0) {
            const bool last_eot = rt[p.cur_len - 1] == p.eot;                       // greedy.rs:37-38
            const int next = last_eot ? p.eot : r.tok[0];                            // :41-45
            p.sum_out[row] = p.sum_in[row] + (last_eot ? 0.0f : r.lp[0]);            // :39
            dst[p.cur_len] = next;
            if (p.dbg_lp) p.dbg_lp[row] = r.lp[0];
            if (p.dbg_tok) p.dbg_tok[row] = r.tok[0];
            if (p.dbg_flags) p.dbg_flags[row] = (uint8_t)r.flags;
            if (r.flags & 128u) atomicAdd(&p.status->nan_rows, 1);
            atomicAdd(&s_rows, 1);
            if (next == p.eot) atomicAdd(&s_eot_rows, 1);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(p.sync + 1, s_eot_rows);
        __threadfence();
        const int arrived = atomicAdd(p.sync, 1);
        if (arrived == (int)gridDim.x - 1) {
            __threadfence();
            const int n_eot = atomicExch(p.sync + 1, 0);
            *p.completed = (n_eot == p.n_rows) ? 1 : 0;
            atomicExch(p.sync, 0);
        }
    }
}

// ---- in-place filters (LogitFilter::apply, src/decode/filter/mod.rs:12-14) ----
// logits.index_fill_(1, idx, -inf)   suppress_tokens.rs:22, suppress_blanks.rs:41
__global__ void filter_index_fill_kernel(float* __restrict__ logits, int64_t row_stride, int n_rows, int n_vocab,
                                         const int32_t* __restrict__ ids, int n_ids) {
    const int64_t total = (int64_t)n_rows * n_ids;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / n_ids), id = ids[i - (int64_t)r * n_ids];
        if (id >= 0 && id < n_vocab) logits[(int64_t)r * row_stride + id] = -CUDART_INF_F;
    }
}

struct TsFillParams {
    float* logits; int64_t row_stride; int n_rows;
    const int64_t* tokens; int64_t tok_stride; int cur_len, sample_begin;
    RuleCfg rc; int no_timestamps;
    const void* partials; int chunks;
    uint8_t* row_flags; DevStatus* status;
};

// timestamp_tokens.rs:37-96 in place: one CTA per row; the statistics come from decode_scan_kernel run on
// the logits as they are (earlier filters already wrote their -inf), so the decision sees what the
// reference's log_softmax sees.
template <int KCAP>
__global__ void __launch_bounds__(kScanThreads) filter_ts_fill_kernel(const TsFillParams p) {
    const int row = blockIdx.x, tid = threadIdx.x;
    __shared__ unsigned s_flags;
    const RowRules rr = row_rules(p.rc, p.tokens + (int64_t)row * p.tok_stride, p.cur_len, p.sample_begin);
    if (tid < 32) {
        const Partial<KCAP>* parts = reinterpret_cast<const Partial<KCAP>*>(p.partials) + (int64_t)row * p.chunks * 2;
        const RowResult<KCAP> r = finish_row<KCAP>(parts, p.chunks, 1, rr.flags, true);
        if (tid == 0) {
            s_flags = r.flags;
            if (p.row_flags) p.row_flags[row] = (uint8_t)r.flags;
        }
    }
    __syncthreads();
    const bool kill_text = (s_flags & 8u) != 0;
    float* rp = p.logits + (int64_t)row * p.row_stride;
    const int tsb = p.rc.ts_begin, V = p.rc.n_vocab;
    // text side: surviving range is [t_lo, t_hi) unless the decision fired
    const int keep_lo = kill_text ? tsb : rr.t_lo, keep_hi = kill_text ? tsb : rr.t_hi;
    for (int i = tid; i < tsb; i += kScanThreads)
        if (i < keep_lo || i >= keep_hi) rp[i] = -CUDART_INF_F;
    for (int i = tsb + tid; i < V; i += kScanThreads)
        if (i < rr.s_lo || i >= rr.s_hi) rp[i] = -CUDART_INF_F;
    if (tid == 0 && p.no_timestamps >= 0 && p.no_timestamps < V) rp[p.no_timestamps] = -CUDART_INF_F;
}

}  // namespace wrs

// =================================================================================================
// host side
// =================================================================================================
namespace {

using namespace wrs;

RuleCfg rule_cfg(const wrs_decode_cfg& c) {
    RuleCfg r;
    r.n_vocab = c.n_vocab; r.eot = c.eot; r.ts_begin = c.ts_begin; r.timestamps = c.timestamps;
    r.max_initial_ts_index = c.max_initial_timestamp_index;
    return r;
}

size_t partial_bytes(int kcap) { return kcap == 8 ? sizeof(Partial<8>) : sizeof(Partial<16>); }

int check_step_args(wrs_decoder* dec, const char* who, const void* logits, int64_t row_stride, int64_t n_rows,
                    const void* tokens, int64_t tok_stride, int64_t cur_len, int64_t sample_begin) {
    if (!dec) return fail(nullptr, WRS_ERR_INVALID, std::string(who) + ": null decoder");
    wrs_ctx* ctx = dec->ctx;
    if (!logits || !tokens) return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": null buffer");
    if (n_rows <= 0 || n_rows > dec->max_rows) return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": n_rows outside 1..max_rows");
    if (row_stride < dec->cfg.n_vocab) return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": row_stride < n_vocab");
    if (cur_len < 1 || tok_stride < cur_len || sample_begin < 1 || sample_begin > cur_len)
        return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": need 1 <= sample_begin <= cur_len <= tok_stride");
    if (cur_len + 1 > dec->cfg.max_len) return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": cur_len + 1 exceeds cfg.max_len");
    if ((reinterpret_cast<uintptr_t>(logits) & 3) != 0) return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": logits not 4-byte aligned");
    return WRS_OK;
}

int launch_scan(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, const int64_t* tokens,
                int64_t tok_stride, int64_t cur_len, int64_t sample_begin, int kout, const uint32_t* bm,
                const uint32_t* bm_first, cudaStream_t stream) {
    ScanParams sp;
    sp.logits = logits; sp.row_stride = row_stride; sp.tokens = tokens; sp.tok_stride = tok_stride;
    sp.cur_len = (int)cur_len; sp.sample_begin = (int)sample_begin; sp.n_rows = (int)n_rows;
    sp.chunks = dec->chunks;
    sp.chunk_len = (((dec->cfg.n_vocab + dec->chunks - 1) / dec->chunks) + 3) & ~3;
    sp.kout = kout; sp.rc = rule_cfg(dec->cfg); sp.bm = bm; sp.bm_first = bm_first; sp.partials = dec->d_partials;
    dim3 grid(dec->chunks, (unsigned)n_rows);
    if (dec->kcap == 8) decode_scan_kernel<8><<<grid, kScanThreads, 0, stream>>>(sp);
    else decode_scan_kernel<16><<<grid, kScanThreads, 0, stream>>>(sp);
    WRS_CUDA(dec->ctx, cudaGetLastError());
    return WRS_OK;
}

}  // namespace

extern "C" {

int wrs_decoder_create(wrs_ctx* ctx, const wrs_decode_cfg* cfg, const int32_t* suppress_ids, int32_t n_suppress,
                       int32_t max_rows, wrs_decoder** out) {
    if (!ctx || !cfg || !out) return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: null argument");
    *out = nullptr;
    const wrs_decode_cfg& c = *cfg;
    if (c.n_vocab < 8 || c.beam_size < 1 || c.beam_size > kMaxBeam || max_rows < c.beam_size || max_rows % c.beam_size != 0)
        return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: need n_vocab >= 8, 1 <= beam_size <= 15, max_rows a multiple of beam_size");
    if (c.eot < 0 || c.eot >= c.n_vocab || (c.timestamps && (c.ts_begin <= c.eot || c.ts_begin > c.n_vocab ||
                                                              c.no_timestamps < 0 || c.no_timestamps >= c.n_vocab)))
        return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: need 0 <= eot < ts_begin <= n_vocab and a valid no_timestamps id");
    if (n_suppress < 0 || (n_suppress > 0 && !suppress_ids)) return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: bad suppress list");
    if (c.max_len < 2 || !(c.patience > 0.0f)) return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: need max_len >= 2 and patience > 0");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    wrs_decoder* d = new wrs_decoder();
    d->ctx = ctx; d->cfg = c; d->max_rows = max_rows; d->max_audio = max_rows / c.beam_size;
    d->max_cand = std::max(1, (int)std::round(c.patience * (float)c.beam_size));   // f32::round, beam_search.rs:122
    d->kcap = (c.beam_size + 1 <= 8) ? 8 : 16;
    d->chunks = 8;
    d->bm_words = (c.n_vocab + 31) / 32 + 2;
    std::vector<uint32_t> base(d->bm_words, 0u), first, ts(d->bm_words, 0u), none(d->bm_words, 0u);
    auto setbit = [&](std::vector<uint32_t>& b, int id) { if (id >= 0 && id < c.n_vocab) b[id >> 5] |= 1u << (id & 31); };
    for (int i = 0; i < n_suppress; ++i) setbit(base, suppress_ids[i]);
    if (c.timestamps) { setbit(base, c.no_timestamps); setbit(ts, c.no_timestamps); }
    first = base;
    if (c.suppress_blank) { setbit(first, c.blank_ids[0]); setbit(first, c.blank_ids[1]); }
    const size_t bmb = sizeof(uint32_t) * d->bm_words;
    auto up = [&](uint32_t** dst, const std::vector<uint32_t>& src) -> cudaError_t {
        cudaError_t e = cudaMalloc(dst, bmb);
        return e != cudaSuccess ? e : cudaMemcpy(*dst, src.data(), bmb, cudaMemcpyHostToDevice);
    };
    cudaError_t e = cudaSuccess;
    if ((e = up(&d->d_bm_base, base)) != cudaSuccess || (e = up(&d->d_bm_first, first)) != cudaSuccess ||
        (e = up(&d->d_bm_ts, ts)) != cudaSuccess || (e = up(&d->d_bm_none, none)) != cudaSuccess) {
        wrs_decoder_destroy(d);
        return cuda_fail(ctx, e, "decoder bitmaps");
    }
    d->n_suppress = n_suppress;
    const size_t fin_tok = sizeof(int64_t) * (size_t)d->max_audio * d->max_cand * c.max_len;
    if ((n_suppress > 0 && ((e = cudaMalloc(&d->d_suppress, sizeof(int32_t) * n_suppress)) != cudaSuccess ||
                            (e = cudaMemcpy(d->d_suppress, suppress_ids, sizeof(int32_t) * n_suppress, cudaMemcpyHostToDevice)) != cudaSuccess)) ||
        (e = cudaMalloc(&d->d_partials, partial_bytes(d->kcap) * 2 * d->chunks * (size_t)max_rows)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_fin_tokens, fin_tok)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_fin_len, sizeof(int32_t) * d->max_audio * d->max_cand)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_fin_score, sizeof(double) * d->max_audio * d->max_cand)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_fin_n, sizeof(int32_t) * d->max_audio)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_sync, sizeof(int32_t) * 2)) != cudaSuccess ||
        (e = cudaMemset(d->d_fin_n, 0, sizeof(int32_t) * d->max_audio)) != cudaSuccess ||
        (e = cudaMemset(d->d_sync, 0, sizeof(int32_t) * 2)) != cudaSuccess) {
        wrs_decoder_destroy(d);
        return cuda_fail(ctx, e, "decoder workspace");
    }
    *out = d;
    return WRS_OK;
}

void wrs_decoder_destroy(wrs_decoder* d) {
    if (!d) return;
    cudaSetDevice(d->ctx->device);
    cudaFree(d->d_bm_base); cudaFree(d->d_bm_first); cudaFree(d->d_bm_ts); cudaFree(d->d_bm_none);
    cudaFree(d->d_suppress); cudaFree(d->d_partials); cudaFree(d->d_fin_tokens); cudaFree(d->d_fin_len);
    cudaFree(d->d_fin_score); cudaFree(d->d_fin_n); cudaFree(d->d_sync);
    delete d;
}

int wrs_decoder_reset(wrs_decoder* d, wrs_stream_t stream) {
    if (!d) return fail(nullptr, WRS_ERR_INVALID, "wrs_decoder_reset: null decoder");
    WRS_CUDA(d->ctx, cudaSetDevice(d->ctx->device));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_fin_n, 0, sizeof(int32_t) * d->max_audio, (cudaStream_t)stream));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_sync, 0, sizeof(int32_t) * 2, (cudaStream_t)stream));
    return WRS_OK;
}

int wrs_filter_apply(wrs_decoder* dec, int32_t which, float* logits, int64_t row_stride, int64_t n_rows,
                     const int64_t* tokens, int64_t tok_stride, int64_t cur_len, int64_t sample_begin,
                     uint8_t* row_flags, wrs_stream_t stream_) {
    int rc = check_step_args(dec, "wrs_filter_apply", logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin);
    if (rc != WRS_OK) return rc;
    wrs_ctx* ctx = dec->ctx;
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    const wrs_decode_cfg& c = dec->cfg;
    if ((which & WRS_FILTER_SUPPRESS_BLANK) && cur_len == sample_begin) {          // suppress_blanks.rs:40
        // blank ids live at the head of a tiny device array: reuse the bitmap-free sparse fill
        int32_t* d_ids = nullptr;
        WRS_CUDA(ctx, cudaMallocAsync(&d_ids, sizeof(int32_t) * 2, stream));
        WRS_CUDA(ctx, cudaMemcpyAsync(d_ids, c.blank_ids, sizeof(int32_t) * 2, cudaMemcpyHostToDevice, stream));
        filter_index_fill_kernel<<<(unsigned)std::min<int64_t>((n_rows * 2 + 255) / 256, 1024), 256, 0, stream>>>(
            logits, row_stride, (int)n_rows, c.n_vocab, d_ids, 2);
        WRS_CUDA(ctx, cudaGetLastError());
        WRS_CUDA(ctx, cudaFreeAsync(d_ids, stream));
    }
    if ((which & WRS_FILTER_SUPPRESS_TOKENS) && dec->n_suppress > 0) {              // suppress_tokens.rs:22
        filter_index_fill_kernel<<<(unsigned)std::min<int64_t>((n_rows * dec->n_suppress + 255) / 256, 1024), 256, 0, stream>>>(
            logits, row_stride, (int)n_rows, c.n_vocab, dec->d_suppress, dec->n_suppress);
        WRS_CUDA(ctx, cudaGetLastError());
    }
    if (which & WRS_FILTER_TIMESTAMPS) {                                             // timestamp_tokens.rs:33-98
        if (!c.timestamps) return fail(ctx, WRS_ERR_INVALID, "wrs_filter_apply: decoder was created with timestamps = 0");
        rc = launch_scan(dec, logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin, 1, dec->d_bm_ts,
                         dec->d_bm_ts, stream);
        if (rc != WRS_OK) return rc;
        TsFillParams fp;
        fp.logits = logits; fp.row_stride = row_stride; fp.n_rows = (int)n_rows; fp.tokens = tokens; fp.tok_stride = tok_stride;
        fp.cur_len = (int)cur_len; fp.sample_begin = (int)sample_begin; fp.rc = rule_cfg(c); fp.no_timestamps = c.no_timestamps;
        fp.partials = dec->d_partials; fp.chunks = dec->chunks; fp.row_flags = row_flags; fp.status = ctx->d_status;
        if (dec->kcap == 8) filter_ts_fill_kernel<8><<<(unsigned)n_rows, kScanThreads, 0, stream>>>(fp);
        else filter_ts_fill_kernel<16><<<(unsigned)n_rows, kScanThreads, 0, stream>>>(fp);
        WRS_CUDA(ctx, cudaGetLastError());
    }
    return WRS_OK;
}

int wrs_beam_step(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, const int64_t* tokens,
                  int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_in, int64_t* next_tokens,
                  int64_t next_stride, int32_t* source_idx, float* sum_out, int32_t* completed, const wrs_step_debug* dbg,
                  wrs_stream_t stream_) {
    int rc = check_step_args(dec, "wrs_beam_step", logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin);
    if (rc != WRS_OK) return rc;
    wrs_ctx* ctx = dec->ctx;
    const wrs_decode_cfg& c = dec->cfg;
    if (!sum_in || !next_tokens || !source_idx || !sum_out || !completed) return fail(ctx, WRS_ERR_INVALID, "wrs_beam_step: null output");
    if (n_rows % c.beam_size != 0) return fail(ctx, WRS_ERR_INVALID, "wrs_beam_step: n_rows not a multiple of beam_size (beam_search.rs:29)");
    if (next_stride < cur_len + 1) return fail(ctx, WRS_ERR_INVALID, "wrs_beam_step: next_stride < cur_len + 1");
    if ((const void*)next_tokens == (const void*)tokens) return fail(ctx, WRS_ERR_INVALID, "wrs_beam_step: next_tokens must not alias tokens");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    const int K1 = c.beam_size + 1;
    rc = launch_scan(dec, logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin, K1, dec->d_bm_base,
                     dec->d_bm_first, stream);
    if (rc != WRS_OK) return rc;
    MergeParams mp;
    mp.partials = dec->d_partials; mp.chunks = dec->chunks; mp.kout = K1; mp.beam = c.beam_size;
    mp.n_audio = (int)(n_rows / c.beam_size);
    mp.tokens = tokens; mp.tok_stride = tok_stride; mp.cur_len = (int)cur_len; mp.sample_begin = (int)sample_begin;
    mp.rc = rule_cfg(c); mp.eot = c.eot; mp.max_cand = dec->max_cand; mp.max_len = c.max_len;
    mp.sum_in = sum_in; mp.sum_out = sum_out; mp.next_tokens = next_tokens; mp.next_stride = next_stride;
    mp.source_idx = source_idx; mp.completed = completed;
    mp.fin_tokens = dec->d_fin_tokens; mp.fin_len = dec->d_fin_len; mp.fin_score = dec->d_fin_score; mp.fin_n = dec->d_fin_n;
    mp.sync = dec->d_sync;
    mp.dbg_lp = dbg ? dbg->cand_logprob : nullptr; mp.dbg_tok = dbg ? dbg->cand_token : nullptr;
    mp.dbg_flags = dbg ? dbg->row_flags : nullptr; mp.dbg_fin_count = dbg ? dbg->fin_count : nullptr;
    mp.dbg_fin_source = dbg ? dbg->fin_source : nullptr; mp.dbg_fin_score = dbg ? dbg->fin_score : nullptr;
    mp.status = ctx->d_status;
    if (dec->kcap == 8) beam_merge_kernel<8><<<mp.n_audio, kMergeThreads, 0, stream>>>(mp);
    else beam_merge_kernel<16><<<mp.n_audio, kMergeThreads, 0, stream>>>(mp);
    WRS_CUDA(ctx, cudaGetLastError());
    return WRS_OK;
}

int wrs_greedy_step(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, const int64_t* tokens,
                    int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_in, int64_t* next_tokens,
                    int64_t next_stride, float* sum_out, int32_t* completed, const wrs_step_debug* dbg, wrs_stream_t stream_) {
    int rc = check_step_args(dec, "wrs_greedy_step", logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin);
    if (rc != WRS_OK) return rc;
    wrs_ctx* ctx = dec->ctx;
    const wrs_decode_cfg& c = dec->cfg;
    if (!sum_in || !next_tokens || !sum_out || !completed) return fail(ctx, WRS_ERR_INVALID, "wrs_greedy_step: null output");
    if (next_stride < cur_len + 1) return fail(ctx, WRS_ERR_INVALID, "wrs_greedy_step: next_stride < cur_len + 1");
    if ((const void*)next_tokens == (const void*)tokens) return fail(ctx, WRS_ERR_INVALID, "wrs_greedy_step: next_tokens must not alias tokens");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    rc = launch_scan(dec, logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin, 1, dec->d_bm_base,
                     dec->d_bm_first, stream);
    if (rc != WRS_OK) return rc;
    GreedyParams gp;
    gp.partials = dec->d_partials; gp.chunks = dec->chunks; gp.n_rows = (int)n_rows; gp.tokens = tokens; gp.tok_stride = tok_stride;
    gp.cur_len = (int)cur_len; gp.sample_begin = (int)sample_begin; gp.rc = rule_cfg(c); gp.eot = c.eot;
    gp.sum_in = sum_in; gp.sum_out = sum_out; gp.next_tokens = next_tokens; gp.next_stride = next_stride;
    gp.completed = completed; gp.sync = dec->d_sync;
    gp.dbg_lp = dbg ? dbg->cand_logprob : nullptr; gp.dbg_tok = dbg ? dbg->cand_token : nullptr;
    gp.dbg_flags = dbg ? dbg->row_flags : nullptr; gp.status = ctx->d_status;
    const unsigned grid = (unsigned)((n_rows + kMergeThreads / 32 - 1) / (kMergeThreads / 32));
    if (dec->kcap == 8) greedy_finish_kernel<8><<<grid, kMergeThreads, 0, stream>>>(gp);
    else greedy_finish_kernel<16><<<grid, kMergeThreads, 0, stream>>>(gp);
    WRS_CUDA(ctx, cudaGetLastError());
    return WRS_OK;
}

}  // extern "C"

namespace wrs {
struct DecoderView {
    wrs_ctx* ctx; int beam, eot, max_len, max_cand, max_audio, max_rows;
    const int64_t* d_fin_tokens; const int32_t* d_fin_len; const double* d_fin_score; const int32_t* d_fin_n;
};
DecoderView decoder_view(const wrs_decoder* d) {
    return DecoderView{d->ctx, d->cfg.beam_size, d->cfg.eot, d->cfg.max_len, d->max_cand, d->max_audio, d->max_rows,
                       d->d_fin_tokens, d->d_fin_len, d->d_fin_score, d->d_fin_n};
}
}  // namespace wrs

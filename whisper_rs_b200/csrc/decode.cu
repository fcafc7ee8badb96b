// K2: the fused per-step logit processing (filters + log-softmax + top-k + beam merge) for sm_100a, ONE launch.
//
//   decode_step_kernel<MODE>  grid (chunks, rows), 512 threads.  Every CTA scans one chunk of one row
//       (decode_scan.cuh: one read of the logits, masks as predicates, {max, sum-exp, top-K} partial to global),
//       then arrives on its group's counter.  The LAST CTA to arrive for a group runs the group's tail:
//   MODE_BEAM    group = audio (beam rows): per row merge of the chunk partials, the timestamp-vs-text rule
//                (src/decode/filter/timestamp_tokens.rs:83-96) on the partition sums, log-softmax of the K
//                candidates; then BeamSearchTokenExtractor::update's host loop
//                (src/decode/extractor/beam_search.rs:49-146) on device
//   MODE_GREEDY  group = row: the same row merge followed by GreedyTokenExtractor::update (greedy.rs:24-50)
//   MODE_FILTER  group = row: the row merge, then the in-place fills of TimestampTokens::apply
//                (LogitFilter::apply equivalents, src/decode/filter/*.rs)
#include <algorithm>
#include <cmath>
#include <cstring>

#include "decode_scan.cuh"
#include "wrs_internal.cuh"

struct wrs_decoder {
    wrs_ctx* ctx = nullptr;
    wrs_decode_cfg cfg{};
    int max_rows = 0, max_audio = 0, max_cand = 0;
    int ts_split = 0, n_text = 0, text_len = 0, n_ts = 0, ts_len = 0, chunks = 0;   // chunk geometry of a row
    int bm_words = 0;
    uint32_t* d_bm_base = nullptr;    // suppress list (+ no_timestamps when timestamps are on)
    uint32_t* d_bm_first = nullptr;   // base + blank ids (first step, suppress_blank)
    uint32_t* d_bm_ts = nullptr;      // only no_timestamps (in-place timestamp filter)
    int32_t* d_suppress = nullptr; int n_suppress = 0;   // raw list for the in-place sparse fill
    int32_t* d_blank = nullptr;       // the two blank ids, for the in-place sparse fill
    wrs::Partial* d_partials = nullptr;   // [max_rows][chunks]
    int32_t* d_arrive = nullptr;      // [max_rows] self-resetting arrival counters
    // finished-sequence store (BeamSearchTokenExtractor::finished_sequences, beam_search.rs:13)
    int64_t* d_fin_tokens = nullptr;  // [max_audio][max_cand][max_len]
    int32_t* d_fin_len = nullptr;     // [max_audio][max_cand]
    double* d_fin_score = nullptr;    // [max_audio][max_cand]
    int32_t* d_fin_n = nullptr;       // [max_audio]
    int32_t* d_sync = nullptr;        // [2] arrive counter, full counter
};

namespace wrs {

enum { MODE_BEAM = 0, MODE_GREEDY = 1, MODE_FILTER = 2 };
constexpr int kMaxBeam = 15;
constexpr int kMaxCandAudio = kMaxBeam * kMaxK;          // 240
constexpr int kRowEntries = kMaxChunks * kMaxK;          // 144 staged candidates per row

struct StepParams {
    float* logits; int64_t row_stride;                   // read-only except MODE_FILTER
    const int64_t* tokens; int64_t tok_stride;
    int cur_len, sample_begin, n_rows, K, beam, n_groups;
    RuleCfg rc;
    int ts_split, n_text, text_len, n_ts, ts_len, chunks;
    const uint32_t* bm; const uint32_t* bm_first;
    Partial* partials; int32_t* arrive;
    int eot, max_cand, max_len, no_timestamps;
    const float* sum_in; float* sum_out;
    int64_t* next_tokens; int64_t next_stride; int32_t* source_idx; int32_t* completed;
    int64_t* fin_tokens; int32_t* fin_len; double* fin_score; int32_t* fin_n; int32_t* sync;
    float* dbg_lp; int32_t* dbg_tok; uint8_t* dbg_flags; int32_t* dbg_fin_count; int32_t* dbg_fin_source; double* dbg_fin_score;
    DevStatus* status;
};

struct TailSmem {
    float cv[kScanWarps][kRowEntries];
    int ci[kScanWarps][kRowEntries];
    float sel_v[kScanWarps][kMaxK];
    int sel_i[kScanWarps][kMaxK];
    float lp[kMaxBeam][kMaxK];
    int tok[kMaxBeam][kMaxK];
    unsigned char same[kMaxBeam][kMaxBeam];
    double c_score[kMaxCandAudio];
    short c_row[kMaxCandAudio], c_shadow[kMaxCandAudio], order[kMaxCandAudio];
    int c_tok[kMaxCandAudio];
    int sel_src[kMaxBeam], sel_tok[kMaxBeam], fin_src[kMaxBeam];
    double sel_score[kMaxBeam], fin_sc[kMaxBeam];
    int n_fin, store_base, n_store, nan_seen;
    unsigned flags;
};

// ---- row finish (one full warp per row): merge chunk partials, the timestamp-vs-text rule, K log-probs.
// Lane i < K returns candidate i; `flags` is the same on every lane.
struct RowResult { float lp; int tok; unsigned flags; };

__device__ __forceinline__ RowResult finish_row(const StepParams& p, const Partial* __restrict__ parts, unsigned rule_flags,
                                                TailSmem& ts) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int K = p.K;
    const bool mine = lane < p.chunks, text = lane < p.n_text;
    const float m_c = mine ? __ldcg(&parts[lane].m) : -CUDART_INF_F;
    const float s_c = mine ? __ldcg(&parts[lane].s) : 0.0f;
    float Tm = (mine && text) ? m_c : -CUDART_INF_F, Sm = (mine && !text) ? m_c : -CUDART_INF_F;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Tm = fmaxf(Tm, __shfl_xor_sync(0xffffffffu, Tm, o));
        Sm = fmaxf(Sm, __shfl_xor_sync(0xffffffffu, Sm, o));
    }
    float Ts = (mine && text) ? s_c * exp_diff(m_c, Tm) : 0.0f, Ss = (mine && !text) ? s_c * exp_diff(m_c, Sm) : 0.0f;
    if (s_c != s_c) { Ts = text ? s_c : Ts; Ss = text ? Ss : s_c; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Ts += __shfl_xor_sync(0xffffffffu, Ts, o);
        Ss += __shfl_xor_sync(0xffffffffu, Ss, o);
    }
    RowResult r;
    r.flags = rule_flags;
    if (Ts != Ts || Ss != Ss) r.flags |= 128u;
    // logsumexp(timestamps) > max(text): both sides carry the same -logsumexp(all), which cancels
    const float lse_s = (Ss > 0.0f) ? Sm + logf(Ss) : -CUDART_INF_F;
    const bool ts_over_text = p.rc.timestamps && (lse_s > Tm);
    if (ts_over_text) { r.flags |= 8u; Tm = -CUDART_INF_F; Ts = 0.0f; }
    // final log-softmax over what survived: (x - M) - log(sum exp(x - M)), the order ATen uses
    const float M = fmaxf(Tm, Sm);
    const float tot = Ts * exp_diff(Tm, M) + Ss * exp_diff(Sm, M);
    const float logtot = logf(tot);
    if (M == -CUDART_INF_F) r.flags |= 128u;   // everything masked: the reference yields NaN log-probs
    // stage the surviving chunks' candidates, then rank-count select the best K
    const int nent = p.chunks * kMaxK;
    float* cv = ts.cv[warp]; int* ci = ts.ci[warp];
    for (int e = lane; e < nent; e += 32) {
        const int c = e / kMaxK, i = e - c * kMaxK;
        const bool dead = (ts_over_text && c < p.n_text) || i >= K;
        cv[e] = dead ? -CUDART_INF_F : __ldcg(&parts[c].v[i]);
        ci[e] = dead ? INT32_MAX : __ldcg(&parts[c].id[i]);
    }
    if (lane < kMaxK) { ts.sel_v[warp][lane] = -CUDART_INF_F; ts.sel_i[warp][lane] = INT32_MAX; }
    __syncwarp();
    for (int e = lane; e < nent; e += 32) {
        const float v = cv[e]; const int id = ci[e];
        if (id == INT32_MAX) continue;
        int rank = 0;
        for (int d = 0; d < nent; ++d) rank += better(cv[d], ci[d], v, id) ? 1 : 0;
        if (rank < K) { ts.sel_v[warp][rank] = v; ts.sel_i[warp][rank] = id; }
    }
    __syncwarp();
    const float x = lane < K ? ts.sel_v[warp][lane] : -CUDART_INF_F;
    r.tok = lane < K ? ts.sel_i[warp][lane] : INT32_MAX;
    r.lp = (x - M) - logtot;
    __syncwarp();
    return r;
}

// ---- MODE_BEAM tail: BeamSearchTokenExtractor::update for one audio (beam_search.rs:49-146), whole CTA ----
__device__ __forceinline__ void beam_tail(const StepParams& p, int audio, TailSmem& ts) {
    const int k = p.beam, K = p.K;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int row0 = audio * k;
    if (tid == 0) ts.nan_seen = 0;
    for (int i = tid; i < k * k; i += kScanThreads) ts.same[i / k][i % k] = 1;
    __syncthreads();
    // 1. which beams share an identical prefix (HashMap key = whole sequence, beam_search.rs:74-77): all pairs,
    //    all positions in parallel
    {
        const int npairs = k * (k - 1) / 2, total = npairs * p.cur_len;
        for (int idx = tid; idx < total; idx += kScanThreads) {
            const int pr = idx / p.cur_len, pos = idx - pr * p.cur_len;
            int a = 0, q = pr;
            while (q >= k - 1 - a) { q -= k - 1 - a; ++a; }
            const int b = a + 1 + q;
            if (p.tokens[(int64_t)(row0 + a) * p.tok_stride + pos] != p.tokens[(int64_t)(row0 + b) * p.tok_stride + pos]) {
                ts.same[a][b] = 0; ts.same[b][a] = 0;
            }
        }
    }
    // 2. per-row finish (one warp per row)
    for (int j = warp; j < k; j += kScanWarps) {
        const int row = row0 + j;
        const RowRules rr = row_rules(p.rc, p.tokens + (int64_t)row * p.tok_stride, p.cur_len, p.sample_begin);
        const RowResult r = finish_row(p, p.partials + (int64_t)row * p.chunks, rr.flags, ts);
        if (lane < K) {
            ts.lp[j][lane] = r.lp; ts.tok[j][lane] = r.tok;
            if (p.dbg_lp) p.dbg_lp[(int64_t)row * K + lane] = r.lp;
            if (p.dbg_tok) p.dbg_tok[(int64_t)row * K + lane] = r.tok;
        }
        if (lane == 0) {
            if (r.flags & 128u) ts.nan_seen = 1;
            if (p.dbg_flags) p.dbg_flags[row] = (uint8_t)r.flags;
        }
    }
    __syncthreads();
    // 3. candidates: score = f64(sum_logprobs[row]) + f64(logprob)   (beam_search.rs:70-71)
    const int n_c = k * K;
    for (int c = tid; c < n_c; c += kScanThreads) {
        const int j = c / K, i = c - j * K;
        const int tok = ts.tok[j][i];
        // a later beam with the same prefix proposing the same token overwrites this map entry (:77)
        bool shadowed = false;
        for (int j2 = j + 1; j2 < k; ++j2)
            if (ts.same[j][j2])
                for (int i2 = 0; i2 < K; ++i2) shadowed |= (ts.tok[j2][i2] == tok);
        ts.c_score[c] = (double)p.sum_in[row0 + j] + (double)ts.lp[j][i];
        ts.c_row[c] = (short)j; ts.c_shadow[c] = (short)((shadowed || tok == INT32_MAX) ? 1 : 0); ts.c_tok[c] = tok;
    }
    __syncthreads();
    // 4. rank sort by (score desc, source row asc, token asc); shadowed entries go last
    for (int c = tid; c < n_c; c += kScanThreads) {
        int rank = 0;
        const double sc = ts.c_score[c];
        for (int d = 0; d < n_c; ++d) {
            if (d == c) continue;
            bool before;
            if (ts.c_shadow[d] != ts.c_shadow[c]) before = ts.c_shadow[d] < ts.c_shadow[c];
            else if (ts.c_score[d] != sc) before = ts.c_score[d] > sc;
            else if (ts.c_row[d] != ts.c_row[c]) before = ts.c_row[d] < ts.c_row[c];
            else before = ts.c_tok[d] < ts.c_tok[c] || (ts.c_tok[d] == ts.c_tok[c] && d < c);
            rank += before ? 1 : 0;
        }
        ts.order[rank] = (short)c;
    }
    __syncthreads();
    // 5. walk (beam_search.rs:87-104): EOT-terminated -> finished, others kept until k are saved
    if (tid == 0) {
        int saved = 0, nf = 0;
        for (int o = 0; o < n_c && saved < k; ++o) {
            const int c = ts.order[o];
            if (ts.c_shadow[c]) break;   // only shadowed entries remain
            if (ts.c_tok[c] == p.eot) { if (nf < kMaxBeam) { ts.fin_src[nf] = ts.c_row[c]; ts.fin_sc[nf] = ts.c_score[c]; ++nf; } }
            else { ts.sel_src[saved] = ts.c_row[c]; ts.sel_tok[saved] = ts.c_tok[c]; ts.sel_score[saved] = ts.c_score[c]; ++saved; }
        }
        // fewer than k live candidates cannot happen with >= k+1 finite logits per row; keep the row count anyway
        for (; saved < k; ++saved) { ts.sel_src[saved] = 0; ts.sel_tok[saved] = p.eot; ts.sel_score[saved] = -CUDART_INF; }
        ts.n_fin = nf;
        // finished store: append in descending score until round(patience*k) entries (:124-138)
        const int have = p.fin_n[audio];
        const int add = min(nf, max(0, p.max_cand - have));
        ts.store_base = have; ts.n_store = add;
        p.fin_n[audio] = have + add;
        if (p.dbg_fin_count) p.dbg_fin_count[audio] = nf;
        for (int i = 0; i < k; ++i) {
            if (p.dbg_fin_source) p.dbg_fin_source[audio * k + i] = i < nf ? row0 + ts.fin_src[i] : -1;
            if (p.dbg_fin_score) p.dbg_fin_score[audio * k + i] = i < nf ? ts.fin_sc[i] : 0.0;
        }
        if (ts.nan_seen) atomicAdd(&p.status->nan_rows, 1);
    }
    __syncthreads();
    // 6. outputs: all rows in parallel
    for (int i = tid; i < k; i += kScanThreads) {
        p.source_idx[row0 + i] = row0 + ts.sel_src[i];
        p.sum_out[row0 + i] = (float)ts.sel_score[i];   // index_put_(.., score as f32) :91-95
    }
    const int L1 = p.cur_len + 1;
    for (int idx = tid; idx < k * L1; idx += kScanThreads) {
        const int i = idx / L1, t = idx - i * L1;
        p.next_tokens[(int64_t)(row0 + i) * p.next_stride + t] =
            t < p.cur_len ? p.tokens[(int64_t)(row0 + ts.sel_src[i]) * p.tok_stride + t] : (int64_t)ts.sel_tok[i];
    }
    for (int idx = tid; idx < ts.n_store * L1; idx += kScanThreads) {
        const int i = idx / L1, t = idx - i * L1;
        const int slot = ts.store_base + i;
        p.fin_tokens[((int64_t)audio * p.max_cand + slot) * p.max_len + t] =
            t < p.cur_len ? p.tokens[(int64_t)(row0 + ts.fin_src[i]) * p.tok_stride + t] : (int64_t)p.eot;
        if (t == 0) {
            p.fin_len[audio * p.max_cand + slot] = L1;
            p.fin_score[audio * p.max_cand + slot] = ts.fin_sc[i];
        }
    }
    // 7. completed = all audios hold max_cand finished sequences (:141-146); last audio to arrive publishes
    if (tid == 0) {
        const int full = (ts.store_base + ts.n_store >= p.max_cand) ? 1 : 0;
        if (full) atomicAdd(p.sync + 1, 1);
        __threadfence();
        const int arrived = atomicAdd(p.sync, 1);
        if (arrived == p.n_groups - 1) {
            __threadfence();
            const int nfull = atomicExch(p.sync + 1, 0);
            *p.completed = (nfull == p.n_groups) ? 1 : 0;
            atomicExch(p.sync, 0);
        }
    }
}

// ---- MODE_GREEDY tail: GreedyTokenExtractor::update for one row (src/decode/extractor/greedy.rs:24-50) ----
__device__ __forceinline__ void greedy_tail(const StepParams& p, int row, TailSmem& ts) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t* rt = p.tokens + (int64_t)row * p.tok_stride;
    int64_t* dst = p.next_tokens + (int64_t)row * p.next_stride;
    for (int t = tid; t < p.cur_len; t += kScanThreads) dst[t] = rt[t];
    if (warp != 0) return;
    const RowRules rr = row_rules(p.rc, rt, p.cur_len, p.sample_begin);
    const RowResult r = finish_row(p, p.partials + (int64_t)row * p.chunks, rr.flags, ts);
    if (lane != 0) return;
    const bool last_eot = rt[p.cur_len - 1] == p.eot;                       // greedy.rs:37-38
    const int next = last_eot ? p.eot : r.tok;                               // :41-45
    p.sum_out[row] = p.sum_in[row] + (last_eot ? 0.0f : r.lp);               // :39
    dst[p.cur_len] = next;
    if (p.dbg_lp) p.dbg_lp[row] = r.lp;
    if (p.dbg_tok) p.dbg_tok[row] = r.tok;
    if (p.dbg_flags) p.dbg_flags[row] = (uint8_t)r.flags;
    if (r.flags & 128u) atomicAdd(&p.status->nan_rows, 1);
    if (next == p.eot) atomicAdd(p.sync + 1, 1);
    __threadfence();
    const int arrived = atomicAdd(p.sync, 1);
    if (arrived == p.n_groups - 1) {                                         // completed: every row's new last token is EOT
        __threadfence();
        const int n_eot = atomicExch(p.sync + 1, 0);
        *p.completed = (n_eot == p.n_rows) ? 1 : 0;
        atomicExch(p.sync, 0);
    }
}

// ---- MODE_FILTER tail: timestamp_tokens.rs:37-96 in place for one row.  The statistics were taken on the
// logits as they are (earlier filters already wrote their -inf), so the decision sees what the reference's
// log_softmax sees.  Every chunk of the row has been read by now: the fills cannot race the scan.
__device__ __forceinline__ void filter_tail(const StepParams& p, int row, TailSmem& ts) {
    const int tid = threadIdx.x;
    const RowRules rr = row_rules(p.rc, p.tokens + (int64_t)row * p.tok_stride, p.cur_len, p.sample_begin);
    if (tid < 32) {
        const RowResult r = finish_row(p, p.partials + (int64_t)row * p.chunks, rr.flags, ts);
        if (tid == 0) {
            ts.flags = r.flags;
            if (p.dbg_flags) p.dbg_flags[row] = (uint8_t)r.flags;
        }
    }
    __syncthreads();
    const bool kill_text = (ts.flags & 8u) != 0;
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

template <int MODE>
__global__ void __launch_bounds__(kScanThreads, 2) decode_step_kernel(const StepParams p) {
    __shared__ ScanSmem sm;
    __shared__ TailSmem ts;
    __shared__ int s_last;
    const int row = blockIdx.y, chunk = blockIdx.x;
    const int tid = threadIdx.x;
    const RowRules rr = row_rules(p.rc, p.tokens + (int64_t)row * p.tok_stride, p.cur_len, p.sample_begin);
    const uint32_t* bm = (p.cur_len == p.sample_begin) ? p.bm_first : p.bm;
    int lo, hi;
    if (chunk < p.n_text) {
        const int c_lo = chunk * p.text_len;
        lo = max(c_lo, rr.t_lo); hi = min(min(p.ts_split, c_lo + p.text_len), rr.t_hi);
    } else {
        const int c_lo = p.ts_split + (chunk - p.n_text) * p.ts_len;
        lo = max(c_lo, rr.s_lo); hi = min(min(p.rc.n_vocab, c_lo + p.ts_len), rr.s_hi);
    }
    scan_chunk(p.logits + (int64_t)row * p.row_stride, lo, hi, bm, p.K, sm, p.partials + (int64_t)row * p.chunks + chunk);
    // arrive: the partial was written by warp 0 (or by threads 0..15 on the empty path)
    const int group = (MODE == MODE_BEAM) ? row / p.beam : row;
    const int expected = ((MODE == MODE_BEAM) ? p.beam : 1) * p.chunks;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const int prev = atomicAdd(p.arrive + group, 1);
        s_last = (prev == expected - 1) ? 1 : 0;
        if (s_last) p.arrive[group] = 0;         // self-reset for the next launch (stream-ordered)
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (MODE == MODE_BEAM) beam_tail(p, group, ts);
    else if (MODE == MODE_GREEDY) greedy_tail(p, group, ts);
    else filter_tail(p, group, ts);
}

// ---- sparse in-place fills: logits.index_fill_(1, idx, -inf)   suppress_tokens.rs:22, suppress_blanks.rs:41
__global__ void filter_index_fill_kernel(float* __restrict__ logits, int64_t row_stride, int n_rows, int n_vocab,
                                         const int32_t* __restrict__ ids, int n_ids) {
    const int64_t total = (int64_t)n_rows * n_ids;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / n_ids), id = ids[i - (int64_t)r * n_ids];
        if (id >= 0 && id < n_vocab) logits[(int64_t)r * row_stride + id] = -CUDART_INF_F;
    }
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

// common part of StepParams
StepParams step_params(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, const int64_t* tokens,
                       int64_t tok_stride, int64_t cur_len, int64_t sample_begin, int K, int beam, const uint32_t* bm,
                       const uint32_t* bm_first) {
    StepParams p;
    memset(&p, 0, sizeof(p));
    p.logits = const_cast<float*>(logits); p.row_stride = row_stride; p.tokens = tokens; p.tok_stride = tok_stride;
    p.cur_len = (int)cur_len; p.sample_begin = (int)sample_begin; p.n_rows = (int)n_rows; p.K = K; p.beam = beam;
    p.n_groups = (int)(n_rows / beam);
    p.rc = rule_cfg(dec->cfg);
    p.ts_split = dec->ts_split; p.n_text = dec->n_text; p.text_len = dec->text_len; p.n_ts = dec->n_ts; p.ts_len = dec->ts_len;
    p.chunks = dec->chunks;
    p.bm = bm; p.bm_first = bm_first; p.partials = dec->d_partials; p.arrive = dec->d_arrive;
    p.eot = dec->cfg.eot; p.max_cand = dec->max_cand; p.max_len = dec->cfg.max_len; p.no_timestamps = dec->cfg.no_timestamps;
    p.fin_tokens = dec->d_fin_tokens; p.fin_len = dec->d_fin_len; p.fin_score = dec->d_fin_score; p.fin_n = dec->d_fin_n;
    p.sync = dec->d_sync; p.status = dec->ctx->d_status;
    return p;
}

void set_debug(StepParams& p, const wrs_step_debug* dbg) {
    if (!dbg) return;
    p.dbg_lp = dbg->cand_logprob; p.dbg_tok = dbg->cand_token; p.dbg_flags = dbg->row_flags;
    p.dbg_fin_count = dbg->fin_count; p.dbg_fin_source = dbg->fin_source; p.dbg_fin_score = dbg->fin_score;
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
    // chunk geometry: text ids [0, ts_split) and timestamp ids [ts_split, V) never share a CTA
    d->ts_split = c.timestamps ? c.ts_begin : c.n_vocab;
    d->n_text = (d->ts_split + kChunkCap - 1) / kChunkCap;
    d->text_len = d->n_text ? ((((d->ts_split + d->n_text - 1) / d->n_text) + 3) & ~3) : 0;
    const int n_tsid = c.n_vocab - d->ts_split;
    d->n_ts = (n_tsid + kChunkCap - 1) / kChunkCap;
    d->ts_len = d->n_ts ? ((((n_tsid + d->n_ts - 1) / d->n_ts) + 3) & ~3) : 0;
    d->chunks = d->n_text + d->n_ts;
    if (d->chunks > kMaxChunks || d->text_len > kChunkCap || d->ts_len > kChunkCap) {
        delete d;
        return fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_decoder_create: vocabulary too large for 9 chunks of 14333 ids per row");
    }
    d->bm_words = (c.n_vocab + 31) / 32 + 2;
    std::vector<uint32_t> base(d->bm_words, 0u), first, ts(d->bm_words, 0u);
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
        (e = up(&d->d_bm_ts, ts)) != cudaSuccess) {
        wrs_decoder_destroy(d);
        return cuda_fail(ctx, e, "decoder bitmaps");
    }
    d->n_suppress = n_suppress;
    const size_t fin_tok = sizeof(int64_t) * (size_t)d->max_audio * d->max_cand * c.max_len;
    if ((n_suppress > 0 && ((e = cudaMalloc(&d->d_suppress, sizeof(int32_t) * n_suppress)) != cudaSuccess ||
                            (e = cudaMemcpy(d->d_suppress, suppress_ids, sizeof(int32_t) * n_suppress, cudaMemcpyHostToDevice)) != cudaSuccess)) ||
        (e = cudaMalloc(&d->d_partials, sizeof(Partial) * d->chunks * (size_t)max_rows)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_arrive, sizeof(int32_t) * max_rows)) != cudaSuccess ||
        (e = cudaMemset(d->d_arrive, 0, sizeof(int32_t) * max_rows)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_blank, sizeof(int32_t) * 2)) != cudaSuccess ||
        (e = cudaMemcpy(d->d_blank, c.blank_ids, sizeof(int32_t) * 2, cudaMemcpyHostToDevice)) != cudaSuccess ||
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
    cudaFree(d->d_bm_base); cudaFree(d->d_bm_first); cudaFree(d->d_bm_ts); cudaFree(d->d_arrive); cudaFree(d->d_blank);
    cudaFree(d->d_suppress); cudaFree(d->d_partials); cudaFree(d->d_fin_tokens); cudaFree(d->d_fin_len);
    cudaFree(d->d_fin_score); cudaFree(d->d_fin_n); cudaFree(d->d_sync);
    delete d;
}

int wrs_decoder_reset(wrs_decoder* d, wrs_stream_t stream) {
    if (!d) return fail(nullptr, WRS_ERR_INVALID, "wrs_decoder_reset: null decoder");
    WRS_CUDA(d->ctx, cudaSetDevice(d->ctx->device));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_fin_n, 0, sizeof(int32_t) * d->max_audio, (cudaStream_t)stream));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_sync, 0, sizeof(int32_t) * 2, (cudaStream_t)stream));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_arrive, 0, sizeof(int32_t) * d->max_rows, (cudaStream_t)stream));
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
        filter_index_fill_kernel<<<(unsigned)std::min<int64_t>((n_rows * 2 + 255) / 256, 1024), 256, 0, stream>>>(
            logits, row_stride, (int)n_rows, c.n_vocab, dec->d_blank, 2);
        WRS_CUDA(ctx, cudaGetLastError());
    }
    if ((which & WRS_FILTER_SUPPRESS_TOKENS) && dec->n_suppress > 0) {              // suppress_tokens.rs:22
        filter_index_fill_kernel<<<(unsigned)std::min<int64_t>((n_rows * dec->n_suppress + 255) / 256, 1024), 256, 0, stream>>>(
            logits, row_stride, (int)n_rows, c.n_vocab, dec->d_suppress, dec->n_suppress);
        WRS_CUDA(ctx, cudaGetLastError());
    }
    if (which & WRS_FILTER_TIMESTAMPS) {                                             // timestamp_tokens.rs:33-98
        if (!c.timestamps) return fail(ctx, WRS_ERR_INVALID, "wrs_filter_apply: decoder was created with timestamps = 0");
        StepParams p = step_params(dec, logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin, 1, 1,
                                   dec->d_bm_ts, dec->d_bm_ts);
        p.dbg_flags = row_flags;
        decode_step_kernel<MODE_FILTER><<<dim3(dec->chunks, (unsigned)n_rows), kScanThreads, 0, stream>>>(p);
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
    StepParams p = step_params(dec, logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin, c.beam_size + 1,
                               c.beam_size, dec->d_bm_base, dec->d_bm_first);
    p.sum_in = sum_in; p.sum_out = sum_out; p.next_tokens = next_tokens; p.next_stride = next_stride;
    p.source_idx = source_idx; p.completed = completed;
    set_debug(p, dbg);
    decode_step_kernel<MODE_BEAM><<<dim3(dec->chunks, (unsigned)n_rows), kScanThreads, 0, stream>>>(p);
    WRS_CUDA(ctx, cudaGetLastError());
    return WRS_OK;
}

int wrs_greedy_step(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, const int64_t* tokens,
                    int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_in, int64_t* next_tokens,
                    int64_t next_stride, float* sum_out, int32_t* completed, const wrs_step_debug* dbg, wrs_stream_t stream_) {
    int rc = check_step_args(dec, "wrs_greedy_step", logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin);
    if (rc != WRS_OK) return rc;
    wrs_ctx* ctx = dec->ctx;
    if (!sum_in || !next_tokens || !sum_out || !completed) return fail(ctx, WRS_ERR_INVALID, "wrs_greedy_step: null output");
    if (next_stride < cur_len + 1) return fail(ctx, WRS_ERR_INVALID, "wrs_greedy_step: next_stride < cur_len + 1");
    if ((const void*)next_tokens == (const void*)tokens) return fail(ctx, WRS_ERR_INVALID, "wrs_greedy_step: next_tokens must not alias tokens");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    StepParams p = step_params(dec, logits, row_stride, n_rows, tokens, tok_stride, cur_len, sample_begin, 1, 1,
                               dec->d_bm_base, dec->d_bm_first);
    p.sum_in = sum_in; p.sum_out = sum_out; p.next_tokens = next_tokens; p.next_stride = next_stride; p.completed = completed;
    set_debug(p, dbg);
    decode_step_kernel<MODE_GREEDY><<<dim3(dec->chunks, (unsigned)n_rows), kScanThreads, 0, stream>>>(p);
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

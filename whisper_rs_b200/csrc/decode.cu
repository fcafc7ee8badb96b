// K2: the fused per-step logit processing (filters + log-softmax + top-k + beam merge) for sm_100a, two launches per
// step, no host round trip.
//
//   decode_scan_kernel        one CTA of 24 warps per SM.  Every row's units (decode_scan.cuh) are cut into runs that never
//       cross the text / timestamp partition; a warp folds a run: bulk copies into its shared-memory stages, suppressed
//       ids poked to -inf in the stage, rule ranges as register masks, then {max, sum exp, the K best CELLS} -- a cell is
//       the 28 ids a lane holds of a unit, valued by their maximum (fold_unit).  One Partial per run; the logits are read
//       once and never written.
//   decode_tail_kernel<MODE>  one CTA per audio (MODE_BEAM) or row, one warp per row.  merge_row: partition sums,
//       timestamp-vs-text rule (src/decode/filter/timestamp_tokens.rs:83-96), k-way merge of the runs' cell lists, the K
//       winning cells read again from the logits (28 ids each), the K best ids and their log-softmax.  Then
//   MODE_BEAM    BeamSearchTokenExtractor::update's host loop (src/decode/extractor/beam_search.rs:49-146) on device,
//                spread over the audio's warps (beam <= 5) or on warp 0 (beam_tail_general)
//   MODE_GREEDY  GreedyTokenExtractor::update for the row (greedy.rs:24-50)
//   MODE_FILTER  the in-place fills of TimestampTokens::apply (LogitFilter::apply equivalents, src/decode/filter/*.rs)
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "async_copy.cuh"
#include "decode_scan.cuh"
#include "decode_proj.cuh"
#include "logmel_fft.cuh"   // packed FP32 helpers (f32x2)
#include "wrs_internal.cuh"

namespace wrs { struct RowResult; }

struct wrs_decoder {
    wrs_ctx* ctx = nullptr;
    wrs_decode_cfg cfg{};
    int max_rows = 0, max_audio = 0, max_cand = 0;
    int ts_split = 0, n_text_units = 0, n_ts_units = 0, upr = 0;   // unit geometry of a row
    int bm_words = 0;
    uint32_t* d_bm_base = nullptr;    // suppress list (+ no_timestamps when timestamps are on)
    uint32_t* d_bm_first = nullptr;   // base + blank ids (first step, suppress_blank)
    uint32_t* d_bm_ts = nullptr;      // only no_timestamps (in-place timestamp filter)
    int32_t* d_suppress = nullptr; int n_suppress = 0;   // raw list for the in-place sparse fill
    int32_t* d_blank = nullptr;       // the two blank ids, for the in-place sparse fill
    wrs::Partial* d_partials = nullptr;   // [max_rows][upr]
    wrs::RowResult* d_rowres = nullptr;   // [max_rows]
    int32_t* d_arrive = nullptr;      // [max_rows + max_audio] self-resetting arrival counters (rows, then audios)
    // finished-sequence store (BeamSearchTokenExtractor::finished_sequences, beam_search.rs:13)
    int64_t* d_fin_tokens = nullptr;  // [max_audio][max_cand][max_len]
    int32_t* d_fin_len = nullptr;     // [max_audio][max_cand]
    double* d_fin_score = nullptr;    // [max_audio][max_cand]
    int32_t* d_fin_n = nullptr;       // [max_audio]
    int32_t* d_sync = nullptr;        // [2] arrive counter, full counter
    bool smem_attr_set = false;
};

// the token embedding of one model packed for proj_kernel, bound to a decoder's vocab geometry
struct wrs_projection {
    wrs_decoder* dec = nullptr;
    int n_state = 0, n_tiles = 0, n_text_tiles = 0, n_kstages = 0, max_mtiles = 0;
    unsigned char* d_wp = nullptr;    // packed weights (f16 hi / lo of w x w_scale)
    float w_scale = 1.0f;             // a power of two: max |w| x w_scale in [512, 1024)
    unsigned char* d_xp = nullptr;    // packed x of the step in flight, [max_mtiles]
    float* d_parts = nullptr;         // [max_rows][2 + 2 (beam + 1)][tiles padded to 4] record planes
    bool attr_set = false;
};

namespace wrs {

enum { MODE_BEAM = 0, MODE_GREEDY = 1, MODE_FILTER = 2 };
constexpr int kMaxBeam = 15;
constexpr int kMaxCandAudio = kMaxBeam * kMaxK;          // 240
#ifndef WRS_STAGES
#define WRS_STAGES 2
#endif
constexpr int kStages = WRS_STAGES;                               // units a warp keeps in flight (bulk copies, one mbarrier each)
constexpr int kStageBytes = kUnit * 4;                    // landing zone of one unit
constexpr int kScratchBytes = 2688;                      // tail launch: rest of a warp's slice behind kStageBytes
constexpr int kTailBytes = kStageBytes + kScratchBytes;  // per-warp slice of the tail launch
constexpr int kTailLists = 2 * 4 * 32 * kMaxK;           // the row merge's cell lists, [entry][lane] values then ids
constexpr int kScanScratch = 1056;                       // scan launch: per-warp scratch behind stage 0 (WarpBuf only)
constexpr int kWarpSmem = kStages * kStageBytes + kScanScratch + 32;   // + the stages' mbarriers
constexpr int kTailRows = kMaxBeam * 144;                // tail launch: the group's row results ...
constexpr int kTailHead = kTailRows + 3200;              // ... and the fast beam tail's tables, in front of the warp slices

// what the row merge leaves for the audio-level tail
struct __align__(16) RowResult {                        // 144 bytes (kTailHead)
    float lp[kMaxK];
    int tok[kMaxK];
    unsigned flags;
    unsigned same;         // bit j2: beam j2 (> this row's beam index) holds an identical token prefix
    int pad[2];
};

static_assert(sizeof(RowResult) * kMaxBeam <= kTailRows, "row results do not fit the head of the tail launch's shared memory");

struct StepParams {
    float* logits; int64_t row_stride;                   // read-only except MODE_FILTER
    const int64_t* tokens; int64_t tok_stride;
    int cur_len, sample_begin, n_rows, K, beam, n_groups;
    // unit dealing of the scan launch (run_in_row), set by launch_step: every row is cut into rt text runs and rs
    // timestamp runs of consecutive units; a run never crosses a partition
    int rt, rs, rpr, n_runs, t_base, t_rem, s_base, s_rem; unsigned long long rpr_magic;
    int npw;                                             // runs per warp of the scan launch: warp w takes runs [w npw, (w + 1) npw)
    RuleCfg rc;
    int ts_split, n_text_units, n_ts_units, upr, bm_words;
    const uint32_t* bm;
    Partial* partials; RowResult* rowres; int32_t* row_arrive; int32_t* audio_arrive;
    int eot, max_cand, max_len, no_timestamps;
    const float* sum_in; float* sum_out;
    int64_t* next_tokens; int64_t next_stride; int32_t* source_idx; int32_t* completed;
    int64_t* fin_tokens; int32_t* fin_len; double* fin_score; int32_t* fin_n; int32_t* sync;
    const float* proj_parts; int proj_tiles, proj_text_tiles, proj_tp;   // projection mode: the tail merges proj_kernel's records (decode_proj.cuh)
    float* dbg_lp; int32_t* dbg_tok; uint8_t* dbg_flags; int32_t* dbg_fin_count; int32_t* dbg_fin_source; double* dbg_fin_score;
    DevStatus* status;
};

#ifdef WRS_TRACE
// investigation aid (tools/debug/dec_trace.py): per-warp event times in ns, [warp][8]
__device__ unsigned long long g_trace[(148 * kStepWarps + 1024) * 8];   // scan warps, then tail rows
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define WRS_TR(slot) do { if (lane == 0) g_trace[gw * 8 + (slot)] = gtime(); } while (0)
#define WRS_TR_ADD(slot, v) do { if (lane == 0) g_trace[gw * 8 + (slot)] += (v); } while (0)
#define WRS_TR_CNT(slot, v) do { if ((threadIdx.x & 31) == 0) g_trace[((threadIdx.x >> 5) * gridDim.x + blockIdx.x) * 8 + (slot)] += (v); } while (0)
#else
#define WRS_TR_CNT(slot, v) do {} while (0)
#define WRS_TR(slot) do {} while (0)
#define WRS_TR_ADD(slot, v) do {} while (0)
#endif

// geometry of one unit as a warp sees it (all warp-uniform)
struct Geom {
    int row, part;      // part 0 = text ids, 1 = timestamp ids
    int lo, hi;         // surviving ids of the unit [lo, hi) after the row's rules (hi <= lo: nothing to read)
    int a0;             // id of the unit's first element slot (16-byte aligned in the row's memory; may be < lo)
};

// SuppressTokens / SuppressBlank (suppress_tokens.rs:22, suppress_blanks.rs:41) on a unit that has landed in its
// shared-memory stage: lane l holds bitmap word word0 + l (bit set = masked) and writes -inf over the ids it names.
// Most words are zero, so this costs a test; the registers are loaded after it.
__device__ __forceinline__ void poke_suppressed(float* st, uint32_t bmw, int word0, int a0) {
    if (bmw) {
        const int base = ((word0 + (int)(threadIdx.x & 31)) << 5) - a0;       // slot of the word's bit 0
        do {
            const unsigned pos = (unsigned)(base + __ffs((int)bmw) - 1);
            bmw &= bmw - 1u;
            if (pos < (unsigned)kUnit) st[pos] = -CUDART_INF_F;
        } while (bmw);
    }
}

// The statistics of one unit (already in registers, masked ids loaded as -inf) folded into the run: the lane maximum is
// returned, the exp-sum goes against the run's reference (no cross-lane traffic unless the reference moves).
__device__ __forceinline__ float fold_stats(float (&x)[kNV][4], Run& run) {
    float tmax = fmaxf(fmaxf(x[0][0], x[0][1]), fmaxf(x[0][2], x[0][3]));
#pragma unroll
    for (int j = 1; j < kNV; ++j) tmax = fmaxf(tmax, fmaxf(fmaxf(x[j][0], x[j][1]), fmaxf(x[j][2], x[j][3])));
    run.lm = fmaxf(run.lm, tmax);
    // the reference lives in the log2 domain and is an INTEGER R = rint(max * log2 e): x * log2 e - R is then one fused
    // operation with one rounding of a small number (a rounded -ref * log2 e would shift every term of the run by the
    // same ~1e-5 relative at |logit| ~ 100), and moving the reference scales the sum by an exact power of two
    if (!run.ref_set || __any_sync(0xffffffffu, fmaf(tmax, kLog2e, -run.ref) > 92.0f)) {
        const float wm = warp_max(tmax);
        if (wm > -CUDART_INF_F) {
            const float R = rintf(wm * kLog2e);
            if (run.ref_set) run.ls *= ex2(run.ref - R);
            run.ref = R; run.ref_set = true;
        }
    }
    // packed FP32 (FFMA2 / FADD2): the four sums of a lane run side by side; NaN inputs keep the sum NaN
    const f32x2 l2e = pack2(kLog2e, kLog2e), nref = pack2(-run.ref, -run.ref);
    f32x2 s01 = pack2(0.0f, 0.0f), s23 = s01;
#pragma unroll
    for (int j = 0; j < kNV; ++j) {
        float a0, a1, a2, a3;
        unpack2(fma_(pack2(x[j][0], x[j][1]), l2e, nref), a0, a1);
        unpack2(fma_(pack2(x[j][2], x[j][3]), l2e, nref), a2, a3);
        s01 = add_(s01, pack2(ex2(a0), ex2(a1)));
        s23 = add_(s23, pack2(ex2(a2), ex2(a3)));
    }
    float t0, t1;
    unpack2(add_(s01, s23), t0, t1);
    run.ls += t0 + t1;
    return tmax;
}

// Fold one unit into the run with a top-K list of IDS: what the tail's rescan_row uses when it reads a whole row again.
__device__ __forceinline__ void reduce_unit(float (&x)[kNV][4], const Geom& g, int K, WarpBuf& wb, Run& run) {
    const int lane = threadIdx.x & 31;
    const float tmax = fold_stats(x, run);
    // gather everything that reaches tau, a value the run's K-th best is known to reach
    if (run.tau == -CUDART_INF_F) run.tau = kth_threshold(tmax, K);      // at least K elements of this unit reach it
    const float tau = run.tau;
    const bool mine = tmax >= tau && tmax > -CUDART_INF_F;
    const unsigned hb = __ballot_sync(0xffffffffu, mine);
    if (!hb) return;
    const int np0 = run.np;
    int c = 0;
    if (mine) {
#pragma unroll
        for (int j = 0; j < kNV; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) c += (x[j][i] >= tau && x[j][i] > -CUDART_INF_F) ? 1 : 0;
    }
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    const int n_tot = np0 + __shfl_sync(0xffffffffu, incl, 31);
    if (n_tot > kCandCap - kMaxK) {
        // massive ties: merge what was pending before this unit, then select exactly straight from the registers
        run_merge(run, wb, np0, K);
        run_exact<4 * kNV>(run, K, [&](int e) { return x[e >> 2][e & 3]; },
                           [&](int e) { return g.a0 + 4 * (lane + 32 * (e >> 2)) + (e & 3); });
    } else {
        if (mine) {
            int w = np0 + incl - c;
#pragma unroll
            for (int j = 0; j < kNV; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (x[j][i] >= tau && x[j][i] > -CUDART_INF_F) {
                        wb.v[w] = x[j][i]; wb.id[w] = g.a0 + 4 * (lane + 32 * j) + i; ++w;
                    }
        }
        __syncwarp();
        // once the list is full its K-th value is an exact bar and later units add only a few stragglers: those wait for a
        // cheap register merge (list + pending <= 32 entries); the first unit merges at once to set that bar
        if (run.n >= K && n_tot + K <= 32) { run.np = n_tot; return; }
        run_merge(run, wb, n_tot, K);
    }
    run.np = 0;
    run.tau = fmaxf(run.tau, __shfl_sync(0xffffffffu, run.rv, K - 1));   // -inf while the list holds fewer than K
    __syncwarp();
}

// Fold one unit into the run, scan launch: the top-K list holds CELLS, not ids.  A cell
// is (unit, lane): the 28 ids a lane holds of a unit, valued by their maximum.  The K best elements of a row lie in its K
// best cells (K cells at or above a value are K distinct elements at or above it), so the scan keeps the best cells only
// -- one candidate per lane and unit, no per-element work -- and the tail reads the few winning cells again.  Cells
// order by (value desc, cell id asc); a dropped cell that equals the list's K-th value raises run.tie, and the tail then
// scans the row itself (exact ties between the best cells of a row: quantised logits, not real ones).
__device__ __forceinline__ void fold_unit(float (&x)[kNV][4], int unit_in_row, int K, WarpBuf& wb, Run& run) {
    const int lane = threadIdx.x & 31;
    const float tmax = fold_stats(x, run);
    // the unit's cells that reach tau, a value the run's K-th best cell is known to reach
    if (run.tau == -CUDART_INF_F) run.tau = kth_threshold(tmax, K);
    bool hit = tmax >= run.tau && tmax > -CUDART_INF_F;
    unsigned hb = __ballot_sync(0xffffffffu, hit);
    if (!hb) return;
    if (run.np > 0 && run.np + __popc(hb) + K > 32) {     // keep the merges one entry per lane
        run_merge(run, wb, run.np, K);
        run.np = 0;
        run.tau = fmaxf(run.tau, __shfl_sync(0xffffffffu, run.rv, K - 1));
        hit = tmax >= run.tau && tmax > -CUDART_INF_F;
        hb = __ballot_sync(0xffffffffu, hit);
        if (!hb) return;
    }
    if (hit) {
        const int pos = run.np + __popc(hb & ((1u << lane) - 1u));
        wb.v[pos] = tmax; wb.id[pos] = unit_in_row * 32 + lane;
    }
    __syncwarp();
    const int n_tot = run.np + __popc(hb);
    // the cells wait for one merge at the end of the run (tau, from the first unit, is a bar K of them reach); a long run
    // merges whenever the next unit's cells might not fit a lane each, and that raises the bar
    if (n_tot + K <= 32) { run.np = n_tot; return; }
    run_merge(run, wb, n_tot, K);
    run.np = 0;
    run.tau = fmaxf(run.tau, __shfl_sync(0xffffffffu, run.rv, K - 1));   // -inf while the list holds fewer than K
    __syncwarp();
}

// End of a run: merge the pending candidates and turn the lane-local statistics into {max, sum exp(x - max)}.
__device__ __forceinline__ void run_finish(Run& run, WarpBuf& wb, int K) {
    if (run.np > 0) run_merge(run, wb, run.np, K);
    run.np = 0;
    run.m = warp_max(run.lm);
    float s = run.ls;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    // sum exp(x - m) = 2^(R - m log2 e) * sum 2^(x log2 e - R); R is within half a unit of a seen maximum's image: no overflow
    run.s = (s != s || !run.ref_set) ? s : s * ex2(fmaf(-run.m, kLog2e, run.ref));
}

// ---- A row's units are cut into RUNS that never cross a partition: rt runs over the n_text_units text units, rs runs
//      over the timestamp units, each as even as possible (launch_step picks rt and rs so that the grid's warps get one
//      run each when the rows allow it).  A run is folded by one warp of the scan launch and leaves ONE partial, in the
//      slot of its first unit.  All constants come from the host: neither launch divides.
struct RunGeo { int part, j0, len; };                 // partition, first unit in row, units
__device__ __forceinline__ RunGeo run_in_row(const StepParams& p, int i) {   // i = run of the row, [text runs | timestamp runs]
    RunGeo r;
    if (i < p.rt) { r.part = 0; r.j0 = i * p.t_base + min(i, p.t_rem); r.len = p.t_base + (i < p.t_rem ? 1 : 0); }
    else {
        i -= p.rt;
        r.part = 1; r.j0 = p.n_text_units + i * p.s_base + min(i, p.s_rem); r.len = p.s_base + (i < p.s_rem ? 1 : 0);
    }
    return r;
}

// ---- the tail reads logits again, cell by cell: the 28 ids that lane L of the scan held of the unit whose first slot id
//      is a0 (7 pieces of 16 bytes), with the masks the scan applied -- the partition's surviving range [lo, hi) and the
//      step's suppress bitmap.  Pieces that would reach outside the logits buffer are read id by id.
__device__ __forceinline__ void load_cell(const StepParams& p, const float* __restrict__ rp, int row, int a0, int L, int lo, int hi,
                                          bool active, float (&x)[kNV][4]) {
    // all loads first, without branches (a piece that is not wanted reads a harmless aligned address): one round trip
    float4 t[kNV]; uint32_t w0[kNV], w1[kNV]; bool need[kNV], inside[kNV];
    const float4* harmless = reinterpret_cast<const float4*>(p.partials);
    const bool first_row = row == 0, last_row = row == p.n_rows - 1;
#pragma unroll
    for (int s = 0; s < kNV; ++s) {
        const int e0 = a0 + 4 * (L + 32 * s);
        need[s] = active && e0 < hi && e0 + 3 >= lo;
        inside[s] = !(first_row && e0 < 0) && !(last_row && e0 + 3 >= p.rc.n_vocab);
        t[s] = __ldcs((need[s] && inside[s]) ? reinterpret_cast<const float4*>(rp + e0) : harmless);
        w0[s] = __ldg(p.bm + min(max(e0, 0) >> 5, p.bm_words - 1));
        w1[s] = __ldg(p.bm + min(max(e0 + 3, 0) >> 5, p.bm_words - 1));
    }
#pragma unroll
    for (int s = 0; s < kNV; ++s) {
        const int e0 = a0 + 4 * (L + 32 * s);
        float v[4] = {t[s].x, t[s].y, t[s].z, t[s].w};
        const bool clean = inside[s] && e0 >= lo && e0 + 3 < hi && (w0[s] | w1[s]) == 0u;
        if (!need[s]) {
            v[0] = v[1] = v[2] = v[3] = -CUDART_INF_F;
        } else if (!clean) {                             // the usual piece is wanted whole and names no suppressed id
            if (need[s] && !inside[s]) {                 // head of row 0 / tail of the last row
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = (e0 + i >= lo && e0 + i < hi) ? __ldcs(rp + e0 + i) : -CUDART_INF_F;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int id = e0 + i;
                const uint32_t w = (id >> 5) == (max(e0, 0) >> 5) ? w0[s] : w1[s];
                if (!need[s] || id < lo || id >= hi || ((w >> (id & 31)) & 1u)) v[i] = -CUDART_INF_F;
            }
        }
        x[s][0] = v[0]; x[s][1] = v[1]; x[s][2] = v[2]; x[s][3] = v[3];
    }
}
// first slot id and surviving id range of unit u of a row (what the scan's walker and rules give)
__device__ __forceinline__ void unit_range(const StepParams& p, const float* rp, const RowRules& rr, int u, int& a0, int& lo, int& hi) {
    if (u < p.n_text_units) {
        a0 = u * kUnit - (int)((reinterpret_cast<uintptr_t>(rp) >> 2) & 3);
        lo = max(max(a0, 0), rr.t_lo); hi = min(min(a0 + kUnit, p.ts_split), rr.t_hi);
    } else {
        a0 = p.ts_split + (u - p.n_text_units) * kUnit - (int)((reinterpret_cast<uintptr_t>(rp + p.ts_split) >> 2) & 3);
        lo = max(max(a0, p.ts_split), rr.s_lo); hi = min(min(a0 + kUnit, p.rc.n_vocab), rr.s_hi);
    }
}

// ---- the row's K best ids the long way: one warp reads the whole row again (reduce_unit).  Taken when the best cells of
//      the row tie exactly (quantised logits) or a winning cell holds more candidates than the merge buffers take.
__device__ __noinline__ void rescan_row(const StepParams& p, int row, const RowRules& rr, bool ts_over_text, WarpBuf& wb, float& out_v,
                                        int& out_i) {
    const int lane = threadIdx.x & 31;
    const float* rp = p.logits + (int64_t)row * p.row_stride;
    Run cand;
    run_reset(cand);
    for (int u = ts_over_text ? p.n_text_units : 0; u < p.upr; ++u) {
        Geom g;
        g.row = row; g.part = u < p.n_text_units ? 0 : 1;
        unit_range(p, rp, rr, u, g.a0, g.lo, g.hi);
        if (g.hi <= g.lo) continue;
        float x[kNV][4];
        load_cell(p, rp, row, g.a0, lane, g.lo, g.hi, true, x);
        reduce_unit(x, g, p.K, wb, cand);
    }
    run_finish(cand, wb, p.K);
    out_v = cand.rv; out_i = cand.ri;
}

// ---- row merge (one full warp): partials of the row -> K log-probs.  Lane i < K ends with candidate i.
// Lane l takes partial l of the row's list [text partials | timestamp partials] (launch_step keeps a row at <= 32 runs);
// header and cell list are fetched with independent 16-byte loads up front.  The partials' lists hold CELLS (fold_unit):
// the K best cells of the row are found by a k-way merge, read again from the logits, and their ids at or above the K-th
// cell's value compete for the K places.
struct RowOut {
    float lp; int tok; unsigned flags;
#ifdef WRS_TRACE
    unsigned long long t_sums, t_cells, t_loaded;      // phase times of the merge (global timer)
#endif
};

__device__ __forceinline__ RowOut merge_row(const StepParams& p, int row, const RowRules& rr, WarpBuf& wb, unsigned char* lists,
                                         int* glist) {
    const int lane = threadIdx.x & 31;
    const int K = p.K, KQ = (K + 3) >> 2;
    const Partial* __restrict__ parts = p.partials + (int64_t)row * p.upr;
    const int cT = p.rt, cnt = p.rpr;
    float4 h0 = make_float4(-CUDART_INF_F, 0.0f, 0.0f, 0.0f);
    float ev0[kMaxK]; int ei0[kMaxK];
#pragma unroll
    for (int e = 0; e < kMaxK; ++e) { ev0[e] = -CUDART_INF_F; ei0[e] = INT32_MAX; }
    if (lane < cnt) {
        // partial i sits in the slot of its run's first unit
        const Partial* pp = parts + run_in_row(p, lane).j0;
        h0 = __ldcg(reinterpret_cast<const float4*>(pp));
#pragma unroll
        for (int q = 0; q < kMaxK / 4; ++q)
            if (q < KQ) {
                const float4 a = __ldcg(reinterpret_cast<const float4*>(pp->v) + q);
                const int4 b = __ldcg(reinterpret_cast<const int4*>(pp->id) + q);
                ev0[4 * q] = a.x; ev0[4 * q + 1] = a.y; ev0[4 * q + 2] = a.z; ev0[4 * q + 3] = a.w;
                ei0[4 * q] = b.x; ei0[4 * q + 1] = b.y; ei0[4 * q + 2] = b.z; ei0[4 * q + 3] = b.w;
            }
    }
    // ---- pass A: partition maxima and sums (text = 0, timestamps = 1)
    float am[2] = {-CUDART_INF_F, -CUDART_INF_F}, as[2] = {0.0f, 0.0f};
    bool nan_seen = false;
    if (lane < cnt) {
        const int q = lane < cT ? 0 : 1;
        if (h0.y != h0.y) nan_seen = true;
        else { as[q] = h0.y; am[q] = h0.x; }
    }
    float M[2], S[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        M[q] = warp_max(am[q]);
        float s = as[q] * exp_diff(am[q], M[q]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        S[q] = s;
    }
    nan_seen = __any_sync(0xffffffffu, nan_seen);
    RowOut r;
    r.flags = 0u;
    if (nan_seen) r.flags |= 128u;
    float Tm = M[0], Ts = S[0];
    const float Sm = M[1], Ss = S[1];
    // logsumexp(timestamps) > max(text): both sides carry the same -logsumexp(all), which cancels
    const float lse_s = (Ss > 0.0f) ? Sm + logf(Ss) : -CUDART_INF_F;
    const bool ts_over_text = p.rc.timestamps && (lse_s > Tm);
    if (ts_over_text) { r.flags |= 8u; Tm = -CUDART_INF_F; Ts = 0.0f; }
    // final log-softmax over what survived: (x - M) - log(sum exp(x - M)), the order ATen uses
    const float Mx = fmaxf(Tm, Sm);
    const float tot = Ts * exp_diff(Tm, Mx) + Ss * exp_diff(Sm, Mx);
    const float logtot = nan_seen ? CUDART_NAN_F : logf(tot);
    if (Mx == -CUDART_INF_F) r.flags |= 128u;   // everything masked: the reference yields NaN log-probs
#ifdef WRS_TRACE
    r.t_sums = gtime(); r.t_cells = r.t_loaded = 0;
#endif
    // ---- pass B: the K best cells of the surviving partitions.  K rounds of "best head wins, its lane moves on to its
    // next entry" (value desc, cell asc) -- a k-way merge of sorted lists, two warp reductions per round.  The lists wait in
    // shared memory ([entry][lane]): a lane that wins reads its next head from there
    float* lv = reinterpret_cast<float*>(lists);
    int* li = reinterpret_cast<int*>(lists) + 32 * kMaxK;
    int pn = (int)__float_as_int(h0.z);
    const bool alive = lane < cnt && (lane >= cT || !ts_over_text);
    pn = alive ? min(pn, K) : 0;
    bool rescan = __any_sync(0xffffffffu, alive && __float_as_int(h0.w) != 0);
#pragma unroll
    for (int e = 0; e < kMaxK; ++e) {
        if (e >= K) break;                               // (warp-uniform: K stores, not kMaxK predicated ones)
        lv[32 * e + lane] = ev0[e]; li[32 * e + lane] = ei0[e];
    }
    // (cells of equal value all come along anyway, so a round needs no tie-break: every head at the round's maximum pops)
    int taken = 0, got = 0;
    float hv = pn > 0 ? ev0[0] : -CUDART_INF_F;
    float vK = -CUDART_INF_F;                            // value of the K-th best cell (-inf: the row has fewer than K cells)
    for (int rnd = 0; rnd < K; ++rnd) {
        const float m = warp_max(hv);
        if (m == -CUDART_INF_F) break;
        const bool pop = hv == m;
        got += __popc(__ballot_sync(0xffffffffu, pop));
        if (pop) {
            ++taken;
            hv = taken < pn ? lv[32 * taken + lane] : -CUDART_INF_F;
        }
        if (got >= K) { vK = m; break; }
    }
    // cells that tie the K-th one come along (a list that dropped such a cell has said so: `tie`)
    while (taken < pn && lv[32 * taken + lane] == vK) ++taken;
    int incl = taken;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    const int nG = __shfl_sync(0xffffffffu, incl, 31);
    if (nG > 32) rescan = true;
#ifdef WRS_TRACE
    r.t_cells = gtime();
#endif
    float ov = -CUDART_INF_F; int oi = INT32_MAX;
    if (!rescan) {
        for (int t = 0; t < taken; ++t) glist[incl - taken + t] = li[32 * t + lane];
        __syncwarp();
        // ---- the winning cells again, one per lane
        const float* rp = p.logits + (int64_t)row * p.row_stride;
        const bool have = lane < nG;
        const int cid = have ? glist[lane] : 0;
        int a0, lo, hi2;
        unit_range(p, rp, rr, cid >> 5, a0, lo, hi2);
        float x[kNV][4];
        load_cell(p, rp, row, a0, cid & 31, lo, hi2, have, x);
#ifdef WRS_TRACE
        if (x[0][0] == 12345.678f) r.flags |= 0x40000000u;   // (wait for the loads)
        r.t_loaded = gtime();
#endif
        const float bar = fmaxf(vK, -CUDART_MAX_NORMAL_F);   // x >= bar: at or above the K-th cell and not -inf
        // the usual cell holds ONE id at or above the bar, its maximum: then the candidates stay in registers, one per lane
        float pm[kNV];
        int np = 0, jj = 0;
#pragma unroll
        for (int s = 0; s < kNV; ++s) {
            pm[s] = fmaxf(fmaxf(x[s][0], x[s][1]), fmaxf(x[s][2], x[s][3]));
            const bool hp = pm[s] >= bar;
            np += hp ? 1 : 0; jj = hp ? s : jj;
        }
        float y[4] = {x[0][0], x[0][1], x[0][2], x[0][3]};
#pragma unroll
        for (int s = 1; s < kNV; ++s)
            if (jj == s) { y[0] = x[s][0]; y[1] = x[s][1]; y[2] = x[s][2]; y[3] = x[s][3]; }
        int ne = 0, ii = 0;
#pragma unroll
        for (int i = 3; i >= 0; --i) { const bool he = y[i] >= bar; ne += he ? 1 : 0; ii = he ? i : ii; }
        float cv = -CUDART_INF_F; int ci = INT32_MAX;
        if (np == 1 && ne == 1) { cv = y[0]; cv = ii == 1 ? y[1] : cv; cv = ii == 2 ? y[2] : cv; cv = ii == 3 ? y[3] : cv; ci = a0 + 4 * ((cid & 31) + 32 * jj) + ii; }
        if (!__any_sync(0xffffffffu, np > 1 || ne > 1)) {
            for (int rnd = 0; rnd < K; ++rnd) {
                const float m = warp_max(cv);
                if (m == -CUDART_INF_F) break;
                const int w2 = __reduce_min_sync(0xffffffffu, cv == m ? ci : INT32_MAX);
                if (cv == m && ci == w2) cv = -CUDART_INF_F;
                if (lane == rnd) { ov = m; oi = w2; }
            }
        } else {
            // several ids of a cell reach the bar: gather them all in shared memory
            int c = 0;
#pragma unroll
            for (int s = 0; s < kNV; ++s)
                if (pm[s] >= bar) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) c += x[s][i] >= bar ? 1 : 0;
                }
            int inc2 = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc2, o); if (lane >= o) inc2 += t; }
            const int nE = __shfl_sync(0xffffffffu, inc2, 31);
            if (nE > kCandCap) rescan = true;
            else {
                int w = inc2 - c;
#pragma unroll
                for (int s = 0; s < kNV; ++s)
                    if (pm[s] >= bar) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (x[s][i] >= bar) { wb.v[w] = x[s][i]; wb.id[w] = a0 + 4 * ((cid & 31) + 32 * s) + i; ++w; }
                    }
                __syncwarp();
                bool dropped = false;
                const int n = warp_compact(wb, nE, K, dropped);
                if (lane < n) { ov = wb.v[lane]; oi = wb.id[lane]; }
                __syncwarp();
            }
        }
    }
    if (rescan) rescan_row(p, row, rr, ts_over_text, wb, ov, oi);
    r.tok = oi;
    r.lp = (ov - Mx) - logtot;
    return r;
}


// Projection mode (decode_proj.cuh): the row's list is one record per vocab tile, stored as planes [plane][tile] -- max,
// sum, then the tile's best K values and their ids in (value desc, id asc) order: exact ids, so the best K of their union
// are the row's best K, ties included, and there is nothing to read again.  The three planes every tile contributes to
// (max, sum, best value) come into the warp's shared-memory slice with one bulk copy each; the loops over them are
// rolled (this code runs once per launch: what it costs is its instruction fetch).
__device__ __forceinline__ RowOut merge_row_proj(const StepParams& p, int row, WarpBuf& wb, unsigned char* slice) {
    const int lane = threadIdx.x & 31;
    const int K = p.K, T = p.proj_tiles, TP = p.proj_tp;
    const float* __restrict__ planes = p.proj_parts + (int64_t)row * (2 + 2 * K) * TP;
    const float* pm = planes; const float* ps = planes + TP; const float* pv0 = planes + 2 * TP;
    if (3 * TP * 4 <= kTailLists - 16) {
        float* sm = reinterpret_cast<float*>(slice);
        unsigned long long* bar = reinterpret_cast<unsigned long long*>(slice + kTailLists - 16);
        if (lane == 0) {
            mbar_init(bar, 1);
            fence_proxy_async();
            mbar_expect_tx(bar, 3u * TP * 4u);
            bulk_g2s(sm, planes, 3u * TP * 4u, bar);      // the three planes are adjacent
        }
        __syncwarp();
        mbar_wait(bar, 0u);
        pm = sm; ps = sm + TP; pv0 = sm + 2 * TP;
    }
    // ---- pass A: partition maxima and sums, folded lane-locally first (ten tiles per lane: exp2 on the fast path, as
    // in the scan -- the tiles' own sums were formed the same way)
    float am0 = -CUDART_INF_F, am1 = -CUDART_INF_F, as0 = 0.0f, as1 = 0.0f, head0 = -CUDART_INF_F, head1 = -CUDART_INF_F;
    bool nan_seen = false;
    auto ediff = [](float a, float b) { return a == -CUDART_INF_F ? 0.0f : ex2((a - b) * kLog2e); };   // exp(a - b), exp(-inf - -inf) = 0
#pragma unroll 1
    for (int t = lane; t < T; t += 32) {
        const float hm = pm[t], hs = ps[t], hv = pv0[t];
        if (t < p.proj_text_tiles) head0 = fmaxf(head0, hv); else head1 = fmaxf(head1, hv);   // (the tiles' best values, for pass B)
        if (hs != hs) { nan_seen = true; continue; }
        if (t < p.proj_text_tiles) {
            if (hm > am0) { as0 = as0 * ediff(am0, hm) + hs; am0 = hm; } else as0 += hs * ediff(hm, am0);
        } else {
            if (hm > am1) { as1 = as1 * ediff(am1, hm) + hs; am1 = hm; } else as1 += hs * ediff(hm, am1);
        }
    }
    float M[2], S[2];
    {
        M[0] = warp_max(am0); M[1] = warp_max(am1);
        float s0 = as0 * exp_diff(am0, M[0]), s1 = as1 * exp_diff(am1, M[1]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
        S[0] = s0; S[1] = s1;
    }
    nan_seen = __any_sync(0xffffffffu, nan_seen);
    RowOut r;
    r.flags = 0u;
    if (nan_seen) r.flags |= 128u;
    float Tm = M[0], Ts = S[0];
    const float Sm = M[1], Ss = S[1];
    const float lse_s = (Ss > 0.0f) ? Sm + logf(Ss) : -CUDART_INF_F;
    const bool ts_over_text = p.rc.timestamps && (lse_s > Tm);          // timestamp_tokens.rs:80-96
    if (ts_over_text) { r.flags |= 8u; Tm = -CUDART_INF_F; Ts = 0.0f; }
    const float Mx = fmaxf(Tm, Sm);
    const float tot = Ts * exp_diff(Tm, Mx) + Ss * exp_diff(Sm, Mx);
    const float logtot = nan_seen ? CUDART_NAN_F : logf(tot);
    if (Mx == -CUDART_INF_F) r.flags |= 128u;
#ifdef WRS_TRACE
    r.t_sums = gtime(); r.t_cells = r.t_loaded = 0;
#endif
    // ---- pass B: a value at least K candidates reach (K rounds over the lanes' best heads), then everything at or above it
    const int t_first = ts_over_text ? p.proj_text_tiles : 0;
    const float head = ts_over_text ? head1 : fmaxf(head0, head1);
    const float tau = kth_threshold(head, K);
    const float* gv = planes + 2 * TP;                                   // value planes and id planes in global memory
    const int* gi = reinterpret_cast<const int*>(planes + (2 + K) * TP);
#pragma unroll 1
    for (int t = lane; t < T; t += 32) {
        if (t < t_first) continue;
        float v = pv0[t];
#pragma unroll 1
        for (int e = 0; e < K; ++e) {                                    // (lists are sorted: stop at the first entry below tau)
            if (!(v >= tau) || v == -CUDART_INF_F) break;
            // the entry's id and the list's next value travel together
            const int id = __ldcg(gi + (int64_t)e * TP + t);
            const float vn = e + 1 < K ? __ldcg(gv + (int64_t)(e + 1) * TP + t) : -CUDART_INF_F;
            const int slot = atomicAdd(&wb.n, 1);
            if (slot < kCandCap) { wb.v[slot] = v; wb.id[slot] = id; }
            v = vn;
        }
    }
    __syncwarp();
    const int n = wb.n;
    float ov = -CUDART_INF_F; int oi = INT32_MAX;
    if (n <= kCandCap) {
        bool dropped = false;
        const int kept = warp_compact(wb, n, K, dropped);
        if (lane < kept) { ov = wb.v[lane]; oi = wb.id[lane]; }
    } else {
        // massive ties: K rounds of "the best entry strictly after the previous pick" over the lane's lists
        float pv = CUDART_INF_F; int pi = -1;
        for (int rnd = 0; rnd < K; ++rnd) {
            float bv = -CUDART_INF_F; int bi = INT32_MAX;
            for (int t = lane; t < T; t += 32) {
                if (t < t_first) continue;
                for (int e = 0; e < K; ++e) {
                    const float v = __ldcg(gv + (int64_t)e * TP + t); const int id = __ldcg(gi + (int64_t)e * TP + t);
                    if (v == -CUDART_INF_F || !(v == v)) break;
                    const bool after = v < pv || (v == pv && id > pi);
                    if (after) { if (better(v, id, bv, bi)) { bv = v; bi = id; } break; }   // the first entry after the pick is the list's best
                }
            }
            warp_pick(bv, bi);
            if (bi == INT32_MAX) break;
            if (lane == rnd) { ov = bv; oi = bi; }
            pv = bv; pi = bi;
        }
    }
    __syncwarp();
    r.tok = oi;
    r.lp = (ov - Mx) - logtot;
    return r;
}


// `completed` of a step: every group adds 1 (+ 1 << 16 when it is "done") to one word; the last group to arrive
// publishes whether all were done and re-arms the word for the next launch.
// (two halves so that a caller can put its output stores between the atomic and the use of its result)
__device__ __forceinline__ int publish_arrive(const StepParams& p, int done) { return atomicAdd(p.sync, 1 + (done << 16)); }
__device__ __forceinline__ void publish_finish(const StepParams& p, int prev, int done) {
    if ((prev & 0xffff) == p.n_groups - 1) {
        *p.completed = ((prev >> 16) + done == p.n_groups) ? 1 : 0;
        atomicExch(p.sync, 0);
    }
}
__device__ __forceinline__ void publish_completed(const StepParams& p, int done) { publish_finish(p, publish_arrive(p, done), done); }

// per-warp tables of the beam tail (alias the WarpBuf: the warp's run has been flushed when a tail starts)
struct TailScratch {
    double score[kMaxCandAudio];
    int tok[kMaxCandAudio];
    unsigned short order[kMaxCandAudio];
    unsigned char shadow[kMaxCandAudio];
    double sel_score[kMaxBeam], fin_sc[kMaxBeam];
    int sel_src[kMaxBeam], sel_tok[kMaxBeam], fin_src[kMaxBeam];
};
static_assert(sizeof(TailScratch) <= kTailBytes, "tail scratch does not fit the per-warp slice");
static_assert(sizeof(WarpBuf) <= kScanScratch, "candidate buffer does not fit the per-warp slice");
static_assert(kTailLists + kScanScratch + 128 <= kTailBytes, "row merge tables do not fit the per-warp slice of the tail launch");

// ---- MODE_BEAM tail: BeamSearchTokenExtractor::update for one audio (beam_search.rs:49-146), ONE warp ----
__device__ __noinline__ void beam_tail_general(const StepParams& p, int audio, TailScratch& ts) {
    const int k = p.beam, K = p.K, n_c = k * K;
    const int lane = threadIdx.x & 31;
    const int row0 = audio * k;
    const unsigned lt = (1u << lane) - 1u;
    // 1. candidates: score = f64(sum_logprobs[row]) + f64(logprob)   (beam_search.rs:70-71)
    bool nan_seen = false;
    for (int c = lane; c < n_c; c += 32) {
        const int j = c / K, i = c - j * K;
        const RowResult* rr = p.rowres + row0 + j;
        const float lp = __ldcg(&rr->lp[i]);
        ts.tok[c] = __ldcg(&rr->tok[i]);
        ts.score[c] = (double)p.sum_in[row0 + j] + (double)lp;
        if (i == 0 && (__ldcg(&rr->flags) & 128u)) nan_seen = true;
    }
    const unsigned same_mine = lane < k ? __ldcg(&p.rowres[row0 + lane].same) : 0u;
    nan_seen = __any_sync(0xffffffffu, nan_seen);
    __syncwarp();
    // 2. a later beam with the same prefix proposing the same token overwrites this map entry (:74-77)
    for (int c0 = 0; c0 < n_c; c0 += 32) {
        const int c = c0 + lane;
        const int j = min(c, n_c - 1) / K;
        const unsigned same = __shfl_sync(0xffffffffu, same_mine, j);
        if (c < n_c) {
            const int tok = ts.tok[c];
            bool shadowed = false;
            for (int j2 = j + 1; j2 < k; ++j2)
                if ((same >> j2) & 1u)
                    for (int i2 = 0; i2 < K; ++i2) shadowed |= (ts.tok[j2 * K + i2] == tok);
            ts.shadow[c] = (unsigned char)((shadowed || tok == INT32_MAX) ? 1 : 0);
        }
    }
    __syncwarp();
    // 3. rank sort by (score desc, source row asc, token asc); shadowed entries go last
    for (int c = lane; c < n_c; c += 32) {
        int rank = 0;
        const double sc = ts.score[c]; const int sh = ts.shadow[c], rw = c / K, tk = ts.tok[c];
        for (int d = 0; d < n_c; ++d) {
            if (d == c) continue;
            const int shd = ts.shadow[d];
            bool before;
            if (shd != sh) before = shd < sh;
            else if (ts.score[d] != sc) before = ts.score[d] > sc;
            else if (d / K != rw) before = d / K < rw;
            else before = ts.tok[d] < tk || (ts.tok[d] == tk && d < c);
            rank += before ? 1 : 0;
        }
        ts.order[rank] = (unsigned short)c;
    }
    __syncwarp();
    // 4. walk (beam_search.rs:87-104): EOT-terminated -> finished, others kept until k are saved.  An entry is
    //    visited iff fewer than k live entries precede it; shadowed entries sort last and end the walk.
    int saved = 0, nf = 0;
    for (int base = 0; base < n_c && saved < k; base += 32) {
        const int o = base + lane;
        const bool valid = o < n_c;
        const int c = valid ? ts.order[o] : 0;
        const bool ok = valid && !ts.shadow[c];
        const bool live = ok && ts.tok[c] != p.eot, fin = ok && ts.tok[c] == p.eot;
        const unsigned lm = __ballot_sync(0xffffffffu, live);
        const int live_before = saved + __popc(lm & lt);
        const bool visited = live_before < k;
        const unsigned vm = __ballot_sync(0xffffffffu, visited);
        const unsigned fm = __ballot_sync(0xffffffffu, fin && visited);
        if (live && visited) { ts.sel_src[live_before] = c / K; ts.sel_tok[live_before] = ts.tok[c]; ts.sel_score[live_before] = ts.score[c]; }
        const int fidx = nf + __popc(fm & lt);
        if (fin && visited && fidx < kMaxBeam) { ts.fin_src[fidx] = c / K; ts.fin_sc[fidx] = ts.score[c]; }
        saved += __popc(lm & vm);
        nf = min(nf + __popc(fm), kMaxBeam);
    }
    __syncwarp();
    // fewer than k live candidates cannot happen with >= k+1 finite logits per row; keep the row count anyway
    if (lane >= saved && lane < k) { ts.sel_src[lane] = 0; ts.sel_tok[lane] = p.eot; ts.sel_score[lane] = -CUDART_INF; }
    // finished store: append in descending score until round(patience*k) entries (:124-138)
    int have = 0;
    if (lane == 0) {
        have = p.fin_n[audio];
        p.fin_n[audio] = have + min(nf, max(0, p.max_cand - have));
        if (p.dbg_fin_count) p.dbg_fin_count[audio] = nf;
        if (nan_seen) atomicAdd(&p.status->nan_rows, 1);
    }
    have = __shfl_sync(0xffffffffu, have, 0);
    const int n_store = min(nf, max(0, p.max_cand - have));
    __syncwarp();
    if (lane < k) {
        if (p.dbg_fin_source) p.dbg_fin_source[audio * k + lane] = lane < nf ? row0 + ts.fin_src[lane] : -1;
        if (p.dbg_fin_score) p.dbg_fin_score[audio * k + lane] = lane < nf ? ts.fin_sc[lane] : 0.0;
        p.source_idx[row0 + lane] = row0 + ts.sel_src[lane];
        p.sum_out[row0 + lane] = (float)ts.sel_score[lane];   // index_put_(.., score as f32) :91-95
    }
    // 5. outputs
    const int L1 = p.cur_len + 1;
    for (int idx = lane; idx < k * L1; idx += 32) {
        const int i = idx / L1, t = idx - i * L1;
        p.next_tokens[(int64_t)(row0 + i) * p.next_stride + t] =
            t < p.cur_len ? p.tokens[(int64_t)(row0 + ts.sel_src[i]) * p.tok_stride + t] : (int64_t)ts.sel_tok[i];
    }
    for (int idx = lane; idx < n_store * L1; idx += 32) {
        const int i = idx / L1, t = idx - i * L1;
        const int slot = have + i;
        p.fin_tokens[((int64_t)audio * p.max_cand + slot) * p.max_len + t] =
            t < p.cur_len ? p.tokens[(int64_t)(row0 + ts.fin_src[i]) * p.tok_stride + t] : (int64_t)p.eot;
        if (t == 0) {
            p.fin_len[audio * p.max_cand + slot] = L1;
            p.fin_score[audio * p.max_cand + slot] = ts.fin_sc[i];
        }
    }
    // 6. completed = all audios hold max_cand finished sequences (:141-146); last audio to arrive publishes
    if (lane == 0) publish_completed(p, (have + n_store >= p.max_cand) ? 1 : 0);
    __syncwarp();
}

// ---- MODE_GREEDY tail: GreedyTokenExtractor::update for one row (src/decode/extractor/greedy.rs:24-50), one warp ----
__device__ __forceinline__ void greedy_tail(const StepParams& p, int row, const RowOut& r) {
    const int lane = threadIdx.x & 31;
    const int64_t* rt = p.tokens + (int64_t)row * p.tok_stride;
    int64_t* dst = p.next_tokens + (int64_t)row * p.next_stride;
    for (int t = lane; t < p.cur_len; t += 32) dst[t] = rt[t];
    if (lane != 0) return;
    const bool last_eot = rt[p.cur_len - 1] == p.eot;                       // greedy.rs:37-38
    const int next = last_eot ? p.eot : r.tok;                               // :41-45
    p.sum_out[row] = p.sum_in[row] + (last_eot ? 0.0f : r.lp);               // :39
    dst[p.cur_len] = next;
    if (r.flags & 128u) atomicAdd(&p.status->nan_rows, 1);
    publish_completed(p, next == p.eot ? 1 : 0);        // completed: every row's new last token is EOT
}

// ---- MODE_FILTER tail: timestamp_tokens.rs:37-96 in place for one row, one warp.  The statistics were taken on
// the logits as they are (earlier filters already wrote their -inf), so the decision sees what the reference's
// log_softmax sees.  Every unit of the row has been read by now: the fills cannot race the scan.
__device__ __forceinline__ void filter_tail(const StepParams& p, int row, const RowRules& rr, unsigned flags) {
    const int lane = threadIdx.x & 31;
    const bool kill_text = (flags & 8u) != 0;
    float* rp = p.logits + (int64_t)row * p.row_stride;
    const int tsb = p.rc.ts_begin, V = p.rc.n_vocab;
    const int keep_lo = kill_text ? tsb : rr.t_lo, keep_hi = kill_text ? tsb : rr.t_hi;
    for (int i = lane; i < keep_lo; i += 32) rp[i] = -CUDART_INF_F;
    for (int i = keep_hi + lane; i < tsb; i += 32) rp[i] = -CUDART_INF_F;
    for (int i = tsb + lane; i < rr.s_lo; i += 32) rp[i] = -CUDART_INF_F;
    for (int i = max(rr.s_hi, tsb) + lane; i < V; i += 32) rp[i] = -CUDART_INF_F;
    if (lane == 0 && p.no_timestamps >= 0 && p.no_timestamps < V) rp[p.no_timestamps] = -CUDART_INF_F;
}

// ---- MODE_BEAM tail, k * (k + 1) <= 32 (beam <= 5): BeamSearchTokenExtractor::update (beam_search.rs:49-146) with one
//      candidate per lane.  After the barrier behind the row merges EVERY warp of the audio sorts the candidates for
//      itself (a 32-key bitonic network on shuffles: no tables, no further barrier), walks them, and writes the output row
//      it stands for: token row r of the next step, finished sequence r, source and sum r.  Warp 0 also keeps the books
//      (finished count, `completed`).
struct FastTail {                                   // CTA-level tables (behind the row results in the tail's shared memory)
    long long tok_stage[256];                       // the audio's k token prefixes when k * cur_len <= 256
    int have;                                       // finished sequences the audio already holds
};
static_assert(sizeof(FastTail) <= kTailHead - kTailRows, "fast tail tables do not fit the head of the tail launch's shared memory");

__device__ __forceinline__ void beam_fast(const StepParams& p, int audio, const FastTail& fs, const RowResult* rr_s, float sum, int warp) {
    const int k = p.beam, K = p.K, n_c = k * K, L = p.cur_len, L1 = L + 1;
    const int lane = threadIdx.x & 31;
    const int row0 = audio * k;
    const bool valid = lane < n_c;
    // ---- lane c < k K is candidate (beam j = c / K, entry i = c % K): score = f64(sum_logprobs[row]) + f64(logprob)   (:70-71)
    const int j = valid ? lane / K : 0, i = valid ? lane - j * K : 0;
    const RowResult* rr = rr_s + j;
    const float lp = rr->lp[i];
    const int tok = rr->tok[i];
    const unsigned fl = rr->flags, same = rr->same;
    const double score = (double)sum + (double)lp;
    // a later beam with the same prefix proposing the same token overwrites this map entry (:74-77)
    bool shadowed = false;
    if (__any_sync(0xffffffffu, valid && same != 0u)) {
        for (int j2 = 1; j2 < k; ++j2)
            for (int i2 = 0; i2 < K; ++i2) {
                const int t2 = __shfl_sync(0xffffffffu, tok, j2 * K + i2);
                if (j2 > j && ((same >> j2) & 1u) && t2 == tok) shadowed = true;
            }
    }
    const bool ok = valid && !shadowed && tok != INT32_MAX;
    // ---- sort by (score desc, source row asc, token asc) == (score desc, lane asc); entries that are not ok go last.
    // Order-preserving 64-bit image of the score (0 for "not ok": below the image of every double), the lane rides along
    long long kb = __double_as_longlong(score);
    kb = kb < 0 ? ~kb : (kb | (long long)0x8000000000000000ull);
    unsigned hi = ok ? (unsigned)((unsigned long long)kb >> 32) : 0u, lo = ok ? (unsigned)kb : 0u;
    int o = lane;
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1)
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            const unsigned phi = __shfl_xor_sync(0xffffffffu, hi, stride), plo = __shfl_xor_sync(0xffffffffu, lo, stride);
            const int po = __shfl_xor_sync(0xffffffffu, o, stride);
            const bool mine_first = hi != phi ? hi > phi : (lo != plo ? lo > plo : o < po);
            const bool want_first = ((lane & stride) == 0) == ((lane & size) == 0 || size == 32);
            if (mine_first != want_first) { hi = phi; lo = plo; o = po; }
        }
    // lane r now holds the r-th entry of the sorted list: candidate o
    const int s_tok = __shfl_sync(0xffffffffu, tok, o), s_j = __shfl_sync(0xffffffffu, j, o);
    const bool s_ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, o) != 0;
    const long long sb = (long long)(((unsigned long long)hi << 32) | lo);
    const double s_score = __longlong_as_double(sb < 0 ? (sb & 0x7fffffffffffffffll) : ~sb);
    const bool nan_seen = __any_sync(0xffffffffu, valid && (fl & 128u));
    // ---- walk (:87-104): EOT-terminated -> finished, others kept until k are saved.  An entry is visited iff fewer than
    //      k live entries precede it in sorted order.
    const bool live = s_ok && s_tok != p.eot, fin = s_ok && s_tok == p.eot;
    const unsigned below = (1u << lane) - 1u;
    const unsigned live_sorted = __ballot_sync(0xffffffffu, live);
    const int live_before = __popc(live_sorted & below);
    const bool visited = live_before < k;
    const unsigned fin_sorted = __ballot_sync(0xffffffffu, fin && visited);
    const int nf = min(__popc(fin_sorted), kMaxBeam), saved = min(__popc(live_sorted), k);
    // finished store: append in descending score until round(patience*k) entries (:124-138)
    const int have = fs.have;
    const int n_store = min(nf, max(0, p.max_cand - have));
    const int done = (have + n_store >= p.max_cand) ? 1 : 0;           // (:141-146)
    int arrived = 0;
    if (warp == 0 && lane == 0) {
        p.fin_n[audio] = have + n_store;
        if (p.dbg_fin_count) p.dbg_fin_count[audio] = nf;
        if (nan_seen) atomicAdd(&p.status->nan_rows, 1);
        arrived = publish_arrive(p, done);                              // its round trip overlaps the output stores
    }
    if (warp == 0 && (p.dbg_fin_source || p.dbg_fin_score)) {
        for (int r = 0; r < k; ++r) {
            const unsigned fm = __ballot_sync(0xffffffffu, fin && visited && __popc(fin_sorted & below) == r);
            const int fl2 = fm ? __ffs((int)fm) - 1 : 0;
            const int fj = __shfl_sync(0xffffffffu, s_j, fl2);
            const double fsr = __shfl_sync(0xffffffffu, s_score, fl2);
            if (lane == 0) {
                if (p.dbg_fin_source) p.dbg_fin_source[audio * k + r] = r < nf ? row0 + fj : -1;
                if (p.dbg_fin_score) p.dbg_fin_score[audio * k + r] = r < nf ? fsr : 0.0;
            }
        }
    }
    const bool staged = L <= 64 && k * L <= 256;          // as decode_tail_kernel staged it
    // ---- output row r = warp: the r-th live entry visited (fewer than k live candidates cannot happen with >= k+1 finite
    //      logits per row; keep the row anyway: source 0, EOT, -inf)
    {
        const unsigned sm = __ballot_sync(0xffffffffu, live && visited && live_before == warp);
        const int sl = sm ? __ffs((int)sm) - 1 : 0;
        int src = __shfl_sync(0xffffffffu, s_j, sl), tk = __shfl_sync(0xffffffffu, s_tok, sl);
        double sc = __shfl_sync(0xffffffffu, s_score, sl);
        if (warp >= saved) { src = 0; tk = p.eot; sc = -CUDART_INF; }
        const int64_t* from = p.tokens + (int64_t)(row0 + src) * p.tok_stride;
        int64_t* to = p.next_tokens + (int64_t)(row0 + warp) * p.next_stride;
        for (int t = lane; t < L1; t += 32) to[t] = t < L ? (staged ? fs.tok_stage[src * L + t] : from[t]) : (long long)tk;
        if (lane == 0) {
            p.source_idx[row0 + warp] = row0 + src;
            p.sum_out[row0 + warp] = (float)sc;                         // index_put_(.., score as f32) :91-95
        }
    }
    // ---- finished sequence r = warp
    if (warp < n_store) {
        const unsigned fm = __ballot_sync(0xffffffffu, fin && visited && __popc(fin_sorted & below) == warp);
        const int fl2 = __ffs((int)fm) - 1;
        const int src = __shfl_sync(0xffffffffu, s_j, fl2);
        const double sc = __shfl_sync(0xffffffffu, s_score, fl2);
        const int slot = have + warp;
        const int64_t* from = p.tokens + (int64_t)(row0 + src) * p.tok_stride;
        int64_t* to = p.fin_tokens + ((int64_t)audio * p.max_cand + slot) * p.max_len;
        for (int t = lane; t < L1; t += 32) to[t] = t < L ? (staged ? fs.tok_stage[src * L + t] : from[t]) : (long long)p.eot;
        if (lane == 0) {
            p.fin_len[audio * p.max_cand + slot] = L1;
            p.fin_score[audio * p.max_cand + slot] = sc;
        }
    }
    if (warp == 0 && lane == 0) publish_finish(p, arrived, done);
}

// ---- where the units of a warp's runs live.  A run's units are consecutive units of one partition of one row: the
//      walker steps through them with one addition and recomputes everything else only when the next run begins
//      (run g + 1 of the same warp; at g_end the walker is exhausted)
struct Walk {
    int g, g_end;        // run, and the run behind the warp's last one
    int row, part, pe;   // the run's row, partition (0 text, 1 timestamps) and the id the partition's ids end at
    int j, j0, j_end;    // current unit in row, the run's first unit and the unit behind its last one
    int a0;              // id of the unit's first element slot: 16-byte aligned in memory, up to 3 ids in front of the partition
    const float* rp;     // the row's logits
};
__device__ __forceinline__ void walk_enter(const StepParams& p, Walk& w) {    // w.g is set and < n_runs
    w.row = (int)(((unsigned long long)(unsigned)w.g * p.rpr_magic) >> 40);
    const RunGeo r = run_in_row(p, w.g - w.row * p.rpr);
    w.part = r.part; w.j = w.j0 = r.j0; w.j_end = r.j0 + r.len;
    w.rp = p.logits + (int64_t)w.row * p.row_stride;
    if (r.part == 0) {
        w.pe = p.ts_split;
        w.a0 = w.j * kUnit - (int)((reinterpret_cast<uintptr_t>(w.rp) >> 2) & 3);
    } else {
        w.pe = p.rc.n_vocab;
        w.a0 = p.ts_split + (w.j - p.n_text_units) * kUnit - (int)((reinterpret_cast<uintptr_t>(w.rp + p.ts_split) >> 2) & 3);
    }
}
__device__ __forceinline__ void walk_step(const StepParams& p, Walk& w) {
    if (++w.j < w.j_end) { w.a0 += kUnit; return; }
    if (++w.g < w.g_end) walk_enter(p, w);
}
// 16-byte chunks [c_lo, c_hi) of the unit that are bulk-copied: those that hold ids of the partition, except a chunk that
// would reach outside the logits buffer (before row 0, behind the last row), whose <= 3 ids are read one by one instead
__device__ __forceinline__ void walk_chunks(const StepParams& p, const Walk& w, int& c_lo, int& c_hi) {
    c_lo = (w.row == 0 && w.a0 < 0) ? 1 : 0;
    c_hi = min(kUnit >> 2, (w.pe - w.a0 + 3) >> 2);
    if (w.row == p.n_rows - 1 && w.a0 + 4 * c_hi > p.rc.n_vocab) --c_hi;
}

// ---- launch 1 of a step: every warp folds its run(s) of units and leaves one partial per run.  No atomics, no fences:
//      the kernel boundary hands the partials to the tail.
__global__ void __launch_bounds__(kStepThreads, 1) decode_scan_kernel(const StepParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // per-warp slice: [stage 0][scratch][stage 1][stage 2][mbarriers]
    unsigned char* slice = smem_raw + warp * kWarpSmem;
    unsigned char* scratch = slice + kStageBytes;
    WarpBuf& wb = *reinterpret_cast<WarpBuf*>(scratch);
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(slice + kStages * kStageBytes + kScanScratch);
    auto stage_of = [&](int s) { return reinterpret_cast<float4*>(s == 0 ? slice : slice + kScanScratch + s * kStageBytes); };

    // Every warp takes a block of npw consecutive runs (one run when the rows allow it); consecutive blocks go to
    // different CTAs: consecutive runs belong to the same row, and rows whose rules mask most of a partition cost next to
    // nothing -- interleaved, every SM gets the same mix of cheap and full runs.  (A block, not a stride: with text and
    // timestamp runs alternating, a stride would hand some warps nothing but long text runs.)
    const int gw = warp * (int)gridDim.x + (int)blockIdx.x;
    if (gw * p.npw >= p.n_runs) return;
#ifdef WRS_TRACE
    if (lane == 0) for (int q = 0; q < 8; ++q) g_trace[gw * 8 + q] = 0;
#endif
    WRS_TR(0);

    Walk wi;                                              // the next unit to request
    wi.g = gw * p.npw;
    wi.g_end = min(wi.g + p.npw, p.n_runs);
    walk_enter(p, wi);
    Walk wf = wi;                                         // the unit being folded
    // start the flight of the next unit not yet requested into the next stage: one bulk copy, completion by byte
    // count on the stage's mbarrier
    int stage_issue = 0;
    auto issue_next = [&]() {
        int c_lo, c_hi;
        walk_chunks(p, wi, c_lo, c_hi);
        if (elect_one()) {
            unsigned long long* bar = bars + stage_issue;
            const int bytes = (c_hi - c_lo) * 16;
            if (bytes > 0) {
                mbar_expect_tx(bar, (unsigned)bytes);
                bulk_g2s(reinterpret_cast<unsigned char*>(stage_of(stage_issue)) + 16 * c_lo, wi.rp + wi.a0 + 4 * c_lo, (unsigned)bytes, bar);
            } else {
                mbar_arrive(bar);
            }
        }
        walk_step(p, wi);
        stage_issue = stage_issue + 1 == kStages ? 0 : stage_issue + 1;
    };
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) mbar_init(bars + s, 1);
    }
    __syncwarp();
    // every warp puts its first unit in flight before anything else
    issue_next();

    // the rules of a run's row come from the row's two newest tokens: those of the first run are fetched now, those of the
    // warp's next run when a run begins
    auto load_tokens = [&](int g, int64_t& last, int64_t& pen) {
        const int r = (int)(((unsigned long long)(unsigned)min(g, p.n_runs - 1) * p.rpr_magic) >> 40);
        const int64_t* rt = p.tokens + (int64_t)r * p.tok_stride;
        last = rt[p.cur_len - 1];
        pen = p.cur_len >= 2 ? rt[p.cur_len - 2] : 0;
    };
    int64_t tk_last, tk_pen;
    load_tokens(wi.g, tk_last, tk_pen);
    // lane l takes bitmap word (first id of the unit >> 5) + l: 29 words cover the unit
    auto bitmap_of = [&](const Walk& q) { return __ldg(p.bm + min((max(q.a0, 0) >> 5) + lane, p.bm_words - 1)); };
    // bring the unit into registers; slot j of a lane holds ids a0 + 4 (lane + 32 j) ..+3, everything outside [lo, hi)
    // and every suppressed id reads as -inf
    auto fetch = [&](int stage, const Walk& q, const Geom& g, uint32_t bmw, float (&x)[kNV][4]) {
        float4* st = stage_of(stage);
        float* stf = reinterpret_cast<float*>(st);
        const bool ragged = g.lo != g.a0 || g.hi != g.a0 + kUnit;
        if (ragged) {
            // the <= 3 ids at the head of row 0 / the tail of the last row that the copy left out
            int c_lo, c_hi;
            walk_chunks(p, q, c_lo, c_hi);
            const int e_tail = g.a0 + 4 * c_hi;
            const bool tail = q.row == p.n_rows - 1 && c_hi < (kUnit >> 2) && e_tail < g.hi;
            if (c_lo > 0 || tail) {
                if (lane < 4) {
                    const int e = g.a0 + lane;
                    if (c_lo > 0 && e >= g.lo && e < g.hi) stf[lane] = __ldcs(q.rp + e);
                } else if (lane < 8) {
                    const int e = e_tail + lane - 4;
                    if (tail && e < g.hi) stf[4 * c_hi + lane - 4] = __ldcs(q.rp + e);
                }
                __syncwarp();
            }
        }
        if (__ballot_sync(0xffffffffu, bmw != 0u)) {
            poke_suppressed(stf, bmw, max(g.a0, 0) >> 5, g.a0);
#ifndef WRS_NO_POKE_FENCE
            fence_proxy_async();                          // the stage is the target of a later bulk copy
#endif
            __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < kNV; ++j) {
            const float4 v = st[lane + 32 * j];
            x[j][0] = v.x; x[j][1] = v.y; x[j][2] = v.z; x[j][3] = v.w;
        }
        if (ragged) {
#pragma unroll
            for (int j = 0; j < kNV; ++j) {
                const int e = g.a0 + 4 * (lane + 32 * j);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (e + i < g.lo || e + i >= g.hi) x[j][i] = -CUDART_INF_F;
            }
        }
    };

    Run run;
    run_reset(run);
    uint32_t bm_c = bitmap_of(wf);
    int stage = 0;                                       // stage and mbarrier phase of the unit being folded
    unsigned phase = 0u;
    bool first = true;
    RowRules rules;
    rules.t_lo = rules.t_hi = rules.s_lo = rules.s_hi = 0; rules.flags = 0;

    const int g_end = wi.g_end;
    while (wf.g < g_end) {
        const Walk q = wf;
        const uint32_t bmw = bm_c;
        if (q.j == q.j0) {                               // a run begins: its row's rules; request the next run's tokens
            rules = row_rules_from(p.rc, tk_last, tk_pen, p.cur_len, p.sample_begin);
            if (q.g + 1 < g_end) load_tokens(q.g + 1, tk_last, tk_pen);
        }
        // ids of the unit that survive the row's rules
        Geom g;
        g.row = q.row; g.part = q.part; g.a0 = q.a0;
        g.lo = max(max(q.a0, q.part ? p.ts_split : 0), q.part ? rules.s_lo : rules.t_lo);
        g.hi = min(min(q.a0 + kUnit, q.pe), q.part ? rules.s_hi : rules.t_hi);
        walk_step(p, wf);
        if (wf.g < g_end) bm_c = bitmap_of(wf);
#ifdef WRS_TRACE
        const unsigned long long tw0 = gtime();
#endif
        mbar_wait(bars + stage, phase);                  // also when nothing survives: the stage is reused
#ifdef WRS_TRACE
        const unsigned long long tw1 = gtime();
        WRS_TR_ADD(4, tw1 - tw0);
#endif
        if (first) {
            // only now the later units: all first units of the grid (a quarter of the step's bytes) land first and
            // the arithmetic starts while the rest streams in behind them
            first = false;
            WRS_TR(1);
#pragma unroll
            for (int k2 = 1; k2 < kStages; ++k2)
                if (wi.g < g_end) issue_next();
        }
        if (g.hi > g.lo) {
            float x[kNV][4];
            fetch(stage, q, g, bmw, x);
            fold_unit(x, q.j, p.K, wb, run);
        }
        __syncwarp();                                    // every lane has taken its share of the stage out
#ifdef WRS_TRACE
        WRS_TR_ADD(5, gtime() - tw1);
#endif
        if (wi.g < g_end) issue_next();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
        WRS_TR(2);
        if (q.j + 1 < q.j_end) continue;

        // ---- end of the run: leave the partial in the slot of the run's first unit
        run_finish(run, wb, p.K);
        Partial* out = p.partials + (int64_t)q.row * p.upr + q.j0;
        if (lane < kMaxK) { out->v[lane] = run.rv; out->id[lane] = run.ri; }
        if (lane == 0) { out->m = run.m; out->s = run.s; out->n = run.n; out->tie = run.tie ? 1 : 0; }
        run_reset(run);
    }

    WRS_TR(3);
}

// ---- launch 2 of a step: one CTA per group (an audio's beams; a single row for greedy / filter), one warp per row.
//      Every warp merges its row's partials into k+1 log-probs; in beam mode warp 0 then runs the audio's beam update.
template <int MODE>
__global__ void __launch_bounds__(kMaxBeam * 32) decode_tail_kernel(const StepParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    RowResult* rr_s = reinterpret_cast<RowResult*>(smem_raw);          // the group's row results
    FastTail& fs = *reinterpret_cast<FastTail*>(smem_raw + kTailRows);
    unsigned char* slice = smem_raw + kTailHead + warp * kTailBytes;   // [merge lists | candidate buffer | winning cells]; the general beam tail's tables overlay it
    WarpBuf& wb = *reinterpret_cast<WarpBuf*>(slice + kTailLists);
    int* glist = reinterpret_cast<int*>(slice + kTailLists + kScanScratch);
    const int row = (int)blockIdx.x * (MODE == MODE_BEAM ? p.beam : 1) + warp;
    const bool fast_tail = MODE == MODE_BEAM && p.beam * p.K <= 32;
#ifdef WRS_TRACE
    const int gw = 148 * kStepWarps + min(row, 1023);
    if (lane == 0) for (int q = 0; q < 8; ++q) g_trace[gw * 8 + q] = 0;
    WRS_TR(0);
    const unsigned long long tm0 = gtime();
#endif
    // everything this row needs from memory is requested before the first use: its two newest tokens (rules), the
    // token prefixes (beam dedup), the running sums and, inside merge_row, the partials
    const int64_t* mine = p.tokens + (int64_t)row * p.tok_stride;
    const int64_t tk_last = mine[p.cur_len - 1], tk_pen = p.cur_len >= 2 ? mine[p.cur_len - 2] : 0;
    // which later beams of this audio hold an identical prefix (HashMap key = whole sequence, beam_search.rs:74-77):
    // the first 32 positions of up to 4 later beams are fetched before the merge so both latencies overlap
    const int bj = MODE == MODE_BEAM ? warp : 0, rbase = row - bj;
    int64_t pre_m = 0, pre_m1 = 0, pre_o[4] = {0, 0, 0, 0};
    float cand_sum = 0.0f;
    int have = 0;
    // (the prefix is staged from two registers per lane, positions lane and lane + 32: prefixes of up to 64 tokens)
    const bool staged = fast_tail && p.cur_len <= 64 && p.beam * p.cur_len <= 256;
    if (MODE == MODE_BEAM) {
        if (lane < p.cur_len) {
            pre_m = mine[lane];
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (bj + 1 + q < p.beam) pre_o[q] = p.tokens[(int64_t)(rbase + bj + 1 + q) * p.tok_stride + lane];
        }
        if (staged && lane + 32 < p.cur_len) pre_m1 = mine[lane + 32];
        if (fast_tail) {
            if (lane < p.beam * p.K) cand_sum = p.sum_in[rbase + lane / p.K];
            if (warp == 0 && lane == 0) have = p.fin_n[blockIdx.x];   // handed to the other warps through fs.have
        }
    }
    if (lane == 0) wb.n = 0;
    __syncwarp();
    const RowRules rr_row = row_rules_from(p.rc, tk_last, tk_pen, p.cur_len, p.sample_begin);
    RowOut r = p.proj_parts ? merge_row_proj(p, row, wb, slice) : merge_row(p, row, rr_row, wb, slice, glist);
    r.flags |= rr_row.flags;
    unsigned same = 0u;
    if (MODE == MODE_BEAM) {
        unsigned diff = 0u;
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (pre_o[q] != pre_m) diff |= 1u << (bj + 1 + q);          // lanes >= cur_len hold 0 == 0
        for (int j2 = bj + 5; j2 < p.beam; ++j2)
            if (lane < p.cur_len && p.tokens[(int64_t)(rbase + j2) * p.tok_stride + lane] != pre_m) diff |= 1u << j2;
        for (int t = lane + 32; t < p.cur_len; t += 32) {
            const int64_t mt = mine[t];
            for (int j2 = bj + 1; j2 < p.beam; ++j2)
                if (p.tokens[(int64_t)(rbase + j2) * p.tok_stride + t] != mt) diff |= 1u << j2;
        }
        diff = __reduce_or_sync(0xffffffffu, diff);
        same = ~diff & ((1u << p.beam) - 1u) & ~((2u << bj) - 1u);
    }
#ifdef WRS_TRACE
    WRS_TR_ADD(4, gtime() - tm0);
    if (lane == 0) { g_trace[gw * 8 + 5] = r.t_sums - tm0; g_trace[gw * 8 + 6] = r.t_cells - tm0; g_trace[gw * 8 + 7] = r.t_loaded ? r.t_loaded - tm0 : 0; }
#endif
    if (lane < p.K) {
        if (p.dbg_lp) p.dbg_lp[(int64_t)row * p.K + lane] = r.lp;
        if (p.dbg_tok) p.dbg_tok[(int64_t)row * p.K + lane] = r.tok;
    }
    if (lane == 0 && p.dbg_flags) p.dbg_flags[row] = (uint8_t)r.flags;
    if (MODE == MODE_GREEDY) { greedy_tail(p, row, r); return; }
    if (MODE == MODE_FILTER) { filter_tail(p, row, rr_row, r.flags); return; }
    {
        RowResult* res = fast_tail ? rr_s + warp : p.rowres + row;
        if (lane < kMaxK) { res->lp[lane] = r.lp; res->tok[lane] = lane < p.K ? r.tok : INT32_MAX; }
        if (lane == 0) { res->flags = r.flags; res->same = same; }
    }
    if (fast_tail && warp == 0 && lane == 0) fs.have = have;
    if (staged) {
        if (lane < p.cur_len) fs.tok_stage[warp * p.cur_len + lane] = pre_m;
        if (lane + 32 < p.cur_len) fs.tok_stage[warp * p.cur_len + lane + 32] = pre_m1;
    }
    __threadfence_block();
    __syncthreads();                                     // the audio's row results are complete (same CTA wrote them)
    if (fast_tail) {
#ifdef WRS_TRACE
        const unsigned long long tt0 = gtime();
#endif
        WRS_TR(1);
        beam_fast(p, (int)blockIdx.x, fs, rr_s, cand_sum, warp);
        WRS_TR(3);
        return;
    }
    if (warp != 0) return;
#ifdef WRS_TRACE
    const unsigned long long tt0 = gtime();
#endif
    beam_tail_general(p, (int)blockIdx.x, *reinterpret_cast<TailScratch*>(slice));
#ifdef WRS_TRACE
    WRS_TR_ADD(6, gtime() - tt0); WRS_TR_ADD(7, 1);
#endif
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
    p.ts_split = dec->ts_split; p.n_text_units = dec->n_text_units; p.n_ts_units = dec->n_ts_units; p.upr = dec->upr;
    p.bm_words = dec->bm_words;
    p.bm = (cur_len == sample_begin) ? bm_first : bm;   // SuppressBlank acts on the first step only (suppress_blanks.rs:40)
    p.partials = dec->d_partials; p.rowres = dec->d_rowres;
    p.row_arrive = dec->d_arrive; p.audio_arrive = dec->d_arrive + dec->max_rows;
    p.eot = dec->cfg.eot; p.max_cand = dec->max_cand; p.max_len = dec->cfg.max_len; p.no_timestamps = dec->cfg.no_timestamps;
    p.fin_tokens = dec->d_fin_tokens; p.fin_len = dec->d_fin_len; p.fin_score = dec->d_fin_score; p.fin_n = dec->d_fin_n;
    p.sync = dec->d_sync; p.status = dec->ctx->d_status;
    return p;
}

// ---- projection mode (decode_proj.cuh): pack x, project + fold, then the same tail launch
int proj_check(wrs_projection* pj, const char* who, const float* x, int64_t x_stride, int64_t n_rows) {
    if (!pj) return fail(nullptr, WRS_ERR_INVALID, std::string(who) + ": null projection");
    wrs_ctx* ctx = pj->dec->ctx;
    if (!x) return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": null x");
    if (n_rows <= 0 || n_rows > pj->dec->max_rows) return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": n_rows outside 1..max_rows");
    if (x_stride < pj->n_state || (x_stride & 3) != 0 || (reinterpret_cast<uintptr_t>(x) & 15) != 0)
        return fail(ctx, WRS_ERR_INVALID, std::string(who) + ": x needs x_stride >= n_state, a multiple of 4, and a 16-byte aligned base");
    return WRS_OK;
}

template <int EPI>
int launch_proj(wrs_projection* pj, ProjParams q, const float* x, int64_t x_stride, cudaStream_t stream) {
    wrs_decoder* dec = pj->dec;
    wrs_ctx* ctx = dec->ctx;
    const size_t smem = sizeof(ProjSmem) + sizeof(uint32_t) * (dec->bm_words + 2);
    if (!pj->attr_set) {
        auto set = [&](auto kern) { return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); };
        WRS_CUDA(ctx, set(proj_kernel<0, false>)); WRS_CUDA(ctx, set(proj_kernel<0, true>));
        WRS_CUDA(ctx, set(proj_kernel<2, false>)); WRS_CUDA(ctx, set(proj_kernel<2, true>));
        WRS_CUDA(ctx, set(proj_kernel<4, false>)); WRS_CUDA(ctx, set(proj_kernel<4, true>));
        WRS_CUDA(ctx, set(proj_kernel<6, false>)); WRS_CUDA(ctx, set(proj_kernel<6, true>));
        WRS_CUDA(ctx, set(proj_kernel<8, false>)); WRS_CUDA(ctx, set(proj_kernel<8, true>));
        WRS_CUDA(ctx, set(proj_kernel<12, false>)); WRS_CUDA(ctx, set(proj_kernel<12, true>));
        WRS_CUDA(ctx, set(proj_kernel<16, false>)); WRS_CUDA(ctx, set(proj_kernel<16, true>));
        pj->attr_set = true;
    }
    q.wp = pj->d_wp; q.xp = pj->d_xp;
    q.n_tiles = pj->n_tiles; q.n_text_tiles = pj->n_text_tiles; q.n_kstages = pj->n_kstages;
    q.n_mtiles = (q.n_rows + kPM - 1) / kPM;
    q.n_vocab = dec->cfg.n_vocab; q.ts_split = dec->ts_split; q.inv_scale = 1.0f / pj->w_scale;
    const int64_t groups = (int64_t)q.n_mtiles * q.n_kstages * 4 * kPM;
    proj_pack_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, stream>>>(x, x_stride, q.n_rows, pj->n_state, q.n_mtiles, kPM, 0, 0, 0, 1.0f, pj->d_xp);
    WRS_CUDA(ctx, cudaGetLastError());
    // (programmatic dependent launch behind the packing kernel -- set-up under the packing -- was measured: 63.0 against
    // 63.5 us per step, within the noise; not kept)
    // Exactly two M tiles (129 .. 256 rows): clusters of two CTAs that share the weights (proj_kernel<.., true>).
    const int pairs = ctx->sm_count / 2;
    if (q.n_mtiles == 2 && pairs >= 1 && ctx->opt_proj_cluster != 0) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(2 * std::min(pairs, pj->n_tiles))); cfg.blockDim = dim3(kPThreads);
        cfg.dynamicSmemBytes = smem; cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        WRS_CUDA(ctx, cudaLaunchKernelEx(&cfg, proj_kernel<EPI, true>, q));
    } else {
        proj_kernel<EPI, false><<<std::min(ctx->sm_count, pj->n_tiles), kPThreads, smem, stream>>>(q);
        WRS_CUDA(ctx, cudaGetLastError());
    }
    return WRS_OK;
}

int ensure_tail_attrs(wrs_decoder* dec);

// Launch 1: one 512-thread CTA per SM at most, every warp gets at least one unit.  Launch 2: one CTA per group, one warp
// per row.
template <int MODE>
int launch_step(wrs_decoder* dec, StepParams p, cudaStream_t stream) {
    wrs_ctx* ctx = dec->ctx;
    const size_t smem = (size_t)kStepWarps * kWarpSmem;
    {
        const int rc_attr = ensure_tail_attrs(dec);
        if (rc_attr != WRS_OK) return rc_attr;
    }
    const int64_t units = (int64_t)p.n_rows * p.upr;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctx->sm_count, (units + kStepWarps - 1) / kStepWarps));
    // the shortest longest run q for which the rows' runs fit the grid's warps (one run per warp); if even one run per
    // partition does not fit, warps take several runs in turn
    {
        const int64_t Wt = (int64_t)grid * kStepWarps;
        const int nt = p.n_text_units, ns = p.n_ts_units;
        int rt = 1, rs = ns > 0 ? 1 : 0;
        for (int q = 1; q <= std::max(nt, ns); ++q) {
            const int a = (nt + q - 1) / q, b2 = ns > 0 ? (ns + q - 1) / q : 0;
            if ((int64_t)(a + b2) * p.n_rows <= Wt && a + b2 <= 32) {     // (the tail merges a row's runs one per lane)
                rs = b2;
                rt = a;      // the fewest runs of that length (more, shorter runs were measured: the per-run work costs more)
                break;
            }
        }
        p.rt = rt; p.rs = rs; p.rpr = rt + rs;
        p.t_base = nt / rt; p.t_rem = nt % rt;
        p.s_base = rs > 0 ? ns / rs : 0; p.s_rem = rs > 0 ? ns % rs : 0;
        const int64_t n_runs = (int64_t)p.rpr * p.n_rows;
        // g / rpr = (g * ceil(2^40 / rpr)) >> 40, exact while g (rpr - 1) < 2^40
        if (n_runs >= (1ll << 31) || n_runs * (p.rpr - 1) >= (1ll << 40))
            return fail(ctx, WRS_ERR_UNSUPPORTED, "decode step: n_rows x runs per row too large for the unit dealing");
        p.n_runs = (int)n_runs;
        p.npw = (int)((n_runs + Wt - 1) / Wt);
        p.rpr_magic = ((1ull << 40) + (uint64_t)p.rpr - 1) / (uint64_t)p.rpr;
    }
#ifdef WRS_TRACE
    // investigation aid (trace builds only, tools/debug/split.sh): WRS_DEBUG_ONLY=scan|tail times one launch alone
    // (results are then meaningless)
    static const char* only = std::getenv("WRS_DEBUG_ONLY");
#else
    const char* only = nullptr;
#endif
    if (!only || only[0] != 't') decode_scan_kernel<<<grid, kStepThreads, smem, stream>>>(p);
    WRS_CUDA(ctx, cudaGetLastError());
    if (only && only[0] == 's') return WRS_OK;
    const int rows_per_group = MODE == MODE_BEAM ? p.beam : 1;
    // (programmatic dependent launch, measured twice: 2 us slower per step -- tail CTAs placed early only wait)
    decode_tail_kernel<MODE><<<p.n_rows / rows_per_group, rows_per_group * 32, (size_t)kTailHead + (size_t)rows_per_group * kTailBytes, stream>>>(p);
    WRS_CUDA(ctx, cudaGetLastError());
    return WRS_OK;
}

int ensure_tail_attrs(wrs_decoder* dec) {
    wrs_ctx* ctx = dec->ctx;
    const size_t smem = (size_t)kStepWarps * kWarpSmem, smem_tail = (size_t)kTailHead + (size_t)kMaxBeam * kTailBytes;
    if (!dec->smem_attr_set) {
        WRS_CUDA(ctx, cudaFuncSetAttribute(decode_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        WRS_CUDA(ctx, cudaFuncSetAttribute(decode_tail_kernel<MODE_BEAM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tail));
        WRS_CUDA(ctx, cudaFuncSetAttribute(decode_tail_kernel<MODE_GREEDY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tail));
        WRS_CUDA(ctx, cudaFuncSetAttribute(decode_tail_kernel<MODE_FILTER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tail));
        dec->smem_attr_set = true;
    }
    return WRS_OK;
}

void set_debug(StepParams& p, const wrs_step_debug* dbg) {
    if (!dbg) return;
    p.dbg_lp = dbg->cand_logprob; p.dbg_tok = dbg->cand_token; p.dbg_flags = dbg->row_flags;
    p.dbg_fin_count = dbg->fin_count; p.dbg_fin_source = dbg->fin_source; p.dbg_fin_score = dbg->fin_score;
}

}  // namespace

extern "C" {

#ifdef WRS_EXP_PJ_TIMES
int wrs_debug_pj_times(unsigned long long* out_host, int n) {
    return (int)cudaMemcpyFromSymbol(out_host, wrs::g_pj_times, sizeof(unsigned long long) * n);
}
int wrs_debug_pj_clk(long long* out_host, int n) { return (int)cudaMemcpyFromSymbol(out_host, wrs::g_pj_clk, sizeof(long long) * n); }
#endif
#ifdef WRS_TRACE
int wrs_debug_trace_read(unsigned long long* out_host, int n) {
    return (int)cudaMemcpyFromSymbol(out_host, wrs::g_trace, sizeof(unsigned long long) * n);
}
#endif

int wrs_decoder_create(wrs_ctx* ctx, const wrs_decode_cfg* cfg, const int32_t* suppress_ids, int32_t n_suppress,
                       int32_t max_rows, wrs_decoder** out) {
    if (!ctx || !cfg || !out) return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: null argument");
    *out = nullptr;
    const wrs_decode_cfg& c = *cfg;
    if (c.n_vocab < 8 || c.beam_size < 1 || c.beam_size > kMaxBeam || max_rows < c.beam_size || max_rows % c.beam_size != 0 ||
        max_rows >= 32768)
        return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: need n_vocab >= 8, 1 <= beam_size <= 15, max_rows a multiple of beam_size below 32768");
    if (c.eot < 0 || c.eot >= c.n_vocab || (c.timestamps && (c.ts_begin <= c.eot || c.ts_begin > c.n_vocab ||
                                                              c.no_timestamps < 0 || c.no_timestamps >= c.n_vocab)))
        return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: need 0 <= eot < ts_begin <= n_vocab and a valid no_timestamps id");
    if (n_suppress < 0 || (n_suppress > 0 && !suppress_ids)) return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: bad suppress list");
    if (c.max_len < 2 || !(c.patience > 0.0f)) return fail(ctx, WRS_ERR_INVALID, "wrs_decoder_create: need max_len >= 2 and patience > 0");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    wrs_decoder* d = new wrs_decoder();
    d->ctx = ctx; d->cfg = c; d->max_rows = max_rows; d->max_audio = max_rows / c.beam_size;
    d->max_cand = std::max(1, (int)std::round(c.patience * (float)c.beam_size));   // f32::round, beam_search.rs:122
    // unit geometry: text ids [0, ts_split) and timestamp ids [ts_split, V) never share a unit
    d->ts_split = c.timestamps ? c.ts_begin : c.n_vocab;
    // units start on 16-byte boundaries of a row's memory: up to 3 slack ids in front of a partition
    d->n_text_units = (d->ts_split + 3 + kUnit - 1) / kUnit;
    d->n_ts_units = c.n_vocab > d->ts_split ? (c.n_vocab - d->ts_split + 3 + kUnit - 1) / kUnit : 0;
    d->upr = d->n_text_units + d->n_ts_units;
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
        (e = cudaMalloc(&d->d_partials, sizeof(Partial) * d->upr * (size_t)max_rows)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_rowres, sizeof(RowResult) * (size_t)max_rows)) != cudaSuccess ||
        (e = cudaMalloc(&d->d_arrive, sizeof(int32_t) * (max_rows + d->max_audio))) != cudaSuccess ||
        (e = cudaMemset(d->d_arrive, 0, sizeof(int32_t) * (max_rows + d->max_audio))) != cudaSuccess ||
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
    cudaFree(d->d_suppress); cudaFree(d->d_partials); cudaFree(d->d_rowres); cudaFree(d->d_fin_tokens); cudaFree(d->d_fin_len);
    cudaFree(d->d_fin_score); cudaFree(d->d_fin_n); cudaFree(d->d_sync);
    delete d;
}

// internal (not in wrs.h): what no_speech.cu borrows from a decoder -- the partial buffer as {max, sum} pairs
// ([max_rows][units per row] x 144 B is far more than [max_rows][8] x 8 B) and the self-resetting row counters
int wrs_decoder_scratch_(wrs_decoder* dec, wrs_ctx** ctx, int32_t* max_rows, int32_t* n_vocab, int32_t* no_speech,
                         void** pairs, int32_t** counters) {
    *ctx = dec->ctx; *max_rows = dec->max_rows; *n_vocab = dec->cfg.n_vocab; *no_speech = dec->cfg.no_speech;
    *pairs = dec->d_partials; *counters = dec->d_arrive;
    return WRS_OK;
}

int wrs_decoder_reset(wrs_decoder* d, wrs_stream_t stream) {
    if (!d) return fail(nullptr, WRS_ERR_INVALID, "wrs_decoder_reset: null decoder");
    WRS_CUDA(d->ctx, cudaSetDevice(d->ctx->device));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_fin_n, 0, sizeof(int32_t) * d->max_audio, (cudaStream_t)stream));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_sync, 0, sizeof(int32_t) * 2, (cudaStream_t)stream));
    WRS_CUDA(d->ctx, cudaMemsetAsync(d->d_arrive, 0, sizeof(int32_t) * (d->max_rows + d->max_audio), (cudaStream_t)stream));
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
        rc = launch_step<MODE_FILTER>(dec, p, stream);
        if (rc != WRS_OK) return rc;
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
    return launch_step<MODE_BEAM>(dec, p, stream);
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
    return launch_step<MODE_GREEDY>(dec, p, stream);
}

// ---- projection + decode step (reference src/model.rs:521 then src/decode/mod.rs:257-270)
int wrs_projection_create(wrs_decoder* dec, const float* token_emb, int64_t emb_stride, int32_t n_state, wrs_stream_t stream_,
                          wrs_projection** out) {
    if (!dec || !out) return fail(dec ? dec->ctx : nullptr, WRS_ERR_INVALID, "wrs_projection_create: null argument");
    *out = nullptr;
    wrs_ctx* ctx = dec->ctx;
    if (!token_emb || n_state < kPK || n_state % kPK != 0 || emb_stride < n_state || (emb_stride & 3) != 0 ||
        (reinterpret_cast<uintptr_t>(token_emb) & 15) != 0)
        return fail(ctx, WRS_ERR_INVALID, "wrs_projection_create: need n_state a positive multiple of 32, emb_stride >= n_state and a multiple of 4, a 16-byte aligned base");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    wrs_projection* pj = new wrs_projection();
    pj->dec = dec; pj->n_state = n_state; pj->n_kstages = n_state / kPK;
    const int V = dec->cfg.n_vocab;
    pj->n_text_tiles = (dec->ts_split + kPN - 1) / kPN;
    pj->n_tiles = pj->n_text_tiles + (V - dec->ts_split + kPN - 1) / kPN;
    pj->max_mtiles = (dec->max_rows + kPM - 1) / kPM;
    const size_t wbytes = (size_t)pj->n_tiles * pj->n_kstages * 2 * kPBBytes;
    const size_t xbytes = (size_t)pj->max_mtiles * pj->n_kstages * 2 * kPABytes;
    const size_t pbytes = sizeof(float) * (size_t)dec->max_rows * (2 + 2 * (dec->cfg.beam_size + 1)) * ((pj->n_tiles + 3) & ~3);
    cudaError_t e;
    if ((e = cudaMalloc(&pj->d_wp, wbytes)) != cudaSuccess || (e = cudaMalloc(&pj->d_xp, xbytes)) != cudaSuccess ||
        (e = cudaMalloc(&pj->d_parts, pbytes)) != cudaSuccess) {
        wrs_projection_destroy(pj);
        return cuda_fail(ctx, e, "projection buffers");
    }
    // the power of two that puts the largest |w| in [512, 1024): one reduction over the embedding, read back here
    // (creation is a load-time call; it synchronises `stream`)
    {
        unsigned* d_bits = reinterpret_cast<unsigned*>(pj->d_parts);      // (scratch: the record buffer is not in use yet)
        unsigned bits = 0u;
        if ((e = cudaMemsetAsync(d_bits, 0, sizeof(unsigned), stream)) == cudaSuccess) {
            proj_absmax_kernel<<<ctx->sm_count * 4, 256, 0, stream>>>(token_emb, emb_stride, V, n_state, d_bits);
            if ((e = cudaGetLastError()) == cudaSuccess &&
                (e = cudaMemcpyAsync(&bits, d_bits, sizeof(unsigned), cudaMemcpyDeviceToHost, stream)) == cudaSuccess)
                e = cudaStreamSynchronize(stream);
        }
        if (e != cudaSuccess) { wrs_projection_destroy(pj); return cuda_fail(ctx, e, "projection scale"); }
        float wmax;
        memcpy(&wmax, &bits, sizeof(float));
        int ex = 0;
        if (wmax > 0.0f) { std::frexp(wmax, &ex); pj->w_scale = std::ldexp(1.0f, 10 - ex); }   // wmax = f x 2^ex, f in [0.5, 1)
    }
    const int64_t groups = (int64_t)pj->n_tiles * pj->n_kstages * 4 * kPN;
    proj_pack_kernel<<<(unsigned)std::min<int64_t>((groups + 255) / 256, 65535), 256, 0, stream>>>(
        token_emb, emb_stride, V, n_state, pj->n_tiles, kPN, pj->n_text_tiles, dec->ts_split, 1, pj->w_scale, pj->d_wp);
    e = cudaGetLastError();
    if (e != cudaSuccess) { wrs_projection_destroy(pj); return cuda_fail(ctx, e, "projection packing"); }
    *out = pj;
    return WRS_OK;
}

void wrs_projection_destroy(wrs_projection* pj) {
    if (!pj) return;
    cudaSetDevice(pj->dec->ctx->device);
    cudaFree(pj->d_wp); cudaFree(pj->d_xp); cudaFree(pj->d_parts);
    delete pj;
}

int wrs_project_logits(wrs_projection* pj, const float* x, int64_t x_stride, int64_t n_rows, float* logits, int64_t row_stride,
                       wrs_stream_t stream_) {
    int rc = proj_check(pj, "wrs_project_logits", x, x_stride, n_rows);
    if (rc != WRS_OK) return rc;
    wrs_ctx* ctx = pj->dec->ctx;
    if (!logits || row_stride < pj->dec->cfg.n_vocab) return fail(ctx, WRS_ERR_INVALID, "wrs_project_logits: need logits and row_stride >= n_vocab");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    ProjParams q;
    memset(&q, 0, sizeof(q));
    q.n_rows = (int)n_rows; q.logits = logits; q.row_stride = row_stride;
    return launch_proj<0>(pj, q, x, x_stride, (cudaStream_t)stream_);
}

int wrs_project_beam_step(wrs_projection* pj, const float* x, int64_t x_stride, int64_t n_rows, const int64_t* tokens,
                          int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_in, int64_t* next_tokens,
                          int64_t next_stride, int32_t* source_idx, float* sum_out, int32_t* completed, const wrs_step_debug* dbg,
                          wrs_stream_t stream_) {
    int rc = proj_check(pj, "wrs_project_beam_step", x, x_stride, n_rows);
    if (rc != WRS_OK) return rc;
    wrs_decoder* dec = pj->dec;
    // (the logits arguments of the common check: x stands in as the non-null, aligned buffer; there is no row_stride)
    rc = check_step_args(dec, "wrs_project_beam_step", x, dec->cfg.n_vocab, n_rows, tokens, tok_stride, cur_len, sample_begin);
    if (rc != WRS_OK) return rc;
    wrs_ctx* ctx = dec->ctx;
    const wrs_decode_cfg& c = dec->cfg;
    if (!sum_in || !next_tokens || !source_idx || !sum_out || !completed) return fail(ctx, WRS_ERR_INVALID, "wrs_project_beam_step: null output");
    if (n_rows % c.beam_size != 0) return fail(ctx, WRS_ERR_INVALID, "wrs_project_beam_step: n_rows not a multiple of beam_size (beam_search.rs:29)");
    if (next_stride < cur_len + 1) return fail(ctx, WRS_ERR_INVALID, "wrs_project_beam_step: next_stride < cur_len + 1");
    if ((const void*)next_tokens == (const void*)tokens) return fail(ctx, WRS_ERR_INVALID, "wrs_project_beam_step: next_tokens must not alias tokens");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    StepParams p = step_params(dec, nullptr, c.n_vocab, n_rows, tokens, tok_stride, cur_len, sample_begin, c.beam_size + 1,
                               c.beam_size, dec->d_bm_base, dec->d_bm_first);
    p.sum_in = sum_in; p.sum_out = sum_out; p.next_tokens = next_tokens; p.next_stride = next_stride;
    p.source_idx = source_idx; p.completed = completed;
    set_debug(p, dbg);
    p.proj_parts = pj->d_parts; p.proj_tiles = pj->n_tiles; p.proj_text_tiles = pj->n_text_tiles; p.proj_tp = (pj->n_tiles + 3) & ~3;
    ProjParams q;
    memset(&q, 0, sizeof(q));
    q.n_rows = (int)n_rows; q.tokens = tokens; q.tok_stride = tok_stride; q.cur_len = p.cur_len; q.sample_begin = p.sample_begin;
    q.K = p.K; q.rc = p.rc; q.bm = p.bm; q.bm_words = p.bm_words; q.parts = pj->d_parts; q.tiles_padded = p.proj_tp;
    rc = ensure_tail_attrs(dec);
    if (rc != WRS_OK) return rc;
    rc = p.K <= 2 ? launch_proj<2>(pj, q, x, x_stride, stream) : p.K <= 4 ? launch_proj<4>(pj, q, x, x_stride, stream)
       : p.K <= 6 ? launch_proj<6>(pj, q, x, x_stride, stream) : p.K <= 8 ? launch_proj<8>(pj, q, x, x_stride, stream)
       : p.K <= 12 ? launch_proj<12>(pj, q, x, x_stride, stream) : launch_proj<16>(pj, q, x, x_stride, stream);
    if (rc != WRS_OK) return rc;
    decode_tail_kernel<MODE_BEAM><<<p.n_rows / p.beam, p.beam * 32, (size_t)kTailHead + (size_t)p.beam * kTailBytes, stream>>>(p);
    WRS_CUDA(ctx, cudaGetLastError());
    return WRS_OK;
}

int wrs_project_greedy_step(wrs_projection* pj, const float* x, int64_t x_stride, int64_t n_rows, const int64_t* tokens,
                            int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_in, int64_t* next_tokens,
                            int64_t next_stride, float* sum_out, int32_t* completed, const wrs_step_debug* dbg, wrs_stream_t stream_) {
    int rc = proj_check(pj, "wrs_project_greedy_step", x, x_stride, n_rows);
    if (rc != WRS_OK) return rc;
    wrs_decoder* dec = pj->dec;
    rc = check_step_args(dec, "wrs_project_greedy_step", x, dec->cfg.n_vocab, n_rows, tokens, tok_stride, cur_len, sample_begin);
    if (rc != WRS_OK) return rc;
    wrs_ctx* ctx = dec->ctx;
    if (!sum_in || !next_tokens || !sum_out || !completed) return fail(ctx, WRS_ERR_INVALID, "wrs_project_greedy_step: null output");
    if (next_stride < cur_len + 1) return fail(ctx, WRS_ERR_INVALID, "wrs_project_greedy_step: next_stride < cur_len + 1");
    if ((const void*)next_tokens == (const void*)tokens) return fail(ctx, WRS_ERR_INVALID, "wrs_project_greedy_step: next_tokens must not alias tokens");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    StepParams p = step_params(dec, nullptr, dec->cfg.n_vocab, n_rows, tokens, tok_stride, cur_len, sample_begin, 1, 1,
                               dec->d_bm_base, dec->d_bm_first);
    p.sum_in = sum_in; p.sum_out = sum_out; p.next_tokens = next_tokens; p.next_stride = next_stride; p.completed = completed;
    set_debug(p, dbg);
    p.proj_parts = pj->d_parts; p.proj_tiles = pj->n_tiles; p.proj_text_tiles = pj->n_text_tiles; p.proj_tp = (pj->n_tiles + 3) & ~3;
    ProjParams q;
    memset(&q, 0, sizeof(q));
    q.n_rows = (int)n_rows; q.tokens = tokens; q.tok_stride = tok_stride; q.cur_len = p.cur_len; q.sample_begin = p.sample_begin;
    q.K = 1; q.rc = p.rc; q.bm = p.bm; q.bm_words = p.bm_words; q.parts = pj->d_parts; q.tiles_padded = p.proj_tp;
    rc = ensure_tail_attrs(dec);
    if (rc != WRS_OK) return rc;
    rc = launch_proj<2>(pj, q, x, x_stride, stream);
    if (rc != WRS_OK) return rc;
    decode_tail_kernel<MODE_GREEDY><<<p.n_rows, 32, (size_t)kTailHead + (size_t)kTailBytes, stream>>>(p);
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

// K2 building blocks: warp-autonomous, single-read scan of logits rows.
//
// Replaces, per (segment, beam) row, the ATen calls of reference src/decode/filter/*.rs
// (index_fill_, slice fill_, log_softmax, logsumexp, max) and src/decode/extractor/beam_search.rs:37,62-65
// (log_softmax, topk) with one read of the row: masks are predicates, never written back.
//
// A row is cut into UNITS of 896 elements that start on 16-byte boundaries of the row's memory (57 text units +
// 2 timestamp units for V = 51865; a unit never straddles ts_begin, the first and last unit of a partition are
// ragged).  A warp owns a run of consecutive units of one partition: the unit being reduced sits in registers (7 x 16
// bytes per lane) while the next one is in flight into the warp's shared-memory stages (cp.async.bulk); every unit is
// folded into the run's {max, sum exp(x - ref), top-K}.  Top-K is threshold based: K rounds of redux.max over the lane
// maxima give a value that at least K candidates reach, so only the handful at or above it are gathered (shared
// memory, per warp) and merged.  The scan launch ranks CELLS (a lane's 28 ids of a unit, decode.cu fold_unit); the
// per-id form (reduce_unit) is what the tail falls back to when cells tie.  No block-level barrier in the scan.
//
// Ordering everywhere is (value desc, id asc) -- north_star's "ties broken by lowest index".
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace wrs {

#ifndef WRS_STEP_WARPS
#define WRS_STEP_WARPS 24
#endif
constexpr int kStepWarps = WRS_STEP_WARPS;               // one CTA per SM: 24 warps, up to 80 registers per thread
constexpr int kStepThreads = kStepWarps * 32;
constexpr int kNV = 7;                                   // float4 slots per lane per unit
constexpr int kUnit = 32 * 4 * kNV;                      // 896 elements per unit, 16-byte aligned in memory
constexpr int kMaxK = 16;                                // beam + 1 <= 16
constexpr int kCandCap = 128;                            // gathered candidates per warp before the exact fallback
constexpr float kLog2e = 1.4426950408889634f;

// what a warp leaves in global memory for one run of units; v / id list the run's best CELLS: value = the cell's
// maximum, id = 32 * unit in row + lane
struct __align__(16) Partial {
    float m, s;            // max and sum of exp(x - m) over the run's surviving ids (m = -inf, s = 0 if none)
    int n, tie;            // valid entries of v/id; tie != 0: an entry equal to v[n-1] was dropped
    float v[kMaxK];        // the run's best values, sorted (value desc, id asc), padded with -inf
    int id[kMaxK];
};

__device__ __forceinline__ bool better(float va, int ia, float vb, int ib) {   // (va,ia) precedes (vb,ib)
    return va > vb || (va == vb && ia < ib);
}
__device__ __forceinline__ float warp_max(float x) {     // CREDUX.MAX.F32 (sm_100a); NaN operands are ignored
    float y;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// exp(a - b) with exp(-inf - -inf) = 0: a fully masked partition is normal
__device__ __forceinline__ float exp_diff(float a, float b) { return a == -CUDART_INF_F ? 0.0f : expf(a - b); }

// per-row description of which ids survive the filters (warp-uniform)
struct RowRules {
    int t_lo, t_hi;   // surviving text range      [t_lo, t_hi)  (ids below ts_begin)
    int s_lo, s_hi;   // surviving timestamp range [s_lo, s_hi)
    unsigned flags;   // WRS_ROW_* bits 0..2
};

struct RuleCfg {
    int n_vocab, eot, ts_begin, timestamps, max_initial_ts_index;
};

// src/decode/filter/timestamp_tokens.rs:47-78 as range arithmetic; `last`/`penult` are the row's two newest tokens
__device__ __forceinline__ RowRules row_rules_from(const RuleCfg& c, int64_t last, int64_t penult, int cur_len, int sample_begin) {
    RowRules r;
    r.flags = 0;
    if (!c.timestamps) { r.t_lo = 0; r.t_hi = c.n_vocab; r.s_lo = r.s_hi = c.n_vocab; return r; }
    r.t_lo = 0; r.t_hi = c.ts_begin; r.s_lo = c.ts_begin; r.s_hi = c.n_vocab;
    const int n = cur_len - sample_begin;
    const bool last_ts = n >= 1 && last >= c.ts_begin;
    const bool penult_ts = n < 2 || penult >= c.ts_begin;
    if (last_ts) {
        if (penult_ts) { r.s_hi = r.s_lo; r.flags |= 1u; }                 // has to be non-timestamp
        else { r.t_lo = min(c.eot, r.t_hi); r.flags |= 2u; }              // cannot be normal text
    }
    if (n == 0) {                                                          // first step
        r.t_lo = r.t_hi; r.flags |= 4u;
        if (c.max_initial_ts_index >= 0) r.s_hi = min(r.s_hi, c.ts_begin + c.max_initial_ts_index + 1);
    }
    return r;
}
__device__ __forceinline__ RowRules row_rules(const RuleCfg& c, const int64_t* __restrict__ row_tokens, int cur_len,
                                              int sample_begin) {
    const int64_t last = row_tokens[cur_len - 1];
    const int64_t penult = cur_len >= 2 ? row_tokens[cur_len - 2] : 0;
    return row_rules_from(c, last, penult, cur_len, sample_begin);
}

// per-warp candidate list in shared memory: entries [0, n) hold the run's current best (sorted) followed by
// whatever the unit being reduced gathered
struct WarpBuf {
    float v[kCandCap];
    int id[kCandCap];
    int n;
};

// A value that at least min(K, #lanes with a finite maximum) elements reach: K rounds of "warp max, winners drop out".
__device__ __forceinline__ float kth_threshold(float lane_max, int K) {
    float t = lane_max, tau = CUDART_INF_F;
    for (int r = 0; r < K; ++r) {
        tau = warp_max(t);
        if (t == tau) t = -CUDART_INF_F;
    }
    return tau;
}

// Rank-sort entries [0, n) of wb (n <= kCandCap) and keep the best min(n, K) in wb[0..): returns the new count.
__device__ __forceinline__ int warp_compact(WarpBuf& wb, int n, int K, bool& tie) {
    const int lane = threadIdx.x & 31;
    constexpr int kPer = kCandCap / 32;
    float mv[kPer]; int mi[kPer], rk[kPer];
#pragma unroll
    for (int q = 0; q < kPer; ++q) {
        const int e = lane + 32 * q;
        mv[q] = e < n ? wb.v[e] : -CUDART_INF_F; mi[q] = e < n ? wb.id[e] : INT32_MAX; rk[q] = 0;
    }
    for (int d = 0; d < n; ++d) {
        const float dv = wb.v[d]; const int di = wb.id[d];
#pragma unroll
        for (int q = 0; q < kPer; ++q) rk[q] += better(dv, di, mv[q], mi[q]) ? 1 : 0;
    }
    __syncwarp();
#pragma unroll
    for (int q = 0; q < kPer; ++q)
        if (lane + 32 * q < n && rk[q] < K) { wb.v[rk[q]] = mv[q]; wb.id[rk[q]] = mi[q]; }
    __syncwarp();
    if (n > K) {                                         // was an entry equal to the K-th value cut off?
        const float vk = wb.v[K - 1];
        bool t = false;
#pragma unroll
        for (int q = 0; q < kPer; ++q) t |= lane + 32 * q < n && rk[q] >= K && mv[q] == vk;
        tie |= __any_sync(0xffffffffu, t);
    }
    return min(n, K);
}

// One warp-wide argmax step of the exact fallback: every lane proposes its best element strictly after the previous
// pick; returns the winner on every lane.
__device__ __forceinline__ void warp_pick(float& bv, int& bi) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float v2 = __shfl_xor_sync(0xffffffffu, bv, o); const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
        if (better(v2, i2, bv, bi)) { bv = v2; bi = i2; }
    }
}

// The running statistics of the (row, partition) run a warp is reducing.
//   softmax: every lane sums 2^(x log2 e - ref) for its own elements against an integer reference fixed by the run's first
//            unit (re-based only if a later unit tops it by more than 2^92), so a unit costs no cross-lane traffic;
//   top-K:   lane r < n holds the run's r-th best candidate (rv, ri) in (value desc, id asc) order, other lanes hold
//            (-inf, INT32_MAX); wb[0, np) holds gathered candidates not merged yet; tau is a value the run's K-th best
//            is known to reach (-inf while unknown).
struct Run { float ref, ls, lm, tau, rv, m, s; int ri, n, np; bool ref_set, tie; };
//   tie:     a candidate equal to the list's K-th value was dropped (the list is then not the whole story at that value)

__device__ __forceinline__ void run_reset(Run& run) {
    run.ref = 0.0f; run.ls = 0.0f; run.lm = -CUDART_INF_F; run.tau = -CUDART_INF_F; run.ref_set = false;
    run.m = -CUDART_INF_F; run.s = 0.0f; run.rv = -CUDART_INF_F; run.ri = INT32_MAX; run.n = 0; run.np = 0; run.tie = false;
}

// Many candidates: rank-sort in shared memory (the old list is appended behind the new candidates).  Out of line and
// by value: it is rare, and inlined three times it made up a sixth of the scan kernel's code.
__device__ __noinline__ Run run_merge_slow(Run run, WarpBuf& wb, int n_new, int K) {
    const int lane = threadIdx.x & 31;
    if (lane < run.n) { wb.v[n_new + lane] = run.rv; wb.id[n_new + lane] = run.ri; }
    __syncwarp();
    const int n = warp_compact(wb, n_new + run.n, K, run.tie);
    run.rv = lane < n ? wb.v[lane] : -CUDART_INF_F; run.ri = lane < n ? wb.id[lane] : INT32_MAX; run.n = n;
    __syncwarp();
    return run;
}

// Merge the n_new gathered candidates wb[0, n_new) into the run's list, keeping the best K.
__device__ __forceinline__ void run_merge(Run& run, WarpBuf& wb, int n_new, int K) {
    const int lane = threadIdx.x & 31;
    if (n_new + K <= 32) {
        // one entry per lane: the old list on lanes [0, K), the new candidates on lanes [K, K + n_new)
        float cv = run.rv; int ci = run.ri;
        if (lane >= K) {
            const int e = lane - K;
            cv = e < n_new ? wb.v[e] : -CUDART_INF_F; ci = e < n_new ? wb.id[e] : INT32_MAX;
        }
        float nv = -CUDART_INF_F, m = -CUDART_INF_F; int ni = INT32_MAX, cnt = 0;
        for (int r = 0; r < K; ++r) {
            m = warp_max(cv);
            if (m == -CUDART_INF_F) break;
            const int w = __reduce_min_sync(0xffffffffu, cv == m ? ci : INT32_MAX);
            if (cv == m && ci == w) cv = -CUDART_INF_F;
            if (lane == r) { nv = m; ni = w; }
            ++cnt;
        }
        if (cnt == K && __any_sync(0xffffffffu, cv == m)) run.tie = true;   // (m > -inf here: extracted lanes hold -inf)
        run.rv = nv; run.ri = ni; run.n = cnt;
        __syncwarp();
        return;
    }
    run = run_merge_slow(run, wb, n_new, K);
}

// Exact fallback for massive ties: K rounds of warp-wide selection over {run list} U {the lane's NE elements}, each
// strictly after the previous pick.  `val(e)` / `idx(e)` give the lane's e-th element (-inf = absent).
template <int NE, class Val, class Idx>
__device__ __forceinline__ void run_exact(Run& run, int K, Val val, Idx idx) {
    const int lane = threadIdx.x & 31;
    const float ov = run.rv; const int oi = run.ri;
    float pv = CUDART_INF_F, nv = -CUDART_INF_F; int pi = -1, ni = INT32_MAX, cnt = 0;
    for (int r = 0; r < K; ++r) {
        float bv = -CUDART_INF_F; int bi = INT32_MAX;
        if ((ov < pv || (ov == pv && oi > pi)) && ov > -CUDART_INF_F) { bv = ov; bi = oi; }
#pragma unroll
        for (int e = 0; e < NE; ++e) {
            const float v = val(e); const int id = idx(e);
            const bool after = v < pv || (v == pv && id > pi);
            if (after && v > -CUDART_INF_F && better(v, id, bv, bi)) { bv = v; bi = id; }
        }
        warp_pick(bv, bi);
        if (bi == INT32_MAX) break;
        if (lane == r) { nv = bv; ni = bi; }
        pv = bv; pi = bi; ++cnt;
    }
    run.rv = nv; run.ri = ni; run.n = cnt;
}

}  // namespace wrs

// K2 building blocks: single-pass masked (max, sum-exp, top-K) scan of a logits row segment.
//
// Replaces, per (segment, beam) row, the ATen calls of reference src/decode/filter/*.rs
// (index_fill_, slice fill_, log_softmax, logsumexp, max) and src/decode/extractor/beam_search.rs:37,62-65
// (log_softmax, topk) with one read of the row: masks are predicates, never written back.
//
// Ordering everywhere is (value desc, token id asc) -- north_star's "ties broken by lowest index".
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace wrs {

constexpr int kScanThreads = 256;
constexpr int kScanWarps = kScanThreads / 32;
constexpr int kMaxChunks = 32;   // chunks per row (one lane per chunk in the merge kernel)

// statistics of one masked range: running max, sum of exp(x - max), and the K best (value, id) pairs
template <int KCAP>
struct Stats {
    float m, s;
    float v[KCAP];
    int id[KCAP];
    __device__ __forceinline__ void init() {
        m = -CUDART_INF_F; s = 0.0f;
#pragma unroll
        for (int i = 0; i < KCAP; ++i) { v[i] = -CUDART_INF_F; id[i] = INT32_MAX; }
    }
};

// what a CTA leaves in global memory for one (row, chunk, partition)
template <int KCAP>
struct Partial {
    float m, s;
    float v[KCAP];
    int id[KCAP];
};

__device__ __forceinline__ bool better(float va, int ia, float vb, int ib) {   // (va,ia) precedes (vb,ib)
    return va > vb || (va == vb && ia < ib);
}

// insert (x, e) into a sorted list; caller guarantees x > v[KCAP-1] or (x == v[KCAP-1] && e < id[KCAP-1])
template <int KCAP>
__device__ __forceinline__ void topk_insert(float (&v)[KCAP], int (&id)[KCAP], float x, int e) {
    v[KCAP - 1] = x; id[KCAP - 1] = e;
#pragma unroll
    for (int j = KCAP - 1; j > 0; --j) {
        if (better(v[j], id[j], v[j - 1], id[j - 1])) {
            const float tv = v[j]; v[j] = v[j - 1]; v[j - 1] = tv;
            const int ti = id[j]; id[j] = id[j - 1]; id[j - 1] = ti;
        }
    }
}

// exp(a - b) with the convention exp(-inf - -inf) = 0 (a fully masked partition is normal)
__device__ __forceinline__ float exp_diff(float a, float b) { return a == -CUDART_INF_F ? 0.0f : __expf(a - b); }

// merge (m2, s2) into (m, s)
__device__ __forceinline__ void ms_merge(float& m, float& s, float m2, float s2) {
    const float M = fmaxf(m, m2);
    s = s * exp_diff(m, M) + s2 * exp_diff(m2, M);
    m = M;
}

// Warp-wide K-way merge: every lane holds a sorted list; afterwards every lane holds the same merged
// top-KCAP list.  KOUT rounds of "find the best head, pop it".
template <int KCAP>
__device__ __forceinline__ void warp_topk_merge(float (&v)[KCAP], int (&id)[KCAP], int kout) {
    float ov[KCAP];
    int oi[KCAP];
#pragma unroll
    for (int r = 0; r < KCAP; ++r) {
        ov[r] = -CUDART_INF_F; oi[r] = INT32_MAX;
        if (r < kout) {
            float bv = v[0]; int bi = id[0];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float v2 = __shfl_xor_sync(0xffffffffu, bv, o);
                const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
                if (better(v2, i2, bv, bi)) { bv = v2; bi = i2; }
            }
            ov[r] = bv; oi[r] = bi;
            if (id[0] == bi && bi != INT32_MAX) {   // ids are unique across lanes: exactly one lane pops
#pragma unroll
                for (int j = 0; j < KCAP - 1; ++j) { v[j] = v[j + 1]; id[j] = id[j + 1]; }
                v[KCAP - 1] = -CUDART_INF_F; id[KCAP - 1] = INT32_MAX;
            }
        }
    }
#pragma unroll
    for (int r = 0; r < KCAP; ++r) { v[r] = ov[r]; id[r] = oi[r]; }
}

template <int KCAP>
__device__ __forceinline__ void warp_stats_merge(Stats<KCAP>& st, int kout) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float m2 = __shfl_xor_sync(0xffffffffu, st.m, o);
        const float s2 = __shfl_xor_sync(0xffffffffu, st.s, o);
        ms_merge(st.m, st.s, m2, s2);
    }
    warp_topk_merge<KCAP>(st.v, st.id, kout);
}

// per-row description of which ids survive the filters (all CTA-uniform)
struct RowRules {
    int t_lo, t_hi;   // surviving text range   [t_lo, t_hi)  (ids below ts_begin)
    int s_lo, s_hi;   // surviving timestamp range [s_lo, s_hi)
    unsigned flags;   // WRS_ROW_* bits 0..2
};

struct RuleCfg {
    int n_vocab, eot, ts_begin, timestamps, max_initial_ts_index;
};

// src/decode/filter/timestamp_tokens.rs:47-78 as range arithmetic
__device__ __forceinline__ RowRules row_rules(const RuleCfg& c, const int64_t* __restrict__ row_tokens, int cur_len,
                                              int sample_begin) {
    RowRules r;
    r.flags = 0;
    if (!c.timestamps) { r.t_lo = 0; r.t_hi = c.n_vocab; r.s_lo = r.s_hi = c.n_vocab; return r; }
    r.t_lo = 0; r.t_hi = c.ts_begin; r.s_lo = c.ts_begin; r.s_hi = c.n_vocab;
    const int n = cur_len - sample_begin;
    const bool last_ts = n >= 1 && row_tokens[cur_len - 1] >= c.ts_begin;
    const bool penult_ts = n < 2 || row_tokens[cur_len - 2] >= c.ts_begin;
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

// Scan ids [lo, hi) of one row (row pointer rp, element 0 = id 0) into `st`.  `bm` is the suppress bitmap
// (bit set = masked).  All threads of the CTA take part; each keeps its own Stats.
template <int KCAP>
__device__ __forceinline__ void scan_range(const float* __restrict__ rp, int lo, int hi,
                                           const uint32_t* __restrict__ bm, Stats<KCAP>& st) {
    if (hi <= lo) return;
    const int tid = threadIdx.x;
    auto one = [&](float x, int e) {
        if ((bm[e >> 5] >> (e & 31)) & 1u) return;
        if (x > st.m) { st.s = st.s * exp_diff(st.m, x) + 1.0f; st.m = x; }
        else if (x == x) st.s += exp_diff(x, st.m);
        else st.s = x;   // NaN poisons the sum, detected later
        if (x > st.v[KCAP - 1]) topk_insert<KCAP>(st.v, st.id, x, e);
    };
    // head: up to 3 scalars until 16-byte alignment
    const int mis = (int)((reinterpret_cast<uintptr_t>(rp + lo) >> 2) & 3);
    const int head = min((4 - mis) & 3, hi - lo);
    if (tid < head) one(rp[lo + tid], lo + tid);
    const int b0 = lo + head;
    const int nvec = (hi - b0) >> 2;
    const float4* __restrict__ vp = reinterpret_cast<const float4*>(rp + b0);
    constexpr int U = 4;
    for (int v0 = tid; v0 < nvec; v0 += kScanThreads * U) {
        float4 x[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int vi = v0 + u * kScanThreads;
            x[u] = vi < nvec ? __ldcs(vp + vi) : make_float4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int vi = v0 + u * kScanThreads;
            if (vi >= nvec) break;
            const int e = b0 + 4 * vi;
            // 4 mask bits starting at bit e (may straddle a word)
            const uint32_t w0 = bm[e >> 5], w1 = bm[(e >> 5) + 1];
            const uint32_t bits = __funnelshift_r(w0, w1, e & 31) & 0xFu;
            float a = (bits & 1u) ? -CUDART_INF_F : x[u].x;
            float b = (bits & 2u) ? -CUDART_INF_F : x[u].y;
            float c = (bits & 4u) ? -CUDART_INF_F : x[u].z;
            float d = (bits & 8u) ? -CUDART_INF_F : x[u].w;
            const float m4 = fmaxf(fmaxf(a, b), fmaxf(c, d));
            if (m4 > st.m) { st.s *= exp_diff(st.m, m4); st.m = m4; }
            if (st.m != -CUDART_INF_F) {
                // NaN inputs: fmaxf drops them from m4 but __expf(NaN - m) keeps the sum NaN
                st.s += (__expf(a - st.m) + __expf(b - st.m)) + (__expf(c - st.m) + __expf(d - st.m));
            } else if (a != a || b != b || c != c || d != d) {
                st.s = CUDART_NAN_F;
            }
            if (m4 > st.v[KCAP - 1]) {
                if (a > st.v[KCAP - 1]) topk_insert<KCAP>(st.v, st.id, a, e);
                if (b > st.v[KCAP - 1]) topk_insert<KCAP>(st.v, st.id, b, e + 1);
                if (c > st.v[KCAP - 1]) topk_insert<KCAP>(st.v, st.id, c, e + 2);
                if (d > st.v[KCAP - 1]) topk_insert<KCAP>(st.v, st.id, d, e + 3);
            }
        }
    }
    // tail
    const int t0 = b0 + 4 * nvec;
    if (tid < hi - t0) one(rp[t0 + tid], t0 + tid);
}

}  // namespace wrs

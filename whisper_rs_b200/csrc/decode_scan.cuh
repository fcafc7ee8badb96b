// K2 building blocks: register-resident, single-read scan of one chunk of a logits row.
//
// Replaces, per (segment, beam) row, the ATen calls of reference src/decode/filter/*.rs
// (index_fill_, slice fill_, log_softmax, logsumexp, max) and src/decode/extractor/beam_search.rs:37,62-65
// (log_softmax, topk) with one read of the row: masks are predicates, never written back.
//
// A CTA of 512 threads owns one chunk (<= 14 333 ids) of one row and one partition (text ids below
// ts_begin, or timestamp ids).  Every thread issues its 7 x 16-byte loads up front (all of the chunk is
// in flight at once), masks in registers, and the CTA leaves {max, sum exp(x - max), top-K} in global
// memory.  Top-K is threshold based: the K-th largest of the 16 warp maxima bounds the K-th largest
// element from below, so only the handful of elements at or above it are gathered and sorted.
//
// Ordering everywhere is (value desc, token id asc) -- north_star's "ties broken by lowest index".
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace wrs {

constexpr int kScanThreads = 512;
constexpr int kScanWarps = kScanThreads / 32;
constexpr int kNV = 7;                                   // float4 slots per thread
constexpr int kChunkCap = kScanThreads * 4 * kNV - 3;    // ids per chunk (3 = worst-case alignment slack)
constexpr int kMaxK = 16;                                // beam + 1 <= 16
constexpr int kMaxChunks = 9;                            // chunks per row (text + timestamp)
constexpr int kCandCap = 256;                            // gathered candidates per CTA before the tie fallback
constexpr float kLog2e = 1.4426950408889634f;

// what a CTA leaves in global memory for one (row, chunk)
struct __align__(16) Partial {
    float m, s;            // max and sum of exp(x - m) over the chunk's surviving ids (m = -inf, s = 0 if none)
    int n, pad;            // valid entries of v/id
    float v[kMaxK];        // the chunk's best values, sorted (value desc, id asc), padded with -inf
    int id[kMaxK];
};

__device__ __forceinline__ bool better(float va, int ia, float vb, int ib) {   // (va,ia) precedes (vb,ib)
    return va > vb || (va == vb && ia < ib);
}

// order-preserving float <-> unsigned key (for redux.sync max)
__device__ __forceinline__ unsigned fkey(float f) {
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float fkey_inv(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// exp(a - b) with exp(-inf - -inf) = 0: a fully masked partition is normal
__device__ __forceinline__ float exp_diff(float a, float b) { return a == -CUDART_INF_F ? 0.0f : expf(a - b); }

// per-row description of which ids survive the filters (all CTA-uniform)
struct RowRules {
    int t_lo, t_hi;   // surviving text range      [t_lo, t_hi)  (ids below ts_begin)
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

// shared scratch of one scanning CTA
struct ScanSmem {
    float wm[kScanWarps], ws[kScanWarps];
    float cv[kCandCap];
    int ci[kCandCap];
    int ncand;
    float red_v[kScanWarps];
    int red_i[kScanWarps];
    float pick_v;
    int pick_i;
};

// Scan ids [lo, hi) of the row at `rp` (element 0 = id 0) with suppress bitmap `bm` (bit set = masked) and
// leave the chunk's statistics in *out.  All 512 threads of the CTA take part; hi - lo <= kChunkCap.
__device__ __forceinline__ void scan_chunk(const float* __restrict__ rp, int lo, int hi, const uint32_t* __restrict__ bm,
                                           int K, ScanSmem& sm, Partial* __restrict__ out) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (hi <= lo) {                                      // nothing survives here (CTA-uniform)
        if (tid < kMaxK) { out->v[tid] = -CUDART_INF_F; out->id[tid] = INT32_MAX; }
        if (tid == 0) { out->m = -CUDART_INF_F; out->s = 0.0f; out->n = 0; }
        return;
    }
    // ---- 1. loads: element base of slot j is a0 + 4 (tid + 512 j), a0 = lo rounded down to 16-byte alignment
    const int mis = (int)((reinterpret_cast<uintptr_t>(rp + lo) >> 2) & 3);
    const int a0 = lo - mis;
    float x[kNV][4];
#pragma unroll
    for (int j = 0; j < kNV; ++j) {
        const int e = a0 + 4 * (tid + kScanThreads * j);
        if (e >= lo && e + 3 < hi) {
            const float4 q = __ldcs(reinterpret_cast<const float4*>(rp + e));
            x[j][0] = q.x; x[j][1] = q.y; x[j][2] = q.z; x[j][3] = q.w;
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) x[j][i] = (e + i >= lo && e + i < hi) ? __ldcs(rp + e + i) : -CUDART_INF_F;
        }
    }
    if (tid == 0) sm.ncand = 0;
    // ---- 2. suppress bitmap -> -inf; thread max
    float tmax = -CUDART_INF_F;
#pragma unroll
    for (int j = 0; j < kNV; ++j) {
        const int e = a0 + 4 * (tid + kScanThreads * j);
        if (e + 3 >= lo && e < hi) {
            const int w = e >> 5;                        // e >= -3; e < 0 only happens with lo = 0: those lanes are -inf already
            const uint32_t w0 = e >= 0 ? __ldg(bm + w) : 0u, w1 = e >= 0 ? __ldg(bm + w + 1) : __ldg(bm);
            const uint32_t bits = e >= 0 ? (__funnelshift_r(w0, w1, e & 31) & 0xFu) : ((w1 << (-e)) & 0xFu);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (bits & (1u << i)) x[j][i] = -CUDART_INF_F;
                tmax = fmaxf(tmax, x[j][i]);
            }
        }
    }
    // ---- 3. warp max (redux), sum of exp(x - warp max), warp sum
    const float wmax = fkey_inv(__reduce_max_sync(0xffffffffu, fkey(tmax)));
    const float ref = (wmax == -CUDART_INF_F) ? 0.0f : wmax;
    const float nref = -ref * kLog2e;
    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
    for (int j = 0; j < kNV; ++j) {
        s0 += ex2(fmaf(x[j][0], kLog2e, nref)) + ex2(fmaf(x[j][1], kLog2e, nref));   // NaN inputs keep the sum NaN
        s1 += ex2(fmaf(x[j][2], kLog2e, nref)) + ex2(fmaf(x[j][3], kLog2e, nref));
    }
    float ssum = s0 + s1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
    if (lane == 0) { sm.wm[warp] = wmax; sm.ws[warp] = ssum; }
    __syncthreads();
    // ---- 4. threshold = K-th largest warp maximum (>= K elements are at or above it)
    float tau;
    {
        const float mine = sm.wm[lane & (kScanWarps - 1)];
        int rank = 0;
#pragma unroll
        for (int w = 0; w < kScanWarps; ++w) {
            const float o = __shfl_sync(0xffffffffu, mine, w);
            rank += (o > mine || (o == mine && w < (lane & (kScanWarps - 1)))) ? 1 : 0;
        }
        const unsigned hit = __ballot_sync(0xffffffffu, lane < kScanWarps && rank == K - 1);
        tau = hit ? __shfl_sync(0xffffffffu, mine, __ffs(hit) - 1) : -CUDART_INF_F;
    }
    // ---- 5. gather everything at or above the threshold
    if (tmax >= tau && tmax > -CUDART_INF_F) {
#pragma unroll
        for (int j = 0; j < kNV; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (x[j][i] >= tau && x[j][i] > -CUDART_INF_F) {
                    const int pos = atomicAdd(&sm.ncand, 1);
                    if (pos < kCandCap) { sm.cv[pos] = x[j][i]; sm.ci[pos] = a0 + 4 * (tid + kScanThreads * j) + i; }
                }
    }
    __syncthreads();
    const int nc = sm.ncand;
    if (nc <= kCandCap) {
        // ---- 6a. warp 0: rank-count sort of the candidates, CTA (max, sum)
        if (warp == 0) {
            if (lane < kMaxK) { out->v[lane] = -CUDART_INF_F; out->id[lane] = INT32_MAX; }
            __syncwarp();
            for (int c = lane; c < nc; c += 32) {
                const float v = sm.cv[c]; const int id = sm.ci[c];
                int rank = 0;
                for (int d = 0; d < nc; ++d) rank += better(sm.cv[d], sm.ci[d], v, id) ? 1 : 0;
                if (rank < K) { out->v[rank] = v; out->id[rank] = id; }
            }
        }
    } else {
        // ---- 6b. massive ties: K rounds of exact CTA-wide selection, each strictly after the previous pick
        float pv = CUDART_INF_F; int pi = -1;
        for (int r = 0; r < K; ++r) {
            float bv = -CUDART_INF_F; int bi = INT32_MAX;
#pragma unroll
            for (int j = 0; j < kNV; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float v = x[j][i]; const int id = a0 + 4 * (tid + kScanThreads * j) + i;
                    const bool after = v < pv || (v == pv && id > pi);
                    if (after && v > -CUDART_INF_F && better(v, id, bv, bi)) { bv = v; bi = id; }
                }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float v2 = __shfl_xor_sync(0xffffffffu, bv, o); const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
                if (better(v2, i2, bv, bi)) { bv = v2; bi = i2; }
            }
            if (lane == 0) { sm.red_v[warp] = bv; sm.red_i[warp] = bi; }
            __syncthreads();
            if (warp == 0) {
                bv = lane < kScanWarps ? sm.red_v[lane] : -CUDART_INF_F; bi = lane < kScanWarps ? sm.red_i[lane] : INT32_MAX;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float v2 = __shfl_xor_sync(0xffffffffu, bv, o); const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (better(v2, i2, bv, bi)) { bv = v2; bi = i2; }
                }
                if (lane == 0) { sm.pick_v = bv; sm.pick_i = bi; out->v[r] = bv; out->id[r] = bi; }
            }
            __syncthreads();
            pv = sm.pick_v; pi = sm.pick_i;
        }
        if (warp == 0 && lane >= K && lane < kMaxK) { out->v[lane] = -CUDART_INF_F; out->id[lane] = INT32_MAX; }
    }
    if (warp == 0) {
        float m = lane < kScanWarps ? sm.wm[lane] : -CUDART_INF_F;
        const float wm_l = m;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        float s = (lane < kScanWarps && wm_l > -CUDART_INF_F) ? sm.ws[lane] * ex2((wm_l - m) * kLog2e) : 0.0f;
        if (lane < kScanWarps && sm.ws[lane] != sm.ws[lane]) s = sm.ws[lane];     // keep NaN visible
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) { out->m = m; out->s = s; out->n = min(nc, K); }
    }
}

}  // namespace wrs

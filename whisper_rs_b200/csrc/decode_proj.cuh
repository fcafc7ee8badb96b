// the logits projection with the decode step's row scan as its epilogue.
//
// Replaces, together with the tail launch of decode.cu, reference src/model.rs:521
// (`x.matmul(&self.token_emb.ws.transpose(0, 1))`, [R, n_state] x [n_state, V] in f32) followed by
// src/decode/mod.rs:257-270 (filters + extractor): the [R, V] logits are never written to HBM.
//
// Arithmetic: tcgen05.mma kind::f16 on f16 operands with f32 accumulators in tensor memory.  f32 operands are split
// v = hi + lo (both f16, lo = f16(v - hi)), and hi*hi + lo*hi + hi*lo is accumulated: 22 mantissa bits per operand, so a
// logit differs from the exact product sum by about 2^-22 of sum |x_i w_i| -- a smaller error than an f32 matmul's own
// rounding.  f16 has 5 exponent bits: the embedding is multiplied by a power of two at packing time that puts its largest
// |w| in [512, 1024) (the lo parts of all but vanishing weights stay normal numbers) and every accumulator value is
// multiplied back; x is taken as it comes (LayerNorm output: |x| must stay below 65504, beyond that the row is reported
// like any non-finite row).  The bars are written in tests/test_gpu_project.py and DESIGN.md.
//
// Geometry: a vocab tile is 176 consecutive ids of ONE partition (text ids below ts_split, timestamp ids from it: a
// tile never straddles, so a tile's statistics belong to one partition); V = 51865 with timestamps gives 287 + 9 = 296
// tiles = two per SM on 148 SMs.  Rows go 128 to an M tile (TMEM lanes).  A CTA (192 threads) takes vocab tiles
// b, b + grid, ... and for each of them every M tile:
//   warp 4 (one lane)  producer: cp.async.bulk of one K stage (32 of n_state) of A (x, 16 KB) and B (weights, 22 KB)
//                      into a 4-deep ring; both operands were packed beforehand into the tensor core's canonical
//                      K-major no-swizzle layout ([16-byte K chunk][row][8 bf16]: 8 x 16 B core matrices, rows 16 B
//                      apart, K chunks `rows x 16 B` apart), so a stage is one contiguous block -- no tensor map.
//   warp 5 (one lane)  issues 6 MMAs per stage (2 K steps of 16 x {hi hi, lo hi, hi lo}) into one of two 176-column
//                      accumulators; tcgen05.commit frees the stage / publishes the accumulator.
//   warps 0-3          epilogue, thread = row: tcgen05.ld 32 columns at a time, masks as predicates (suppress bitmap,
//                      timestamp rules as id ranges), online {max, sum exp}, exact top-K of the tile in registers
//                      (value desc, id asc), one record per (row, tile) to global memory for the tail launch.
#pragma once

#include <cuda_fp16.h>

#include "async_copy.cuh"
#include "decode_scan.cuh"
#include "decode_proj_select.cuh"

namespace wrs {



constexpr int kPK = 32;                                   // elements of n_state per pipeline stage (two MMA K steps)
constexpr int kPStages = 4;
constexpr int kPEpiWarps = 16;                            // warps 0-7 select (top-K), warps 8-15 sum (max, sum exp); of each
                                                          // eight, warps {4g .. 4g+3} drain accumulator g
constexpr int kPThreads = (kPEpiWarps + 2) * 32;
constexpr int kPABytes = kPM * kPK * 2;                   // one part (hi or lo) of a stage of x: 8 KB
constexpr int kPBBytes = kPN * kPK * 2;                   // ... of the weights: 11 KB
constexpr int kPStageBytes = 2 * kPABytes + 2 * kPBBytes; // [A hi | A lo | B hi | B lo] = 38 912 B
constexpr int kPAccStride = 256;                          // TMEM columns between the two accumulators
constexpr int kPTmemCols = 512;

struct ProjParams {
    const unsigned char* wp;     // packed weights [tile][stage][hi | lo][4][176][8] bf16
    const unsigned char* xp;     // packed x       [M tile][stage][hi | lo][4][128][8] bf16
    int n_tiles, n_text_tiles, n_mtiles, n_kstages, n_rows, n_vocab, ts_split;
    float inv_scale;             // 1 / (the power of two the packed weights were multiplied by): the epilogue works on the scaled
                                 // accumulators and multiplies what it writes (a power of two: order and ties are untouched)
    // EPI 0: the plain projection
    float* logits; int64_t row_stride;
    // EPI 1: fold
    const int64_t* tokens; int64_t tok_stride; int cur_len, sample_begin, K;
    RuleCfg rc; const uint32_t* bm; int bm_words;
    // records, one per (row, vocab tile), as planes of 4-byte words: parts[(row * (2 + 2K) + plane) * tiles_padded + tile];
    // planes: 0 max, 1 sum, 2 .. 2+K-1 the tile's best values (value desc, id asc), 2+K .. their ids
    float* parts; int tiles_padded;
};

// ---- operand packing.  One thread per 16-byte group (8 consecutive elements of n_state of one row).
__device__ __forceinline__ void split8(const float* v, float scale, uint4& hi, uint4& lo) {
    unsigned h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float a0 = v[2 * i] * scale, a1 = v[2 * i + 1] * scale;          // (scale is a power of two: exact)
        const __half h0 = __float2half_rn(a0), h1 = __float2half_rn(a1);
        const __half l0 = __float2half_rn(a0 - __half2float(h0)), l1 = __float2half_rn(a1 - __half2float(h1));
        h[i] = (unsigned)__half_as_ushort(h0) | ((unsigned)__half_as_ushort(h1) << 16);
        l[i] = (unsigned)__half_as_ushort(l0) | ((unsigned)__half_as_ushort(l1) << 16);
    }
    hi = make_uint4(h[0], h[1], h[2], h[3]); lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// largest |w| of the embedding (for the power-of-two scale of its packed copy)
__global__ void proj_absmax_kernel(const float* __restrict__ src, int64_t src_stride, int n_rows, int n_cols, unsigned* __restrict__ out_bits) {
    float m = 0.0f;
    const int64_t total = (int64_t)n_rows * n_cols;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = fabsf(src[(i / n_cols) * src_stride + (i % n_cols)]);
        m = (v == v && v < CUDART_INF_F) ? fmaxf(m, v) : m;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));      // (non-negative floats order like their bits)
}

// rows_per_tile = 176 (weights; src row = the tile's id) or 128 (x; src row = tile * 128 + r)
__global__ void proj_pack_kernel(const float* __restrict__ src, int64_t src_stride, int n_src_rows, int n_state, int n_tiles,
                                 int rows_per_tile, int n_text_tiles, int ts_split, int is_w, float scale, unsigned char* __restrict__ dst) {
    const int n_kstages = n_state / kPK;
    const int64_t groups = (int64_t)n_tiles * n_kstages * 4 * rows_per_tile;
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (int64_t)gridDim.x * blockDim.x) {
        // consecutive threads take consecutive K chunks of one row: a warp reads 8 rows x 128 contiguous bytes
        const int kc = (int)(g & 3);
        int64_t rest = g >> 2;
        const int s = (int)(rest % n_kstages); rest /= n_kstages;
        const int r = (int)(rest % rows_per_tile);
        const int tile = (int)(rest / rows_per_tile);
        int row, end, rows_here = rows_per_tile;
        if (is_w) { row = proj_tile_base(tile, n_text_tiles, ts_split) + r; end = proj_tile_end(tile, n_text_tiles, ts_split, n_src_rows); }
        else {
            // an M tile with fewer than 128 rows is packed compactly (its rows rounded up to 8): proj_rows8()
            row = tile * rows_per_tile + r; end = n_src_rows;
            rows_here = proj_rows8(n_src_rows, tile);
            if (r >= rows_here) continue;
        }
        float v[8];
        if (row < end) {
            const float4 a = *reinterpret_cast<const float4*>(src + (int64_t)row * src_stride + s * kPK + kc * 8);
            const float4 b = *reinterpret_cast<const float4*>(src + (int64_t)row * src_stride + s * kPK + kc * 8 + 4);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.0f;
        }
        uint4 hi, lo;
        split8(v, scale, hi, lo);
        const int64_t slot_bytes = (int64_t)rows_per_tile * kPK * 2 * 2;       // a (tile, stage) slot; a short tile uses its front
        const int64_t part_bytes = (int64_t)rows_here * kPK * 2;
        unsigned char* blk = dst + ((int64_t)tile * n_kstages + s) * slot_bytes;
        const int64_t off = ((int64_t)kc * rows_here + r) * 16;
        *reinterpret_cast<uint4*>(blk + off) = hi;
        *reinterpret_cast<uint4*>(blk + part_bytes + off) = lo;
    }
}

// mbarrier wait with a bound: a pipeline that stops (a bug, never a data condition) ends the launch with an error
// instead of hanging the device.  The hardware parks the warp (suspend-time hint) and, between tries, the warp sleeps:
// ten of a CTA's eighteen warps are waiting at any time, and a waiting warp that spins takes issue slots from the one
// epilogue warp of its scheduler that is working (measured: 5 cycles per instruction there, before).
template <int SLEEP_NS>
__device__ __forceinline__ void mbar_wait_b(unsigned long long* bar, unsigned parity) {
    for (unsigned spins = 0;; ++spins) {
        unsigned done;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(smem_addr(bar)), "r"(parity), "r"(0x989680u) : "memory");
        if (done) return;
        if (SLEEP_NS > 0) __nanosleep(SLEEP_NS);
        if (spins > (1u << 22)) __trap();
    }
}

// ---- tcgen05 plumbing (PTX as the CUDA 12.9 assembler takes it for sm_100a)
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// two-CTA clusters: rank, barrier across the pair, a bulk copy and a tcgen05.commit that reach both CTAs (the shared-memory
// offsets of destination and mbarrier are the issuing CTA's; the same offsets are used in every CTA of the mask)
__device__ __forceinline__ unsigned cluster_rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bulk_g2s_mc(void* dst, const void* src, unsigned bytes, unsigned long long* bar, unsigned short mask) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar)), "h"(mask)
                 : "memory");
}
__device__ __forceinline__ void tc_commit_mc(unsigned long long* bar, unsigned short mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_addr(bar)), "h"(mask) : "memory");
}

// shared-memory operand descriptor, K-major, no swizzle: start address, LBO = bytes between the two 16-byte K chunks of
// an MMA, SBO = bytes between 8-row groups; bit 46 = descriptor version 1 (Blackwell)
__device__ __forceinline__ unsigned long long tc_desc(unsigned saddr, unsigned lbo, unsigned sbo) {
    return (unsigned long long)((saddr & 0x3ffffu) >> 4) | ((unsigned long long)(lbo >> 4) << 16) |
           ((unsigned long long)(sbo >> 4) << 32) | (1ull << 46);
}
// instruction descriptor: D f32 (bits 4-5 = 1), A and B f16 (bits 7-9, 10-12 = 0), both K-major (bits 15, 16 = 0),
// N / 8 at bit 17, M / 16 at bit 24
constexpr unsigned kPIdesc = (1u << 4) | ((unsigned)(kPN >> 3) << 17) | ((unsigned)(kPM >> 4) << 24);
__device__ __forceinline__ void tc_mma(unsigned d_tmem, unsigned long long a, unsigned long long b, unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem), "l"(a), "l"(b), "r"(kPIdesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_ld32(unsigned taddr, float (&x)[32]) {
    unsigned r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, "
        "%19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        "tcgen05.wait::ld.sync.aligned;\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tc_ld16(unsigned taddr, float (&x)[32]) {
    unsigned r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        "tcgen05.wait::ld.sync.aligned;\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = __uint_as_float(r[i]);
#pragma unroll
    for (int i = 16; i < 32; ++i) x[i] = -CUDART_INF_F;
}

#ifdef WRS_EXP_PJ_TIMES
// investigation aid (tools/proj_variants.py): per CTA, globaltimer stamps of [unit][mma start, mma issued, epi start, epi end]
__device__ unsigned long long g_pj_times[148 * 32];
__device__ long long g_pj_clk[148 * 32];       // clock64 stamps inside proj_topk32 of unit 2, warp 0: [chunk][after max, after staging, after loop, iterations]
__device__ int g_pj_unit;
__device__ __forceinline__ unsigned long long pj_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define WRS_PJ_T(u, k) do { if ((u) < 8) g_pj_times[blockIdx.x * 32 + (u) * 4 + (k)] = pj_now(); } while (0)
#else
#define WRS_PJ_T(u, k) do {} while (0)
#endif

// 32 columns of the row's accumulator (ids id0 .. id0 + 31; `valid` bit i: id0 + i survives the filters): masked ids
// become -inf, the chunk maximum is returned (NaN operands are ignored by fmaxf: the callers look for them)
__device__ __forceinline__ float proj_mask32(float (&x)[32], unsigned valid) {
    if (valid != 0xffffffffu) {
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = (valid & (1u << i)) ? x[i] : -CUDART_INF_F;
    }
    float t[8];                                          // (a tree, not a chain: one warp per scheduler lives on latency)
#pragma unroll
    for (int i = 0; i < 8; ++i) t[i] = fmaxf(fmaxf(x[4 * i], x[4 * i + 1]), fmaxf(x[4 * i + 2], x[4 * i + 3]));
    return fmaxf(fmaxf(fmaxf(t[0], t[1]), fmaxf(t[2], t[3])), fmaxf(fmaxf(t[4], t[5]), fmaxf(t[6], t[7])));
}

// the statistics half of the epilogue: online {max, sum exp(x - max)} of the survivors
__device__ __forceinline__ void proj_stats32(float& m, float& s, float (&x)[32], unsigned valid, float c) {   // c = log2(e) / scale
    const float cm = proj_mask32(x, valid);
    if (cm == -CUDART_INF_F || cm != cm) {
        // nothing finite: the sum is unchanged unless a NaN is among the survivors (then it is NaN, as in ATen)
        bool bad = false;
#pragma unroll
        for (int i = 0; i < 32; ++i) bad |= x[i] != x[i];
        if (bad) s = CUDART_NAN_F;
        return;
    }
    if (cm > m) { s *= ex2((m - cm) * c); m = cm; }
#ifndef WRS_EXP_PJ_NOSUM
    float acc0 = 0.0f, acc1 = 0.0f, acc2 = 0.0f, acc3 = 0.0f;
    const float mb = -m * c;
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
        acc0 += ex2(fmaf(x[i], c, mb));
        acc1 += ex2(fmaf(x[i + 1], c, mb));
        acc2 += ex2(fmaf(x[i + 2], c, mb));
        acc3 += ex2(fmaf(x[i + 3], c, mb));
    }
    s += (acc0 + acc1) + (acc2 + acc3);
#endif
}

// the selection half: the chunk's survivors that beat the row's K-th best so far enter its list.  xs: the thread's
// column of its warp's [32][32] staging tile in shared memory (element i at xs[32 i]): the few candidates are picked out
// of it by index -- one copy of the insertion code, run about K ln(176 / K) times per fresh tile, instead of one
// predicated copy per element.
template <int KM>
__device__ __forceinline__ void proj_topk32(ProjRow<KM>& st, float (&x)[32], unsigned valid, int id0, float* xs, int pj_trace_unit = -1) {
#ifdef WRS_EXP_PJ_TIMES
    const bool tr = (threadIdx.x == 0) && pj_trace_unit >= 0;
    long long* tc = g_pj_clk + blockIdx.x * 32 + 4 * ((id0 / 32) % 6);
    const long long c0 = clock64();
    int iters = 0;
#endif
    const float cm = proj_mask32(x, valid);
#ifdef WRS_EXP_PJ_TIMES
    if (tr) tc[0] = clock64() - c0;
#endif
#ifndef WRS_EXP_PJ_NOINS
    if (cm > st.kth) {
        unsigned c4[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            c4[i & 3] |= (x[i] > st.kth) ? (1u << i) : 0u;
            xs[32 * i] = x[i];
        }
        unsigned cand = (c4[0] | c4[1]) | (c4[2] | c4[3]);
#ifdef WRS_EXP_PJ_TIMES
        if (tr) tc[1] = clock64() - c0;
#endif
        // one candidate per round, the next one's value already on its way from shared memory
        int i = __ffs((int)cand) - 1;
        float v = xs[32 * i];
        cand &= cand - 1u;
        for (;;) {
            const int i_now = i; const float v_now = v;
            const bool more = cand != 0u;
            if (more) { i = __ffs((int)cand) - 1; v = xs[32 * i]; cand &= cand - 1u; }
            proj_insert<KM>(st, v_now, id0 + i_now);
#ifdef WRS_EXP_PJ_TIMES
            ++iters;
#endif
            if (!more) break;
        }
#ifdef WRS_EXP_PJ_TIMES
        if (tr) { tc[2] = clock64() - c0; tc[3] = iters; }
#endif
    }
#endif
}

struct ProjSmem {
    unsigned char stage[kPStages][kPStageBytes];
    unsigned long long full[kPStages], empty[kPStages], tfull[2], tempty[2];
    unsigned tmem_base;
    unsigned pad[3];
    float xstage[kPEpiWarps / 2][32 * 32];
    float tile_floor[kPEpiWarps / 2][32];                // sum warp -> select warp of the pair: a floor for the row's list in this tile                            // epilogue warps: [element][lane] staging of one chunk
    uint32_t bm[1];                                      // the suppress bitmap, bm_words + 2 words (EPI 1)
};

// EPI 0: store the logits; EPI 2, 4, 6, 8, 12, 16: fold, with room for K <= EPI candidates per (row, tile)
// CL: launched as clusters of two CTAs, for batches of exactly two M tiles (129 .. 256 rows): CTA r of a pair takes M tile r
// of the pair's vocab tiles (four each for V = 51865), and the pair shares the weights -- each CTA fetches one half of a
// stage's weights (rank 0 the hi parts, rank 1 the lo parts) and multicasts it into both CTAs' stage, so a weight leaves
// L2 once per pair instead of once per M tile.  A stage is free when BOTH CTAs' MMAs have read it (their commits reach
// both CTAs' barriers).  What an SM takes IN per stage is unchanged, and that, with the MMAs, is what paces the main loop:
// the gain is the L2 traffic and that a CTA meets the same rows in all its tiles (floors carry over three times).
template <int EPI, bool CL>
__global__ void __launch_bounds__(kPThreads, 1) proj_kernel(const ProjParams p) {
    constexpr int KM = EPI == 0 ? 1 : EPI;
    extern __shared__ __align__(1024) unsigned char proj_smem_raw[];
    ProjSmem& sm = *reinterpret_cast<ProjSmem*>(proj_smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kPStages; ++s) { mbar_init(&sm.full[s], 1); mbar_init(&sm.empty[s], CL ? 2 : 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&sm.tfull[b], 1); mbar_init(&sm.tempty[b], 8); }
    }
    if (EPI != 0)
        for (int i = threadIdx.x; i < p.bm_words + 2; i += kPThreads) sm.bm[i] = i < p.bm_words ? p.bm[i] : 0u;
    __syncwarp();
    if (warp == kPEpiWarps + 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&sm.tmem_base)), "r"(kPTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (CL) cluster_sync_all();                          // the peer's barriers exist before anything of ours reaches them
    tc_fence_after();
    const unsigned tmem = sm.tmem_base;
    // the vocab tiles of this CTA (of this pair): first, first + stride, ...
    const int crank = CL ? (int)cluster_rank() : 0;
    const int first = CL ? (int)blockIdx.x >> 1 : (int)blockIdx.x, stride = CL ? (int)gridDim.x >> 1 : (int)gridDim.x;
    const int my_tiles = (first < p.n_tiles) ? (p.n_tiles - 1 - first) / stride + 1 : 0;
    const int mt_per_tile = CL ? 1 : p.n_mtiles;         // M tiles this CTA takes of each of its vocab tiles
    const int n_units = my_tiles * mt_per_tile;          // accumulations of this CTA: unit u = (tile u / mt_per_tile, M tile)

    if (warp == kPEpiWarps) {
        // ================================================= producer
        if (lane == 0) {
            unsigned it = 0;
            for (int u = 0; u < n_units; ++u) {
                const int tile = first + (u / mt_per_tile) * stride, mt = CL ? crank : u % p.n_mtiles;
                const unsigned char* a = p.xp + (int64_t)mt * p.n_kstages * (2 * kPABytes);
                const unsigned char* b = p.wp + (int64_t)tile * p.n_kstages * (2 * kPBBytes);
                const unsigned a_bytes = (unsigned)proj_rows8(p.n_rows, mt) * (kPK * 2 * 2);      // hi and lo of the tile's rows
                for (int ks = 0; ks < p.n_kstages; ++ks, ++it) {
                    const unsigned slot = it % kPStages, round = it / kPStages;
                    mbar_wait_b<0>(&sm.empty[slot], (round & 1u) ^ 1u);      // (a fresh barrier passes the first round)
                    mbar_expect_tx(&sm.full[slot], a_bytes + 2u * kPBBytes);
                    bulk_g2s(sm.stage[slot], a + (int64_t)ks * (2 * kPABytes), a_bytes, &sm.full[slot]);
                    if (CL) bulk_g2s_mc(sm.stage[slot] + 2 * kPABytes + crank * kPBBytes, b + (int64_t)ks * (2 * kPBBytes) + crank * kPBBytes, kPBBytes, &sm.full[slot], 3);
                    else bulk_g2s(sm.stage[slot] + 2 * kPABytes, b + (int64_t)ks * (2 * kPBBytes), 2 * kPBBytes, &sm.full[slot]);
                }
            }
        }
    } else if (warp == kPEpiWarps + 1) {
        // ================================================= MMA issuer
        if (lane == 0) {
            unsigned it = 0;
            for (int u = 0; u < n_units; ++u) {
                const unsigned buf = (unsigned)u & 1u, use = (unsigned)u >> 1;
                mbar_wait_b<100>(&sm.tempty[buf], (use & 1u) ^ 1u);            // the epilogue has drained this accumulator
                tc_fence_after();
                WRS_PJ_T(u, 0);
                const unsigned d = tmem + buf * kPAccStride;
                const unsigned a_rows = (unsigned)proj_rows8(p.n_rows, CL ? crank : u % p.n_mtiles);
                for (int ks = 0; ks < p.n_kstages; ++ks, ++it) {
                    const unsigned slot = it % kPStages, round = it / kPStages;
                    mbar_wait_b<0>(&sm.full[slot], round & 1u);
                    tc_fence_after();
                    const unsigned a_hi = smem_addr(sm.stage[slot]), a_lo = a_hi + a_rows * (kPK * 2);
                    const unsigned b_hi = a_hi + 2 * kPABytes, b_lo = b_hi + kPBBytes;
#pragma unroll
                    for (int j = 0; j < kPK / 16; ++j) {
                        const unsigned ao = j * 2 * (a_rows * 16), bo = j * 2 * (kPN * 16);
                        const unsigned long long dah = tc_desc(a_hi + ao, a_rows * 16, 128), dal = tc_desc(a_lo + ao, a_rows * 16, 128);
                        const unsigned long long dbh = tc_desc(b_hi + bo, kPN * 16, 128), dbl = tc_desc(b_lo + bo, kPN * 16, 128);
#if defined(WRS_EXP_PJ_NOMMA)
                        (void)dal; (void)dbl; (void)dah; (void)dbh; (void)d;
#elif defined(WRS_EXP_PJ_ONEPROD)
                        (void)dal; (void)dbl;
                        tc_mma(d, dah, dbh, (ks | j) != 0);
#else
                        tc_mma(d, dal, dbh, (ks | j) != 0);             // the small products first
                        tc_mma(d, dah, dbl, 1u);
                        tc_mma(d, dah, dbh, 1u);
#endif
                    }
                    if (CL) tc_commit_mc(&sm.empty[slot], 3);           // (both CTAs' producers wait for both CTAs' MMAs)
                    else tc_commit(&sm.empty[slot]);                    // the stage is free once these MMAs have read it
                }
                tc_commit(&sm.tfull[buf]);                              // the accumulator is complete
                WRS_PJ_T(u, 1);
            }
        }
    } else {
        // ================================================= epilogue: thread = row of the M tile.  Each accumulator is read by
        // two warps per TMEM lane quarter (a quarter belongs to warp % 4): a SELECT warp (0-7) keeps the row's best K of
        // the tile, a SUM warp (8-15) its {max, sum exp}; of each eight, warps 0-3 take the even units (accumulator 0),
        // warps 4-7 the odd ones.  The two never talk: each writes its own planes of the row's record.
        // A select thread that meets the same row again in a later tile of the same partition (always, with one or two M
        // tiles) carries a floor over: the K-th best value it has seen there bounds the row's K-th best from below, so only
        // values at or above it (ties can still win on the id) need to enter the new tile's list.
        const bool sums = warp >= kPEpiWarps / 2;
        const int ew = warp & (kPEpiWarps / 2 - 1);
        const unsigned lane_base = (unsigned)((ew & 3) * 32) << 16;
        float* const xmine = sm.xstage[ew] + lane;
        int prev_row = -1, prev_part = -1;
        float tau_run = -CUDART_INF_F;
        for (int u = ew >> 2; u < n_units; u += 2) {
            const int tile = first + (u / mt_per_tile) * stride, mt = CL ? crank : u % p.n_mtiles;
            const unsigned buf = (unsigned)u & 1u, use = (unsigned)u >> 1;
            const int row = mt * kPM + (int)(threadIdx.x & 127);
            const int base = proj_tile_base(tile, p.n_text_tiles, p.ts_split);
            const int end = proj_tile_end(tile, p.n_text_tiles, p.ts_split, p.n_vocab);
            ProjRow<KM> st;
            float sm_m = -CUDART_INF_F, sm_s = 0.0f;
            RowRules rr;
            if (EPI != 0) {
                const int part = tile < p.n_text_tiles ? 0 : 1;
                if (row != prev_row || part != prev_part) tau_run = -CUDART_INF_F;
                prev_row = row; prev_part = part;
                st.kth = -CUDART_INF_F; st.floor = -CUDART_INF_F;       // (set below, once the tile's floor is known)
#pragma unroll
                for (int e = 0; e < KM; ++e) { st.v[e] = e < KM - p.K ? CUDART_INF_F : -CUDART_INF_F; st.id[e] = INT32_MAX; }
                rr.t_lo = rr.t_hi = rr.s_lo = rr.s_hi = 0; rr.flags = 0u;
                if (row < p.n_rows) rr = row_rules(p.rc, p.tokens + (int64_t)row * p.tok_stride, p.cur_len, p.sample_begin);
            }
            mbar_wait_b<200>(&sm.tfull[buf], use & 1u);
            tc_fence_after();
            if (warp == (int)(buf * 4) && lane == 0) WRS_PJ_T(u, 2);
            const unsigned t0 = tmem + buf * kPAccStride + lane_base;
            if (EPI != 0) {
                // A floor for the tile before any list is built: the sum warp runs over the accumulator once for the maxima
                // of its twelve 16-column groups (masks applied); the K-th largest of them is reached by at least K ids of
                // the tile, so nothing below it can be among the row's best K.  It hands the value to the select warp of its
                // pair (named barrier, 64 threads), which would otherwise insert about K (1 + ln(176 / K)) times into a
                // fresh list; with the floor it is about a third of that.
                if (sums) {
                    float h[12];
#pragma unroll
                    for (int g = 0; g < 12; ++g) h[g] = -CUDART_INF_F;
                    if (p.K <= 12) {
#pragma unroll
                        for (int c = 0; c < (kPN + 31) / 32; ++c) {
                            float x[32];
                            if (32 * c + 32 <= kPN) tc_ld32(t0 + 32 * c, x);
                            else tc_ld16(t0 + 32 * c, x);
                            const int id0 = base + 32 * c;
                            if (row < p.n_rows && id0 < end) {
                                const unsigned w0 = sm.bm[id0 >> 5], w1 = sm.bm[(id0 >> 5) + 1];
                                const unsigned supp = __funnelshift_r(w0, w1, id0 & 31);
                                const unsigned ok = (range_bits(rr.t_lo - id0, min(rr.t_hi, end) - id0) | range_bits(rr.s_lo - id0, min(rr.s_hi, end) - id0)) & ~supp;
                                if (ok != 0xffffffffu) {
#pragma unroll
                                    for (int i = 0; i < 32; ++i) x[i] = (ok & (1u << i)) ? x[i] : -CUDART_INF_F;
                                }
                                float a[8];
#pragma unroll
                                for (int i = 0; i < 8; ++i) a[i] = fmaxf(fmaxf(x[4 * i], x[4 * i + 1]), fmaxf(x[4 * i + 2], x[4 * i + 3]));
                                h[2 * c] = fmaxf(fmaxf(a[0], a[1]), fmaxf(a[2], a[3]));
                                h[2 * c + 1] = fmaxf(fmaxf(a[4], a[5]), fmaxf(a[6], a[7]));
                            }
                        }
                    }
                    sm.tile_floor[ew][lane] = p.K <= 12 ? kth_of_12(h, p.K) : -CUDART_INF_F;
                    asm volatile("bar.arrive %0, 64;" ::"r"(1 + ew) : "memory");
                } else {
                    asm volatile("bar.sync %0, 64;" ::"r"(1 + ew) : "memory");
                    float fl = fmaxf(tau_run, sm.tile_floor[ew][lane]);
                    // the largest float below it: "x > floor" is "x >= the bound"
                    if (fl > -CUDART_INF_F) fl = fl == 0.0f ? -1.4e-45f : __int_as_float(__float_as_int(fl) + (fl > 0.0f ? -1 : 1));
                    st.kth = fl; st.floor = fl;
                }
            }
#pragma unroll 1
            for (int c = 0; c < (kPN + 31) / 32; ++c) {
                if (EPI == 0 && ((c & 1) != 0) != sums) continue;      // (plain projection: the two warps share the chunks)
                float x[32];
                if (32 * c + 32 <= kPN) tc_ld32(t0 + 32 * c, x);
                else tc_ld16(t0 + 32 * c, x);
                const int id0 = base + 32 * c;
#ifdef WRS_EXP_PJ_TIMES
                if (warp == 0 && lane == 0 && u == 0) g_pj_times[blockIdx.x * 32 + 16 + 2 * c] = pj_now();
#endif
#ifdef WRS_EXP_PJ_NOEPI
                if (x[0] == 123.456f && p.logits) p.logits[0] = 0.0f;
                continue;
#endif
                if (EPI == 0) {
                    if (row < p.n_rows) {
                        float* o = p.logits + (int64_t)row * p.row_stride + id0;
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (id0 + i < end) o[i] = x[i] * p.inv_scale;
                    }
                } else if (row < p.n_rows && id0 < end) {
                    // suppress bits of ids id0 .. id0 + 31 (two words of the bitmap), the row's ranges, the tile's end
                    const unsigned w0 = sm.bm[id0 >> 5], w1 = sm.bm[(id0 >> 5) + 1];
                    const unsigned supp = __funnelshift_r(w0, w1, id0 & 31);
                    const unsigned ok = (range_bits(rr.t_lo - id0, min(rr.t_hi, end) - id0) | range_bits(rr.s_lo - id0, min(rr.s_hi, end) - id0)) & ~supp;
                    if (sums) proj_stats32(sm_m, sm_s, x, ok, kLog2e * p.inv_scale);
                    else proj_topk32<KM>(st, x, ok, id0, xmine, u == 2 ? 2 : -1);
                }
#ifdef WRS_EXP_PJ_TIMES
                if (warp == 0 && lane == 0 && u == 0) g_pj_times[blockIdx.x * 32 + 17 + 2 * c] = pj_now();
#endif
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.tempty[buf]);
            if (warp == (int)(buf * 4) && lane == 0) WRS_PJ_T(u, 3);
            if (EPI != 0 && row < p.n_rows) {
                float* rec = p.parts + (int64_t)row * (2 + 2 * p.K) * p.tiles_padded + tile;
                if (sums) {
                    rec[0] = sm_m * p.inv_scale; rec[p.tiles_padded] = sm_s;
                } else {
                    tau_run = fmaxf(tau_run, st.v[KM - 1]);
                    float* rv = rec + (int64_t)(2 - (KM - p.K)) * p.tiles_padded;      // plane of list entry e: 2 + e - (KM - K)
                    int* ri = reinterpret_cast<int*>(rv + (int64_t)p.K * p.tiles_padded);
#pragma unroll
                    for (int e = 0; e < KM; ++e)
                        if (e >= KM - p.K) { rv[(int64_t)e * p.tiles_padded] = st.v[e] * p.inv_scale; ri[(int64_t)e * p.tiles_padded] = st.id[e]; }
                }
            }
        }
    }
    // ---- teardown: every tcgen05 operation of this CTA has completed (the epilogue waited for the last commit)
    tc_fence_before();
    __syncthreads();
    if (CL) cluster_sync_all();                          // nothing of the peer's is still on its way into this CTA
    if (warp == kPEpiWarps + 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kPTmemCols) : "memory");
    }
}

}  // namespace wrs

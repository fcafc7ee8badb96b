// K1: fused log-mel front-end for batches of audio segments (sm_100a).
//
// Replaces the ATen op sequence of reference src/audio.rs:230-259 (hann_window, stft, abs, pow, matmul,
// clamp_min, log10, max, maximum, add, div) with one persistent, warp-specialised kernel.  A CTA (16 warps, one per
// SM) takes tiles of 64 frames; every lane works on TWO frames held side by side in 64-bit register pairs, so the
// FFT and the mel projection run on Blackwell's packed FP32 instructions (FADD2 / FMUL2 / FFMA2).
//
//   warps 0-7  "A"  : real 25-point DFTs (FMA-pipe bound).  A tile is two passes of 32 frames (16 frame pairs
//                     (f, f+16)); pass h reads PCM half-tile buffer h (34 rows of 160 samples, pitch 164, filled by
//                     bulk async copies that the A warps issue themselves two passes ahead), every lane takes the
//                     taps 16m + t of its frame pair at compile-time offsets, multiplies by the Hann window
//                     (src/audio.rs:233), runs the 25-point DFT and the twiddles W400^{tq}, and stores
//                     Y[t][q][frame pair] (float4: re2, im2).
//   warps 8-15 "BC" : complex 16-point DFTs over t, |X|^2 -> P[bin][frame pair]; then the banded mel projection
//                     (src/audio.rs:252), clamp + log10 (:254), raw rows -> HBM, slice maximum (:256); plus the
//                     deferred floor / scale (:256-258) of an earlier slice of this CTA (LSU / memory bound).
//
// The two roles meet only through mbarriers (Y full / Y empty, PCM full / PCM empty per buffer): A of tile k+1 runs
// while BC works on tile k, so the FMA-bound and the LSU-bound halves of the algorithm share the SM's issue slots
// instead of alternating behind CTA barriers.  BC warps keep one named barrier between their two stages.
//
//   slice end       : raw log-mel -> HBM, the slice's max key -> slice_max[segment][slice] (written once)
//   one slice later : once all slices of a segment have posted, its max is final and floor (max-8), (x+4)/4 are
//                     applied in place on the slice's own 20 KB (still in L2)
#include <algorithm>
#include <cmath>

#include "async_copy.cuh"
#include "logmel_fft.cuh"
#include "wrs_internal.cuh"

namespace wrs {

constexpr int kThreads = 512;
constexpr int kAWarps = 8;           // stage A (producers of Y)
constexpr int kBWarps = 8;           // stages B and C (consumers)
constexpr int kBThreads = kBWarps * 32;
constexpr int kTW = 64;              // frames per tile (two per lane in stages B and C)
constexpr int kHalf = 32;            // frames per stage-A pass
constexpr int kHRows = kHalf + 2;    // rows of 160 samples a pass reads: frame i uses rows i .. i+2
constexpr int kPitch = WRS_HOP_LENGTH + 4;   // words per tile row: 16-byte aligned rows, frame stride == 4 banks
constexpr float kLog10Of2 = 0.30102999566398120f;
constexpr int kSlotEmpty = (int)0x80808080;         // memset pattern; as a key it decodes to -3.4e38, never a log10 value

__constant__ float c_win[WRS_N_FFT];                 // periodic Hann, src/audio.rs:233
__constant__ float2 c_tw[16 * 13];                   // W400^{tq} = (cos, sin)(-2 pi t q / 400)

struct LogmelParams {
    const float* pcm;
    float* out;
    int64_t seg_stride;
    int n_segments;
    int n_samples;
    int n_frames;
    int slices_per_seg;
    int total_slices;
    int vec_in;
    int vec_out;
    const MelBank* bank;
    int* seg_max;   // two-launch mode: [n_segments] order-preserving int key of the per-segment max of log10(mel)
    int* slice_max; // fused mode: [n_segments][slices_per_seg] the same key per slice, kSlotEmpty until the slice has posted
    DevStatus* status;
};

constexpr int kPRows = WRS_N_BINS + 3;               // 201 bins + rows the zero-padded band tails may touch

struct LogmelSmem {
    float4 Y[16 * 13 * 32];                          // 106 496 B  [t][q][frame pair]
    float2 P[kPRows * 32];                           // 52 224 B   [bin][frame pair]
    float pcm[2][kHRows * kPitch];                   // 44 608 B   half-tile buffers
    float w[MelBank::kMaxPadded];                    // 3 072 B: in-band filter weights, bands zero-padded to even counts
    int4 desc[WRS_N_MELS];                           // 1 280 B: {first bin, padded count, weight offset, 0} per mel row
    float win[WRS_N_FFT];                            // 1 600 B
    float2 tw[16 * 13];                              // 1 664 B
    unsigned long long pcm_full[2], pcm_empty[2];    // mbarriers: half-tile buffer h filled / read by all A warps
    unsigned long long y_full, y_empty;              // mbarriers: Y written by both passes / read by all BC warps
    int row_begin[kBWarps + 1];                      // mel rows [row_begin[w], row_begin[w+1]) of BC warp w
    int red_max;                                     // max key of the slice in progress (reset by the last warp to add to it)
    int warps_done;
    int fin_ready;                                   // the oldest pending slice's segment is complete: seg_floor is valid
    float seg_floor;
};
static_assert(sizeof(LogmelSmem) <= 227 * 1024, "log-mel working set exceeds the 227 KB a CTA can own");

__device__ __forceinline__ int float_key(float f) {
    const int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }
// max that keeps a NaN (the reference's clamp_min / max / maximum all propagate it)
__device__ __forceinline__ float fmax_nan(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
// log2 of a normal, positive number (the argument has been clamped to >= 1e-10): one MUFU, no denormal fix-up
__device__ __forceinline__ float log2_pos(float a) {
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
}
__device__ __forceinline__ float floor_scale(float x, float fl) { return (fmax_nan(x, fl) + 4.0f) * 0.25f; }

// one sample of the (virtually reflect-padded) segment; i is relative to the segment start
__device__ __forceinline__ float load_reflect(const float* __restrict__ seg, int i, int n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    return (i >= 0 && i < n) ? __ldg(seg + i) : 0.0f;
}

__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) { mbar_expect_tx(bar, bytes); }
__device__ __forceinline__ void bar_bc() { asm volatile("bar.sync 1, %0;" ::"n"(kBThreads) : "memory"); }

// Half-tile buffer `buf` <- rows f0 .. f0+33 of segment `seg` cut into 160-sample rows that start 200 samples before
// the frame they are named after (row R = samples [160 R - 200, 160 R - 40) of the segment, so that frame f reads rows
// f, f+1, f+2 and tap n sits at a compile-time offset).  A warp w of the eight takes rows w, w+8, ...: rows that lie
// entirely inside the segment move as 640-byte bulk copies (16-byte aligned PCM), issued by one elected lane from
// uniform registers; rows in front of the segment (reflection, src/audio.rs:236 center = true) and at or behind its end
// are filled element-wise.  Every A warp arrives once on pcm_full[buf], announcing the bytes of its own copies.
__device__ __forceinline__ void stage_half(LogmelSmem& sm, const LogmelParams& p, int seg, int f0, int buf, int aw, int lane) {
    const float* __restrict__ segp = p.pcm + (int64_t)seg * p.seg_stride;
    float* tile = sm.pcm[buf];
    const int r_plain_hi = p.vec_in != 0 ? (p.n_samples + 40) / WRS_HOP_LENGTH : 0;      // rows R in [2, r_plain_hi] are plain
    int nplain = 0;
    for (int r = aw; r < kHRows; r += kAWarps) {
        const int R = f0 + r;
        if (R >= 2 && R <= r_plain_hi) { ++nplain; continue; }
        const int base = WRS_HOP_LENGTH * R - 200;
#pragma unroll
        for (int e = 0; e < WRS_HOP_LENGTH / 32; ++e) tile[r * kPitch + lane + 32 * e] = load_reflect(segp, base + lane + 32 * e, p.n_samples);
    }
    __syncwarp();
    if (elect_one()) {
        if (nplain != 0) {
            fence_proxy_async();                     // the buffer was last touched by ordinary loads / stores
            mbar_arrive_expect_tx(&sm.pcm_full[buf], (unsigned)(nplain * WRS_HOP_LENGTH * sizeof(float)));
            for (int r = aw; r < kHRows; r += kAWarps) {
                const int R = f0 + r;
                if (R >= 2 && R <= r_plain_hi)
                    bulk_g2s(tile + r * kPitch, segp + (WRS_HOP_LENGTH * R - 200), WRS_HOP_LENGTH * (unsigned)sizeof(float), &sm.pcm_full[buf]);
            }
        } else {
            mbar_arrive(&sm.pcm_full[buf]);
        }
    }
}

// Per-slice maxima are single words written once (relaxed, gpu scope) and read by other CTAs the same way: a word is
// either still kSlotEmpty or final, so no ordering between different words is ever needed -- no fences, no counters.
__device__ __forceinline__ int ld_relaxed(const int* p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(int* p, int v) { asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// One warp: have all slices of `seg` posted?  Folds the slots this lane owns beyond the two it may already hold
// (a = slot lane, b = slot lane + 32, fetched earlier); returns the segment's max key on every lane, `complete` likewise.
__device__ __forceinline__ int fold_slots(const LogmelParams& p, int seg, int a, int b, int lane, bool& complete) {
    const int* slots = p.slice_max + (int64_t)seg * p.slices_per_seg;
    bool miss = (lane < p.slices_per_seg && a == kSlotEmpty) || (lane + 32 < p.slices_per_seg && b == kSlotEmpty);
    int mx = max(lane < p.slices_per_seg ? a : INT32_MIN, lane + 32 < p.slices_per_seg ? b : INT32_MIN);
    for (int i = lane + 64; i < p.slices_per_seg; i += 32) {
        const int v = ld_relaxed(slots + i);
        miss = miss || v == kSlotEmpty;
        mx = max(mx, v);
    }
    complete = !__any_sync(0xffffffffu, miss);
    return __reduce_max_sync(0xffffffffu, mx);
}

// out = (max(raw, segmax - 8) + 4) / 4 (src/audio.rs:256-258) in place on one [80][nf] slice, by NT threads (t = 0..NT-1).
// Aligned rows: all of a thread's 16-byte loads go out before the first store.
template <int NT>
__device__ __forceinline__ void floor_scale_slice(float* __restrict__ outp, int64_t n_frames, int nf, float fl, int t, bool vec) {
    if (vec) {
        constexpr int kPer = (WRS_N_MELS * (kTW / 4) + NT - 1) / NT, kBatch = kPer < 5 ? kPer : 5;
        const int nf4 = nf >> 2;
#pragma unroll 1
        for (int k0 = 0; k0 < kPer; k0 += kBatch) {
            float4 v[kBatch];
#pragma unroll
            for (int k = 0; k < kBatch; ++k) {
                const int idx = t + (k0 + k) * NT;
                if (idx < WRS_N_MELS * (kTW / 4) && (idx & 15) < nf4)
                    v[k] = __ldcg(reinterpret_cast<const float4*>(outp + (int64_t)(idx >> 4) * n_frames + 4 * (idx & 15)));
            }
#pragma unroll
            for (int k = 0; k < kBatch; ++k) {
                const int idx = t + (k0 + k) * NT;
                if (idx < WRS_N_MELS * (kTW / 4) && (idx & 15) < nf4) {
                    float4 x = v[k];
                    x.x = floor_scale(x.x, fl); x.y = floor_scale(x.y, fl); x.z = floor_scale(x.z, fl); x.w = floor_scale(x.w, fl);
                    *reinterpret_cast<float4*>(outp + (int64_t)(idx >> 4) * n_frames + 4 * (idx & 15)) = x;
                }
            }
        }
    } else {
        const int d = max(nf, 1);
        for (int idx = t; idx < WRS_N_MELS * nf; idx += NT) {
            const int m = idx / d, c = idx - m * d;
            float* q = outp + (int64_t)m * n_frames + c;
            *q = floor_scale(__ldcg(q), fl);
        }
    }
}

// Blocking floor + scale of one slice in place, by the eight BC warps (bt = 0..255).  Only the drain at the end of the
// kernel (and a full pending queue) comes here; the steady state hides it inside the tile loop.
__device__ __forceinline__ void finalize_slice_blocking(LogmelSmem& sm, const LogmelParams& p, int seg, int fs, int bt) {
    const int nf = min(kTW, p.n_frames - fs);
    if (bt < 32) {
        // every slice of the segment is owned by a CTA that is running or has finished (persistent grid no larger than
        // what is co-resident; a CTA posts its slice before it ever waits), so this wait ends; the bound is a safety net
        const int* slots = p.slice_max + (int64_t)seg * p.slices_per_seg;
        int spins = 0, key;
        for (;;) {
            const int a = bt < p.slices_per_seg ? ld_relaxed(slots + bt) : 0;
            const int b = bt + 32 < p.slices_per_seg ? ld_relaxed(slots + bt + 32) : 0;
            bool complete;
            key = fold_slots(p, seg, a, b, bt, complete);
            if (complete) break;
            __nanosleep(64);
            if (++spins > (1 << 22)) { if (bt == 0) atomicExch(&p.status->timeout, 1); break; }
        }
        if (bt == 0) sm.seg_floor = key_float(key) - 8.0f;
    }
    bar_bc();
    floor_scale_slice<kBThreads>(p.out + ((int64_t)seg * WRS_N_MELS) * p.n_frames + fs, p.n_frames, nf, sm.seg_floor, bt, p.vec_out != 0);
    bar_bc();
}

constexpr int kPend = 4;              // slices a CTA may have waiting for their segment's maximum

// ---- stage A, one pass: lane (i8, tl) of A warp (g, tq) takes frames i and i+16 of the half-tile (i = 8g + i8) and the
// taps 16m + t (t = 4tq + tl).  8 frames x 4 consecutive taps per load = 32 distinct banks at pitch 164.
__device__ __forceinline__ void stage_a_pass(LogmelSmem& sm, const float* __restrict__ tile, float4* __restrict__ ydst,
                                             const float* __restrict__ winp, const float2* __restrict__ twp,
                                             unsigned long long* pcm_empty, unsigned long long* y_wait, unsigned y_parity, int lane) {
    f32x2 v[25];
#pragma unroll
    for (int m = 0; m < 25; ++m) {
        constexpr int kHi = 16 * kPitch;
        const int off = (m / 10) * kPitch + 16 * (m % 10);           // tap n = 16m + t: row n / 160 = m / 10 (t < 16)
        const float wv = winp[16 * m];
        v[m] = mul_(pack2(tile[off], tile[off + kHi]), pack2(wv, wv));
    }
    cx<f32x2> y[13];
    rdft25<true>(v, y);
    // every tap has been consumed: the buffer may be refilled once all eight A warps say so
    __syncwarp();
    if (lane == 0) mbar_arrive(pcm_empty);
    cx<f32x2> z[13];
    z[0] = y[0];
#pragma unroll
    for (int q = 1; q < 13; ++q) {
        const float2 w = twp[q];
        const cx<f32x2> wq{pack2(w.x, w.x), pack2(w.y, w.y)};
        z[q] = rdft25_conj_slot(q) ? cmul_conj(y[q], wq) : cmul(y[q], wq);
    }
    if (y_wait) mbar_wait(y_wait, y_parity);         // BC has read the previous tile's Y
#pragma unroll
    for (int q = 0; q < 13; ++q) {
        float a, b, c, d;
        unpack2(z[q].re, a, b); unpack2(z[q].im, c, d);
        ydst[q * 32] = make_float4(a, b, c, d);
    }
}

// ---- stage B, one item: q, lane = frame pair.  Complex 16-point DFT over t, |X|^2 -> P.
__device__ __forceinline__ void stage_b_load(const LogmelSmem& sm, int q, int lane, cx<f32x2> (&z)[16]) {
#pragma unroll
    for (int t = 0; t < 16; ++t) {
        const float4 a = sm.Y[(t * 13 + q) * 32 + lane];
        z[t].re = pack2(a.x, a.y); z[t].im = pack2(a.z, a.w);
    }
    cdft16(z);
}
__device__ __forceinline__ void stage_b_store(LogmelSmem& sm, int q, int lane, const cx<f32x2> (&z)[16]) {
    // bins q + 25 s for s < 8; s >= 8 gives the conjugates of bins 400 - q - 25 s (q = 0: bin 200 at s = 8, nothing beyond)
    float2* lo = sm.P + q * 32 + lane;
    float2* hi = sm.P + (400 - q) * 32 + lane;                             // hi[-25 s * 32] = bin 400 - q - 25 s
#pragma unroll
    for (int s = 0; s < 16; ++s) {
        float a, b;
        unpack2(power(z[s]), a, b);
        if (s < 8) lo[25 * s * 32] = make_float2(a, b);
        else if (q != 0) hi[-25 * s * 32] = make_float2(a, b);
        else if (s == 8) lo[25 * s * 32] = make_float2(a, b);
    }
}

template <bool FUSED>
__global__ void __launch_bounds__(kThreads, 1) logmel_kernel(const LogmelParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LogmelSmem& sm = *reinterpret_cast<LogmelSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // tables -> smem (once per CTA); the rows of P behind bin 200 stay zero (zero-padded band tails read them)
    for (int i = tid; i < WRS_N_MELS; i += kThreads) sm.desc[i] = make_int4(p.bank->lo[i], p.bank->n2[i], p.bank->off2[i], 0);
    for (int i = tid; i < MelBank::kMaxPadded; i += kThreads) sm.w[i] = p.bank->wp[i];
    for (int i = tid; i < WRS_N_FFT; i += kThreads) sm.win[i] = c_win[i];
    for (int i = tid; i < 16 * 13; i += kThreads) sm.tw[i] = c_tw[i];
    for (int i = tid; i < (kPRows - WRS_N_BINS) * 32; i += kThreads) sm.P[WRS_N_BINS * 32 + i] = make_float2(0.0f, 0.0f);
    if (tid <= kBWarps) sm.row_begin[tid] = p.bank->row_begin[tid];
    if (tid == 0) {
        mbar_init(&sm.pcm_full[0], kAWarps); mbar_init(&sm.pcm_full[1], kAWarps);
        mbar_init(&sm.pcm_empty[0], kAWarps); mbar_init(&sm.pcm_empty[1], kAWarps);
        mbar_init(&sm.y_full, 2 * kAWarps); mbar_init(&sm.y_empty, kBWarps);
        sm.red_max = float_key(-INFINITY); sm.warps_done = 0; sm.fin_ready = 0;
    }
    __syncthreads();                                     // the only CTA-wide barrier: the roles part here

    // tiles are dealt round-robin; (seg, sl) of the current tile advance without divisions
    const int step_seg = (int)gridDim.x / p.slices_per_seg, step_sl = (int)gridDim.x % p.slices_per_seg;
    int seg = (int)blockIdx.x / p.slices_per_seg, sl = (int)blockIdx.x % p.slices_per_seg;

    if (warp < kAWarps) {
        // =================================================== A warps
        const int i = 8 * (warp & 1) + (lane & 7), t = 4 * (warp >> 1) + (lane >> 3);
        const float* winp = sm.win + t;
        const float2* twp = sm.tw + t * 13;
        if ((int)blockIdx.x < p.total_slices) {
            stage_half(sm, p, seg, sl * kTW, 0, warp, lane);
            stage_half(sm, p, seg, sl * kTW + kHalf, 1, warp, lane);
        }
        unsigned k = 0;
        for (int slice = blockIdx.x; slice < p.total_slices; slice += gridDim.x, ++k) {
            int nseg = seg + step_seg, nsl = sl + step_sl;   // the tile after this one
            if (nsl >= p.slices_per_seg) { nsl -= p.slices_per_seg; ++nseg; }
            const bool has_next = slice + (int)gridDim.x < p.total_slices;
#pragma unroll 1
            for (int h = 0; h < 2; ++h) {
                mbar_wait(&sm.pcm_full[h], k & 1u);
                stage_a_pass(sm, sm.pcm[h] + i * kPitch + t, sm.Y + (t * 13) * 32 + 16 * h + i, winp, twp, &sm.pcm_empty[h],
                             (h == 0 && k > 0) ? &sm.y_empty : nullptr, (k - 1u) & 1u, lane);
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.y_full);
                if (has_next) {
                    mbar_wait(&sm.pcm_empty[h], k & 1u);     // all eight A warps have taken their taps of this pass
                    stage_half(sm, p, nseg, nsl * kTW + kHalf * h, h, warp, lane);
                }
            }
            seg = nseg; sl = nsl;
        }
        return;
    }

    // ======================================================= BC warps
    const int bw = warp - kAWarps, bt = tid - kAWarps * 32;
    // slices whose floor/scale is still to be applied, oldest first (fs < 0: empty)
    int pend_seg[kPend], pend_fs[kPend];
#pragma unroll
    for (int k = 0; k < kPend; ++k) { pend_seg[k] = 0; pend_fs[k] = -1; }
    const bool ovec = p.vec_out != 0;                    // output rows 16-byte aligned, n_frames % 4 == 0
    const int m_begin = sm.row_begin[bw], m_end = sm.row_begin[bw + 1];
    // lane = frame pair: frames fl and fl + 16 of the tile
    const int fl = 32 * (lane >> 4) + (lane & 15);
    unsigned k = 0;
    for (int slice = blockIdx.x; slice < p.total_slices; slice += gridDim.x, ++k) {
        const int fs = sl * kTW;                         // first frame of the tile
        const int nf = min(kTW, p.n_frames - fs);
        int nseg = seg + step_seg, nsl = sl + step_sl;
        if (nsl >= p.slices_per_seg) { nsl -= p.slices_per_seg; ++nseg; }
        // has the oldest pending slice's segment completed?  The answer is asked for here and looked at after stage B,
        // so that the round trip to L2 is off every warp's critical path.
        int poll_a = 0, poll_b = 0;
        if (FUSED && bw == kBWarps - 1 && pend_fs[0] >= 0) {
            const int* slots = p.slice_max + (int64_t)pend_seg[0] * p.slices_per_seg;
            if (lane < p.slices_per_seg) poll_a = ld_relaxed(slots + lane);
            if (lane + 32 < p.slices_per_seg) poll_b = ld_relaxed(slots + lane + 32);
        }
        mbar_wait(&sm.y_full, k & 1u);

        // ---- stage B: 13 items (q) over 8 warps: warps 0-4 take q and q + 8, warps 5-7 take one
        {
            cx<f32x2> z[16];
            stage_b_load(sm, bw, lane, z);
            if (bw >= 5) { __syncwarp(); if (lane == 0) mbar_arrive(&sm.y_empty); }
            bar_bc();                                    // every BC warp has finished stage C of the previous tile: P is free
            stage_b_store(sm, bw, lane, z);
            if (bw < 5) {
                stage_b_load(sm, bw + 8, lane, z);
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.y_empty);
                stage_b_store(sm, bw + 8, lane, z);
            }
        }
        if (FUSED && bw == kBWarps - 1) {
            bool ready = false;
            int key = 0;
            if (pend_fs[0] >= 0) key = fold_slots(p, pend_seg[0], poll_a, poll_b, lane, ready);
            if (lane == 0) {
                if (ready) sm.seg_floor = key_float(key) - 8.0f;
                sm.fin_ready = ready ? 1 : 0;
            }
        }
        bar_bc();                                        // P complete, fin_ready / seg_floor visible

        // ---- stage C: banded mel projection (src/audio.rs:252), clamp + log10 (:254), raw rows -> HBM
        const bool fin = FUSED && sm.fin_ready != 0;
        const int fseg = pend_seg[0], ffs = pend_fs[0];
        const int fnf = fin ? min(kTW, p.n_frames - ffs) : 0;
        float* __restrict__ foutp = p.out + ((int64_t)fseg * WRS_N_MELS) * p.n_frames + ffs;
        // the pending slice whose floor has just become known: its raw rows come back from L2 behind the mel arithmetic
        constexpr int kPer = (WRS_N_MELS * (kTW / 4) + kBThreads - 1) / kBThreads;   // float4 of a [80][64] slice per thread: 5
        const int fnf4 = fnf >> 2;
        float4 raw[kPer];
        if (fin && ovec) {
#pragma unroll
            for (int j = 0; j < kPer; ++j) {
                const int idx = bt + j * kBThreads;
                if ((idx & 15) < fnf4)
                    raw[j] = __ldcg(reinterpret_cast<const float4*>(foutp + (int64_t)(idx >> 4) * p.n_frames + 4 * (idx & 15)));
            }
        }
        float lmax = -INFINITY;
        {
            const bool ok0 = fl < nf, ok1 = fl + 16 < nf;
            float* __restrict__ orow = p.out + ((int64_t)seg * WRS_N_MELS + m_begin) * p.n_frames + fs + fl;
            int4 d = sm.desc[m_begin < WRS_N_MELS ? m_begin : 0];
#pragma unroll 1
            for (int m = m_begin; m < m_end; ++m) {
                const float2* pp = sm.P + d.x * 32 + lane;
                const float2* wv = reinterpret_cast<const float2*>(sm.w + d.z);
                const int n2 = d.y;
                d = sm.desc[min(m + 1, WRS_N_MELS - 1)];     // the next row's description is fetched behind this row's bins
                float2 w0 = wv[0];
                float2 p0 = pp[0], p1 = pp[32];
                f32x2 acc0 = mul_(pack2(p0.x, p0.y), pack2(w0.x, w0.x));
                f32x2 acc1 = mul_(pack2(p1.x, p1.y), pack2(w0.y, w0.y));
#pragma unroll 2
                for (int j = 2; j < n2; j += 2) {
                    const float2 wj = wv[j >> 1];
                    const float2 q0 = pp[j * 32], q1 = pp[j * 32 + 32];
                    acc0 = fma_(pack2(q0.x, q0.y), pack2(wj.x, wj.x), acc0);
                    acc1 = fma_(pack2(q1.x, q1.y), pack2(wj.y, wj.y), acc1);
                }
                float a0, a1;
                unpack2(add_(acc0, acc1), a0, a1);
                const float v0 = log2_pos(fmax_nan(a0, 1e-10f)) * kLog10Of2, v1 = log2_pos(fmax_nan(a1, 1e-10f)) * kLog10Of2;
                // raw row -> HBM straight from the registers (src/audio.rs:254)
                if (ok0) { orow[0] = v0; lmax = fmax_nan(lmax, v0); }
                if (ok1) { orow[16] = v1; lmax = fmax_nan(lmax, v1); }
                orow += p.n_frames;
            }
        }
        if (fin) {
            const float fl_ = sm.seg_floor;              // out = (max(raw, segmax - 8) + 4) / 4   (src/audio.rs:256-258)
            if (ovec) {
#pragma unroll
                for (int j = 0; j < kPer; ++j) {
                    const int idx = bt + j * kBThreads;
                    if ((idx & 15) < fnf4) {
                        float4 v = raw[j];
                        v.x = floor_scale(v.x, fl_); v.y = floor_scale(v.y, fl_); v.z = floor_scale(v.z, fl_); v.w = floor_scale(v.w, fl_);
                        *reinterpret_cast<float4*>(foutp + (int64_t)(idx >> 4) * p.n_frames + 4 * (idx & 15)) = v;
                    }
                }
            } else {
                floor_scale_slice<kBThreads>(foutp, p.n_frames, fnf, fl_, bt, false);   // unaligned output
            }
#pragma unroll
            for (int j = 0; j + 1 < kPend; ++j) { pend_seg[j] = pend_seg[j + 1]; pend_fs[j] = pend_fs[j + 1]; }
            pend_fs[kPend - 1] = -1;
        }
        // ---- per-slice max -> per-segment max (src/audio.rs:256 `log_spec.max()`); a NaN wins (its key is above +inf)
        {
            int kmax = float_key(lmax);
            kmax = __reduce_max_sync(0xffffffffu, kmax);
            // The last warp to add its maximum posts the slice: once the maximum is out the slice has posted (its raw
            // rows are only ever re-read by this CTA).
            if (lane == 0) {
                atomicMax(&sm.red_max, kmax);
                __threadfence_block();
                if (atomicAdd(&sm.warps_done, 1) == kBWarps - 1) {
                    __threadfence_block();
                    const int key = atomicExch(&sm.red_max, float_key(-INFINITY));
                    atomicExch(&sm.warps_done, 0);
                    if (FUSED) st_relaxed(p.slice_max + (int64_t)seg * p.slices_per_seg + sl, key);
                    else atomicMax(p.seg_max + seg, key);
                }
            }
        }
        if (FUSED) {
            if (pend_fs[kPend - 1] >= 0) {               // queue full (a straggling segment): wait for the oldest
                bar_bc();
                finalize_slice_blocking(sm, p, pend_seg[0], pend_fs[0], bt);
#pragma unroll
                for (int j = 0; j + 1 < kPend; ++j) { pend_seg[j] = pend_seg[j + 1]; pend_fs[j] = pend_fs[j + 1]; }
                pend_fs[kPend - 1] = -1;
            }
            bool placed = false;
#pragma unroll
            for (int j = 0; j < kPend; ++j)
                if (!placed && pend_fs[j] < 0) { pend_seg[j] = seg; pend_fs[j] = fs; placed = true; }
        }
        seg = nseg; sl = nsl;
    }
    if (FUSED) {
        bar_bc();
#pragma unroll
        for (int j = 0; j < kPend; ++j)
            if (pend_fs[j] >= 0) finalize_slice_blocking(sm, p, pend_seg[j], pend_fs[j], bt);
    }
}

// second launch of the two-pass mode: in-place max(x, segmax-8), (x+4)/4
__global__ void __launch_bounds__(256) logmel_finalize_kernel(float* __restrict__ out, const int* __restrict__ seg_max,
                                                              int64_t per_seg, int64_t total, int vec) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (vec) {
        const int64_t per4 = per_seg >> 2, total4 = total >> 2;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += stride) {
            const float fl = key_float(seg_max[i / per4]) - 8.0f;
            float4 v = reinterpret_cast<float4*>(out)[i];
            v.x = (fmaxf(v.x, fl) + 4.0f) * 0.25f; v.y = (fmaxf(v.y, fl) + 4.0f) * 0.25f;
            v.z = (fmaxf(v.z, fl) + 4.0f) * 0.25f; v.w = (fmaxf(v.w, fl) + 4.0f) * 0.25f;
            reinterpret_cast<float4*>(out)[i] = v;
        }
    } else {
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
            const float fl = key_float(seg_max[i / per_seg]) - 8.0f;
            out[i] = (fmaxf(out[i], fl) + 4.0f) * 0.25f;
        }
    }
}

__global__ void pad_or_trim_kernel(const float* __restrict__ in, int64_t rows, int64_t in_len, int64_t in_stride,
                                   float* __restrict__ out, int64_t out_len) {
    const int64_t total = rows * out_len, stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int64_t r = i / out_len, c = i - r * out_len;
        out[i] = c < in_len ? in[r * in_stride + c] : 0.0f;
    }
}

static bool g_tables_ready[64] = {false};

static int init_tables(wrs_ctx* ctx) {
    if (ctx->device < 64 && g_tables_ready[ctx->device]) return WRS_OK;
    const double PI = 3.14159265358979323846;
    static float win[WRS_N_FFT];
    static float2 tw[16 * 13];
    for (int n = 0; n < WRS_N_FFT; ++n) win[n] = (float)(0.5 - 0.5 * std::cos(2.0 * PI * n / WRS_N_FFT));
    for (int t = 0; t < 16; ++t)
        for (int q = 0; q < 13; ++q) {
            const double a = -2.0 * PI * (double)(t * q) / 400.0;
            tw[t * 13 + q] = make_float2((float)std::cos(a), (float)std::sin(a));
        }
    WRS_CUDA(ctx, cudaMemcpyToSymbol(c_win, win, sizeof(win)));
    WRS_CUDA(ctx, cudaMemcpyToSymbol(c_tw, tw, sizeof(tw)));
    if (ctx->device < 64) g_tables_ready[ctx->device] = true;
    return WRS_OK;
}

}  // namespace wrs

extern "C" {


int wrs_logmel_set_filter(wrs_ctx* ctx, const float* filter_host) {
    if (!ctx) return wrs::fail(nullptr, WRS_ERR_INVALID, "wrs_logmel_set_filter: null ctx");
    std::vector<float> builtin;
    if (!filter_host) {
        builtin.resize(WRS_N_MELS * WRS_N_BINS);
        wrs::mel_filter_slaney(builtin.data());
        filter_host = builtin.data();
    }
    wrs::MelBank bank;
    if (!wrs::build_mel_bank(filter_host, &bank))
        return wrs::fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_logmel_set_filter: filterbank is not banded (more than 768 in-band weights after padding every band to an even count)");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    WRS_CUDA(ctx, cudaMemcpy(ctx->d_bank, &bank, sizeof(bank), cudaMemcpyHostToDevice));
    return WRS_OK;
}

}  // extern "C"

namespace wrs {

// ---- scratch of wrs_logmel_batch_noalloc: the per-slice (or per-segment) max words of ONE call, resident in every
// GPU that loads the library until exit.  64 Ki words = 256 KB: up to 1394 segments of 30 s (47 slices each).
constexpr int64_t kNoAllocWords = 65536;
static __device__ __align__(256) int g_noalloc_seg_state[kNoAllocWords];
static int* g_noalloc_ptr[64] = {nullptr};            // device address of the array, per device (host code must not use the symbol itself)

// `scratch` == nullptr: the context's own buffer, grown with cudaMalloc when a call needs more (wrs_logmel_batch);
// otherwise the caller-independent words [0, scratch_cap) and no allocation of any kind (wrs_logmel_batch_noalloc).
static int logmel_batch_impl(wrs_ctx* ctx, const char* who, const float* pcm, int64_t n_segments, int64_t n_samples,
                             int64_t seg_stride, float* out, wrs_stream_t stream_, int* scratch, int64_t scratch_cap);

}  // namespace wrs

extern "C" {

int wrs_logmel_batch(wrs_ctx* ctx, const float* pcm, int64_t n_segments, int64_t n_samples, int64_t seg_stride,
                     float* out, wrs_stream_t stream_) {
    return wrs::logmel_batch_impl(ctx, "wrs_logmel_batch", pcm, n_segments, n_samples, seg_stride, out, stream_, nullptr, 0);
}

int wrs_logmel_batch_noalloc(wrs_ctx* ctx, const float* pcm, int64_t n_segments, int64_t n_samples, int64_t seg_stride,
                             float* out, wrs_stream_t stream_) {
    using namespace wrs;
    if (!ctx) return fail(nullptr, WRS_ERR_INVALID, "wrs_logmel_batch_noalloc: null ctx");
    if (ctx->device < 0 || ctx->device >= 64) return fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_logmel_batch_noalloc: device index beyond 63");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!g_noalloc_ptr[ctx->device]) {                // once per device, with that device current
        int* ptr = nullptr;
        WRS_CUDA(ctx, cudaGetSymbolAddress((void**)&ptr, g_noalloc_seg_state));
        g_noalloc_ptr[ctx->device] = ptr;
    }
    return logmel_batch_impl(ctx, "wrs_logmel_batch_noalloc", pcm, n_segments, n_samples, seg_stride, out, stream_,
                             g_noalloc_ptr[ctx->device], kNoAllocWords);
}

}  // extern "C"

namespace wrs {

static int logmel_batch_impl(wrs_ctx* ctx, const char* who_, const float* pcm, int64_t n_segments, int64_t n_samples,
                             int64_t seg_stride, float* out, wrs_stream_t stream_, int* scratch, int64_t scratch_cap) {
    const std::string who(who_);
    if (!ctx) return fail(nullptr, WRS_ERR_INVALID, who + ": null ctx");
    if (n_segments == 0) return WRS_OK;
    if (!pcm || !out) return fail(ctx, WRS_ERR_INVALID, who + ": null buffer");
    if (n_segments < 0 || n_samples < WRS_N_FFT || seg_stride < n_samples)
        return fail(ctx, WRS_ERR_INVALID, who + ": need n_segments >= 0, n_samples >= 400 (reflect padding needs > 200), seg_stride >= n_samples");
    if (n_samples > (int64_t)1 << 30 || n_segments > (int64_t)1 << 24)
        return fail(ctx, WRS_ERR_UNSUPPORTED, who + ": segment longer than 2^30 samples or more than 2^24 segments");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = init_tables(ctx);
    if (rc != WRS_OK) return rc;

    constexpr int TW = kTW;
    LogmelParams p;
    p.pcm = pcm; p.out = out; p.seg_stride = seg_stride;
    p.n_segments = (int)n_segments; p.n_samples = (int)n_samples; p.n_frames = (int)(n_samples / WRS_HOP_LENGTH);
    p.slices_per_seg = (p.n_frames + TW - 1) / TW;
    const int64_t total = (int64_t)p.slices_per_seg * n_segments;
    if (total > INT32_MAX) return fail(ctx, WRS_ERR_UNSUPPORTED, who + ": too many slices");
    p.total_slices = (int)total;
    p.vec_in = ((reinterpret_cast<uintptr_t>(pcm) & 15) == 0 && (seg_stride & 3) == 0) ? 1 : 0;
    p.vec_out = ((reinterpret_cast<uintptr_t>(out) & 15) == 0 && (p.n_frames & 3) == 0) ? 1 : 0;
    p.bank = ctx->d_bank;
    p.status = ctx->d_status;

    const size_t smem = sizeof(LogmelSmem);
    if (ctx->logmel_max_ctas_per_sm == 0) {
        WRS_CUDA(ctx, cudaFuncSetAttribute(logmel_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        WRS_CUDA(ctx, cudaFuncSetAttribute(logmel_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        WRS_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, logmel_kernel<true>, kThreads, smem));
        if (occ < 1) return fail(ctx, WRS_ERR_CUDA, "logmel kernel does not fit on an SM");
        ctx->logmel_max_ctas_per_sm = occ;
    }
    const int resident = ctx->sm_count * ctx->logmel_max_ctas_per_sm;
    const int grid = (int)std::min<int64_t>(total, resident);
    // The single fused launch needs a segment's slices to be dealt within one round of the grid: then every slice is
    // posted by a CTA that has not had to wait yet, and no wait can be circular.  Longer segments (a whole clip as one
    // "segment": src/transcribe.rs:112 takes the maximum over the clip) take the two-launch form.
    const bool fused = ctx->opt_logmel_fused != 0 && p.slices_per_seg <= grid;
    const int64_t need = fused ? total : n_segments;     // fused: one word per slice; two launches: one per segment
    int* seg_state = scratch;
    if (scratch) {
        if (need > scratch_cap)
            return fail(ctx, WRS_ERR_UNSUPPORTED, who + ": the call needs " + std::to_string(need) + " scratch words (one per 64-frame slice, or one per segment for a clip longer than one round of the grid); the fixed scratch holds " +
                                                      std::to_string(scratch_cap) + " -- use wrs_logmel_batch or split the batch");
    } else {
        if (need > ctx->seg_cap) {
            WRS_CUDA(ctx, cudaStreamSynchronize(stream));
            cudaFree(ctx->d_seg_state);
            ctx->d_seg_state = nullptr;
            const int64_t cap = std::max<int64_t>(need, 4096);
            WRS_CUDA(ctx, cudaMalloc(&ctx->d_seg_state, sizeof(int) * cap));
            ctx->seg_cap = cap;
        }
        seg_state = ctx->d_seg_state;
    }
    p.seg_max = seg_state;
    p.slice_max = seg_state;
    // 0x80808080: kSlotEmpty for the per-slice words, and (decoding to about -3.4e38) "minus infinity" for atomicMax
    // (the shared scratch keeps the last call's words: every call clears what it is going to use)
    WRS_CUDA(ctx, cudaMemsetAsync(seg_state, 0x80, sizeof(int) * need, stream));

    // fused: the floor/scale of a slice is applied one slice later by the CTA that produced it, once the segment's
    // other slices (all owned by co-resident CTAs: cooperative launch) have posted their maxima
    ctx->stat_last_logmel_mode = fused ? 1 : 0;
    ctx->stat_last_logmel_grid = grid;
    if (fused) {
        void* args[] = {(void*)&p};
        WRS_CUDA(ctx, cudaLaunchCooperativeKernel((void*)logmel_kernel<true>, dim3(grid), dim3(kThreads), args, smem, stream));
    } else {
        logmel_kernel<false><<<grid, kThreads, smem, stream>>>(p);
        WRS_CUDA(ctx, cudaGetLastError());
        const int64_t per_seg = (int64_t)WRS_N_MELS * p.n_frames, tot = per_seg * n_segments;
        const int vec = ((reinterpret_cast<uintptr_t>(out) & 15) == 0 && (per_seg & 3) == 0) ? 1 : 0;
        const int fgrid = (int)std::min<int64_t>((tot / (vec ? 4 : 1) + 255) / 256, (int64_t)ctx->sm_count * 8);
        logmel_finalize_kernel<<<std::max(fgrid, 1), 256, 0, stream>>>(out, p.seg_max, per_seg, tot, vec);
        WRS_CUDA(ctx, cudaGetLastError());
    }
    return WRS_OK;
}

}  // namespace wrs

extern "C" {

int wrs_logmel_batch_host(wrs_ctx* ctx, const float* pcm_host, int64_t n_segments, int64_t n_samples,
                          int64_t seg_stride, float* out_host) {
    using namespace wrs;
    if (!ctx) return fail(nullptr, WRS_ERR_INVALID, "wrs_logmel_batch_host: null ctx");
    if (n_segments == 0) return WRS_OK;
    if (!pcm_host || !out_host || n_segments < 0 || n_samples < WRS_N_FFT || seg_stride < n_samples)
        return fail(ctx, WRS_ERR_INVALID, "wrs_logmel_batch_host: bad arguments");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t n_frames = n_samples / WRS_HOP_LENGTH;
    const size_t in_bytes = sizeof(float) * ((n_segments - 1) * seg_stride + n_samples);
    const size_t out_bytes = sizeof(float) * n_segments * WRS_N_MELS * n_frames;
    if (in_bytes > ctx->stage_in_bytes) {
        cudaFree(ctx->d_stage_in); ctx->d_stage_in = nullptr; ctx->stage_in_bytes = 0;
        WRS_CUDA(ctx, cudaMalloc(&ctx->d_stage_in, in_bytes));
        ctx->stage_in_bytes = in_bytes;
    }
    if (out_bytes > ctx->stage_out_bytes) {
        cudaFree(ctx->d_stage_out); ctx->d_stage_out = nullptr; ctx->stage_out_bytes = 0;
        WRS_CUDA(ctx, cudaMalloc(&ctx->d_stage_out, out_bytes));
        ctx->stage_out_bytes = out_bytes;
    }
    // Three streams: chunk c is copied in while chunk c-1 is computed and chunk c-2 is copied out, so that a large
    // batch costs max(copy-in, copy-out) over PCIe (full duplex) instead of their sum; pinned host buffers make the
    // copies truly asynchronous.  Chunks are whole segments.
    if (!ctx->h2d_stream) {
        WRS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->h2d_stream, cudaStreamNonBlocking));
        WRS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking));
        for (int i = 0; i < wrs_ctx::kHostChunks; ++i) {
            WRS_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_in[i], cudaEventDisableTiming));
            WRS_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
        }
    }
    const int64_t per_chunk = std::max<int64_t>(4, (n_segments + wrs_ctx::kHostChunks - 1) / wrs_ctx::kHostChunks);
    const int64_t out_per_seg = (int64_t)WRS_N_MELS * n_frames;
    cudaStream_t s = ctx->own_stream;
    int c = 0;
    for (int64_t s0 = 0; s0 < n_segments; s0 += per_chunk, ++c) {
        const int64_t ns = std::min(per_chunk, n_segments - s0);
        const size_t bytes_in = sizeof(float) * ((ns - 1) * seg_stride + n_samples);
        WRS_CUDA(ctx, cudaMemcpyAsync(ctx->d_stage_in + s0 * seg_stride, pcm_host + s0 * seg_stride, bytes_in,
                                      cudaMemcpyHostToDevice, ctx->h2d_stream));
        WRS_CUDA(ctx, cudaEventRecord(ctx->ev_in[c], ctx->h2d_stream));
        WRS_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_in[c], 0));
        int rc = wrs_logmel_batch(ctx, ctx->d_stage_in + s0 * seg_stride, ns, n_samples, seg_stride,
                                  ctx->d_stage_out + s0 * out_per_seg, s);
        if (rc != WRS_OK) { cudaDeviceSynchronize(); return rc; }
        WRS_CUDA(ctx, cudaEventRecord(ctx->ev_done[c], s));
        WRS_CUDA(ctx, cudaStreamWaitEvent(ctx->d2h_stream, ctx->ev_done[c], 0));
        WRS_CUDA(ctx, cudaMemcpyAsync(out_host + s0 * out_per_seg, ctx->d_stage_out + s0 * out_per_seg,
                                      sizeof(float) * ns * out_per_seg, cudaMemcpyDeviceToHost, ctx->d2h_stream));
    }
    WRS_CUDA(ctx, cudaStreamSynchronize(ctx->d2h_stream));
    WRS_CUDA(ctx, cudaStreamSynchronize(ctx->h2d_stream));
    WRS_CUDA(ctx, cudaStreamSynchronize(s));
    return WRS_OK;
}

int wrs_pad_or_trim(wrs_ctx* ctx, const float* in, int64_t rows, int64_t in_len, int64_t in_stride, float* out,
                    int64_t out_len, wrs_stream_t stream) {
    using namespace wrs;
    if (!ctx) return fail(nullptr, WRS_ERR_INVALID, "wrs_pad_or_trim: null ctx");
    if (rows < 0 || in_len < 0 || out_len < 0 || in_stride < in_len) return fail(ctx, WRS_ERR_INVALID, "wrs_pad_or_trim: bad sizes");
    if (rows * out_len == 0) return WRS_OK;
    if (!in && in_len > 0) return fail(ctx, WRS_ERR_INVALID, "wrs_pad_or_trim: null input");
    if (!out) return fail(ctx, WRS_ERR_INVALID, "wrs_pad_or_trim: null output");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    const int grid = (int)std::min<int64_t>((rows * out_len + 255) / 256, (int64_t)ctx->sm_count * 8);
    pad_or_trim_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, rows, in_len, in_stride, out, out_len);
    WRS_CUDA(ctx, cudaGetLastError());
    return WRS_OK;
}

}  // extern "C"

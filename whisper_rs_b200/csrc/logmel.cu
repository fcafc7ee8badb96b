// K1: fused log-mel front-end for batches of audio segments (sm_100a).
//
// Replaces the ATen op sequence of reference src/audio.rs:230-259 (hann_window, stft, abs, pow, matmul,
// clamp_min, log10, max, maximum, add, div) with one persistent kernel.  A CTA (16 warps) takes slices of 64
// frames; every lane works on TWO frames held side by side in 64-bit register pairs, so the whole FFT runs on
// Blackwell's packed FP32 instructions (FADD2 / FMUL2 / FFMA2: one issue slot, two frames):
//
//   PCM (HBM) --bulk async copy--> smem tile, row-major [67 rows of 160 samples], pitch 164     (no registers, one
//            mbarrier; issued by an idle warp as soon as stage A has released the tile, lands behind stages B and C)
//   stage A  warp (g, tq), lane (f8, tl): frames 8g+f8 and 8g+f8+32, t = 4tq+tl:
//            real 25-point DFT over the taps 16m+t, times W400^{tq}   -> smem Y[t][q][frame pair]  (float4: re2, im2)
//            (8 frames x 4 consecutive taps per load = 32 distinct banks at pitch 164)
//   stage B  warp q, lane = frame pair: complex 16-point DFT, |X|^2   -> smem P[bin][frame pair]   (float2)
//            warp 14 issues the next slice's PCM copy; warp 15 posts the previous slice and polls the pending one
//   stage C  warp w, lane = frame pair: 5 mel rows (banded FP32 FMA), log10 -> smem tile[mel][frame]
//   slice end       : raw log-mel -> HBM, block max -> atomicMax per segment
//   one slice later : the slice is counted as posted; once all slices of a segment have posted, its max is final
//                     and floor (max-8), (x+4)/4 are applied in place on the slice's own 20 KB (still in L2)
#include <algorithm>
#include <cmath>

#include "async_copy.cuh"
#include "logmel_fft.cuh"
#include "wrs_internal.cuh"

namespace wrs {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kTW = 64;              // frames per slice (two per lane)
constexpr int kRows = kTW + 3;       // 160-sample rows f0-2 .. f0+64 needed by 64 frames
constexpr int kPitch = WRS_HOP_LENGTH + 4;   // words per tile row: 16-byte aligned rows, frame stride == 4 banks
constexpr float kLog10Of2 = 0.30102999566398120f;
constexpr int kSlotEmpty = (int)0x80808080;         // memset pattern; as a key it decodes to -3.4e38, never a log10 value

struct TabEntry { int off; float win; };             // tap n of frame 0: word offset in the tile, periodic Hann weight
__constant__ TabEntry c_tab[WRS_N_FFT];              // src/audio.rs:233
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
    float pcm[kRows * kPitch];                       // 43 952 B
    float4 Y[16 * 13 * 32];                          // 106 496 B
    float2 P[kPRows * 32];                           // 52 224 B
    float w[MelBank::kMaxPadded];                    // 3 072 B: in-band filter weights, bands zero-padded to multiples of 4
    TabEntry tab[WRS_N_FFT];                         // 3 200 B
    float2 tw[16 * 13];                              // 1 664 B
    int lo[WRS_N_MELS], n4[WRS_N_MELS], off4[WRS_N_MELS];
    unsigned long long pcm_bar;                      // mbarrier: bytes of the PCM tile in flight
    int red_max;                                     // max key of the slice in progress (reset by the last warp to add to it)
    int warps_done;
#ifdef WRS_TRACE
    long long t_arrive[4];                           // barrier-B arrival times of warps 5, 13, 15 (investigation aid)
#endif
    int fin_ready;                                   // the oldest pending slice's segment is complete: seg_floor is valid
    float seg_floor;
};
static_assert(sizeof(LogmelSmem) <= 227 * 1024, "log-mel working set exceeds the 227 KB a CTA can own");

__device__ __forceinline__ int float_key(float f) {
    const int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

// one sample of the (virtually reflect-padded) segment; i is relative to the segment start
__device__ __forceinline__ float load_reflect(const float* __restrict__ seg, int i, int n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    return (i >= 0 && i < n) ? __ldg(seg + i) : 0.0f;
}

// The tile of slice (seg, fs) is rows fs-2 .. fs+64 of the segment cut into 160-sample rows.  Rows that lie entirely
// inside the segment are one contiguous range of the input and move as bulk copies (16-byte aligned PCM); the rows
// in front of the segment (reflection, first slice) and at or behind its end (last slice) are filled element-wise.
struct TilePlan { int r_lo, r_hi; };                 // plain rows [r_lo, r_hi)
__device__ __forceinline__ TilePlan tile_plan(const LogmelParams& p, int fs) {
    TilePlan t;
    t.r_lo = max(0, 2 - fs);
    t.r_hi = p.vec_in != 0 ? max(t.r_lo, min(kRows, p.n_samples / WRS_HOP_LENGTH - fs + 2)) : t.r_lo;
    return t;
}
// Row copies of 640 B, completion counted in bytes on the tile's mbarrier.  Two warps share the issue (part 0 also makes
// the phase's single arrival and announces the byte count; a copy that completes before that only drives the pending
// byte count negative for a moment).
__device__ __forceinline__ void stage_bulk(LogmelSmem& sm, const LogmelParams& p, int seg, int fs, TilePlan t, int lane, int part) {
    const float* __restrict__ src = p.pcm + (int64_t)seg * p.seg_stride + (int64_t)WRS_HOP_LENGTH * (fs - 2);
    const int nb = t.r_hi - t.r_lo, half = (nb + 1) >> 1;
    fence_proxy_async();                             // the tile was last touched by ordinary loads/stores
    if (part == 0 && lane == 0) mbar_expect_tx(&sm.pcm_bar, (unsigned)(nb * WRS_HOP_LENGTH * sizeof(float)));
    const int r0 = t.r_lo + part * half, r1 = min(t.r_hi, r0 + half);
    // warp-uniform loop, one elected lane issues: the addresses stay in uniform registers (a per-lane loop would be
    // serialised lane by lane through R2UR at about 80 cycles per copy)
    if (elect_one()) {
#pragma unroll 4
        for (int r = r0; r < r1; ++r)
            bulk_g2s(sm.pcm + r * kPitch, src + r * WRS_HOP_LENGTH, WRS_HOP_LENGTH * (unsigned)sizeof(float), &sm.pcm_bar);
    }
}
// the rows outside [r_lo, r_hi): element-wise with reflection at both ends (src/audio.rs:236 center = true), whole CTA
__device__ __forceinline__ void stage_general(LogmelSmem& sm, const LogmelParams& p, int seg, int fs, TilePlan t, int tid) {
    const float* __restrict__ segp = p.pcm + (int64_t)seg * p.seg_stride;
    const int base = WRS_HOP_LENGTH * (fs - 2);
    const int n1 = t.r_lo * WRS_HOP_LENGTH, total = n1 + (kRows - t.r_hi) * WRS_HOP_LENGTH;
#pragma unroll 4
    for (int idx = tid; idx < total; idx += kThreads) {
        const int e = idx < n1 ? idx : idx - n1 + t.r_hi * WRS_HOP_LENGTH;
        const int r = e / WRS_HOP_LENGTH, c = e - r * WRS_HOP_LENGTH;
        sm.pcm[r * kPitch + c] = load_reflect(segp, base + e, p.n_samples);
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
        constexpr int kPer = (WRS_N_MELS * (kTW / 4) + NT - 1) / NT, kBatch = kPer < 7 ? kPer : 7;
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
                    x.x = (fmaxf(x.x, fl) + 4.0f) * 0.25f; x.y = (fmaxf(x.y, fl) + 4.0f) * 0.25f;
                    x.z = (fmaxf(x.z, fl) + 4.0f) * 0.25f; x.w = (fmaxf(x.w, fl) + 4.0f) * 0.25f;
                    *reinterpret_cast<float4*>(outp + (int64_t)(idx >> 4) * n_frames + 4 * (idx & 15)) = x;
                }
            }
        }
    } else {
        const int d = max(nf, 1);
        for (int idx = t; idx < WRS_N_MELS * nf; idx += NT) {
            const int m = idx / d, c = idx - m * d;
            float* q = outp + (int64_t)m * n_frames + c;
            *q = (fmaxf(__ldcg(q), fl) + 4.0f) * 0.25f;
        }
    }
}

// Blocking floor + scale of one slice in place: out = (max(raw, segmax - 8) + 4) / 4 (src/audio.rs:256-258).  Only the
// drain at the end of the kernel (and a full pending queue) comes here; the steady state hides it inside the slice loop.
__device__ __forceinline__ void finalize_slice_blocking(LogmelSmem& sm, const LogmelParams& p, int seg, int fs, int tid) {
    const int nf = min(kTW, p.n_frames - fs);
    if (tid < 32) {
        // every slice of the segment is owned by a CTA that is running or has finished (persistent grid no larger than
        // what is co-resident; a CTA posts its slice before it ever waits), so this wait ends; the bound is a safety net
        const int* slots = p.slice_max + (int64_t)seg * p.slices_per_seg;
        int spins = 0, key;
        for (;;) {
            const int a = tid < p.slices_per_seg ? ld_relaxed(slots + tid) : 0;
            const int b = tid + 32 < p.slices_per_seg ? ld_relaxed(slots + tid + 32) : 0;
            bool complete;
            key = fold_slots(p, seg, a, b, tid, complete);
            if (complete) break;
            __nanosleep(64);
            if (++spins > (1 << 22)) { if (tid == 0) atomicExch(&p.status->timeout, 1); break; }
        }
        if (tid == 0) sm.seg_floor = key_float(key) - 8.0f;
    }
    __syncthreads();
    floor_scale_slice<kThreads>(p.out + ((int64_t)seg * WRS_N_MELS) * p.n_frames + fs, p.n_frames, nf, sm.seg_floor, tid, p.vec_out != 0);
    __syncthreads();
}


constexpr int kPend = 4;              // slices a CTA may have waiting for their segment's maximum

#ifdef WRS_TRACE
// investigation aid (tools/debug/logmel_trace.py): per-CTA cycles spent in each phase of the slice loop
__device__ unsigned long long g_lm_trace[148 * 12];
#define LM_T(k) do { if (tid == 0) { const long long now_ = clock64(); acc_[k] += now_ - t_; t_ = now_; } } while (0)
#else
#define LM_T(k) do {} while (0)
#endif

template <bool FUSED>
__global__ void __launch_bounds__(kThreads, 1) logmel_kernel(const LogmelParams p) {
#ifdef WRS_TRACE
    long long acc_[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, t_ = clock64();
#endif
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LogmelSmem& sm = *reinterpret_cast<LogmelSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // tables -> smem (once per CTA); the rows of P behind bin 200 stay zero (zero-padded band tails read them)
    for (int i = tid; i < WRS_N_MELS; i += kThreads) { sm.lo[i] = p.bank->lo[i]; sm.n4[i] = p.bank->n4[i]; sm.off4[i] = p.bank->off4[i]; }
    for (int i = tid; i < MelBank::kMaxPadded; i += kThreads) sm.w[i] = p.bank->wp[i];
    for (int i = tid; i < WRS_N_FFT; i += kThreads) sm.tab[i] = c_tab[i];
    for (int i = tid; i < 16 * 13; i += kThreads) sm.tw[i] = c_tw[i];
    for (int i = tid; i < (kPRows - WRS_N_BINS) * 32; i += kThreads) sm.P[WRS_N_BINS * 32 + i] = make_float2(0.0f, 0.0f);
    if (tid == 0) { mbar_init(&sm.pcm_bar, 1); sm.red_max = float_key(-INFINITY); sm.warps_done = 0; }
    __syncthreads();

    // slices are dealt round-robin; (seg, fs) of the current slice advance without divisions
    const int step_seg = (int)gridDim.x / p.slices_per_seg, step_sl = (int)gridDim.x % p.slices_per_seg;
    int seg = (int)blockIdx.x / p.slices_per_seg, sl = (int)blockIdx.x % p.slices_per_seg;
    // slices whose floor/scale is still to be applied, oldest first (fs < 0: empty)
    int pend_seg[kPend], pend_fs[kPend];
#pragma unroll
    for (int k = 0; k < kPend; ++k) { pend_seg[k] = 0; pend_fs[k] = -1; }
    unsigned pcm_parity = 0;                             // phase of pcm_bar the next bulk-copied tile completes
    bool cur_bulk = false;                               // the current slice's tile arrives through pcm_bar
    const bool ovec = p.vec_out != 0;                    // output rows 16-byte aligned, n_frames % 4 == 0: rows move as bulk copies
    if ((int)blockIdx.x < p.total_slices) {
        const TilePlan t = tile_plan(p, sl * kTW);
        cur_bulk = t.r_hi > t.r_lo;
        if (cur_bulk && (warp == 13 || warp == 14)) stage_bulk(sm, p, seg, sl * kTW, t, lane, warp - 13);
        stage_general(sm, p, seg, sl * kTW, t, tid);
    }
    __syncthreads();

    for (int slice = blockIdx.x; slice < p.total_slices; slice += gridDim.x) {
        const int fs = sl * kTW;                         // first frame of the slice
        const int nf = min(kTW, p.n_frames - fs);
        int nseg = seg + step_seg, nsl = sl + step_sl;   // the slice after this one
        if (nsl >= p.slices_per_seg) { nsl -= p.slices_per_seg; ++nseg; }
        const bool has_next = slice + (int)gridDim.x < p.total_slices;
        const TilePlan nplan = tile_plan(p, nsl * kTW);
        const bool next_bulk = has_next && nplan.r_hi > nplan.r_lo;
        const bool next_edge = has_next && nplan.r_hi - nplan.r_lo < kRows;
        // has the oldest pending slice's segment completed?  The answer is asked for here and looked at in stage B, so
        // that the round trip to L2 is off every warp's critical path.
        int poll_a = 0, poll_b = 0;
        if (FUSED && warp == 15 && pend_fs[0] >= 0) {
            const int* slots = p.slice_max + (int64_t)pend_seg[0] * p.slices_per_seg;
            if (lane < p.slices_per_seg) poll_a = ld_relaxed(slots + lane);
            if (lane + 32 < p.slices_per_seg) poll_b = ld_relaxed(slots + lane + 32);
        }
        LM_T(10);
        if (cur_bulk) { mbar_wait(&sm.pcm_bar, pcm_parity); pcm_parity ^= 1u; }
        LM_T(7);

        // ---- stage A: 4 frame groups x 4 tap quads.  Lane (f8, tl): frames 8g+f8 and 8g+f8+32, taps 16m + 4tq + tl.
        {
            // opaque copy: keeps the compiler from hoisting this thread's 25 tap offsets / window values / twiddles out
            // of the slice loop, where they would sit in registers (and spill) through stages B and C
            int wv = warp;
            asm volatile("" : "+r"(wv));
            const int g = wv & 3, t = (wv >> 2) * 4 + (lane >> 3), fp = 8 * g + (lane & 7);
            const float* src = sm.pcm + fp * kPitch;
            f32x2 v[25];
#pragma unroll
            for (int m = 0; m < 25; ++m) {
                const TabEntry e = sm.tab[16 * m + t];
                v[m] = mul_(pack2(src[e.off], src[e.off + 32 * kPitch]), pack2(e.win, e.win));
            }
            cx<f32x2> y[13];
            rdft25(v, y);
            float4* dst = sm.Y + (t * 13) * 32 + fp;
            {
                float a, b, c, d;
                unpack2(y[0].re, a, b); unpack2(y[0].im, c, d);
                dst[0] = make_float4(a, b, c, d);
            }
#pragma unroll
            for (int q = 1; q < 13; ++q) {
                const float2 w = sm.tw[t * 13 + q];
                const f32x2 wr = pack2(w.x, w.x), wi = pack2(w.y, w.y), nwi = pack2(-w.y, -w.y);
                const f32x2 re = fma_(y[q].im, nwi, mul_(y[q].re, wr));
                const f32x2 im = fma_(y[q].re, wi, mul_(y[q].im, wr));
                float a, b, c, d;
                unpack2(re, a, b); unpack2(im, c, d);
                dst[q * 32] = make_float4(a, b, c, d);
            }
        }
        __syncthreads();
        LM_T(0);

        // ---- stage B: warp = q < 13, lane = frame pair (lane, lane + 32).  The tile is free now: warp 14 starts the
        //      next slice's copy.  One lane of warp 15 counts the previous slice as posted and asks whether the oldest
        //      pending slice's segment has completed (all its slices have posted their maxima).
        if (warp < 13) {
            const int q = warp;
            cx<f32x2> z[16];
#pragma unroll
            for (int t = 0; t < 16; ++t) {
                const float4 a = sm.Y[(t * 13 + q) * 32 + lane];
                z[t].re = pack2(a.x, a.y); z[t].im = pack2(a.z, a.w);
            }
            cdft16(z);
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                if (q != 0 || s < 9) {
                    float a, b;
                    unpack2(power(z[s]), a, b);
                    sm.P[stageb_bin(q, s) * 32 + lane] = make_float2(a, b);
                }
            }
        } else {
            // the three warps without a bin: 13 and 14 start the next slice's copy (the tile is free now); 15 folds the
            // per-slice maxima it asked for at the top of the slice: is the oldest pending slice's segment complete?
            if (warp < 15) {
                if (next_bulk) stage_bulk(sm, p, nseg, nsl * kTW, nplan, lane, warp - 13);
            } else if (FUSED) {
                bool ready = false;
                int key = 0;
                if (pend_fs[0] >= 0) key = fold_slots(p, pend_seg[0], poll_a, poll_b, lane, ready);
                if (lane == 0) {
                    if (ready) sm.seg_floor = key_float(key) - 8.0f;
                    sm.fin_ready = ready ? 1 : 0;
                }
            }
        }
#ifdef WRS_TRACE
        if (lane == 0 && (warp == 5 || warp == 13 || warp == 15)) sm.t_arrive[warp == 5 ? 0 : (warp == 13 ? 1 : 2)] = clock64();
        const long long tB0_ = t_;
#endif
        __syncthreads();
        LM_T(1);
#ifdef WRS_TRACE
        if (tid == 0) { acc_[9] += sm.t_arrive[0] - tB0_; acc_[11] += sm.t_arrive[1] - tB0_; acc_[8] += sm.t_arrive[2] - tB0_; }
#endif

        // ---- stage C: banded mel projection (src/audio.rs:252), clamp + log10 (:254), raw rows -> HBM
        const bool fin = FUSED && sm.fin_ready != 0;
        const int fseg = pend_seg[0], ffs = pend_fs[0];
        const int fnf = fin ? min(kTW, p.n_frames - ffs) : 0;
        float* __restrict__ foutp = p.out + ((int64_t)fseg * WRS_N_MELS) * p.n_frames + ffs;
        // the pending slice whose floor has just become known: its raw rows come back from L2 behind the mel arithmetic
        constexpr int kPer = (WRS_N_MELS * (kTW / 4) + kThreads - 1) / kThreads;   // float4 of a [80][64] slice per thread: 3
        const int fnf4 = fnf >> 2;
        float4 raw[kPer];
        if (fin && ovec) {
#pragma unroll
            for (int k = 0; k < kPer; ++k) {
                const int idx = tid + k * kThreads;
                if (idx < WRS_N_MELS * (kTW / 4) && (idx & 15) < fnf4)
                    raw[k] = __ldcg(reinterpret_cast<const float4*>(foutp + (int64_t)(idx >> 4) * p.n_frames + 4 * (idx & 15)));
            }
        }
        LM_T(2);
        if (next_edge) {
            stage_general(sm, p, nseg, nsl * kTW, nplan, tid);

        }
        float lmax = -INFINITY;
        float* __restrict__ outp = p.out + ((int64_t)seg * WRS_N_MELS) * p.n_frames + fs;
        {
            const bool ok0 = lane < nf, ok1 = lane + 32 < nf;
            // software pipeline over the warp's 5 rows: a row's band description and its first four bins are fetched
            // while the previous row is still in its log / store tail (most bands are 1..8 bins wide, so a row is
            // mostly latency otherwise)
            int n4 = sm.n4[warp];
            const float4* w = reinterpret_cast<const float4*>(sm.w + sm.off4[warp]);
            const float2* pp = sm.P + sm.lo[warp] * 32 + lane;
            float4 wa = w[0];
            float2 p0 = pp[0], p1 = pp[32], p2 = pp[64], p3 = pp[96];
#pragma unroll 1
            for (int i = 0; i < WRS_N_MELS / kWarps; ++i) {
                const int m = i * kWarps + warp;
                const int mn = min(m + kWarps, WRS_N_MELS - 1);      // next row (the last iteration re-reads its own)
                const int n4n = sm.n4[mn];
                const float4* wn = reinterpret_cast<const float4*>(sm.w + sm.off4[mn]);
                const float2* ppn = sm.P + sm.lo[mn] * 32 + lane;
                f32x2 acc0 = mul_(pack2(p0.x, p0.y), pack2(wa.x, wa.x));
                f32x2 acc1 = mul_(pack2(p1.x, p1.y), pack2(wa.y, wa.y));
                acc0 = fma_(pack2(p2.x, p2.y), pack2(wa.z, wa.z), acc0);
                acc1 = fma_(pack2(p3.x, p3.y), pack2(wa.w, wa.w), acc1);
                for (int j = 4; j < n4; j += 4) {
                    const float4 wb = w[j >> 2];
                    const float2 q0 = pp[j * 32], q1 = pp[j * 32 + 32], q2 = pp[j * 32 + 64], q3 = pp[j * 32 + 96];
                    acc0 = fma_(pack2(q0.x, q0.y), pack2(wb.x, wb.x), acc0);
                    acc1 = fma_(pack2(q1.x, q1.y), pack2(wb.y, wb.y), acc1);
                    acc0 = fma_(pack2(q2.x, q2.y), pack2(wb.z, wb.z), acc0);
                    acc1 = fma_(pack2(q3.x, q3.y), pack2(wb.w, wb.w), acc1);
                }
                wa = wn[0];
                p0 = ppn[0]; p1 = ppn[32]; p2 = ppn[64]; p3 = ppn[96];
                n4 = n4n; w = wn; pp = ppn;
                float a0, a1;
                unpack2(add_(acc0, acc1), a0, a1);
                const float v0 = __log2f(fmaxf(a0, 1e-10f)) * kLog10Of2, v1 = __log2f(fmaxf(a1, 1e-10f)) * kLog10Of2;
                // raw row -> HBM straight from the registers, 128 contiguous bytes per store (src/audio.rs:254)
                float* __restrict__ orow = outp + (int64_t)m * p.n_frames + lane;
                if (ok0) { orow[0] = v0; lmax = fmaxf(lmax, v0); }
                if (ok1) { orow[32] = v1; lmax = fmaxf(lmax, v1); }
            }
        }
        LM_T(3);
        if (fin) {
            const float fl = sm.seg_floor;               // out = (max(raw, segmax - 8) + 4) / 4   (src/audio.rs:256-258)
            if (ovec) {
#pragma unroll
                for (int k = 0; k < kPer; ++k) {
                    const int idx = tid + k * kThreads;
                    if (idx < WRS_N_MELS * (kTW / 4) && (idx & 15) < fnf4) {
                        float4 v = raw[k];
                        v.x = (fmaxf(v.x, fl) + 4.0f) * 0.25f; v.y = (fmaxf(v.y, fl) + 4.0f) * 0.25f;
                        v.z = (fmaxf(v.z, fl) + 4.0f) * 0.25f; v.w = (fmaxf(v.w, fl) + 4.0f) * 0.25f;
                        *reinterpret_cast<float4*>(foutp + (int64_t)(idx >> 4) * p.n_frames + 4 * (idx & 15)) = v;
                    }
                }
            } else {
                floor_scale_slice<kThreads>(foutp, p.n_frames, fnf, fl, tid, false);   // unaligned output
            }
#pragma unroll
            for (int k = 0; k + 1 < kPend; ++k) { pend_seg[k] = pend_seg[k + 1]; pend_fs[k] = pend_fs[k + 1]; }
            pend_fs[kPend - 1] = -1;
        }
        // ---- per-slice max -> per-segment max (src/audio.rs:256 `log_spec.max()`)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
        // No CTA barrier closes stage C: a warp goes straight on to stage A of the next slice (which touches nothing
        // stage C reads).  The last warp to add its maximum posts the slice: once the maximum is out the slice has
        // posted (its raw rows are only ever re-read by this CTA).
        if (lane == 0) {
            atomicMax(&sm.red_max, float_key(lmax));
            __threadfence_block();
            if (atomicAdd(&sm.warps_done, 1) == kWarps - 1) {
                __threadfence_block();
                const int key = atomicExch(&sm.red_max, float_key(-INFINITY));
                atomicExch(&sm.warps_done, 0);
                if (FUSED) st_relaxed(p.slice_max + (int64_t)seg * p.slices_per_seg + sl, key);
                else atomicMax(p.seg_max + seg, key);
            }
        }
        LM_T(4);
        if (next_edge) __syncthreads();                  // the element-wise rows of the next tile came from every warp
        LM_T(5);
        if (FUSED) {
            if (pend_fs[kPend - 1] >= 0) {               // queue full (a straggling segment): wait for the oldest
                __syncthreads();
                finalize_slice_blocking(sm, p, pend_seg[0], pend_fs[0], tid);
#pragma unroll
                for (int k = 0; k + 1 < kPend; ++k) { pend_seg[k] = pend_seg[k + 1]; pend_fs[k] = pend_fs[k + 1]; }
                pend_fs[kPend - 1] = -1;
            }
            bool placed = false;
#pragma unroll
            for (int k = 0; k < kPend; ++k)
                if (!placed && pend_fs[k] < 0) { pend_seg[k] = seg; pend_fs[k] = fs; placed = true; }
        }
        LM_T(6);
        seg = nseg; sl = nsl; cur_bulk = next_bulk;
    }
#ifdef WRS_TRACE
    if (tid == 0) for (int k = 0; k < 12; ++k) g_lm_trace[blockIdx.x * 12 + k] = acc_[k];
#endif
    if (FUSED) {
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kPend; ++k)
            if (pend_fs[k] >= 0) finalize_slice_blocking(sm, p, pend_seg[k], pend_fs[k], tid);
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
    static TabEntry tab[WRS_N_FFT];
    static float2 tw[16 * 13];
    for (int n = 0; n < WRS_N_FFT; ++n) {
        const int u = n + 120;   // frame f, tap n  ->  row f + u/160 of the tile (row 0 = frame row fs-2), column u%160
        tab[n].off = (u / WRS_HOP_LENGTH) * kPitch + (u % WRS_HOP_LENGTH);
        tab[n].win = (float)(0.5 - 0.5 * std::cos(2.0 * PI * n / WRS_N_FFT));
    }
    for (int t = 0; t < 16; ++t)
        for (int q = 0; q < 13; ++q) {
            const double a = -2.0 * PI * (double)(t * q) / 400.0;
            tw[t * 13 + q] = make_float2((float)std::cos(a), (float)std::sin(a));
        }
    WRS_CUDA(ctx, cudaMemcpyToSymbol(c_tab, tab, sizeof(tab)));
    WRS_CUDA(ctx, cudaMemcpyToSymbol(c_tw, tw, sizeof(tw)));
    if (ctx->device < 64) g_tables_ready[ctx->device] = true;
    return WRS_OK;
}

}  // namespace wrs

extern "C" {

#ifdef WRS_TRACE
int wrs_debug_logmel_trace_read(unsigned long long* out_host, int n) {
    return (int)cudaMemcpyFromSymbol(out_host, wrs::g_lm_trace, sizeof(unsigned long long) * n);
}
#endif

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
        return wrs::fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_logmel_set_filter: filterbank is not banded (more than 768 in-band weights after padding every band to a multiple of four)");
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    WRS_CUDA(ctx, cudaMemcpy(ctx->d_bank, &bank, sizeof(bank), cudaMemcpyHostToDevice));
    return WRS_OK;
}

int wrs_logmel_batch(wrs_ctx* ctx, const float* pcm, int64_t n_segments, int64_t n_samples, int64_t seg_stride,
                     float* out, wrs_stream_t stream_) {
    using namespace wrs;
    if (!ctx) return fail(nullptr, WRS_ERR_INVALID, "wrs_logmel_batch: null ctx");
    if (n_segments == 0) return WRS_OK;
    if (!pcm || !out) return fail(ctx, WRS_ERR_INVALID, "wrs_logmel_batch: null buffer");
    if (n_segments < 0 || n_samples < WRS_N_FFT || seg_stride < n_samples)
        return fail(ctx, WRS_ERR_INVALID, "wrs_logmel_batch: need n_segments >= 0, n_samples >= 400 (reflect padding needs > 200), seg_stride >= n_samples");
    if (n_samples > (int64_t)1 << 30 || n_segments > (int64_t)1 << 24)
        return fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_logmel_batch: segment longer than 2^30 samples or more than 2^24 segments");
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
    if (total > INT32_MAX) return fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_logmel_batch: too many slices");
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
    if (need > ctx->seg_cap) {
        WRS_CUDA(ctx, cudaStreamSynchronize(stream));
        cudaFree(ctx->d_seg_state);
        ctx->d_seg_state = nullptr;
        const int64_t cap = std::max<int64_t>(need, 4096);
        WRS_CUDA(ctx, cudaMalloc(&ctx->d_seg_state, sizeof(int) * cap));
        ctx->seg_cap = cap;
    }
    p.seg_max = ctx->d_seg_state;
    p.slice_max = ctx->d_seg_state;
    // 0x80808080: kSlotEmpty for the per-slice words, and (decoding to about -3.4e38) "minus infinity" for atomicMax
    WRS_CUDA(ctx, cudaMemsetAsync(ctx->d_seg_state, 0x80, sizeof(int) * need, stream));

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

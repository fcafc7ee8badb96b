// K1: fused log-mel front-end for batches of audio segments (sm_100a).
//
// Replaces the ATen op sequence of reference src/audio.rs:230-259 (hann_window, stft, abs, pow, matmul,
// clamp_min, log10, max, maximum, add, div) with one persistent kernel:
//
//   PCM (HBM) --LDG.128--> smem, transposed [sample mod 160][frame row]     (conflict-free, lane = frame)
//   stage A  warp t : 32 frames x real 25-point DFT + W400^{tq}       -> smem Y[t][q][lane]
//   stage B  warp q : 32 frames x complex 16-point DFT, |X|^2         -> smem P[bin][lane]
//   stage C  warp w : 32 frames x 5 mel rows (banded FP32 FMA), log10 -> smem tile[mel][frame]
//   slice end       : block max -> atomicMax per segment; wait for the segment's other slices
//                     (co-resident CTAs, bounded spin); floor (max-8), (x+4)/4, float4 stores.
//
// Within a warp every lane works on a different frame and the same (t | q | mel): window values,
// twiddles and filter weights are warp-uniform (constant bank / smem broadcast) and no lane ever reads
// another lane's column of Y or P.
#include <algorithm>
#include <cmath>

#include "logmel_fft.cuh"
#include "wrs_internal.cuh"

namespace wrs {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kNFP = 37;            // padded row count of the transposed PCM tile (== 5 mod 8 -> conflict-free stores)
constexpr int kRows = 35;           // frame rows f0-2 .. f0+32 needed by 32 frames
constexpr int kStageTasks = 45;     // 9 row blocks x 5 column blocks of (4 rows x 8 float4)
constexpr int kStagePerWarp = 3;    // ceil(45 / 16)
constexpr float kLog10Of2 = 0.30102999566398120f;

__constant__ float c_win[WRS_N_FFT];        // periodic Hann, src/audio.rs:233
__constant__ int c_aoff[WRS_N_FFT];         // sample n of a frame -> offset in the transposed PCM tile
__constant__ cf c_tw[16 * 13];              // W400^{tq}

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
    int* seg_max;   // [n_segments] order-preserving int key of the per-segment max of log10(mel)
    int* seg_cnt;   // [n_segments] slices that have posted
    DevStatus* status;
};

template <int NG>
struct LogmelSmem {
    static constexpr int TW = NG * 32;
    float pcmT[WRS_HOP_LENGTH * kNFP];
    cf Y[16 * 13 * 32];
    float P[(WRS_N_BINS + 1) * 32];
    float tile[WRS_N_MELS * TW];
    MelBank bank;
    int red_max;
    float seg_floor;
};

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

__device__ __forceinline__ float4 load_stage4(const float* __restrict__ seg, int i0, int n, bool vec) {
    if (vec && i0 >= 0 && i0 + 3 < n) return __ldg(reinterpret_cast<const float4*>(seg + i0));
    return make_float4(load_reflect(seg, i0, n), load_reflect(seg, i0 + 1, n), load_reflect(seg, i0 + 2, n),
                       load_reflect(seg, i0 + 3, n));
}

template <int NG, bool FUSED>
__global__ void __launch_bounds__(kThreads, 1) logmel_kernel(const LogmelParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LogmelSmem<NG>& sm = *reinterpret_cast<LogmelSmem<NG>*>(smem_raw);
    constexpr int TW = LogmelSmem<NG>::TW;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // filterbank -> smem (once per CTA)
    {
        const int* src = reinterpret_cast<const int*>(p.bank);
        int* dst = reinterpret_cast<int*>(&sm.bank);
        for (int i = tid; i < (int)(sizeof(MelBank) / 4); i += kThreads) dst[i] = src[i];
    }
    // staging geometry of this lane: 4 rows x 8 float4 per warp task
    const int st_rr = lane >> 3, st_jj = lane & 7;

    for (int slice = blockIdx.x; slice < p.total_slices; slice += gridDim.x) {
        const int seg = slice / p.slices_per_seg;
        const int fs = (slice - seg * p.slices_per_seg) * TW;   // first frame of the slice
        const int nf = min(TW, p.n_frames - fs);
        const int ngroups = (nf + 31) >> 5;
        const float* __restrict__ segp = p.pcm + (int64_t)seg * p.seg_stride;
        float lmax = -INFINITY;
        if (tid == 0) sm.red_max = float_key(-INFINITY);

        float4 pre[kStagePerWarp];
        auto stage_load = [&](int f0) {
#pragma unroll
            for (int k = 0; k < kStagePerWarp; ++k) {
                const int task = warp + k * kWarps;
                const int r = (task / 5) * 4 + st_rr, j = (task % 5) * 8 + st_jj;
                if (task < kStageTasks && r < kRows)
                    pre[k] = load_stage4(segp, WRS_HOP_LENGTH * (f0 - 2 + r) + 4 * j, p.n_samples, p.vec_in != 0);
            }
        };
        auto stage_store = [&]() {
#pragma unroll
            for (int k = 0; k < kStagePerWarp; ++k) {
                const int task = warp + k * kWarps;
                const int r = (task / 5) * 4 + st_rr, j = (task % 5) * 8 + st_jj;
                if (task < kStageTasks && r < kRows) {
                    float* d = sm.pcmT + (4 * j) * kNFP + r;
                    d[0] = pre[k].x; d[kNFP] = pre[k].y; d[2 * kNFP] = pre[k].z; d[3 * kNFP] = pre[k].w;
                }
            }
        };

        stage_load(fs);
        stage_store();
        __syncthreads();

        for (int g = 0; g < ngroups; ++g) {
            const int f0 = fs + 32 * g;
            const bool more = (g + 1 < ngroups);
            if (more) stage_load(f0 + 32);

            // ---- stage A: warp = t
            {
                const int t = warp;
                float v[25];
#pragma unroll
                for (int m = 0; m < 25; ++m) {
                    const int n = 16 * m + t;
                    v[m] = sm.pcmT[c_aoff[n] + lane] * c_win[n];
                }
                cf y[13];
                rdft25(v, y);
                cf* dst = sm.Y + (t * 13) * 32 + lane;
                dst[0] = y[0];
#pragma unroll
                for (int q = 1; q < 13; ++q) dst[q * 32] = cmul(y[q], c_tw[t * 13 + q]);
            }
            __syncthreads();
            if (more) stage_store();

            // ---- stage B: warp = q
            if (warp < 13) {
                const int q = warp;
                cf z[16];
#pragma unroll
                for (int t = 0; t < 16; ++t) z[t] = sm.Y[(t * 13 + q) * 32 + lane];
                cdft16(z);
#pragma unroll
                for (int s = 0; s < 16; ++s) {
                    if (q != 0 || s < 9) sm.P[stageb_bin(q, s) * 32 + lane] = z[s].re * z[s].re + z[s].im * z[s].im;
                }
            }
            __syncthreads();

            // ---- stage C: banded mel projection (src/audio.rs:252), clamp + log10 (:254)
            {
                const bool valid = (f0 + lane) < p.n_frames;
#pragma unroll
                for (int i = 0; i < WRS_N_MELS / kWarps; ++i) {
                    const int m = i * kWarps + warp;
                    const int lo = sm.bank.lo[m], n = sm.bank.n[m];
                    const float* w = sm.bank.w + sm.bank.off[m];
                    const float* pp = sm.P + lo * 32 + lane;
                    float acc = 0.0f;
                    for (int j = 0; j < n; ++j) acc = fmaf(w[j], pp[j * 32], acc);
                    const float val = __log2f(fmaxf(acc, 1e-10f)) * kLog10Of2;
                    sm.tile[m * TW + g * 32 + lane] = val;
                    if (valid) lmax = fmaxf(lmax, val);
                }
            }
        }

        // ---- per-slice max -> per-segment max (src/audio.rs:256 `log_spec.max()`)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
        if (lane == 0) atomicMax(&sm.red_max, float_key(lmax));
        __syncthreads();
        if (tid == 0) {
            atomicMax(p.seg_max + seg, sm.red_max);
            if (FUSED) {
                __threadfence();
                atomicAdd(p.seg_cnt + seg, 1);
                // every slice of a segment is owned by a co-resident CTA (cooperative launch, slices_per_seg <=
                // gridDim.x, slices taken in index order) so this wait always ends; the bound is a safety net.
                volatile int* cnt = p.seg_cnt + seg;
                int spins = 0;
                while (*cnt < p.slices_per_seg) {
                    __nanosleep(64);
                    if (++spins > (1 << 22)) { atomicExch(&p.status->timeout, 1); break; }
                }
                __threadfence();
                sm.seg_floor = key_float(*(volatile int*)(p.seg_max + seg)) - 8.0f;
            }
        }
        __syncthreads();

        // ---- floor, scale (src/audio.rs:256-258) and coalesced store of the [80][nf] tile
        const float floor_v = FUSED ? sm.seg_floor : -INFINITY;
        float* __restrict__ outp = p.out + ((int64_t)seg * WRS_N_MELS) * p.n_frames + fs;
        if (p.vec_out && (nf & 3) == 0) {
            const int nf4 = nf >> 2;
            for (int idx = tid; idx < WRS_N_MELS * nf4; idx += kThreads) {
                const int m = idx / nf4, c4 = idx - m * nf4;
                float4 v = *reinterpret_cast<const float4*>(sm.tile + m * TW + 4 * c4);
                if (FUSED) {
                    v.x = (fmaxf(v.x, floor_v) + 4.0f) * 0.25f; v.y = (fmaxf(v.y, floor_v) + 4.0f) * 0.25f;
                    v.z = (fmaxf(v.z, floor_v) + 4.0f) * 0.25f; v.w = (fmaxf(v.w, floor_v) + 4.0f) * 0.25f;
                }
                *reinterpret_cast<float4*>(outp + (int64_t)m * p.n_frames + 4 * c4) = v;
            }
        } else {
            for (int idx = tid; idx < WRS_N_MELS * nf; idx += kThreads) {
                const int m = idx / nf, c = idx - m * nf;
                float v = sm.tile[m * TW + c];
                if (FUSED) v = (fmaxf(v, floor_v) + 4.0f) * 0.25f;
                outp[(int64_t)m * p.n_frames + c] = v;
            }
        }
        __syncthreads();
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
    float win[WRS_N_FFT];
    int aoff[WRS_N_FFT];
    cf tw[16 * 13];
    for (int n = 0; n < WRS_N_FFT; ++n) {
        win[n] = (float)(0.5 - 0.5 * std::cos(2.0 * PI * n / WRS_N_FFT));
        const int u = n + 120;   // frame f, sample n  ->  row f-2+u/160 of the tile, column u%160
        aoff[n] = (u % WRS_HOP_LENGTH) * kNFP + (u / WRS_HOP_LENGTH);
    }
    for (int t = 0; t < 16; ++t)
        for (int q = 0; q < 13; ++q) {
            const double a = -2.0 * PI * (double)(t * q) / 400.0;
            tw[t * 13 + q] = cf{(float)std::cos(a), (float)std::sin(a)};
        }
    WRS_CUDA(ctx, cudaMemcpyToSymbol(c_win, win, sizeof(win)));
    WRS_CUDA(ctx, cudaMemcpyToSymbol(c_aoff, aoff, sizeof(aoff)));
    WRS_CUDA(ctx, cudaMemcpyToSymbol(c_tw, tw, sizeof(tw)));
    if (ctx->device < 64) g_tables_ready[ctx->device] = true;
    return WRS_OK;
}

constexpr int kNG = 4;   // 128-frame slices

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
        return wrs::fail(ctx, WRS_ERR_UNSUPPORTED, "wrs_logmel_set_filter: filterbank is not banded (more than 1024 in-band weights)");
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

    constexpr int TW = kNG * 32;
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

    if (n_segments > ctx->seg_cap) {
        WRS_CUDA(ctx, cudaStreamSynchronize(stream));
        cudaFree(ctx->d_seg_state);
        ctx->d_seg_state = nullptr;
        const int64_t cap = std::max<int64_t>(n_segments, 256);
        WRS_CUDA(ctx, cudaMalloc(&ctx->d_seg_state, sizeof(int) * 2 * cap));
        ctx->seg_cap = cap;
    }
    p.seg_max = ctx->d_seg_state;
    p.seg_cnt = ctx->d_seg_state + ctx->seg_cap;
    // 0x80808080 decodes to about -3.4e38: "minus infinity" for the max; counters start at 0
    WRS_CUDA(ctx, cudaMemsetAsync(p.seg_max, 0x80, sizeof(int) * n_segments, stream));
    WRS_CUDA(ctx, cudaMemsetAsync(p.seg_cnt, 0, sizeof(int) * n_segments, stream));

    const size_t smem = sizeof(LogmelSmem<kNG>);
    if (ctx->logmel_max_ctas_per_sm == 0) {
        WRS_CUDA(ctx, cudaFuncSetAttribute(logmel_kernel<kNG, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        WRS_CUDA(ctx, cudaFuncSetAttribute(logmel_kernel<kNG, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        WRS_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, logmel_kernel<kNG, true>, kThreads, smem));
        if (occ < 1) return fail(ctx, WRS_ERR_CUDA, "logmel kernel does not fit on an SM");
        ctx->logmel_max_ctas_per_sm = occ;
    }
    const int resident = ctx->sm_count * ctx->logmel_max_ctas_per_sm;
    const int grid = (int)std::min<int64_t>(total, resident);
    bool fused = p.slices_per_seg <= grid;
    if (ctx->opt_logmel_fused == 0) fused = false;
    ctx->stat_last_logmel_mode = fused ? 1 : 0;
    ctx->stat_last_logmel_grid = grid;
    if (fused) {
        void* args[] = {(void*)&p};
        WRS_CUDA(ctx, cudaLaunchCooperativeKernel((void*)logmel_kernel<kNG, true>, dim3(grid), dim3(kThreads), args, smem, stream));
    } else {
        logmel_kernel<kNG, false><<<grid, kThreads, smem, stream>>>(p);
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
    cudaStream_t s = ctx->own_stream;
    WRS_CUDA(ctx, cudaMemcpyAsync(ctx->d_stage_in, pcm_host, in_bytes, cudaMemcpyHostToDevice, s));
    int rc = wrs_logmel_batch(ctx, ctx->d_stage_in, n_segments, n_samples, seg_stride, ctx->d_stage_out, s);
    if (rc != WRS_OK) return rc;
    WRS_CUDA(ctx, cudaMemcpyAsync(out_host, ctx->d_stage_out, out_bytes, cudaMemcpyDeviceToHost, s));
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

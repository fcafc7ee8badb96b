// K3: beam reorder of the decoder's cached keys / values (sm_100a).
//
// Replaces reference src/model.rs:53-62 (`Cached::update_cache`: cache.index(&[Some(sources_indices)]) = an
// index_select along dim 0) for EVERY cached tensor of the decoder -- self- and cross-attention K and V of all
// blocks (src/model.rs:208-211, 321-327, 524-528, 591-594), one ATen call each in the reference -- with one launch:
// dst[i][r] = src[i][source_idx[r]] for all tensors i.  source_idx is the device array wrs_beam_step writes
// (beam_search.rs:111-114), so the step never returns to the host in between.
//
// Pure HBM copy: every byte is read once and written once (algorithmic bytes = 2 x the caches' size).  A row is cut
// into pieces of 16 KB; the persistent grid walks the flat list of pieces, a CTA moves one piece with eight 16-byte
// loads in flight per thread before the first store.
#include <algorithm>

#include "wrs_internal.cuh"

namespace wrs {

constexpr int kKvThreads = 256;
constexpr int kKvUnroll = 8;
constexpr int kKvPiece = kKvThreads * kKvUnroll * 16;    // 32 KB per CTA iteration
constexpr int kKvMaxTensors = 96;                        // descriptors ride in the kernel parameters (~3 KB)

struct KvParams {
    const unsigned char* src[kKvMaxTensors];
    unsigned char* dst[kKvMaxTensors];
    long long row_bytes[kKvMaxTensors];
    long long piece_begin[kKvMaxTensors + 1];            // flat index of a tensor's first piece
    int pieces_per_row[kKvMaxTensors];
    int n_tensors, n_rows;
    const int* source_idx;
    DevStatus* status;
};

__global__ void __launch_bounds__(kKvThreads) kv_reorder_kernel(const __grid_constant__ KvParams p) {
    const long long total = p.piece_begin[p.n_tensors];
    for (long long w = blockIdx.x; w < total; w += gridDim.x) {
        // which tensor: the list is short, a linear scan over the prefix table is warp-uniform and cheap next to 32 KB
        int t = 0;
        while (t + 1 < p.n_tensors && p.piece_begin[t + 1] <= w) ++t;
        const long long local = w - p.piece_begin[t];
        const int ppr = p.pieces_per_row[t];
        const int row = (int)(local / ppr), piece = (int)(local - (long long)row * ppr);
        int srow = p.source_idx[row];
        if (srow < 0 || srow >= p.n_rows) {              // a bad index must not read outside the cache
            if (threadIdx.x == 0) atomicExch(&p.status->bad_index, 1);
            srow = row;
        }
        const long long rb = p.row_bytes[t], off = (long long)piece * kKvPiece;
        const long long n = min((long long)kKvPiece, rb - off);
        const unsigned char* s = p.src[t] + (long long)srow * rb + off;
        unsigned char* d = p.dst[t] + (long long)row * rb + off;
        if ((((uintptr_t)s | (uintptr_t)d) & 15) == 0) {
            const long long n16 = n >> 4;
            const int4* s4 = reinterpret_cast<const int4*>(s);
            int4* d4 = reinterpret_cast<int4*>(d);
            int4 v[kKvUnroll];
#pragma unroll
            for (int k = 0; k < kKvUnroll; ++k) {
                const long long i = threadIdx.x + (long long)k * kKvThreads;
                if (i < n16) v[k] = __ldcs(s4 + i);      // streaming: nothing here is read twice
            }
#pragma unroll
            for (int k = 0; k < kKvUnroll; ++k) {
                const long long i = threadIdx.x + (long long)k * kKvThreads;
                if (i < n16) __stcs(d4 + i, v[k]);
            }
            for (long long i = (n16 << 4) + threadIdx.x; i < n; i += kKvThreads) d[i] = s[i];
        } else {
            for (long long i = threadIdx.x; i < n; i += kKvThreads) d[i] = s[i];
        }
    }
}

// Debug check of a tensor flagged group-invariant (option kv_verify_group): every row equals the first row of its beam
// group, and every source index stays inside its row's group.
__global__ void __launch_bounds__(256) kv_verify_kernel(const unsigned char* __restrict__ t, long long row_bytes, int n_rows, int group,
                                                        const int* __restrict__ source_idx, DevStatus* status) {
    bool bad = false;
    for (int r = blockIdx.y; r < n_rows; r += gridDim.y) {
        const int lead = r - r % group;
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            const int s = source_idx[r];
            bad |= s < lead || s >= lead + group;
        }
        if (r == lead) continue;
        const unsigned char* a = t + (long long)r * row_bytes;
        const unsigned char* b = t + (long long)lead * row_bytes;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < row_bytes; i += (long long)gridDim.x * blockDim.x) bad |= a[i] != b[i];
    }
    if (bad) atomicExch(&status->group_violation, 1);
}

}  // namespace wrs

extern "C" int wrs_kv_reorder(wrs_ctx* ctx, const wrs_kv_tensor* tensors_host, int32_t n_tensors, int64_t n_rows,
                              const int32_t* source_idx, wrs_stream_t stream_) {
    using namespace wrs;
    if (!ctx) return fail(nullptr, WRS_ERR_INVALID, "wrs_kv_reorder: null ctx");
    if (n_tensors == 0 || n_rows == 0) return WRS_OK;
    if (!tensors_host || !source_idx || n_tensors < 0 || n_rows < 0 || n_rows > INT32_MAX)
        return fail(ctx, WRS_ERR_INVALID, "wrs_kv_reorder: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    for (int32_t i = 0; i < n_tensors; ++i) {
        const wrs_kv_tensor& t = tensors_host[i];
        if (t.row_bytes < 0 || (t.row_bytes > 0 && !t.src))
            return fail(ctx, WRS_ERR_INVALID, "wrs_kv_reorder: tensor with null pointer or negative row size");
        if (t.group_invariant) {                          // skipped: the gather would reproduce src bit for bit
            const int g = ctx->opt_kv_verify_group;
            if (g > 0 && t.row_bytes > 0) {
                if (n_rows % g != 0) return fail(ctx, WRS_ERR_INVALID, "wrs_kv_reorder: kv_verify_group does not divide n_rows");
                const dim3 grid((unsigned)std::min<int64_t>(64, (t.row_bytes + 255) / 256), (unsigned)std::min<int64_t>(n_rows, 1024));
                kv_verify_kernel<<<grid, 256, 0, stream>>>(static_cast<const unsigned char*>(t.src), t.row_bytes, (int)n_rows, g, source_idx, ctx->d_status);
                WRS_CUDA(ctx, cudaGetLastError());
            }
            continue;
        }
        if (t.row_bytes > 0 && !t.dst) return fail(ctx, WRS_ERR_INVALID, "wrs_kv_reorder: tensor with null dst");
        if (t.row_bytes > 0 && t.src == t.dst)
            return fail(ctx, WRS_ERR_INVALID, "wrs_kv_reorder: src and dst must be different buffers (the reference's index() allocates too)");
    }
    // the tensors that do move
    std::vector<wrs_kv_tensor> moving;
    moving.reserve(n_tensors);
    for (int32_t i = 0; i < n_tensors; ++i)
        if (!tensors_host[i].group_invariant) moving.push_back(tensors_host[i]);
    tensors_host = moving.data();
    n_tensors = (int32_t)moving.size();
    for (int32_t base = 0; base < n_tensors; base += kKvMaxTensors) {
        KvParams p;
        p.n_tensors = std::min<int32_t>(kKvMaxTensors, n_tensors - base);
        p.n_rows = (int)n_rows; p.source_idx = source_idx; p.status = ctx->d_status;
        long long total = 0;
        for (int i = 0; i < p.n_tensors; ++i) {
            const wrs_kv_tensor& t = tensors_host[base + i];
            p.src[i] = static_cast<const unsigned char*>(t.src); p.dst[i] = static_cast<unsigned char*>(t.dst);
            p.row_bytes[i] = t.row_bytes;
            p.pieces_per_row[i] = (int)std::max<int64_t>(1, (t.row_bytes + kKvPiece - 1) / kKvPiece);
            p.piece_begin[i] = total;
            total += t.row_bytes > 0 ? (long long)p.pieces_per_row[i] * n_rows : 0;
            if (t.row_bytes == 0) p.pieces_per_row[i] = 1;
        }
        p.piece_begin[p.n_tensors] = total;
        if (total == 0) continue;
        const int grid = (int)std::min<long long>(total, (long long)ctx->sm_count * 8);
        kv_reorder_kernel<<<grid, kKvThreads, 0, stream>>>(p);
        WRS_CUDA(ctx, cudaGetLastError());
    }
    return WRS_OK;
}

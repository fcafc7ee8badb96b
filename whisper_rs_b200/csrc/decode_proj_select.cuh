// The pure, per-thread part of the projection kernel's epilogue and geometry (decode_proj.cuh): vocab tile bounds, the
// compact row count of a short M tile, mask ranges, the branch-free insertion into a row's best-K list, the K-th largest
// of the twelve group maxima.  No CUDA intrinsics: this header also compiles for the host, where
// tests/host/proj_select_check.cpp checks it against brute force (tests/test_proj_select_host.py, no GPU).
#pragma once

#include <limits.h>
#include <math.h>

#ifdef __CUDACC__
#define WRS_PHD __host__ __device__ __forceinline__
#else
#define WRS_PHD inline
#endif

namespace wrs {

constexpr int kPM = 128;                                  // rows per M tile
constexpr int kPN = 176;                                  // ids per vocab tile (a multiple of 16, at most 256)

WRS_PHD float proj_inf() { return INFINITY; }

WRS_PHD int proj_tile_base(int tile, int n_text_tiles, int ts_split) {
    return tile < n_text_tiles ? tile * kPN : ts_split + (tile - n_text_tiles) * kPN;
}
WRS_PHD int proj_tile_end(int tile, int n_text_tiles, int ts_split, int n_vocab) {
    return tile < n_text_tiles ? (tile * kPN + kPN < ts_split ? tile * kPN + kPN : ts_split) : (ts_split + (tile - n_text_tiles + 1) * kPN < n_vocab ? ts_split + (tile - n_text_tiles + 1) * kPN : n_vocab);
}

// rows of M tile `mt` as the operands hold them: the real rows rounded up to 8 (a core matrix), at most 128.  The MMA
// always spans 128 rows; with fewer in memory its K chunks lie closer together (LBO = rows x 16 B) and the rows beyond them
// alias the next chunk: they only reach accumulator rows that belong to no row of the batch and are never read.
WRS_PHD int proj_rows8(int n_rows, int mt) {
    const int left = n_rows - mt * kPM;
    return left >= kPM ? kPM : ((left + 7) & ~7);
}

// ---- epilogue (EPI 1): what a thread keeps for its row within one vocab tile
template <int KM>
struct ProjRow {
    // the best so far, (value desc, id asc), right-aligned: the K live entries are v[KM - K .. KM) (-inf / INT32_MAX =
    // empty), in front of them +inf sentinels that nothing displaces -- so the K-th best is always v[KM - 1]
    float v[KM]; int id[KM];
    float kth;                                           // max(v[KM - 1], floor)
    float floor;                                         // nothing at or below it can be among the row's best K (see proj_kernel)
};

WRS_PHD unsigned range_bits(int lo, int hi) {   // bits i of a 32-bit word with lo <= i < hi
    lo = lo > 0 ? lo : 0; hi = hi < 32 ? hi : 32;
    if (hi <= lo) return 0u;
    return (0xffffffffu << lo) & (hi >= 32 ? 0xffffffffu : (1u << hi) - 1u);
}

// ids arrive in ascending order, so a new entry goes behind every entry of equal value: strict comparisons only.
// Branch-free: every slot picks among {its own entry, its left neighbour's, the new one} from two comparisons against the
// OLD list, so the whole insertion is one layer of selects (a value that does not beat the K-th best changes nothing).
template <int KM>
WRS_PHD void proj_insert(ProjRow<KM>& st, float x, int nid) {
    x = x > st.kth ? x : -proj_inf();                  // (also keeps out what a carried floor excludes)
    bool gt[KM];
#pragma unroll
    for (int e = 0; e < KM; ++e) gt[e] = x > st.v[e];
#pragma unroll
    for (int e = KM - 1; e >= 1; --e) {
        st.v[e] = gt[e - 1] ? st.v[e - 1] : (gt[e] ? x : st.v[e]);
        st.id[e] = gt[e - 1] ? st.id[e - 1] : (gt[e] ? nid : st.id[e]);
    }
    st.v[0] = gt[0] ? x : st.v[0];
    st.id[0] = gt[0] ? nid : st.id[0];
    st.kth = (st.v[KM - 1] > st.floor ? st.v[KM - 1] : st.floor);
}

// The K-th largest of 12 values (K <= 12), by rank counting: 66 independent comparisons, ties ranked by position.
WRS_PHD float kth_of_12(const float (&h)[12], int K) {
    int rk[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) rk[i] = 0;
#pragma unroll
    for (int i = 0; i < 12; ++i)
#pragma unroll
        for (int j = i + 1; j < 12; ++j) {
            const bool jw = h[j] > h[i];                 // j ranks before i only if strictly greater
            rk[i] += jw ? 1 : 0; rk[j] += jw ? 0 : 1;
        }
    float t = -proj_inf();
#pragma unroll
    for (int i = 0; i < 12; ++i) t = rk[i] == K - 1 ? h[i] : t;
    return t;
}


}  // namespace wrs

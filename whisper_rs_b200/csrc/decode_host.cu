// Host-side epilogue of a decode window: BeamSearchTokenExtractor::finalize and MaximumLikelihoodRanker::rank.
// O(beam) work per audio (reference src/decode/extractor/beam_search.rs:151-206,
// src/decode/sequencer/max_likelihood.rs:10-35); nothing here is data-parallel.
#include <algorithm>
#include <cmath>
#include <vector>

#include "wrs_internal.cuh"

// layout of wrs_decoder is private to decode.cu; finalize reaches it through these accessors
namespace wrs {
struct DecoderView {
    wrs_ctx* ctx; int beam, eot, max_len, max_cand, max_audio, max_rows;
    const int64_t* d_fin_tokens; const int32_t* d_fin_len; const double* d_fin_score; const int32_t* d_fin_n;
};
DecoderView decoder_view(const wrs_decoder* d);
}  // namespace wrs

extern "C" {

int wrs_beam_finalize(wrs_decoder* dec, const int64_t* tokens, int64_t tok_stride, int64_t cur_len, int64_t n_rows,
                      const float* sum_logprobs, int32_t n_cand_cap, int64_t* out_tokens, int32_t* out_len,
                      double* out_scores, int32_t* out_n_cand, wrs_stream_t stream_) {
    using namespace wrs;
    if (!dec) return fail(nullptr, WRS_ERR_INVALID, "wrs_beam_finalize: null decoder");
    const DecoderView v = decoder_view(dec);
    wrs_ctx* ctx = v.ctx;
    if (!tokens || !sum_logprobs || !out_tokens || !out_len || !out_scores || !out_n_cand)
        return fail(ctx, WRS_ERR_INVALID, "wrs_beam_finalize: null buffer");
    if (n_rows <= 0 || n_rows > v.max_rows || n_rows % v.beam != 0 || cur_len < 1 || tok_stride < cur_len || cur_len + 1 > v.max_len)
        return fail(ctx, WRS_ERR_INVALID, "wrs_beam_finalize: bad sizes");
    const int n_audio = (int)(n_rows / v.beam), k = v.beam;
    if (n_cand_cap < std::max(k, v.max_cand)) return fail(ctx, WRS_ERR_INVALID, "wrs_beam_finalize: n_cand_cap < max(beam, round(patience*beam))");
    cudaStream_t stream = (cudaStream_t)stream_;
    WRS_CUDA(ctx, cudaSetDevice(ctx->device));
    std::vector<int64_t> h_tok((size_t)n_rows * cur_len), h_fin((size_t)n_audio * v.max_cand * v.max_len);
    std::vector<float> h_sum(n_rows);
    std::vector<int32_t> h_fin_len((size_t)n_audio * v.max_cand), h_fin_n(n_audio);
    std::vector<double> h_fin_score((size_t)n_audio * v.max_cand);
    WRS_CUDA(ctx, cudaMemcpy2DAsync(h_tok.data(), cur_len * 8, tokens, tok_stride * 8, cur_len * 8, n_rows, cudaMemcpyDeviceToHost, stream));
    WRS_CUDA(ctx, cudaMemcpyAsync(h_sum.data(), sum_logprobs, n_rows * 4, cudaMemcpyDeviceToHost, stream));
    WRS_CUDA(ctx, cudaMemcpyAsync(h_fin.data(), v.d_fin_tokens, h_fin.size() * 8, cudaMemcpyDeviceToHost, stream));
    WRS_CUDA(ctx, cudaMemcpyAsync(h_fin_len.data(), v.d_fin_len, h_fin_len.size() * 4, cudaMemcpyDeviceToHost, stream));
    WRS_CUDA(ctx, cudaMemcpyAsync(h_fin_score.data(), v.d_fin_score, h_fin_score.size() * 8, cudaMemcpyDeviceToHost, stream));
    WRS_CUDA(ctx, cudaMemcpyAsync(h_fin_n.data(), v.d_fin_n, n_audio * 4, cudaMemcpyDeviceToHost, stream));
    WRS_CUDA(ctx, cudaStreamSynchronize(stream));

    struct Cand { std::vector<int64_t> seq; double score; };
    for (int a = 0; a < n_audio; ++a) {
        std::vector<Cand> cands;
        for (int c = 0; c < h_fin_n[a]; ++c) {
            const size_t slot = (size_t)a * v.max_cand + c;
            const int64_t* s = h_fin.data() + slot * v.max_len;
            cands.push_back(Cand{std::vector<int64_t>(s, s + h_fin_len[slot]), h_fin_score[slot]});
        }
        if ((int)cands.size() < k) {   // beam_search.rs:162-178: top up with live beams, best sum_logprobs first
            std::vector<int> ord(k);
            for (int j = 0; j < k; ++j) ord[j] = j;
            std::stable_sort(ord.begin(), ord.end(), [&](int x, int y) { return h_sum[a * k + x] > h_sum[a * k + y]; });
            for (int j : ord) {
                const int64_t* s = h_tok.data() + (size_t)(a * k + j) * cur_len;
                Cand cd{std::vector<int64_t>(s, s + cur_len), (double)h_sum[a * k + j]};
                cd.seq.push_back(v.eot);
                bool dup = false;   // HashMap insert of an existing key only overwrites the score
                for (auto& e : cands) if (e.seq == cd.seq) { e.score = cd.score; dup = true; break; }
                if (!dup) cands.push_back(std::move(cd));
                if ((int)cands.size() >= k) break;
            }
        }
        std::stable_sort(cands.begin(), cands.end(), [](const Cand& x, const Cand& y) {
            return x.score != y.score ? x.score > y.score : x.seq < y.seq;   // canonical order
        });
        if ((int)cands.size() > n_cand_cap) return fail(ctx, WRS_ERR_INVALID, "wrs_beam_finalize: n_cand_cap too small");
        out_n_cand[a] = (int32_t)cands.size();
        for (int c = 0; c < n_cand_cap; ++c) {
            int64_t* dst = out_tokens + ((size_t)a * n_cand_cap + c) * v.max_len;
            std::fill(dst, dst + v.max_len, (int64_t)v.eot);
            out_len[a * n_cand_cap + c] = 0;
            out_scores[a * n_cand_cap + c] = -INFINITY;
            if (c < (int)cands.size()) {
                std::copy(cands[c].seq.begin(), cands[c].seq.end(), dst);
                out_len[a * n_cand_cap + c] = (int32_t)cands[c].seq.size();
                out_scores[a * n_cand_cap + c] = cands[c].score;
            }
        }
    }
    return WRS_OK;
}

int wrs_rank_max_likelihood_host(const double* sum_logprobs, const int32_t* lengths, int32_t n_audio, int32_t n_cand,
                                 int32_t has_penalty, double alpha, int32_t* selected) {
    if (!sum_logprobs || !lengths || !selected || n_audio < 0 || n_cand < 1)
        return wrs::fail(nullptr, WRS_ERR_INVALID, "wrs_rank_max_likelihood_host: bad arguments");
    for (int a = 0; a < n_audio; ++a) {
        int best = 0; double best_score = 0.0; bool have = false;
        for (int c = 0; c < n_cand; ++c) {
            const double len = (double)lengths[a * n_cand + c];
            const double penalty = has_penalty ? std::pow((5.0 + len) / 6.0, alpha) : len;   // max_likelihood.rs:22-26
            const double score = sum_logprobs[a * n_cand + c] / penalty;                     // len 0 -> inf/NaN as in Rust
            // ATen argmax semantics: first maximum; a NaN counts as the maximum
            if (!have || (score > best_score && !(best_score != best_score)) || (score != score && !(best_score != best_score))) {
                best = c; best_score = score; have = true;
            }
        }
        selected[a] = best;
    }
    return WRS_OK;
}

}  // extern "C"

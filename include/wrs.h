/* wrs.h -- C ABI of the B200-native hot paths of laptou/whisper-rs.
 *
 * The reference has no FFI of its own: its operator surface is Rust (free functions in src/audio.rs and
 * the LogitFilter / TokenExtractor / SequenceRanker traits in src/decode/).  Every entry point below
 * names the reference item it replaces (path:line relative to the reference root); INTEGRATION.md shows
 * the Rust `extern "C"` block + trait impls that bind them.
 *
 * Conventions: plain pointers and sizes; *device* pointers unless a name ends in `_host`; `stream` is a
 * cudaStream_t passed as void*; calls enqueue work and return without synchronising unless documented;
 * no per-call allocation once the context has grown to the problem size; return 0 (WRS_OK) or an error
 * code, with text in wrs_last_error().  There is NO CPU fallback: without a CUDA device every compute call
 * fails with WRS_ERR_CUDA.
 */
#ifndef WRS_H
#define WRS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WRS_ABI_VERSION 2

enum {
    WRS_OK = 0,
    WRS_ERR_INVALID = 1,      /* bad argument (what a tch shape/arg panic is in the reference) */
    WRS_ERR_CUDA = 2,         /* CUDA runtime error, or no device */
    WRS_ERR_UNSUPPORTED = 3,  /* valid in the reference, outside what the kernels cover */
    WRS_ERR_NUMERIC = 4,      /* NaN met where the reference panics (beam_search.rs:86 partial_cmp().unwrap()) */
    WRS_ERR_TIMEOUT = 5       /* bounded device-side wait expired (fused log-mel max exchange) */
};

/* audio hyper-parameters, src/audio.rs:12-18 */
#define WRS_SAMPLE_RATE 16000
#define WRS_N_FFT 400
#define WRS_N_MELS 80
#define WRS_HOP_LENGTH 160
#define WRS_N_SAMPLES 480000
#define WRS_N_FRAMES 3000
#define WRS_N_BINS 201

typedef struct wrs_ctx wrs_ctx;
typedef void* wrs_stream_t;

int wrs_abi_version(void);
/* One context per device/thread (the reference's tasks are !Send, SURVEY.md §8b "Threading"). */
int wrs_ctx_create(int device, wrs_ctx** out);
void wrs_ctx_destroy(wrs_ctx* ctx);
/* Text of the last error on this context (ctx may be NULL: last error of the calling thread). */
const char* wrs_last_error(const wrs_ctx* ctx);
/* Host-side check that a device-side bounded wait or NaN check tripped since the last call; synchronises
 * `stream`.  Returns WRS_OK / WRS_ERR_TIMEOUT / WRS_ERR_NUMERIC. */
int wrs_ctx_check(wrs_ctx* ctx, wrs_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Log-mel front-end
 * ---------------------------------------------------------------------------------------------- */

/* audio::mel_filter()  src/audio.rs:175-188 -- writes the 80x201 slaney filterbank to HOST memory. */
int wrs_mel_filter_host(float* out_80x201_host);

/* Select the filterbank used by wrs_logmel_*: HOST pointer to [80,201] row-major f32, or NULL for the
 * built-in one.  Mirrors the `mel_filter: &Tensor` argument of log_mel_spectrogram (src/audio.rs:230). */
int wrs_logmel_set_filter(wrs_ctx* ctx, const float* filter_80x201_host);

/* audio::log_mel_spectrogram  src/audio.rs:230-259, batched per segment:
 *   pcm  [n_segments][seg_stride] f32, the first n_samples of each row are used (n_samples >= 400)
 *   out  [n_segments][80][n_samples/160] f32
 * Each segment gets its own reflect padding and its own max-8 floor, i.e. the reference function applied
 * to every row.  n_segments = 1 with a long clip is the reference's whole-clip call (src/transcribe.rs:112). */
int wrs_logmel_batch(wrs_ctx* ctx, const float* pcm, int64_t n_segments, int64_t n_samples, int64_t seg_stride,
                     float* out, wrs_stream_t stream);

/* wrs_logmel_batch for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same kernels, same results; the per-slice max words (the kernel's only scratch) live in
 * a fixed 256 KB (65 536 words) static device array instead of the context's growable buffer, so the call makes no
 * cudaMalloc / cudaFree / allocating library call.  A batch needs one word per 64-frame slice (47 per 30 s segment:
 * up to 1394 segments), or one word per segment for clips longer than one round of the grid; more than that is
 * rejected with WRS_ERR_UNSUPPORTED.  Every call on a device shares the array (it keeps the last call's words; a
 * call clears what it uses, on `stream`): calls that overlap on the GPU race -- serialise them on one stream. */
int wrs_logmel_batch_noalloc(wrs_ctx* ctx, const float* pcm, int64_t n_segments, int64_t n_samples,
                             int64_t seg_stride, float* out, wrs_stream_t stream);

/* Same call for a CPU-resident caller (the reference runs on Device::Cpu, examples/transcribe.rs:19):
 * HOST buffers; H2D, kernel, D2H and a stream synchronise happen inside. */
int wrs_logmel_batch_host(wrs_ctx* ctx, const float* pcm_host, int64_t n_segments, int64_t n_samples,
                          int64_t seg_stride, float* out_host);

/* The step in front of the front-end, as far as the reference pins it in its own tree (src/audio.rs:116-131, SURVEY.md
 * 8f row 4): 16-bit PCM frames (`channels` interleaved samples each) -> mono f32, device pointers.  Each sample becomes
 * s / 32768 (symphonia's SampleBuffer<f32> conversion, exact in f32); several channels are averaged per frame, summed left
 * to right in f32 and divided by the count (`c.iter().sum::<f32>() / c.len() as f32`, :127-131).  The resampler behind it
 * (rubato::FftFixedIn, :73-86, 133-166) is a third-party crate absent from the reference tree: not restated. */
int wrs_pcm_i16_to_f32(wrs_ctx* ctx, const int16_t* pcm_i16, int64_t n_frames, int32_t channels, float* out,
                       wrs_stream_t stream);

/* wrs_logmel_batch_host for a caller that holds the decoder's 16-bit mono PCM (the reference converts to f32 on the host,
 * src/audio.rs:116-122): HOST int16 in, host log-mel out; the conversion above runs on the device, so the copy in moves
 * half the bytes.  Results equal wrs_logmel_batch_host fed the converted samples, bit for bit. */
int wrs_logmel_batch_host_i16(wrs_ctx* ctx, const int16_t* pcm_host, int64_t n_segments, int64_t n_samples,
                              int64_t seg_stride, float* out_host);

/* audio::pad_or_trim  src/audio.rs:213-228 on the last dimension of a [rows][in_len] f32 array:
 * out [rows][out_len] = in[:, :out_len] or in right-padded with 0.0 (src/transcribe.rs:152 pads MEL frames). */
int wrs_pad_or_trim(wrs_ctx* ctx, const float* in, int64_t rows, int64_t in_len, int64_t in_stride, float* out,
                    int64_t out_len, wrs_stream_t stream);

/* Tuning / diagnostics.  key "logmel_fused": 1 = single fused launch (default when a segment's slices
 * fit co-resident), 0 = two launches (raw log-mel, then floor+scale).  key "proj_cluster": 0 = never launch the projection
 * kernel as clusters of two CTAs that share the weights (the default, 1, does for 129 .. 256 rows; A/B timing). */
int wrs_ctx_set_option(wrs_ctx* ctx, const char* key, int64_t value);
int64_t wrs_ctx_get_stat(wrs_ctx* ctx, const char* key);

/* ------------------------------------------------------------------------------------------------
 * Decode step: LogitFilter (src/decode/filter/), TokenExtractor (src/decode/extractor/),
 * SequenceRanker (src/decode/sequencer/)
 * ---------------------------------------------------------------------------------------------- */

typedef struct {
    int32_t n_vocab;          /* 51864 (*.en) / 51865 (multilingual) */
    int32_t eot;              /* Tokenizer::token_id_eot            src/tokenize.rs:122 */
    int32_t sot;              /* Tokenizer::token_id_sot            src/tokenize.rs:116 */
    int32_t no_timestamps;    /* Tokenizer::token_id_no_timestamps  src/tokenize.rs:124 */
    int32_t ts_begin;         /* Tokenizer::token_id_ts_begin       src/tokenize.rs:130 */
    int32_t no_speech;        /* Tokenizer::token_id_no_speech      src/tokenize.rs:126 */
    int32_t blank_ids[2];     /* {encode(" ")[0], eot}              src/decode/filter/suppress_blanks.rs:21-33 */
    int32_t suppress_blank;   /* DecodeOptions::suppress_blank      src/decode/mod.rs:158 */
    int32_t timestamps;       /* DecodeOptions::timestamps          src/decode/mod.rs:176 */
    int32_t max_initial_timestamp_index; /* -1 = None; round(mit / (30 / n_audio_ctx)) src/decode/mod.rs:178-182 */
    int32_t beam_size;        /* TokenExtractMode::BeamSearch{beam_size} or Greedy{group_sze}; 1..15 */
    float patience;           /* TokenExtractMode::BeamSearch{patience} */
    int32_t max_len;          /* capacity of a token row: n_text_ctx + 2 */
} wrs_decode_cfg;

/* bits of row_flags[] -- which branch of src/decode/filter/timestamp_tokens.rs fired for a row */
#define WRS_ROW_TS_MASKED 1u         /* last and second-last sampled tokens were timestamps  (:54-59) */
#define WRS_ROW_TEXT_LT_EOT_MASKED 2u/* last was a timestamp, second-last was not             (:60-63) */
#define WRS_ROW_FIRST_STEP 4u        /* tokens.size()[1] == sample_begin                      (:68-78) */
#define WRS_ROW_TS_OVER_TEXT 8u      /* logsumexp(timestamps) > max(text): text masked        (:83-96) */
#define WRS_ROW_NAN 128u             /* a NaN reached the softmax (reference panics) */

#define WRS_FILTER_SUPPRESS_BLANK 1   /* src/decode/filter/suppress_blanks.rs:38-44 */
#define WRS_FILTER_SUPPRESS_TOKENS 2  /* src/decode/filter/suppress_tokens.rs:20-24 */
#define WRS_FILTER_TIMESTAMPS 4       /* src/decode/filter/timestamp_tokens.rs:33-98 */
#define WRS_FILTER_ALL 7

typedef struct wrs_decoder wrs_decoder;

/* DecodeTask::new  src/decode/mod.rs:122-204 (the part that builds filters/extractor):
 * suppress_ids_host = DecodeOptions::suppress_tokens ++ tokenizer.non_speech_tokens() (duplicates allowed),
 * n_suppress = 0 when neither option is set.  max_rows = n_audio * beam_size upper bound. */
int wrs_decoder_create(wrs_ctx* ctx, const wrs_decode_cfg* cfg, const int32_t* suppress_ids_host, int32_t n_suppress,
                       int32_t max_rows, wrs_decoder** out);
void wrs_decoder_destroy(wrs_decoder* dec);

/* TokenExtractor::reset  src/decode/extractor/mod.rs:22, beam_search.rs:18-20 */
int wrs_decoder_reset(wrs_decoder* dec, wrs_stream_t stream);

/* LogitFilter::apply(&self, logits: &mut Tensor, tokens: &Tensor)  src/decode/filter/mod.rs:12-14, in place.
 * `which` is a bit-set of WRS_FILTER_*; set bits are applied in the reference's order Blank -> Suppress ->
 * Timestamp (src/decode/mod.rs:156-190).  logits [n_rows][row_stride] f32 (first n_vocab of each row),
 * tokens [n_rows][tok_stride] i64 (first cur_len of each row), sample_begin = DecodePrompt::sample_begin.
 * row_flags (optional, [n_rows] u8) receives WRS_ROW_* when WRS_FILTER_TIMESTAMPS is applied. */
int wrs_filter_apply(wrs_decoder* dec, int32_t which, float* logits, int64_t row_stride, int64_t n_rows,
                     const int64_t* tokens, int64_t tok_stride, int64_t cur_len, int64_t sample_begin,
                     uint8_t* row_flags, wrs_stream_t stream);

/* Optional per-step diagnostics (device pointers, any may be NULL) */
typedef struct {
    float* cand_logprob;    /* [n_rows][beam+1] log_softmax values of the per-row top-(beam+1), sorted */
    int32_t* cand_token;    /* [n_rows][beam+1] their token ids, ties lowest id first */
    uint8_t* row_flags;     /* [n_rows] WRS_ROW_* */
    int32_t* fin_count;     /* [n_audio]        candidates that ended in EOT this step (<= beam) */
    int32_t* fin_source;    /* [n_audio][beam]  their source rows, descending score */
    double* fin_score;      /* [n_audio][beam]  their f64 scores */
} wrs_step_debug;

/* One fused pass = for f in logit_filters { f.apply(..) } + BeamSearchTokenExtractor::update
 * (src/decode/mod.rs:257-270, src/decode/extractor/beam_search.rs:22-149) without a host round trip:
 *   logits read once, never written; tokens/sum_logprobs_in read; outputs
 *   next_tokens [n_rows][next_stride] i64 = selected prefix ++ token (first cur_len+1 entries),
 *   source_idx  [n_rows] i32            = rows of `tokens` the new rows continue (-> Whisper::update_cache),
 *   sum_logprobs_out [n_rows] f32       (may alias sum_logprobs_in),
 *   completed   [1] i32                 = 1 when every audio holds round(patience*beam) finished sequences.
 * Finished (EOT-terminated) candidates are appended to the decoder's device-side store.
 * Ties are canonical: lowest token id, then lowest source row. */
int wrs_beam_step(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, const int64_t* tokens,
                  int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_logprobs_in,
                  int64_t* next_tokens, int64_t next_stride, int32_t* source_idx, float* sum_logprobs_out,
                  int32_t* completed, const wrs_step_debug* dbg, wrs_stream_t stream);

/* Filters + the intent of GreedyTokenExtractor::update (src/decode/extractor/greedy.rs:24-50; dead code in
 * the reference): argmax (lowest id on ties), log_softmax gather, masked accumulate,
 * EOT propagation, append.  completed = every row's new last token is EOT. */
int wrs_greedy_step(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, const int64_t* tokens,
                    int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_logprobs_in,
                    int64_t* next_tokens, int64_t next_stride, float* sum_logprobs_out, int32_t* completed,
                    const wrs_step_debug* dbg, wrs_stream_t stream);

/* ---- the logits projection with the decode step as its epilogue.
 * Reference: `x.matmul(&self.token_emb.ws.transpose(0, 1))` (src/model.rs:521), whose [n_rows, n_vocab] f32 result
 * is what src/decode/mod.rs:257-270 then filters and extracts from.  Here the product runs on the tensor cores
 * (tcgen05, f16 hi/lo splits of both f32 operands -- 22 bits each --, three products, tensor-core f32 accumulation) and
 * the decode step's row scan is the epilogue of each tile: the logits never reach HBM.  Accuracy: measured up to 1.2e-6
 * of sum |x_i w_i| per logit (4e-5 absolute on logits of sigma 3 at n_state 512): the tensor core's accumulation is not
 * IEEE f32, so this is four to seven times the error of an f32 matmul -- NOT f32-faithful (tests/test_gpu_project.py
 * states the bars).  |x| must stay below 65504 (f16 range; LayerNorm output does).
 *
 * wrs_projection_create packs `token_emb` (DEVICE [n_vocab][emb_stride] f32, the rows of TextDecoder::token_emb,
 * src/model.rs:463) for `dec`'s vocabulary geometry; n_state a multiple of 32.  The packed copy (4 bytes per weight)
 * belongs to the projection; token_emb may be freed once the call returns (it synchronises `stream`: one reduction over
 * the embedding is read back for the scale).  A projection owns per-step scratch (packed x, per-tile records): calls on one
 * projection must be ordered on one stream, like calls on its decoder. */
typedef struct wrs_projection wrs_projection;
int wrs_projection_create(wrs_decoder* dec, const float* token_emb, int64_t emb_stride, int32_t n_state, wrs_stream_t stream,
                          wrs_projection** out);
void wrs_projection_destroy(wrs_projection* proj);

/* The projection alone: logits[r][v] = sum_i x[r][i] token_emb[v][i]  (src/model.rs:521).  x DEVICE [n_rows][x_stride]
 * f32 (the output of TextDecoder::ln, src/model.rs:519), 16-byte aligned, x_stride a multiple of 4. */
int wrs_project_logits(wrs_projection* proj, const float* x, int64_t x_stride, int64_t n_rows, float* logits,
                       int64_t row_stride, wrs_stream_t stream);

/* src/model.rs:521 + wrs_beam_step in one pass: same outputs and semantics as wrs_beam_step on the projected logits. */
int wrs_project_beam_step(wrs_projection* proj, const float* x, int64_t x_stride, int64_t n_rows, const int64_t* tokens,
                          int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_logprobs_in,
                          int64_t* next_tokens, int64_t next_stride, int32_t* source_idx, float* sum_logprobs_out,
                          int32_t* completed, const wrs_step_debug* dbg, wrs_stream_t stream);

/* src/model.rs:521 + wrs_greedy_step in one pass (TokenExtractMode::Greedy, src/decode/extractor/greedy.rs:14-51). */
int wrs_project_greedy_step(wrs_projection* proj, const float* x, int64_t x_stride, int64_t n_rows, const int64_t* tokens,
                            int64_t tok_stride, int64_t cur_len, int64_t sample_begin, const float* sum_logprobs_in,
                            int64_t* next_tokens, int64_t next_stride, float* sum_logprobs_out, int32_t* completed,
                            const wrs_step_debug* dbg, wrs_stream_t stream);

/* BeamSearchTokenExtractor::finalize  src/decode/extractor/beam_search.rs:151-206.  Synchronises `stream`,
 * copies the finished store and the live beams to the host and returns, per audio, exactly
 * max(beam, stored) candidates ordered (score desc, sequence lexicographic):
 *   out_tokens_host [n_audio][n_cand_cap][max_len] i64, out_len_host [n_audio][n_cand_cap] i32 (length incl. EOT),
 *   out_scores_host [n_audio][n_cand_cap] f64, out_n_cand_host [n_audio] i32.
 * WRS_ERR_INVALID if n_cand_cap is too small. */
int wrs_beam_finalize(wrs_decoder* dec, const int64_t* tokens, int64_t tok_stride, int64_t cur_len, int64_t n_rows,
                      const float* sum_logprobs, int32_t n_cand_cap, int64_t* out_tokens_host, int32_t* out_len_host,
                      double* out_scores_host, int32_t* out_n_cand_host, wrs_stream_t stream);

/* MaximumLikelihoodRanker::rank  src/decode/sequencer/max_likelihood.rs:10-35 (pure host, O(n_cand)):
 * lengths = tokens between sample_begin and the first EOT; has_penalty=0 -> logprob/len, else GNMT
 * ((5+len)/6)^alpha.  selected_host [n_audio] = argmax (first maximum). */
int wrs_rank_max_likelihood_host(const double* sum_logprobs_host, const int32_t* lengths_host, int32_t n_audio,
                                 int32_t n_cand, int32_t has_penalty, double alpha, int32_t* selected_host);

/* Step-0 no-speech probability  src/decode/mod.rs:250-255:
 *   probs_at_sot = logits.i((.., sot_idx)).softmax(-1); no_speech_probs = probs_at_sot[.., token_id_no_speech]
 * `logits` is the UNFILTERED row at sot_idx (the last position of the first forward pass: sot is the last initial token,
 * src/decode/mod.rs:46-66), [n_rows][row_stride] f32 on the device; out_probs [n_rows] f32 on the device. */
int wrs_no_speech_probs(wrs_decoder* dec, const float* logits, int64_t row_stride, int64_t n_rows, float* out_probs,
                        wrs_stream_t stream);

/* avg_logprob of the selected candidates  src/decode/mod.rs:389-397 (pure host): sum_logprob / (n_tokens + 1). */
int wrs_avg_logprob_host(const double* sum_logprobs_host, const int32_t* n_tokens_host, int32_t n, double* out_host);

/* ------------------------------------------------------------------------------------------------
 * KV-cache beam reorder: Cached::update_cache  src/model.rs:53-62  (cache.index(&[Some(sources_indices)]))
 * applied to every cached tensor of the decoder (src/model.rs:208-211, 321-327, 524-528, 591-594) in one launch.
 * tensors_host: n_tensors descriptors of device arrays [n_rows][row_bytes] with contiguous rows;
 * dst[i][r] = src[i][source_idx[r]], source_idx a DEVICE array [n_rows] (the one wrs_beam_step writes).
 * src != dst (ping-pong the two buffers between steps).  An index outside [0, n_rows) copies the row's own
 * source and is reported by wrs_ctx_check (WRS_ERR_INVALID).
 *
 * group_invariant != 0 marks a tensor whose rows are identical within every beam group -- the cross-attention K / V,
 * projections of audio_features.repeat_interleave(k) (src/decode/mod.rs:311-313).  A beam step's source_idx never
 * leaves a row's group (beam_search.rs:91-114), so gathering such a tensor is a bit-exact no-op (the reference does it
 * anyway, src/model.rs:321-327): it is skipped, `dst` is ignored (may be NULL or == src) and the caller keeps using
 * `src`.  Option "kv_verify_group" = k > 0 (wrs_ctx_set_option) checks on the device, before the gather, that every
 * flagged tensor really has identical rows within groups of k and that source_idx stays inside them; a violation is
 * reported by wrs_ctx_check (WRS_ERR_INVALID).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    const void* src;
    void* dst;
    int64_t row_bytes;
    int32_t group_invariant;
    int32_t reserved;
} wrs_kv_tensor;

int wrs_kv_reorder(wrs_ctx* ctx, const wrs_kv_tensor* tensors_host, int32_t n_tensors, int64_t n_rows,
                   const int32_t* source_idx, wrs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* WRS_H */

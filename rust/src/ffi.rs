// UNVERIFIED: written without a Rust toolchain (none exists in the build image); kept in step with INTEGRATION.md
// by tools/sync_rust.py.  See INTEGRATION.md for where each piece goes in the reference tree.
// UNVERIFIED source (no rustc here).  One-to-one with include/wrs.h.
use std::os::raw::{c_char, c_int, c_void};
#[repr(C)] pub struct WrsCtx { _p: [u8; 0] }
#[repr(C)] pub struct WrsDecoder { _p: [u8; 0] }
#[repr(C)] pub struct WrsProjection { _p: [u8; 0] }
#[repr(C)] #[derive(Clone, Copy)]
pub struct WrsDecodeCfg {
    pub n_vocab: i32, pub eot: i32, pub sot: i32, pub no_timestamps: i32, pub ts_begin: i32, pub no_speech: i32,
    pub blank_ids: [i32; 2], pub suppress_blank: i32, pub timestamps: i32, pub max_initial_timestamp_index: i32,
    pub beam_size: i32, pub patience: f32, pub max_len: i32,
}
extern "C" {
    pub fn wrs_ctx_create(device: c_int, out: *mut *mut WrsCtx) -> c_int;
    pub fn wrs_ctx_destroy(ctx: *mut WrsCtx);
    pub fn wrs_last_error(ctx: *const WrsCtx) -> *const c_char;
    pub fn wrs_logmel_set_filter(ctx: *mut WrsCtx, filter_host: *const f32) -> c_int;
    pub fn wrs_logmel_batch(ctx: *mut WrsCtx, pcm: *const f32, n_segments: i64, n_samples: i64, seg_stride: i64,
                            out: *mut f32, stream: *mut c_void) -> c_int;
    pub fn wrs_logmel_batch_noalloc(ctx: *mut WrsCtx, pcm: *const f32, n_segments: i64, n_samples: i64, seg_stride: i64,
                                    out: *mut f32, stream: *mut c_void) -> c_int;
    pub fn wrs_logmel_batch_host(ctx: *mut WrsCtx, pcm_host: *const f32, n_segments: i64, n_samples: i64, seg_stride: i64,
                                 out_host: *mut f32) -> c_int;
    pub fn wrs_logmel_batch_host_i16(ctx: *mut WrsCtx, pcm_host: *const i16, n_segments: i64, n_samples: i64,
                                     seg_stride: i64, out_host: *mut f32) -> c_int;
    pub fn wrs_pcm_i16_to_f32(ctx: *mut WrsCtx, pcm_i16: *const i16, n_frames: i64, channels: i32, out: *mut f32,
                              stream: *mut c_void) -> c_int;
    pub fn wrs_pad_or_trim(ctx: *mut WrsCtx, inp: *const f32, rows: i64, in_len: i64, in_stride: i64,
                           out: *mut f32, out_len: i64, stream: *mut c_void) -> c_int;
    pub fn wrs_decoder_create(ctx: *mut WrsCtx, cfg: *const WrsDecodeCfg, suppress: *const i32, n_suppress: i32,
                              max_rows: i32, out: *mut *mut WrsDecoder) -> c_int;
    pub fn wrs_decoder_destroy(dec: *mut WrsDecoder);
    pub fn wrs_decoder_reset(dec: *mut WrsDecoder, stream: *mut c_void) -> c_int;
    pub fn wrs_filter_apply(dec: *mut WrsDecoder, which: i32, logits: *mut f32, row_stride: i64, n_rows: i64,
                            tokens: *const i64, tok_stride: i64, cur_len: i64, sample_begin: i64,
                            row_flags: *mut u8, stream: *mut c_void) -> c_int;
    pub fn wrs_beam_step(dec: *mut WrsDecoder, logits: *const f32, row_stride: i64, n_rows: i64, tokens: *const i64,
                         tok_stride: i64, cur_len: i64, sample_begin: i64, sum_in: *const f32, next_tokens: *mut i64,
                         next_stride: i64, source_idx: *mut i32, sum_out: *mut f32, completed: *mut i32,
                         dbg: *const c_void, stream: *mut c_void) -> c_int;
    pub fn wrs_projection_create(dec: *mut WrsDecoder, token_emb: *const f32, emb_stride: i64, n_state: i32,
                                 stream: *mut c_void, out: *mut *mut WrsProjection) -> c_int;
    pub fn wrs_projection_destroy(proj: *mut WrsProjection);
    pub fn wrs_project_logits(proj: *mut WrsProjection, x: *const f32, x_stride: i64, n_rows: i64, logits: *mut f32,
                              row_stride: i64, stream: *mut c_void) -> c_int;
    pub fn wrs_project_beam_step(proj: *mut WrsProjection, x: *const f32, x_stride: i64, n_rows: i64, tokens: *const i64,
                                 tok_stride: i64, cur_len: i64, sample_begin: i64, sum_in: *const f32,
                                 next_tokens: *mut i64, next_stride: i64, source_idx: *mut i32, sum_out: *mut f32,
                                 completed: *mut i32, dbg: *const c_void, stream: *mut c_void) -> c_int;
    pub fn wrs_project_greedy_step(proj: *mut WrsProjection, x: *const f32, x_stride: i64, n_rows: i64, tokens: *const i64,
                                   tok_stride: i64, cur_len: i64, sample_begin: i64, sum_in: *const f32,
                                   next_tokens: *mut i64, next_stride: i64, sum_out: *mut f32, completed: *mut i32,
                                   dbg: *const c_void, stream: *mut c_void) -> c_int;
    pub fn wrs_beam_finalize(dec: *mut WrsDecoder, tokens: *const i64, tok_stride: i64, cur_len: i64, n_rows: i64,
                             sum_logprobs: *const f32, n_cand_cap: i32, out_tokens: *mut i64, out_len: *mut i32,
                             out_scores: *mut f64, out_n_cand: *mut i32, stream: *mut c_void) -> c_int;
    pub fn wrs_rank_max_likelihood_host(sum: *const f64, lens: *const i32, n_audio: i32, n_cand: i32,
                                        has_penalty: i32, alpha: f64, selected: *mut i32) -> c_int;
    pub fn wrs_no_speech_probs(dec: *mut WrsDecoder, logits: *const f32, row_stride: i64, n_rows: i64,
                               out_probs: *mut f32, stream: *mut c_void) -> c_int;
    pub fn wrs_avg_logprob_host(sum: *const f64, n_tokens: *const i32, n: i32, out: *mut f64) -> c_int;
    pub fn wrs_kv_reorder(ctx: *mut WrsCtx, tensors: *const WrsKvTensor, n_tensors: i32, n_rows: i64,
                          source_idx: *const i32, stream: *mut c_void) -> c_int;
}
#[repr(C)] pub struct WrsKvTensor { pub src: *const c_void, pub dst: *mut c_void, pub row_bytes: i64,
                                    pub group_invariant: i32, pub reserved: i32 }
/// The reference panics on ATen errors (tch non-`f_` methods unwrap); keep that behaviour.
pub fn check(ctx: *const WrsCtx, rc: c_int) {
    if rc != 0 {
        let msg = unsafe { std::ffi::CStr::from_ptr(wrs_last_error(ctx)) }.to_string_lossy().into_owned();
        panic!("wrs_b200 error {rc}: {msg}");
    }
}

// UNVERIFIED: written without a Rust toolchain (none exists in the build image); kept in step with INTEGRATION.md
// by tools/sync_rust.py.  See INTEGRATION.md for where each piece goes in the reference tree.
// Replacements with the reference's own signatures: audio::log_mel_spectrogram, the LogitFilter / TokenExtractor /
// SequenceRanker impls, Whisper::update_cache.
// src/audio.rs -- same signature as :230.  UNVERIFIED.
pub fn log_mel_spectrogram(audio: &Tensor, mel_filter: &Tensor) -> Tensor {
    assert!(audio.device().is_cuda() && audio.kind() == Kind::Float && audio.dim() == 1);
    let ctx = crate::ffi::thread_ctx(audio.device());                 // one WrsCtx per thread+device
    let filt = Vec::<f32>::from(&mel_filter.to_device(Device::Cpu).contiguous().view(-1));
    unsafe { check(ctx, wrs_logmel_set_filter(ctx, filt.as_ptr())) };
    let n = audio.size()[0];
    let audio = audio.contiguous();
    let out = Tensor::empty(&[N_MELS, n / HOP_LENGTH], (Kind::Float, audio.device()));
    unsafe { check(ctx, wrs_logmel_batch(ctx, audio.data_ptr() as *const f32, 1, n, n,
                                         out.data_ptr() as *mut f32, current_cuda_stream())) };
    out
}

// src/decode/filter/timestamp_tokens.rs -- same trait, same struct name.  UNVERIFIED.
impl LogitFilter for TimestampTokens {
    fn apply(&self, logits: &mut Tensor, tokens: &Tensor) {
        let (r, l) = (logits.size()[0], tokens.size()[1]);
        let tokens = tokens.contiguous();
        unsafe { check(self.ctx, wrs_filter_apply(self.dec, 4 /*WRS_FILTER_TIMESTAMPS*/,
            logits.data_ptr() as *mut f32, logits.stride()[0], r, tokens.data_ptr() as *const i64,
            tokens.stride()[0], l, self.prompt.borrow().sample_begin as i64, std::ptr::null_mut(),
            current_cuda_stream())) };
    }
}
// SuppressBlank (which = 1) and SuppressTokens (which = 2) are the same three lines.

// src/decode/extractor/beam_search.rs -- TokenExtractor::update.  UNVERIFIED.
fn update(&mut self, model: &Whisper, tokens: Tensor, logits: Tensor, sum_logprobs: &mut Tensor) -> (Tensor, bool) {
    let (r, l) = (tokens.size()[0], tokens.size()[1]);
    let next = Tensor::empty(&[r, l + 1], (Kind::Int64, tokens.device()));
    let src = Tensor::empty(&[r], (Kind::Int, tokens.device()));
    let done = Tensor::zeros(&[1], (Kind::Int, tokens.device()));
    unsafe { check(self.ctx, wrs_beam_step(self.dec, logits.data_ptr() as *const f32, logits.stride()[0], r,
        tokens.data_ptr() as *const i64, tokens.stride()[0], l, self.sample_begin,
        sum_logprobs.data_ptr() as *const f32, next.data_ptr() as *mut i64, l + 1, src.data_ptr() as *mut i32,
        sum_logprobs.data_ptr() as *mut f32, done.data_ptr() as *mut i32, std::ptr::null(), current_cuda_stream())) };
    model.update_cache(&src.to_kind(Kind::Int64));      // beam_search.rs:111-114, stays on the device
    (next, i64::from(&done) != 0)                        // the only host read per step, as in the reference loop
}

// model.rs:519-522: let TextDecoder::forward_hidden return `self.ln.forward(&x)` ([rows, n_ctx, n_state]) and keep the
// matmul for callers that want logits.  Once per model (the embedding does not change):
//     wrs_projection_create(dec, token_emb.ws.data_ptr(), token_emb.ws.stride()[0], n_state, stream, &mut proj)
// decode/mod.rs:257-270: instead of `logits = model.logits(..).i((.., -1))` + filters + extractor.update
pub fn update_from_hidden(&mut self, tokens: &Tensor, hidden_last: &Tensor /* [rows, n_state] f32, contiguous */,
                          sum_logprobs: &mut Tensor) -> (Tensor, bool) {
    let (r, l) = (tokens.size()[0], tokens.size()[1]);
    let next = Tensor::empty(&[r, l + 1], (Kind::Int64, tokens.device()));
    let done = Tensor::zeros(&[1], (Kind::Int, tokens.device()));
    unsafe { check(self.ctx, wrs_project_beam_step(self.proj, hidden_last.data_ptr() as *const f32, hidden_last.stride()[0], r,
        tokens.data_ptr() as *const i64, tokens.stride()[0], l, self.sample_begin, sum_logprobs.data_ptr() as *const f32,
        next.data_ptr() as *mut i64, l + 1, self.source_idx.data_ptr() as *mut i32, sum_logprobs.data_ptr() as *mut f32,
        done.data_ptr() as *mut i32, std::ptr::null(), stream())); }
    (next, i64::from(&done) != 0)
}

// model.rs: instead of walking the blocks and calling cache.index() per tensor (24 ATen launches for base.en),
// collect every `Cached<T>::cache` once and reorder them together; `spare` holds the second buffer set.
pub fn update_cache(&self, source_idx_dev: *const i32, rows: i64) {
    let cur: Vec<&Tensor> = self.all_caches();                 // self/cross K and V of every block, in a fixed order
    // cross-attention K / V are projections of audio_features.repeat_interleave(k) (decode/mod.rs:311-313): identical rows
    // within a beam group, so their gather is a no-op -- flagged, the library skips them and they are not swapped
    let descs: Vec<WrsKvTensor> = cur.iter().zip(self.spare.iter()).zip(self.is_cross.iter()).map(|((c, s), &x)| WrsKvTensor {
        src: c.data_ptr(), dst: s.data_ptr(), row_bytes: (c.numel() as i64 / rows) * c.kind().elt_size_in_bytes() as i64,
        group_invariant: x as i32, reserved: 0,
    }).collect();
    check(self.ctx, unsafe { wrs_kv_reorder(self.ctx, descs.as_ptr(), descs.len() as i32, rows, source_idx_dev, stream()) });
    self.swap_with_spare_except_cross();                       // Cell::set(Some(spare_i)) per moved Cached, old buffers become spare
}

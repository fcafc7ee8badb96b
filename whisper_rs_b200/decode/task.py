"""Mirror of the reference's `DecodeTask` (src/decode/mod.rs:122-418): the caller of the per-step hot path.

The model forward stays whatever the caller brings (the reference's tch forward is untouched): `model` is duck-typed,

    model.encoder(mel [n_audio, 80, 3000])                    -> audio_features [n_audio, n_audio_ctx, d]
    model.decoder.forward_ext(tokens, audio_features, step)   -> logits [rows, T, V]  (CUDA f32)
    model.update_cache(source_indices) / model.clear_cache()  (src/model.rs:591-600)
    model.dims.n_text_ctxs / model.dims.n_audio_ctxs

and everything between "logits come out of the forward" and "tokens go back in" runs in the CUDA library: the
step-0 no-speech probability, the three filters, log-softmax, top-(k+1), the per-audio beam merge and -- through
`model.update_cache` -- the KV-cache reorder, with one 4-byte host read per step (`completed`, which the reference
reads too, decode/mod.rs:273).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import torch

from ..constants import N_FRAMES, TokenIds
from .core import DecodeOptions, DecodePrompt, avg_logprobs
from .extractor import BeamSearchTokenExtractor, FusedDecodeStep, GreedyTokenExtractor
from .filter import build_filters
from .sequencer import MaximumLikelihoodRanker


@dataclass
class DecodeOutput:
    """src/decode/mod.rs:108-119"""
    audio_features: torch.Tensor
    tokens: List[int]
    text: Optional[str]
    avg_logprob: float
    no_speech_prob: float


class DecodeTask:
    """src/decode/mod.rs:24-37,122-204.

    fused=True (default) runs `for f in logit_filters {f.apply}; token_extractor.update(..)` as the fused step
    (FusedDecodeStep); fused=False keeps the reference's trait-level sequence -- three in-place filters, then the
    extractor -- on the same kernels.  `detokenize` stands in for `Tokenizer::decode` (text is not on the path).
    """

    def __init__(self, model, ids: TokenIds, options: DecodeOptions, *, prompt: Optional[Sequence[int]] = None,
                 max_audio: int = 32, fused: bool = True, detokenize: Optional[Callable[[List[int]], str]] = None,
                 device: int = 0):
        self.model, self.ids, self.options, self.device = model, ids, options, device
        dims = model.dims
        if options.n_text_ctxs != dims.n_text_ctxs or options.n_audio_ctxs != dims.n_audio_ctxs:
            options = DecodeOptions(**{**options.__dict__, "n_text_ctxs": dims.n_text_ctxs, "n_audio_ctxs": dims.n_audio_ctxs})
            self.options = options
        self.sample_len = options.sample_len if options.sample_len is not None else dims.n_text_ctxs // 2   # :129-131
        self.prompt = [DecodePrompt(prompt, ids)]          # one-element list = Rc<RefCell<DecodePrompt>> (:150-154)
        mode = options.token_extract_mode
        self.n_group = mode.group_sze if mode.kind == "greedy" else mode.beam_size
        self.max_rows = max_audio * self.n_group
        self.fused = fused
        self.sequence_ranker = MaximumLikelihoodRanker(options.len_penalty)
        self.detokenize = detokenize
        if fused:
            self.step = FusedDecodeStep(options, ids, self.prompt, self.max_rows, device)
            self.logit_filters, self.token_extractor = [], None
            self._decoder = self.step.decoder
        else:
            self.step = None
            self.logit_filters = build_filters(options, ids, self.prompt, self.max_rows, device)
            if mode.kind == "greedy":
                self.token_extractor = GreedyTokenExtractor(ids, mode.group_sze, self.max_rows, dims.n_text_ctxs, device)
            else:
                self.token_extractor = BeamSearchTokenExtractor(ids, mode.beam_size, mode.patience, self.max_rows,
                                                                dims.n_text_ctxs, device)
            self._decoder = self.token_extractor.decoder

    def set_prompt(self, prompt: Optional[Sequence[int]]):
        """src/decode/mod.rs:206-218 (an empty prompt counts as none)."""
        self.prompt[0] = DecodePrompt(prompt if prompt is not None and len(prompt) > 0 else None, self.ids)

    # ---- src/decode/mod.rs:220-282
    def main_loop(self, audio_features: torch.Tensor, tokens: torch.Tensor):
        n_batch = tokens.shape[0]
        assert audio_features.shape[0] == n_batch
        sum_logprobs = torch.zeros(n_batch, dtype=torch.float32, device=tokens.device)
        no_speech_probs = torch.full((n_batch,), float("nan"), dtype=torch.float32, device=tokens.device)
        prompt = self.prompt[0]
        for i in range(self.sample_len):
            t_in = tokens[:, -1:] if tokens.shape[-1] > prompt.sample_begin else tokens    # :237-240
            logits = self.model.decoder.forward_ext(t_in, audio_features, i)
            if i == 0:                                                                      # :250-255
                no_speech_probs = self._decoder.no_speech_probs(logits[:, prompt.sot_idx])
            last = logits[:, -1]                                                            # strided view, never copied
            if self.fused:
                out = self.step.step(tokens, last, sum_logprobs)
                tokens, sum_logprobs = out["next_tokens"], out["sum_logprobs"]
                if "source_idx" in out and hasattr(self.model, "update_cache"):
                    self.model.update_cache(out["source_idx"])                              # beam_search.rs:111-114
                self._decoder.ctx.check(self._decoder._stream())
                completed = bool(out["completed"].item())
            else:
                for f in self.logit_filters:                                                # :261-264
                    f.apply(last, tokens)
                tokens, completed = self.token_extractor.update(self.model, tokens, last, sum_logprobs)
            if completed or tokens.shape[-1] > self.model.dims.n_text_ctxs:                 # :273-275
                break
        self.model.clear_cache()
        return tokens, sum_logprobs, no_speech_probs

    # ---- src/decode/mod.rs:284-418
    @torch.no_grad()
    def run(self, mel_audio: torch.Tensor) -> List[DecodeOutput]:
        if mel_audio.dim() == 2:
            mel_audio = mel_audio.unsqueeze(0)
        n_audio, _, n_frames = mel_audio.shape
        assert n_frames == N_FRAMES
        if n_audio * self.n_group > self.max_rows:
            raise ValueError(f"{n_audio} audios x group {self.n_group} exceed max_audio given at construction")
        audio_features = self.model.encoder(mel_audio)
        prompt = self.prompt[0]
        init = torch.tensor(prompt.initial_tokens, dtype=torch.int64, device=mel_audio.device)
        n_group = self.n_group
        feats = audio_features.repeat_interleave(n_group, 0)
        tokens = init.repeat(n_audio, 1).repeat_interleave(n_group, 0).contiguous()
        self.model.clear_cache()
        (self.step if self.fused else self.token_extractor).reset()
        tokens, sum_logprobs, no_speech = self.main_loop(feats, tokens)
        no_speech = no_speech[::n_group].double().cpu().numpy()
        tokens = tokens.reshape(n_audio, n_group, -1)
        sum_logprobs = sum_logprobs.reshape(n_audio, n_group)
        cands, scores = (self.step if self.fused else self.token_extractor).finalize(tokens, sum_logprobs)
        eot, sb = self.ids.eot, prompt.sample_begin
        sliced = [[t[sb:t.index(eot)] for t in group] for group in cands]                  # :356-370 (first EOT)
        selected = self.sequence_ranker.rank(sliced, scores)
        chosen = [group[s] for group, s in zip(sliced, selected)]
        sums = [float(scores[a][s]) for a, s in enumerate(selected)]
        avg = avg_logprobs(sums, [len(t) for t in chosen])                                  # :389-397
        return [DecodeOutput(audio_features[a], chosen[a], self.detokenize(chosen[a]) if self.detokenize else None,
                             float(avg[a]), float(no_speech[a])) for a in range(n_audio)]

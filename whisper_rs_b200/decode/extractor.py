"""Mirror of reference src/decode/extractor/ plus the fused single-launch step."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ..constants import TokenIds
from .core import DecodeOptions, DecodePrompt, Decoder, TokenExtractMode, build_decoder_cfg, suppress_list


class TokenExtractor:
    """src/decode/extractor/mod.rs:20-73.  `model` is whatever owns the KV caches; if it has
    `update_cache(source_indices)` it is called like beam_search.rs:114 does."""

    def reset(self):
        raise NotImplementedError

    def update(self, model, tokens, logits, sum_logprobs):
        raise NotImplementedError

    def finalize(self, tokens, sum_logprobs):
        raise NotImplementedError

    def group_size(self) -> int:
        raise NotImplementedError


def _plain_cfg(ids: TokenIds, mode: TokenExtractMode, n_text_ctxs: int):
    opt = DecodeOptions(token_extract_mode=mode, timestamps=False, suppress_blank=False, suppress_non_speech=False,
                        n_text_ctxs=n_text_ctxs)
    return build_decoder_cfg(opt, ids, filters=False)


class BeamSearchTokenExtractor(TokenExtractor):
    """src/decode/extractor/beam_search.rs:9-211.  `update` takes logits the filters have already masked
    in place (trait-level drop-in); the whole of its host loop runs in one device pass."""

    def __init__(self, ids: TokenIds, beam_size: int, patience: float = 1.0, max_rows: int = 0, n_text_ctxs: int = 448,
                 device: int = 0):
        self.token_id_eot = ids.eot
        self.beam_size = beam_size
        self.patience = patience
        self.decoder = Decoder(_plain_cfg(ids, TokenExtractMode.BeamSearch(beam_size, patience), n_text_ctxs), [],
                               max_rows or beam_size * 64, device)
        self.last_source_indices: Optional[torch.Tensor] = None
        self.last: Optional[dict] = None

    def reset(self):
        self.decoder.reset()

    def group_size(self) -> int:
        return self.beam_size

    def update(self, model, tokens: torch.Tensor, logits: torch.Tensor, sum_logprobs: torch.Tensor, debug: bool = False
               ) -> Tuple[torch.Tensor, bool]:
        # sample_begin is irrelevant without filters; any value in 1..L is accepted
        out = self.decoder.beam_step(logits, tokens, 1, sum_logprobs, sum_out=sum_logprobs, debug=debug)
        self.last_source_indices = out["source_idx"]
        self.last = out
        if model is not None and hasattr(model, "update_cache"):
            model.update_cache(out["source_idx"])
        self.decoder.ctx.check(self.decoder._stream())        # NaN => the reference panics (beam_search.rs:86)
        return out["next_tokens"], bool(out["completed"].item())

    def finalize(self, tokens: torch.Tensor, sum_logprobs: torch.Tensor):
        """tokens [n_audio, k, L], sum_logprobs [n_audio, k] -> (candidates, scores [n_audio, n_cand])."""
        n_audio, k, L = tokens.shape
        return self.decoder.finalize(tokens.reshape(n_audio * k, L).contiguous(), sum_logprobs)


class GreedyTokenExtractor(TokenExtractor):
    """Intent of src/decode/extractor/greedy.rs:8-66 (dead code in the reference)."""

    def __init__(self, ids: TokenIds, group_size: int = 1, max_rows: int = 64, n_text_ctxs: int = 448, device: int = 0):
        self.token_id_eot = ids.eot
        self._group_size = group_size
        self.decoder = Decoder(_plain_cfg(ids, TokenExtractMode.Greedy(group_size), n_text_ctxs), [], max_rows, device)
        self.last: Optional[dict] = None

    def reset(self):
        pass

    def group_size(self) -> int:
        return self._group_size

    def update(self, model, tokens, logits, sum_logprobs, debug: bool = False):
        out = self.decoder.greedy_step(logits, tokens, 1, sum_logprobs, sum_out=sum_logprobs, debug=debug)
        self.last = out
        self.decoder.ctx.check(self.decoder._stream())
        return out["next_tokens"], bool(out["completed"].item())

    def finalize(self, tokens: torch.Tensor, sum_logprobs: torch.Tensor):
        eot = self.token_id_eot
        return [[seq.tolist() + [eot] for seq in group] for group in tokens.cpu()], sum_logprobs.double().cpu().numpy()


class FusedDecodeStep:
    """One launch pair for `for f in logit_filters { f.apply } ; token_extractor.update(..)`
    (src/decode/mod.rs:257-270): logits are read once and never written, tokens stay on the device."""

    def __init__(self, opt: DecodeOptions, ids: TokenIds, prompt_ref: List[DecodePrompt], max_rows: int, device: int = 0):
        self.opt, self.ids, self.prompt = opt, ids, prompt_ref
        self.greedy = opt.token_extract_mode.kind == "greedy"
        self.decoder = Decoder(build_decoder_cfg(opt, ids, filters=True), suppress_list(opt, ids), max_rows, device)

    def reset(self):
        self.decoder.reset()

    def group_size(self) -> int:
        m = self.opt.token_extract_mode
        return m.group_sze if self.greedy else m.beam_size

    def step(self, tokens: torch.Tensor, logits: torch.Tensor, sum_logprobs: torch.Tensor, debug: bool = False) -> dict:
        fn = self.decoder.greedy_step if self.greedy else self.decoder.beam_step
        return fn(logits, tokens, self.prompt[0].sample_begin, sum_logprobs, debug=debug)

    def finalize(self, tokens: torch.Tensor, sum_logprobs: torch.Tensor):
        n_audio, k, L = tokens.shape
        if self.greedy:
            # the decoder knows rows, not groups (its beam size is 1): group here, as GreedyTokenExtractor.finalize does
            # (src/decode/extractor/greedy.rs:52-65: every sequence of the group, EOT appended)
            eot = self.ids.eot
            cands = [[seq.tolist() + [eot] for seq in group] for group in tokens.cpu()]
            return cands, sum_logprobs.reshape(n_audio, k).double().cpu().numpy()
        return self.decoder.finalize(tokens.reshape(n_audio * k, L).contiguous(), sum_logprobs)

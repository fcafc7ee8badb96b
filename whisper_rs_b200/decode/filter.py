"""Mirror of reference src/decode/filter/: `LogitFilter.apply(logits, tokens)` mutates logits in place."""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from .. import _lib
from ..constants import TokenIds
from .core import DecodeOptions, DecodePrompt, Decoder, build_decoder_cfg, max_initial_timestamp_index, suppress_list


class LogitFilter:
    """src/decode/filter/mod.rs:12-14"""

    def apply(self, logits: torch.Tensor, tokens: torch.Tensor) -> None:   # pragma: no cover - interface
        raise NotImplementedError


class _DecoderBacked(LogitFilter):
    which = 0

    def __init__(self, decoder: Decoder, prompt_ref: List[DecodePrompt]):
        self.decoder = decoder
        self.prompt = prompt_ref    # one-element list standing in for Rc<RefCell<DecodePrompt>> (decode/mod.rs:150)
        self.last_row_flags: Optional[torch.Tensor] = None

    def apply(self, logits: torch.Tensor, tokens: torch.Tensor) -> None:
        self.decoder.filter_apply(self.which, logits, tokens, self.prompt[0].sample_begin)


class SuppressTokens(_DecoderBacked):
    """src/decode/filter/suppress_tokens.rs:6-24"""
    which = _lib.FILTER_SUPPRESS_TOKENS


class SuppressBlank(_DecoderBacked):
    """src/decode/filter/suppress_blanks.rs:10-44"""
    which = _lib.FILTER_SUPPRESS_BLANK


class TimestampTokens(_DecoderBacked):
    """src/decode/filter/timestamp_tokens.rs:10-98"""
    which = _lib.FILTER_TIMESTAMPS

    def apply(self, logits: torch.Tensor, tokens: torch.Tensor) -> None:
        flags = torch.zeros(tokens.shape[0], dtype=torch.uint8, device=logits.device)
        self.decoder.filter_apply(self.which, logits, tokens, self.prompt[0].sample_begin, flags)
        self.last_row_flags = flags


def build_filters(opt: DecodeOptions, ids: TokenIds, prompt_ref: List[DecodePrompt], max_rows: int,
                  device: int = 0) -> List[LogitFilter]:
    """src/decode/mod.rs:156-190 -- Blank, then Suppress, then Timestamp; all share one Decoder."""
    cfg = build_decoder_cfg(opt, ids, filters=True)
    dec = Decoder(cfg, suppress_list(opt, ids), max_rows, device)
    out: List[LogitFilter] = []
    if opt.suppress_blank:
        out.append(SuppressBlank(dec, prompt_ref))
    if opt.suppress_tokens is not None or opt.suppress_non_speech:
        out.append(SuppressTokens(dec, prompt_ref))
    if opt.timestamps:
        out.append(TimestampTokens(dec, prompt_ref))
    return out

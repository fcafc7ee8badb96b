"""Mirror of reference src/decode/sequencer/."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from .. import _lib


class SequenceRanker:
    """src/decode/sequencer/mod.rs:8-12"""

    def rank(self, tokens, sum_logprobs) -> List[int]:   # pragma: no cover - interface
        raise NotImplementedError


class MaximumLikelihoodRanker(SequenceRanker):
    """src/decode/sequencer/max_likelihood.rs:6-35: logprob / len, or the GNMT penalty ((5+len)/6)^alpha."""

    def __init__(self, length_penalty: Optional[float] = None):
        self.length_penalty = length_penalty

    def rank(self, tokens: Sequence[Sequence[Sequence[int]]], sum_logprobs) -> List[int]:
        n_audio = len(tokens)
        if n_audio == 0:
            return []
        n_cand = len(tokens[0])
        lengths = np.asarray([[len(t) for t in group] for group in tokens], dtype=np.int32)
        scores = np.ascontiguousarray(np.asarray(sum_logprobs, dtype=np.float64).reshape(n_audio, n_cand))
        sel = np.empty(n_audio, dtype=np.int32)
        _lib.check(_lib.lib().wrs_rank_max_likelihood_host(
            scores.ctypes.data_as(C.c_void_p), lengths.ctypes.data_as(C.c_void_p), n_audio, n_cand,
            0 if self.length_penalty is None else 1, float(self.length_penalty or 0.0), sel.ctypes.data_as(C.c_void_p)))
        return [int(i) for i in sel]

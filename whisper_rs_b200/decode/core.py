"""Options, prompt state and the low-level `Decoder` handle (wrs_decoder) shared by filters and extractors."""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _lib
from ..constants import CHUNK_LENGTH, TokenIds


@dataclass(frozen=True)
class TokenExtractMode:
    """src/decode/extractor/mod.rs:12-16.  kind = "beam" -> BeamSearch{beam_size, patience};
    kind = "greedy" -> Greedy{group_sze} (sic)."""
    kind: str = "beam"
    beam_size: int = 5
    patience: float = 1.0
    group_sze: int = 1

    @staticmethod
    def BeamSearch(beam_size: int, patience: float = 1.0) -> "TokenExtractMode":
        return TokenExtractMode("beam", beam_size, patience)

    @staticmethod
    def Greedy(group_sze: int = 1) -> "TokenExtractMode":
        return TokenExtractMode("greedy", 1, 1.0, group_sze)


@dataclass
class DecodeOptions:
    """src/decode/mod.rs:95-106 (fields that shape the hot path) + the two ModelDims it reads."""
    token_extract_mode: TokenExtractMode = TokenExtractMode()
    sample_len: Optional[int] = None
    len_penalty: Optional[float] = None
    max_initial_timestamp: Optional[float] = 1.0
    timestamps: bool = True
    suppress_blank: bool = True
    suppress_non_speech: bool = True
    suppress_tokens: Optional[List[int]] = None
    n_audio_ctxs: int = 1500
    n_text_ctxs: int = 448


class DecodePrompt:
    """src/decode/mod.rs:40-92: initial_tokens = [startofprev, prompt.., sot] or [sot]."""

    def __init__(self, prompt: Optional[Sequence[int]], ids: TokenIds):
        seq_sot = list(ids.sequence_sot())
        if prompt is not None and len(prompt) > 0:
            self.initial_tokens = [ids.startofprev, *[int(t) for t in prompt], *seq_sot]
        else:
            self.initial_tokens = seq_sot
        self.sample_begin = len(self.initial_tokens)
        self.sot_idx = self.initial_tokens.index(ids.sot)


def max_initial_timestamp_index(opt: DecodeOptions) -> Optional[int]:
    """src/decode/mod.rs:176-182 (f64::round = half away from zero)."""
    if opt.max_initial_timestamp is None:
        return None
    v = opt.max_initial_timestamp / (CHUNK_LENGTH / opt.n_audio_ctxs)
    return int(math.copysign(math.floor(abs(v) + 0.5), v))


def suppress_list(opt: DecodeOptions, ids: TokenIds) -> List[int]:
    """src/decode/mod.rs:166-174"""
    if opt.suppress_tokens is None and not opt.suppress_non_speech:
        return []
    lst = list(opt.suppress_tokens or [])
    if opt.suppress_non_speech:
        lst.extend(ids.non_speech_tokens)
    return lst


def build_decoder_cfg(opt: DecodeOptions, ids: TokenIds, *, filters: bool = True) -> _lib.DecodeCfg:
    """wrs_decode_cfg from options.  filters=False builds the plain extractor configuration (no masks):
    what BeamSearchTokenExtractor::update sees after the filters already ran in place."""
    mode = opt.token_extract_mode
    mit = max_initial_timestamp_index(opt)
    cfg = _lib.DecodeCfg()
    cfg.n_vocab, cfg.eot, cfg.sot = ids.n_vocab, ids.eot, ids.sot
    cfg.no_timestamps, cfg.ts_begin, cfg.no_speech = ids.no_timestamps, ids.ts_begin, ids.no_speech
    cfg.blank_ids[0], cfg.blank_ids[1] = ids.blank, ids.eot
    cfg.suppress_blank = int(filters and opt.suppress_blank)
    cfg.timestamps = int(filters and opt.timestamps)
    cfg.max_initial_timestamp_index = -1 if mit is None else mit
    cfg.beam_size = mode.beam_size if mode.kind == "beam" else 1
    cfg.patience = float(mode.patience)
    cfg.max_len = opt.n_text_ctxs + 2
    return cfg


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class Decoder:
    """Owning handle of a wrs_decoder (device-side bitmaps, workspace and finished-sequence store)."""

    def __init__(self, cfg: _lib.DecodeCfg, suppress_ids: Sequence[int], max_rows: int, device: int = 0):
        self.ctx = _lib.context(device)
        self.cfg = cfg
        self.device = torch.device("cuda", device)
        self.max_rows = max_rows
        self.handle = C.c_void_p()
        arr = np.asarray(list(suppress_ids), dtype=np.int32)
        _lib.check(_lib.lib().wrs_decoder_create(self.ctx.handle, C.byref(cfg), arr.ctypes.data_as(C.c_void_p) if len(arr) else None,
                                                 len(arr), max_rows, C.byref(self.handle)), self.ctx.handle)
        self.max_cand = max(1, int(math.floor(float(np.float32(cfg.patience) * np.float32(cfg.beam_size)) + 0.5)))

    def close(self):
        if getattr(self, "handle", None):
            _lib.lib().wrs_decoder_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    @staticmethod
    def _check_io(logits: torch.Tensor, tokens: torch.Tensor):
        if logits.dtype != torch.float32 or logits.dim() != 2 or logits.stride(1) != 1 or not logits.is_cuda:
            raise TypeError("logits must be a CUDA f32 [rows, vocab] tensor with unit inner stride")
        if tokens.dtype != torch.int64 or tokens.dim() != 2 or tokens.stride(1) != 1 or not tokens.is_cuda:
            raise TypeError("tokens must be a CUDA i64 [rows, len] tensor with unit inner stride")
        if logits.shape[0] != tokens.shape[0]:
            raise ValueError("logits and tokens disagree on the number of rows")

    def reset(self):
        _lib.check(_lib.lib().wrs_decoder_reset(self.handle, self._stream()), self.ctx.handle)

    def filter_apply(self, which: int, logits: torch.Tensor, tokens: torch.Tensor, sample_begin: int,
                     row_flags: Optional[torch.Tensor] = None):
        self._check_io(logits, tokens)
        _lib.check(_lib.lib().wrs_filter_apply(self.handle, which, logits.data_ptr(), logits.stride(0), logits.shape[0],
                                               tokens.data_ptr(), tokens.stride(0), tokens.shape[1], sample_begin,
                                               _ptr(row_flags), self._stream()), self.ctx.handle)

    def _debug(self, n_rows: int, debug: bool):
        if not debug:
            return None, None
        k = self.cfg.beam_size
        n_audio = max(1, n_rows // k)
        bufs = dict(cand_logprob=torch.zeros(n_rows, k + 1, dtype=torch.float32, device=self.device),
                    cand_token=torch.zeros(n_rows, k + 1, dtype=torch.int32, device=self.device),
                    row_flags=torch.zeros(n_rows, dtype=torch.uint8, device=self.device),
                    fin_count=torch.zeros(n_audio, dtype=torch.int32, device=self.device),
                    fin_source=torch.zeros(n_audio, k, dtype=torch.int32, device=self.device),
                    fin_score=torch.zeros(n_audio, k, dtype=torch.float64, device=self.device))
        dbg = _lib.StepDebug(*[bufs[n].data_ptr() for n, _ in _lib.StepDebug._fields_])
        return dbg, bufs

    def beam_step(self, logits, tokens, sample_begin, sum_logprobs, next_tokens=None, sum_out=None, debug=False):
        """-> dict(next_tokens [R, L+1], source_idx [R], sum_logprobs [R], completed [1] (device), debug buffers)."""
        self._check_io(logits, tokens)
        R, L = tokens.shape
        if next_tokens is None:
            next_tokens = torch.empty(R, L + 1, dtype=torch.int64, device=self.device)
        if sum_out is None:
            sum_out = torch.empty(R, dtype=torch.float32, device=self.device)
        source = torch.empty(R, dtype=torch.int32, device=self.device)
        completed = torch.zeros(1, dtype=torch.int32, device=self.device)
        dbg, bufs = self._debug(R, debug)
        _lib.check(_lib.lib().wrs_beam_step(self.handle, logits.data_ptr(), logits.stride(0), R, tokens.data_ptr(),
                                            tokens.stride(0), L, sample_begin, sum_logprobs.data_ptr(),
                                            next_tokens.data_ptr(), next_tokens.stride(0), source.data_ptr(),
                                            sum_out.data_ptr(), completed.data_ptr(),
                                            C.byref(dbg) if dbg is not None else None, self._stream()), self.ctx.handle)
        out = dict(next_tokens=next_tokens, source_idx=source, sum_logprobs=sum_out, completed=completed)
        if bufs:
            out.update(bufs)
        return out

    def no_speech_probs(self, logits: torch.Tensor) -> torch.Tensor:
        """src/decode/mod.rs:250-255: softmax of the unfiltered step-0 logits row at sot_idx, probability of
        token_id_no_speech -> [rows] f32 on the device."""
        if logits.dtype != torch.float32 or logits.dim() != 2 or logits.stride(1) != 1 or not logits.is_cuda:
            raise TypeError("logits must be a CUDA f32 [rows, vocab] tensor with unit inner stride")
        out = torch.empty(logits.shape[0], dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().wrs_no_speech_probs(self.handle, logits.data_ptr(), logits.stride(0), logits.shape[0],
                                                  out.data_ptr(), self._stream()), self.ctx.handle)
        return out

    def greedy_step(self, logits, tokens, sample_begin, sum_logprobs, next_tokens=None, sum_out=None, debug=False):
        self._check_io(logits, tokens)
        R, L = tokens.shape
        if next_tokens is None:
            next_tokens = torch.empty(R, L + 1, dtype=torch.int64, device=self.device)
        if sum_out is None:
            sum_out = torch.empty(R, dtype=torch.float32, device=self.device)
        completed = torch.zeros(1, dtype=torch.int32, device=self.device)
        dbg, bufs = self._debug(R, debug)
        _lib.check(_lib.lib().wrs_greedy_step(self.handle, logits.data_ptr(), logits.stride(0), R, tokens.data_ptr(),
                                              tokens.stride(0), L, sample_begin, sum_logprobs.data_ptr(),
                                              next_tokens.data_ptr(), next_tokens.stride(0), sum_out.data_ptr(),
                                              completed.data_ptr(), C.byref(dbg) if dbg is not None else None,
                                              self._stream()), self.ctx.handle)
        out = dict(next_tokens=next_tokens, sum_logprobs=sum_out, completed=completed)
        if bufs:
            out.update(bufs)
        return out

    def finalize(self, tokens: torch.Tensor, sum_logprobs: torch.Tensor):
        """-> (list per audio of candidate token lists (incl. trailing EOT), scores ndarray [n_audio, n_cand] f64)."""
        R, L = tokens.shape
        k = self.cfg.beam_size
        n_audio = R // k
        cap = max(k, self.max_cand)
        max_len = self.cfg.max_len
        out_tok = np.empty((n_audio, cap, max_len), dtype=np.int64)
        out_len = np.empty((n_audio, cap), dtype=np.int32)
        out_sc = np.empty((n_audio, cap), dtype=np.float64)
        out_n = np.empty(n_audio, dtype=np.int32)
        sums = sum_logprobs.reshape(-1).contiguous()
        _lib.check(_lib.lib().wrs_beam_finalize(self.handle, tokens.data_ptr(), tokens.stride(0), L, R, sums.data_ptr(), cap,
                                                out_tok.ctypes.data_as(C.c_void_p), out_len.ctypes.data_as(C.c_void_p),
                                                out_sc.ctypes.data_as(C.c_void_p), out_n.ctypes.data_as(C.c_void_p),
                                                self._stream()), self.ctx.handle)
        cands = [[out_tok[a, c, :out_len[a, c]].tolist() for c in range(out_n[a])] for a in range(n_audio)]
        if len({int(n) for n in out_n}) != 1:
            raise RuntimeError("ragged candidate lists (the reference panics in Tensor::stack, beam_search.rs:203)")
        return cands, out_sc[:, :int(out_n[0])].copy()


def avg_logprobs(sum_logprobs: Sequence[float], n_tokens: Sequence[int]) -> np.ndarray:
    """src/decode/mod.rs:389-397: sum_logprob / (len(tokens) + 1) per selected candidate (host arithmetic, f64)."""
    lp = np.ascontiguousarray(sum_logprobs, dtype=np.float64)
    nt = np.ascontiguousarray(n_tokens, dtype=np.int32)
    out = np.empty(len(lp), dtype=np.float64)
    _lib.check(_lib.lib().wrs_avg_logprob_host(lp.ctypes.data_as(C.c_void_p), nt.ctypes.data_as(C.c_void_p), len(lp),
                                               out.ctypes.data_as(C.c_void_p)))
    return out


class Projection:
    """`x.matmul(token_emb.T)` (reference src/model.rs:521) on the tensor cores, alone
    (`logits`) or with the beam step as its epilogue (`beam_step`: same outputs as Decoder.beam_step, the logits never
    reach HBM).  Bound to one Decoder (vocabulary geometry, masks, finished store)."""

    def __init__(self, decoder: Decoder, token_emb: torch.Tensor):
        if token_emb.dtype != torch.float32 or token_emb.dim() != 2 or token_emb.stride(1) != 1 or not token_emb.is_cuda:
            raise TypeError("token_emb must be a CUDA f32 [vocab, n_state] tensor with unit inner stride")
        if token_emb.shape[0] != decoder.cfg.n_vocab:
            raise ValueError("token_emb rows != n_vocab")
        self.dec = decoder
        self.n_state = token_emb.shape[1]
        self.handle = C.c_void_p()
        _lib.check(_lib.lib().wrs_projection_create(decoder.handle, token_emb.data_ptr(), token_emb.stride(0), self.n_state,
                                                    decoder._stream(), C.byref(self.handle)), decoder.ctx.handle)
        torch.cuda.current_stream(decoder.device).synchronize()     # token_emb may go away after this

    def close(self):
        if getattr(self, "handle", None):
            _lib.lib().wrs_projection_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check_x(self, x: torch.Tensor):
        if x.dtype != torch.float32 or x.dim() != 2 or x.stride(1) != 1 or not x.is_cuda or x.shape[1] != self.n_state:
            raise TypeError("x must be a CUDA f32 [rows, n_state] tensor with unit inner stride")

    def logits(self, x: torch.Tensor) -> torch.Tensor:
        self._check_x(x)
        out = torch.empty(x.shape[0], self.dec.cfg.n_vocab, dtype=torch.float32, device=self.dec.device)
        _lib.check(_lib.lib().wrs_project_logits(self.handle, x.data_ptr(), x.stride(0), x.shape[0], out.data_ptr(),
                                                 out.stride(0), self.dec._stream()), self.dec.ctx.handle)
        return out

    def beam_step(self, x, tokens, sample_begin, sum_logprobs, next_tokens=None, sum_out=None, debug=False):
        self._check_x(x)
        d = self.dec
        R, L = tokens.shape
        if x.shape[0] != R:
            raise ValueError("x and tokens disagree on the number of rows")
        if next_tokens is None:
            next_tokens = torch.empty(R, L + 1, dtype=torch.int64, device=d.device)
        if sum_out is None:
            sum_out = torch.empty(R, dtype=torch.float32, device=d.device)
        source = torch.empty(R, dtype=torch.int32, device=d.device)
        completed = torch.zeros(1, dtype=torch.int32, device=d.device)
        dbg, bufs = d._debug(R, debug)
        _lib.check(_lib.lib().wrs_project_beam_step(self.handle, x.data_ptr(), x.stride(0), R, tokens.data_ptr(),
                                                    tokens.stride(0), L, sample_begin, sum_logprobs.data_ptr(),
                                                    next_tokens.data_ptr(), next_tokens.stride(0), source.data_ptr(),
                                                    sum_out.data_ptr(), completed.data_ptr(),
                                                    C.byref(dbg) if dbg is not None else None, d._stream()), d.ctx.handle)
        out = dict(next_tokens=next_tokens, source_idx=source, sum_logprobs=sum_out, completed=completed)
        if bufs:
            out.update(bufs)
        return out

    def greedy_step(self, x, tokens, sample_begin, sum_logprobs, next_tokens=None, sum_out=None, debug=False):
        """src/model.rs:521 + Decoder.greedy_step in one pass (the decoder must have been built for a greedy mode: K = 1)."""
        self._check_x(x)
        d = self.dec
        R, L = tokens.shape
        if x.shape[0] != R:
            raise ValueError("x and tokens disagree on the number of rows")
        if next_tokens is None:
            next_tokens = torch.empty(R, L + 1, dtype=torch.int64, device=d.device)
        if sum_out is None:
            sum_out = torch.empty(R, dtype=torch.float32, device=d.device)
        completed = torch.zeros(1, dtype=torch.int32, device=d.device)
        dbg, bufs = d._debug(R, debug)
        _lib.check(_lib.lib().wrs_project_greedy_step(self.handle, x.data_ptr(), x.stride(0), R, tokens.data_ptr(),
                                                      tokens.stride(0), L, sample_begin, sum_logprobs.data_ptr(),
                                                      next_tokens.data_ptr(), next_tokens.stride(0), sum_out.data_ptr(),
                                                      completed.data_ptr(), C.byref(dbg) if dbg is not None else None,
                                                      d._stream()), d.ctx.handle)
        out = dict(next_tokens=next_tokens, sum_logprobs=sum_out, completed=completed)
        if bufs:
            out.update(bufs)
        return out

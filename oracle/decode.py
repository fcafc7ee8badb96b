"""Oracle for the per-step logit processing -- restates reference src/decode/** (TEST INFRASTRUCTURE).

Same ATen operators in the same order as the reference issues through tch; host-side bookkeeping
(HashMap merge, finished lists, finalize, rank) restated with a CANONICAL order wherever the reference's
order is hash-random:
  * per-row top-(k+1): ties broken by lowest token id (north_star), where ATen topk is arbitrary
  * candidate sort:    (score desc, source row asc, token asc), where the reference sorts a HashMap dump
  * finished inserts:  descending score, where the reference iterates a HashMap
  * finalize/rank:     candidates ordered (score desc, sequence lexicographic)

Parity status: UNPINNED by the reference (no test touches src/decode/**); see oracle/__init__.py.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

NEG_INF = float("-inf")


# --------------------------------------------------------------------------------------------------
# DecodePrompt  (src/decode/mod.rs:40-92)
# --------------------------------------------------------------------------------------------------
@dataclass
class DecodePrompt:
    sample_begin: int
    sot_idx: int
    initial_tokens: torch.Tensor

    @staticmethod
    def new(prompt: Optional[torch.Tensor], ids) -> "DecodePrompt":
        seq_sot = [int(t) for t in ids.sequence_sot()]
        if prompt is not None and prompt.numel() > 0:                       # :56-77, set_prompt :206-218
            initial = torch.cat([torch.tensor([ids.startofprev], dtype=torch.int64), prompt.to(torch.int64),
                                 torch.tensor(seq_sot, dtype=torch.int64)])
        else:
            initial = torch.tensor(seq_sot, dtype=torch.int64)              # :79
        sample_begin = initial.shape[0]                                      # :81
        sot_idx = int((initial == ids.sot).nonzero()[0, 0])                  # :82-85
        return DecodePrompt(sample_begin, sot_idx, initial)


# --------------------------------------------------------------------------------------------------
# LogitFilter impls  (src/decode/filter/*.rs)
# --------------------------------------------------------------------------------------------------
class SuppressTokens:
    """src/decode/filter/suppress_tokens.rs:6-24"""

    def __init__(self, token_ids):
        self.suppress_indices = torch.tensor([int(i) for i in token_ids], dtype=torch.int64)

    def apply(self, logits: torch.Tensor, tokens: torch.Tensor) -> None:
        if self.suppress_indices.numel():
            logits.index_fill_(1, self.suppress_indices, NEG_INF)            # :22


class SuppressBlank:
    """src/decode/filter/suppress_blanks.rs:10-44"""

    def __init__(self, ids, prompt_ref):
        self.suppress_indices = torch.tensor([ids.blank, ids.eot], dtype=torch.int64)   # :21-33
        self.prompt = prompt_ref                                                         # shared cell

    def apply(self, logits: torch.Tensor, tokens: torch.Tensor) -> None:
        if tokens.shape[1] == self.prompt[0].sample_begin:                   # :40
            logits.index_fill_(1, self.suppress_indices, NEG_INF)            # :41


class TimestampTokens:
    """src/decode/filter/timestamp_tokens.rs:10-98"""

    def __init__(self, ids, prompt_ref, max_initial_timestamp_index: Optional[int]):
        self.ids = ids
        self.prompt = prompt_ref
        self.max_initial_timestamp_index = max_initial_timestamp_index

    def apply(self, logits: torch.Tensor, tokens: torch.Tensor) -> None:
        sample_begin = self.prompt[0].sample_begin                            # :35
        self.rule_flags = [0] * tokens.shape[0]   # oracle-only record of which rules fired (bit layout = wrs.h)
        logits.index_fill_(1, torch.tensor([self.ids.no_timestamps]), NEG_INF)  # :37-41
        tsb = self.ids.ts_begin
        for k in range(tokens.shape[0]):                                      # :47-66
            seq = tokens[k, sample_begin:]
            n = seq.shape[0]
            last_was_timestamp = n >= 1 and int(seq[-1]) >= tsb
            second_last_was_timestamp = n < 2 or int(seq[-2]) >= tsb
            if last_was_timestamp:
                if second_last_was_timestamp:
                    logits[k, tsb:].fill_(NEG_INF)                            # :57-59
                    self.rule_flags[k] |= 1
                else:
                    logits[k, :self.ids.eot].fill_(NEG_INF)                   # :61-63
                    self.rule_flags[k] |= 2
        if tokens.shape[1] == sample_begin:                                   # :68-78
            logits[:, :tsb].fill_(NEG_INF)
            self.rule_flags = [f | 4 for f in self.rule_flags]
            if self.max_initial_timestamp_index is not None:
                last_allowed = tsb + self.max_initial_timestamp_index
                logits[:, last_allowed + 1:].fill_(NEG_INF)
        logprobs = logits.log_softmax(-1, dtype=torch.float32)                # :82
        for k in range(tokens.shape[0]):                                      # :83-96
            timestamp_logprob = float(logprobs[k, tsb:].logsumexp(-1))        # f32 result widened to f64
            max_text_token_logprob = float(logprobs[k, :tsb].max())
            if timestamp_logprob > max_text_token_logprob:
                logits[k, :tsb].fill_(NEG_INF)
                self.rule_flags[k] |= 8
            self.margins = getattr(self, "margins", {})
            self.margins[k] = timestamp_logprob - max_text_token_logprob


@dataclass
class DecodeOptions:
    """The subset of src/decode/mod.rs:95-106 that shapes the hot path."""
    beam_size: int = 5
    patience: float = 1.0
    len_penalty: Optional[float] = None
    max_initial_timestamp: Optional[float] = 1.0
    timestamps: bool = True
    suppress_blank: bool = True
    suppress_non_speech: bool = True
    suppress_tokens: Optional[List[int]] = None
    n_audio_ctxs: int = 1500
    n_text_ctxs: int = 448
    sample_len: Optional[int] = None
    greedy: bool = False
    greedy_group: int = 1     # TokenExtractMode::Greedy(group_sze), src/decode/extractor/mod.rs:14-18


def max_initial_timestamp_index(opt: DecodeOptions) -> Optional[int]:
    """src/decode/mod.rs:176-182: round(mit / (CHUNK_LENGTH / n_audio_ctxs)); Rust f64::round is
    half-away-from-zero."""
    if opt.max_initial_timestamp is None:
        return None
    precision = 30.0 / opt.n_audio_ctxs
    v = opt.max_initial_timestamp / precision
    return int(math.floor(abs(v) + 0.5) * (1 if v >= 0 else -1))


def build_filters(opt: DecodeOptions, ids, prompt_ref) -> list:
    """src/decode/mod.rs:156-190 -- order matters: Blank -> Suppress -> Timestamp."""
    filters = []
    if opt.suppress_blank:
        filters.append(SuppressBlank(ids, prompt_ref))
    if opt.suppress_tokens is not None or opt.suppress_non_speech:
        lst = list(opt.suppress_tokens or [])
        if opt.suppress_non_speech:
            lst.extend(ids.non_speech_tokens)
        filters.append(SuppressTokens(lst))
    if opt.timestamps:
        filters.append(TimestampTokens(ids, prompt_ref, max_initial_timestamp_index(opt)))
    return filters


# --------------------------------------------------------------------------------------------------
# TokenExtractor impls  (src/decode/extractor/*.rs)
# --------------------------------------------------------------------------------------------------
def canonical_topk(row: torch.Tensor, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """`row.topk(k, -1, largest=True, sorted=True)` (beam_search.rs:62-65) with ties canonicalised to
    lowest index first (ATen leaves tied indices in arbitrary order)."""
    vals = torch.topk(row, k, dim=-1, largest=True, sorted=True).values
    thr = vals[-1]
    idx = (row >= thr).nonzero()[:, 0]
    v = row[idx]
    order = np.lexsort((idx.numpy(), -v.numpy().astype(np.float64)))[:k]   # primary -value, secondary index
    idx = idx[torch.from_numpy(order)]
    return row[idx], idx


def rust_round_f32(x: float) -> int:
    """f32::round -- half away from zero (beam_search.rs:122)."""
    x = float(np.float32(x))
    return int(math.floor(abs(x) + 0.5)) * (1 if x >= 0 else -1)


class BeamSearchTokenExtractor:
    """src/decode/extractor/beam_search.rs:9-211"""

    def __init__(self, beam_size: int, eot: int, patience: float = 1.0):
        self.token_id_eot = eot
        self.beam_size = beam_size
        self.patience = float(np.float32(patience))
        self.finished_sequences: Optional[List[Dict[tuple, float]]] = None

    def reset(self):                                                            # :18-20
        self.finished_sequences = None

    def group_size(self):
        return self.beam_size

    def update(self, tokens: torch.Tensor, logits: torch.Tensor, sum_logprobs: torch.Tensor):
        """:22-149.  `sum_logprobs` [R] f32 is mutated in place.  Returns (tokens [R,L+1], completed).
        Oracle-only records: self.last_sources, self.last_candidates, self.last_finished."""
        k = self.beam_size
        assert tokens.shape[0] % k == 0
        n_audio = tokens.shape[0] // k
        if self.finished_sequences is None:
            self.finished_sequences = [dict() for _ in range(n_audio)]
        logprobs = logits.log_softmax(-1, dtype=torch.float32)                  # :37
        if torch.isnan(logprobs).any():
            raise FloatingPointError("NaN log-probability (reference panics in partial_cmp().unwrap(), :86)")
        next_tokens, source_indices, finished_sequences = [], [], []
        prefixes = tokens.tolist()
        cand_lp = torch.empty(tokens.shape[0], k + 1, dtype=torch.float32)
        cand_tok = torch.empty(tokens.shape[0], k + 1, dtype=torch.int64)
        for i in range(n_audio):
            scores_sources: Dict[tuple, Tuple[float, int]] = {}
            finished: Dict[tuple, float] = {}
            for j in range(k):                                                  # STEP 1 :55-79
                idx = i * k + j
                lp, tk = canonical_topk(logprobs[idx], k + 1)
                cand_lp[idx], cand_tok[idx] = lp, tk
                for c in range(k + 1):
                    new_logprob = float(sum_logprobs[idx]) + float(lp[c])       # f64(f32)+f64(f32) :70-71
                    sequence = tuple(prefixes[idx]) + (int(tk[c]),)
                    scores_sources[sequence] = (new_logprob, idx)               # later insert overrides :77
            saved = 0                                                            # STEP 2 :81-104
            items = sorted(scores_sources.items(), key=lambda kv: (-kv[1][0], kv[1][1], kv[0][-1]))
            for sequence, (score, source) in items:
                if sequence[-1] == self.token_id_eot:
                    finished[sequence] = score
                else:
                    sum_logprobs[len(next_tokens)] = float(np.float32(score))   # index_put_ :91-95
                    next_tokens.append(list(sequence))
                    source_indices.append(source)
                    saved += 1
                    if saved == k:
                        break
            finished_sequences.append(finished)
        new_tokens = torch.tensor(next_tokens, dtype=torch.int64)               # :109
        self.last_sources = torch.tensor(source_indices, dtype=torch.int64)     # -> model.update_cache :114
        self.last_candidates = (cand_lp, cand_tok)
        self.last_finished = finished_sequences
        max_candidates = rust_round_f32(np.float32(self.patience) * np.float32(k))   # :122
        for prev, new in zip(self.finished_sequences, finished_sequences):      # :124-138
            for seq, score in sorted(new.items(), key=lambda kv: (-kv[1], kv[0])):   # canonical: score desc
                if len(prev) >= max_candidates:
                    break
                prev[seq] = score
        completed = all(len(s) >= max_candidates for s in self.finished_sequences)   # :141-146
        return new_tokens, completed

    def finalize(self, tokens: torch.Tensor, sum_logprobs: torch.Tensor):
        """:151-206.  tokens [n_audio, k, L], sum_logprobs [n_audio, k].  Returns (list of per-audio
        candidate lists (python lists of ints), scores [n_audio, n_cand] f64)."""
        for i, seqs in enumerate(self.finished_sequences):
            if len(seqs) < self.beam_size:                                      # :162
                s = sum_logprobs[i].to(torch.float64).numpy()
                order = np.lexsort((np.arange(len(s)), -s))                     # canonical: score desc, j asc
                for j in order:
                    seq = tuple(tokens[i, j].tolist()) + (self.token_id_eot,)
                    seqs[seq] = float(sum_logprobs[i, j])
                    if len(seqs) >= self.beam_size:
                        break
        out_tokens, out_scores = [], []
        for seqs in self.finished_sequences:
            items = sorted(seqs.items(), key=lambda kv: (-kv[1], kv[0]))       # canonical order
            out_tokens.append([list(kv[0]) for kv in items])
            out_scores.append([kv[1] for kv in items])
        n = {len(s) for s in out_scores}
        if len(n) != 1:
            raise RuntimeError("Tensor::stack of ragged candidate lists (reference panics, beam_search.rs:203)")
        return out_tokens, torch.tensor(out_scores, dtype=torch.float64)


class GreedyTokenExtractor:
    """Intent of src/decode/extractor/greedy.rs:14-66 (dead code in the reference):
    temperature 0 only; `completed` tests the last column."""

    def __init__(self, eot: int, group_size: int = 1):
        self.token_id_eot = eot
        self._group_size = group_size

    def reset(self):
        pass

    def group_size(self):
        return self._group_size

    def update(self, tokens: torch.Tensor, logits: torch.Tensor, sum_logprobs: torch.Tensor):
        next_tokens = logits.argmax(-1)                                         # :25 (first max wins on CPU)
        logprobs = logits.log_softmax(-1, dtype=torch.float32)                  # :31
        current = logprobs[torch.arange(logprobs.shape[0]), next_tokens]        # :32-35
        last_eot = tokens[:, -1] == self.token_id_eot                           # :37-38
        sum_logprobs += current * (~last_eot)                                   # :39
        next_tokens = next_tokens.clone()
        next_tokens[last_eot] = self.token_id_eot                               # :41-45
        tokens = torch.cat([tokens, next_tokens[:, None]], dim=-1)              # :47
        completed = bool((tokens[:, -1] == self.token_id_eot).all())            # intent of :48
        return tokens, completed

    def finalize(self, tokens: torch.Tensor, sum_logprobs: torch.Tensor):
        """Commented-out intent at :54-58: every sequence gets one trailing EOT."""
        out = [[t.tolist() + [self.token_id_eot] for t in group] for group in tokens]
        return out, sum_logprobs.to(torch.float64)


# --------------------------------------------------------------------------------------------------
# SequenceRanker  (src/decode/sequencer/max_likelihood.rs:6-35)
# --------------------------------------------------------------------------------------------------
class MaximumLikelihoodRanker:
    def __init__(self, length_penalty: Optional[float] = None):
        self.length_penalty = length_penalty

    def rank(self, tokens: List[List[List[int]]], sum_logprobs: torch.Tensor) -> List[int]:
        out = []
        for i, group in enumerate(tokens):
            scores = []
            for logprob, seq in zip(sum_logprobs[i].tolist(), group):
                length = float(len(seq))
                penalty = ((5.0 + length) / 6.0) ** self.length_penalty if self.length_penalty is not None else length
                with np.errstate(divide="ignore", invalid="ignore"):
                    scores.append(float(np.float64(logprob) / np.float64(penalty)))   # len 0 => inf/NaN, as in Rust
            out.append(int(torch.tensor(scores, dtype=torch.float64).argmax()))       # :30-31
        return out


def slice_candidates(tokens: List[List[List[int]]], sample_begin: int, eot: int) -> List[List[List[int]]]:
    """src/decode/mod.rs:356-370: tokens[sample_begin .. first EOT]."""
    return [[seq[sample_begin:seq.index(eot)] for seq in group] for group in tokens]


# --------------------------------------------------------------------------------------------------
# one decode step end to end: the section of main_loop the fused kernel replaces (decode/mod.rs:257-270)
# --------------------------------------------------------------------------------------------------
def decode_step_reference(case: dict, ids, opt: Optional[DecodeOptions] = None, extractor=None):
    """Runs filters (in reference order) then `update` on a synth.decode_case dict.  Returns numpy
    results keyed like the C ABI outputs."""
    opt = opt or DecodeOptions(beam_size=case["beam"])
    logits = torch.from_numpy(case["logits"]).clone()
    tokens = torch.from_numpy(case["tokens"]).clone()
    sums = torch.from_numpy(case["sum_logprobs"]).clone()
    prompt = DecodePrompt(case["sample_begin"], 0, tokens[0, :case["sample_begin"]])
    prompt_ref = [prompt]
    filters = build_filters(opt, ids, prompt_ref)
    for f in filters:
        f.apply(logits, tokens)
    flags = np.zeros(tokens.shape[0], dtype=np.uint8)
    margins = np.full(tokens.shape[0], np.nan)
    for f in filters:
        if isinstance(f, TimestampTokens):
            flags = np.asarray(f.rule_flags, dtype=np.uint8)
            margins = np.asarray([f.margins[k] for k in range(tokens.shape[0])])
    res = dict(n_finite=torch.isfinite(logits).sum(-1).numpy().astype(np.int64), row_flags=flags, margins=margins,
               masked_logits=logits.numpy())
    if opt.greedy:
        ex = extractor or GreedyTokenExtractor(ids.eot)
        new_tokens, completed = ex.update(tokens, logits, sums)
        res.update(next_tokens=new_tokens.numpy(), sum_logprobs=sums.numpy(), completed=np.array(completed))
        return res
    ex = extractor or BeamSearchTokenExtractor(opt.beam_size, ids.eot, opt.patience)
    new_tokens, completed = ex.update(tokens, logits, sums)
    k = opt.beam_size
    n_audio = tokens.shape[0] // k
    fin_count = np.zeros(n_audio, dtype=np.int32)
    fin_score = np.full((n_audio, k), np.nan)
    fin_last_prefix_tok = np.full((n_audio, k), -1, dtype=np.int64)
    for i, fin in enumerate(ex.last_finished):
        items = sorted(fin.items(), key=lambda kv: (-kv[1], kv[0]))
        fin_count[i] = len(items)
        for c, (seq, score) in enumerate(items):
            fin_score[i, c] = score
            fin_last_prefix_tok[i, c] = seq[-2]
    res.update(cand_logprob=ex.last_candidates[0].numpy(), cand_token=ex.last_candidates[1].numpy(),
               next_tokens=new_tokens.numpy(), source_idx=ex.last_sources.numpy(), sum_logprobs=sums.numpy(),
               fin_count=fin_count, fin_score=fin_score, completed=np.array(completed),
               stored_finished=np.array([len(s) for s in ex.finished_sequences], dtype=np.int32))
    return res


def no_speech_probs(logits: torch.Tensor, no_speech: int) -> torch.Tensor:
    """src/decode/mod.rs:250-255: `logits.i((.., sot_idx)).softmax(-1)` then column token_id_no_speech.
    `logits` is the [rows, V] slice at sot_idx (unfiltered)."""
    return logits.softmax(-1)[:, no_speech]


def avg_logprobs(sum_logprobs, n_tokens):
    """src/decode/mod.rs:389-397: lp / (t.size()[0] as f64 + 1.)."""
    return [float(lp) / (float(n) + 1.0) for lp, n in zip(sum_logprobs, n_tokens)]


# --------------------------------------------------------------------------------------------------
# DecodeTask  (src/decode/mod.rs:220-418): main_loop + run_inner around a caller-supplied model
# --------------------------------------------------------------------------------------------------
@dataclass
class DecodeOutput:
    tokens: List[int]
    avg_logprob: float
    no_speech_prob: float
    sum_logprob: float
    n_steps: int


def decode_task_reference(model, ids, opt: DecodeOptions, mel: torch.Tensor, prompt: Optional[torch.Tensor] = None,
                          trace: Optional[list] = None) -> List[DecodeOutput]:
    """`DecodeTask::run` with the filters / extractor / ranker of this file on CPU copies of the logits.  `model` is
    the duck-typed forward of whisper_rs_b200.decode.task (it may live on any device); its `update_cache` receives
    the extractor's source rows like beam_search.rs:111-114.  `trace`, if given, collects the token tensor of every
    step."""
    dev = mel.device
    p = DecodePrompt.new(prompt, ids)
    prompt_ref = [p]
    filters = build_filters(opt, ids, prompt_ref)
    ex = GreedyTokenExtractor(ids.eot, opt.greedy_group) if opt.greedy else BeamSearchTokenExtractor(opt.beam_size, ids.eot, opt.patience)
    n_group = ex.group_size()
    sample_len = opt.sample_len if opt.sample_len is not None else model.dims.n_text_ctxs // 2      # :129-131
    with torch.no_grad():
        n_audio = mel.shape[0]
        feats = model.encoder(mel).repeat_interleave(n_group, 0)                                     # :293,307-311
        tokens = p.initial_tokens.repeat(n_audio, 1).repeat_interleave(n_group, 0)
        model.clear_cache()
        ex.reset()
        # ---- main_loop :220-282
        sums = torch.zeros(tokens.shape[0])
        no_speech = torch.full((tokens.shape[0],), float("nan"))
        n_steps = 0
        for i in range(sample_len):
            t_in = tokens[:, -1:] if tokens.shape[-1] > p.sample_begin else tokens
            logits = model.decoder.forward_ext(t_in.to(dev), feats, i).float().cpu()
            if i == 0:
                no_speech = no_speech_probs(logits[:, p.sot_idx], ids.no_speech)
            last = logits[:, -1].clone()
            for f in filters:
                f.apply(last, tokens)
            tokens, completed = ex.update(tokens, last, sums)
            if not opt.greedy and hasattr(model, "update_cache"):
                model.update_cache(ex.last_sources.to(dev))
            n_steps += 1
            if trace is not None:
                trace.append(tokens.clone())
            if completed or tokens.shape[-1] > model.dims.n_text_ctxs:
                break
        model.clear_cache()
    # ---- run_inner :335-418
    L = tokens.shape[1]
    cands, scores = ex.finalize(tokens.reshape(n_audio, n_group, L), sums.reshape(n_audio, n_group))
    cands = [[[int(t) for t in seq] for seq in group] for group in cands]
    sliced = slice_candidates(cands, p.sample_begin, ids.eot)
    scores = torch.as_tensor(np.asarray(scores), dtype=torch.float64)
    sel = MaximumLikelihoodRanker(opt.len_penalty).rank(sliced, scores)
    chosen = [group[s] for group, s in zip(sliced, sel)]
    lps = [float(scores[a][s]) for a, s in enumerate(sel)]
    avg = avg_logprobs(lps, [len(t) for t in chosen])
    return [DecodeOutput(chosen[a], avg[a], float(no_speech[a * n_group]), lps[a], n_steps) for a in range(n_audio)]

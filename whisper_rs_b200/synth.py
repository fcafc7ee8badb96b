"""Seeded synthetic inputs for both hot paths.

Used by the tests (oracle and CUDA path see identical bytes), by bench.py and by smoke().
Pure numpy; nothing here is on the product path.
"""
from __future__ import annotations

import numpy as np

from . import constants as C

AUDIO_KINDS = ("gaussian", "am_noise", "chirp", "tone", "burst", "zeros")


def audio_segment(kind: str, seed: int, n_samples: int = C.N_SAMPLES) -> np.ndarray:
    """One f32 mono 16 kHz clip of `n_samples` samples."""
    rng = np.random.default_rng([seed, AUDIO_KINDS.index(kind)])
    t = np.arange(n_samples, dtype=np.float64) / C.SAMPLE_RATE
    if kind == "gaussian":
        x = 0.1 * rng.standard_normal(n_samples)
    elif kind == "am_noise":  # "speech-like": noise with a 3 Hz syllabic envelope
        x = 0.3 * rng.standard_normal(n_samples) * (0.5 + 0.5 * np.sin(2 * np.pi * 3.0 * t)) ** 4
    elif kind == "chirp":     # 100 Hz -> 7 kHz linear chirp over the clip, -40 dB noise floor
        dur = n_samples / C.SAMPLE_RATE
        phase = 2 * np.pi * (100.0 * t + 0.5 * (7000.0 - 100.0) / dur * t * t)
        x = 0.5 * np.sin(phase) + 0.005 * rng.standard_normal(n_samples)
    elif kind == "tone":      # pure 440 Hz: exercises the max-8 floor far from the peak
        x = 0.5 * np.sin(2 * np.pi * 440.0 * t)
    elif kind == "burst":     # digital silence with a 1 s noise burst in the middle
        x = np.zeros(n_samples)
        a = n_samples // 2
        b = min(n_samples, a + C.SAMPLE_RATE)
        x[a:b] = 0.2 * rng.standard_normal(b - a)
    elif kind == "zeros":
        x = np.zeros(n_samples)
    else:
        raise ValueError(kind)
    return x.astype(np.float32)


def audio_batch(n_segments: int, seed: int = 0, n_samples: int = C.N_SAMPLES) -> np.ndarray:
    """Config-2 mixture: 1/3 gaussian, 1/3 AM-noise, 1/6 chirp, 1/12 tone, 1/12 burst -> [S, n_samples] f32."""
    pattern = ["gaussian"] * 4 + ["am_noise"] * 4 + ["chirp"] * 2 + ["tone", "burst"]
    out = np.empty((n_segments, n_samples), dtype=np.float32)
    for s in range(n_segments):
        out[s] = audio_segment(pattern[s % len(pattern)], seed + s, n_samples)
    return out


def suppress_list(ids: C.TokenIds, suppress_non_speech: bool = True, user=None):
    """decode/mod.rs:166-174: user list followed by tokenizer.non_speech_tokens()."""
    out = list(user or [])
    if suppress_non_speech:
        out.extend(ids.non_speech_tokens)
    return out


def decode_case(ids: C.TokenIds, n_audio: int, beam: int, step_kind: str, seed: int, prompt_len: int = 0,
                hist_len: int = 6, planted_ties: bool = False, sigma: float = 3.0, quantum: float = 0.0):
    """One decode step's inputs.

    step_kind:
      "first"  L == sample_begin (all beams of an audio share the prompt, sum_logprobs 0)
      "second" one sampled token per row (penultimate defaults to "timestamp")
      "mixed"  hist_len sampled tokens; rows cycle through the four (penult, last) in {text,ts}^2 cases
    quantum > 0 rounds the logits to multiples of it: thousands of exact ties per row, also among the best ids.
    Returns a dict of numpy arrays + scalars; logits are f32 [R, V] contiguous.
    """
    rng = np.random.default_rng([seed, n_audio, beam, prompt_len])
    R, V = n_audio * beam, ids.n_vocab
    if prompt_len > 0:
        prompt = rng.integers(0, 50000, size=prompt_len)
        initial = np.concatenate([[ids.startofprev], prompt, ids.sequence_sot()]).astype(np.int64)
    else:
        initial = np.asarray(ids.sequence_sot(), dtype=np.int64)
    sample_begin = len(initial)
    n_ts = V - ids.ts_begin

    if step_kind == "first":
        tokens = np.tile(initial, (R, 1))
        sums = np.zeros(R, dtype=np.float32)
    else:
        h = 1 if step_kind == "second" else hist_len
        hist = rng.integers(0, 50000, size=(R, h)).astype(np.int64)
        for r in range(R):
            case = r % 4
            ts_a = ids.ts_begin + int(rng.integers(0, n_ts))
            ts_b = ids.ts_begin + int(rng.integers(0, n_ts))
            if h >= 2:
                if case == 1:    # (ts, text)
                    hist[r, -2] = ts_a
                elif case == 2:  # (text, ts) -> text below eot is masked
                    hist[r, -1] = ts_b
                elif case == 3:  # (ts, ts)   -> timestamps masked
                    hist[r, -2], hist[r, -1] = ts_a, ts_b
            elif case >= 2:      # single sampled token that is a timestamp
                hist[r, -1] = ts_b
        tokens = np.concatenate([np.tile(initial, (R, 1)), hist], axis=1)
        sums = (-rng.random(R) * 3.0 * h).astype(np.float32)

    logits = (sigma * rng.standard_normal((R, V))).astype(np.float32)
    offsets = np.array([-4.0, 0.0, 2.0, 6.0], dtype=np.float32)
    row_off = offsets[(np.arange(R) // 4) % 4]
    logits[:, ids.ts_begin:] += row_off[:, None]

    if quantum > 0.0:
        logits = (np.round(logits / np.float32(quantum)) * np.float32(quantum)).astype(np.float32)

    if planted_ties:
        # several exact ties for the row maximum, in both partitions; lowest index must win
        for r in range(R):
            top = np.float32(logits[r].max() + 1.0)
            idx = rng.choice(ids.eot - 300, size=4, replace=False) + 300   # clear of the suppress list head
            logits[r, idx] = top
            tsi = ids.ts_begin + rng.choice(n_ts, size=3, replace=False)
            logits[r, tsi] = top

    # keep the timestamp-vs-text decision clear of fp32 rounding: |lse(ts) - max(text)| >= 5e-3
    static = np.zeros(V, dtype=bool)
    static[np.asarray(suppress_list(ids), dtype=np.int64)] = True
    static[ids.no_timestamps] = True
    if step_kind == "first":
        static[[ids.blank, ids.eot]] = True
    for r in range(R):
        for _ in range(8):
            x = logits[r].astype(np.float64)
            x[static] = -np.inf
            ts = x[ids.ts_begin:]
            m = ts.max()
            lse_ts = m + np.log(np.exp(ts - m).sum())
            d = lse_ts - x[:ids.ts_begin].max()
            if abs(d) >= 5e-3:
                break
            logits[r, ids.ts_begin:] += np.float32(0.0625)

    return dict(logits=logits, tokens=tokens, sum_logprobs=sums, sample_begin=sample_begin, n_audio=n_audio,
                beam=beam, step_kind=step_kind)

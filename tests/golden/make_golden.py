#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/ from the reference tree.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

Outputs (all small, committed):
  reference_kat.json      -- the reference's own known-answer data for the hot path:
                             * mel-filter 8x8 corner   (reference test/data/mel-filter-8x8.csv,
                               asserted by src/audio.rs:266-279 `test_mel_filter`)
                             * invariants of the 80x1100 log-mel golden
                               (reference test/data/mel-spectrogram.csv, src/audio.rs:281-301;
                               its input WAV is not in the reference repo, so only
                               shape / floor / peak / range can be pinned)
  tokenizer_constants.json -- ids derived from reference tokenizer/gpt2.json by replaying
                             src/tokenize.rs:142-193 (special tokens) and :215-307
                             (non_speech_tokens) with the HF `tokenizers` package.
  oracle_vectors.npz      -- seeded inputs -> oracle outputs (log-mel excerpts and one decode
                             step), so that the oracle itself is regression-pinned.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def reference_kat():
    corner = np.loadtxt(os.path.join(REF, "test/data/mel-filter-8x8.csv"), delimiter=",")
    spec = np.loadtxt(os.path.join(REF, "test/data/mel-spectrogram.csv"), delimiter=",", dtype=np.float32)
    assert corner.shape == (8, 8)
    kat = {
        "mel_filter_8x8": [[float(v) for v in row] for row in corner],
        "mel_filter_8x8_rtol": 1e-5,
        "mel_filter_8x8_atol": 1e-8,
        "mel_spectrogram_shape": list(spec.shape),
        "mel_spectrogram_min": float(spec.min()),
        "mel_spectrogram_max": float(spec.max()),
        # first six frames of every mel row sit on the max-8 floor (leading silence)
        "mel_spectrogram_floor_count": int((spec == spec.min()).sum()),
    }
    with open(os.path.join(HERE, "reference_kat.json"), "w") as f:
        json.dump(kat, f, indent=1)
    print("reference_kat.json:", kat["mel_spectrogram_shape"], kat["mel_spectrogram_min"], kat["mel_spectrogram_max"])


def tokenizer_constants():
    import tokenizers

    tok = tokenizers.Tokenizer.from_file(os.path.join(REF, "tokenizer/gpt2.json"))
    base_vocab = tok.get_vocab_size()
    # src/tokenize.rs:154-172 -- sot, 99 languages, then six fixed control tokens.  The language
    # order is HashMap-random in the reference; only the count matters for the ids that follow.
    n_lang = 99
    specials = ["<|startoftranscript|>"] + [f"<|lang{i}|>" for i in range(n_lang)] + [
        "<|translate|>", "<|transcribe|>", "<|startoflm|>", "<|startofprev|>", "<|nospeech|>", "<|notimestamps|>"]
    tok.add_special_tokens([tokenizers.AddedToken(s, special=True) for s in specials])
    ids = {s: tok.token_to_id(s) for s in specials}

    def enc(s):
        return tok.encode(s, add_special_tokens=True).ids

    # src/tokenize.rs:215-307
    symbols = ['"', "#", "(", ")", "*", "+", "/", ":", ";", "<", "=", ">", "@", "[", "\\", "]", "^", "_", "`",
               "{", "|", "}", "~", "「", "」", "『", "』", "<<", ">>", "<<<", ">>>", "--", "---", "-(", "-[",
               "('", '("', "((", "))", "(((", ")))", "[[", "]]", "{{", "}}", "♪♪", "♪♪♪"]
    result = [enc(" -")[0], enc(" '")[0]]
    for s in symbols:
        for cand in (s, " " + s):
            t = enc(cand)
            if len(t) == 1:
                result.append(t[0])
    for s in ["♩", "♪", "♫", "♬", "♭", "♮", "♯"]:
        result.append(enc(s)[0])
        result.append(enc(" " + s)[0])
    result.sort()

    out = {
        "base_vocab": base_vocab,
        "eot": tok.token_to_id("<|endoftext|>"),
        "sot": ids["<|startoftranscript|>"],
        "translate": ids["<|translate|>"],
        "transcribe": ids["<|transcribe|>"],
        "startoflm": ids["<|startoflm|>"],
        "startofprev": ids["<|startofprev|>"],
        "no_speech": ids["<|nospeech|>"],
        "no_timestamps": ids["<|notimestamps|>"],
        "ts_begin": ids["<|notimestamps|>"] + 1,
        "blank": enc(" ")[0],
        "non_speech_tokens": result,          # sorted, NOT deduped (src/tokenize.rs:305)
        "n_vocab": ids["<|notimestamps|>"] + 1 + 1501,
    }
    with open(os.path.join(HERE, "tokenizer_constants.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("tokenizer_constants.json: eot", out["eot"], "sot", out["sot"], "ts_begin", out["ts_begin"],
          "n_vocab", out["n_vocab"], "non_speech", len(result), "unique", len(set(result)))


def oracle_vectors():
    import torch
    from oracle import audio as oa
    from oracle import decode as od
    from whisper_rs_b200 import synth, constants

    torch.manual_seed(0)
    out = {}
    filt = oa.mel_filter()
    # log-mel: one short clip per signal family (2 s each keeps the fixture small)
    for kind in synth.AUDIO_KINDS:
        x = synth.audio_segment(kind, seed=7, n_samples=32000)
        y = oa.log_mel_spectrogram(torch.from_numpy(x), filt).numpy()
        out[f"logmel_{kind}_in_seed"] = np.array([7, 32000])
        out[f"logmel_{kind}_out"] = y.astype(np.float32)
    # decode step: small-R slice of the config-3 generator
    ids = constants.MULTILINGUAL
    case = synth.decode_case(ids, n_audio=3, beam=5, step_kind="mixed", seed=11)
    res = od.decode_step_reference(case, ids)
    for k, v in res.items():
        if k != "masked_logits":      # R x V floats: regenerated from the seed instead of stored
            out[f"decode_{k}"] = np.asarray(v)
    out["decode_case"] = np.array([3, 5, 11])   # n_audio, beam, seed of synth.decode_case(step_kind="mixed")
    np.savez_compressed(os.path.join(HERE, "oracle_vectors.npz"), **out)
    print("oracle_vectors.npz:", sorted(out.keys()))


if __name__ == "__main__":
    which = sys.argv[1:] or ["kat", "tok", "vec"]
    if "kat" in which:
        reference_kat()
    if "tok" in which:
        tokenizer_constants()
    if "vec" in which:
        oracle_vectors()

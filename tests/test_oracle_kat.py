"""The oracle against the reference's own known answers and against the committed regression vectors
(CPU only).  reference: src/audio.rs:266-301, test/data/*.csv."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import audio as oa
from oracle import decode as od
from whisper_rs_b200 import constants as C
from whisper_rs_b200 import synth

KAT = json.load(open(os.path.join(GOLDEN, "reference_kat.json")))
VEC = np.load(os.path.join(GOLDEN, "oracle_vectors.npz"))


def test_mel_filter_matches_reference_8x8():
    # reference test_mel_filter: allclose(rtol 1e-5, atol 1e-8) on the top-left 8x8
    f = oa.mel_filter_np()
    assert f.shape == (80, 201) and f.dtype == np.float32
    assert np.allclose(f[:8, :8], np.array(KAT["mel_filter_8x8"]), rtol=KAT["mel_filter_8x8_rtol"], atol=KAT["mel_filter_8x8_atol"])


def test_mel_filter_structure():
    f = oa.mel_filter_np()
    nz = f != 0
    assert nz.sum() == 391 and nz.sum(0).max() == 2 and nz.sum(1).max() == 14 and not nz[:, 200].any()
    for m in range(80):   # banded: non-zeros of a row are contiguous
        idx = np.nonzero(nz[m])[0]
        assert (np.diff(idx) == 1).all()


def test_log_mel_invariants_of_reference_golden():
    # the reference golden (80x1100) pins: range exactly 8/4 = 2.0 between floor and peak
    assert KAT["mel_spectrogram_shape"] == [80, 1100]
    assert abs((KAT["mel_spectrogram_max"] - KAT["mel_spectrogram_min"]) - 2.0) < 1e-6
    filt = oa.mel_filter()
    for kind in ("am_noise", "burst", "tone"):
        y = oa.log_mel_spectrogram(torch.from_numpy(synth.audio_segment(kind, 5, 11 * 16000)), filt).numpy()
        assert y.shape == (80, 1100)
        assert abs((y.max() - y.min()) - 2.0) < 1e-6     # floor is active for these signals
    z = oa.log_mel_spectrogram(torch.zeros(16000), filt)
    assert torch.all(z == -1.5)                            # log10(1e-10) = -10 -> (-10 + 4) / 4


@pytest.mark.parametrize("kind", synth.AUDIO_KINDS)
def test_log_mel_regression_vectors(kind):
    seed, n = VEC[f"logmel_{kind}_in_seed"]
    y = oa.log_mel_spectrogram(torch.from_numpy(synth.audio_segment(kind, int(seed), int(n))), oa.mel_filter()).numpy()
    assert np.abs(y - VEC[f"logmel_{kind}_out"]).max() <= 2e-5   # same op sequence; slack for BLAS/FFT build differences


@pytest.mark.parametrize("kind", ["gaussian", "am_noise", "chirp", "tone", "burst"])
def test_fp32_error_budget(kind):
    # how far the f32 op sequence sits from exact arithmetic: must stay inside north_star's 1e-4
    x = synth.audio_segment(kind, 3, 160000)
    y32 = oa.log_mel_spectrogram(torch.from_numpy(x), oa.mel_filter()).numpy()
    y64 = oa.log_mel_spectrogram_f64(x, oa.mel_filter_np())
    assert np.abs(y32 - y64).max() < 1e-4


def test_pad_or_trim():
    a = torch.arange(12, dtype=torch.float32).reshape(3, 4)
    assert torch.equal(oa.pad_or_trim(a, 2), a[:, :2])
    p = oa.pad_or_trim(a, 6)
    assert p.shape == (3, 6) and torch.equal(p[:, :4], a) and torch.all(p[:, 4:] == 0)
    assert torch.equal(oa.pad_or_trim(a, 4), a)


def test_decode_regression_vectors():
    n_audio, beam, seed = [int(v) for v in VEC["decode_case"]]
    case = synth.decode_case(C.MULTILINGUAL, n_audio, beam, "mixed", seed)
    res = od.decode_step_reference(case, C.MULTILINGUAL)
    for k in ("cand_token", "next_tokens", "source_idx", "row_flags", "n_finite", "fin_count"):
        assert np.array_equal(res[k], VEC[f"decode_{k}"]), k
    assert np.allclose(res["cand_logprob"], VEC["decode_cand_logprob"], rtol=1e-5, atol=0)
    assert np.allclose(res["sum_logprobs"], VEC["decode_sum_logprobs"], rtol=1e-5, atol=0)


def test_tokenizer_constants_match_package():
    tk = json.load(open(os.path.join(GOLDEN, "tokenizer_constants.json")))
    ids = C.GPT2_EN
    for k in ("eot", "sot", "translate", "transcribe", "startoflm", "startofprev", "no_speech", "no_timestamps",
              "ts_begin", "blank", "n_vocab"):
        assert getattr(ids, k) == tk[k], k
    assert list(ids.non_speech_tokens) == tk["non_speech_tokens"]
    assert len(ids.non_speech_tokens) == 96 and len(set(ids.non_speech_tokens)) == 84


def test_oracle_breaks_exact_ties_by_lowest_id():
    """Quantised logits (synth quantum): many ids share the best values; the oracle's canonical order must list equal
    log-probs by ascending id -- the bar the CUDA step is held to (tests/test_gpu_decode.py)."""
    ids = C.MULTILINGUAL
    case = synth.decode_case(ids, 1, 5, "mixed", seed=11, quantum=2.0)
    want = od.decode_step_reference(case, ids)
    lp, tok = want["cand_logprob"], want["cand_token"]
    tied = lp[:, :-1] == lp[:, 1:]
    assert tied.any()
    assert np.all(tok[:, :-1][tied] < tok[:, 1:][tied])

"""`DecodeTask` (src/decode/mod.rs:220-418) around a random-init Whisper-shaped forward: the caller of the hot path.

CPU part: BASELINE config 1 in the oracle (one 30 s clip, log-mel + greedy decode, tiny.en-shaped random-init weights,
V = 51864).  GPU part: the same loop with the CUDA library doing everything between "logits out" and "tokens in"
(no-speech probability, filters, log-softmax, top-k, beam merge, KV-cache reorder), checked step by step against the
oracle on the logits of the SAME forward; BASELINE config 4 (1 h = 120 segments, base-shaped, timestamp mode) bounded
to what the CPU oracle finishes in seconds, plus grammar properties over all 120 segments.
"""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host"))
import toy_whisper as tw                                              # noqa: E402

from oracle import audio as oa                                        # noqa: E402
from oracle import decode as od                                       # noqa: E402
from whisper_rs_b200 import constants as C                            # noqa: E402
from whisper_rs_b200 import synth                                     # noqa: E402


def timestamp_grammar_ok(seq, ids, max_initial=50):
    """What the three filters guarantee about a sampled sequence (timestamp_tokens.rs:47-78, suppress_blanks.rs)."""
    if not seq:
        return True
    if not (ids.ts_begin <= seq[0] <= ids.ts_begin + max_initial):    # first sampled token: a timestamp <= 1.0 s
        return False
    for i in range(1, len(seq)):
        last_ts = seq[i - 1] >= ids.ts_begin
        penult_ts = i < 2 or seq[i - 2] >= ids.ts_begin
        if last_ts and penult_ts and seq[i] >= ids.ts_begin:          # after a pair: text (or EOT) only
            return False
        if last_ts and not penult_ts and seq[i] < ids.eot:            # after text + one timestamp: no plain text
            return False
    return ids.no_timestamps not in seq


def test_config1_cpu_oracle_pipeline():
    """BASELINE config 1: examples/transcribe.rs on CPU, greedy (intent of greedy.rs), tiny.en-shaped random init."""
    ids = C.GPT2_EN
    torch.manual_seed(0)
    x = torch.from_numpy(synth.audio_segment("gaussian", 0))
    mel = oa.pad_or_trim(oa.log_mel_spectrogram(x, oa.mel_filter()), C.N_FRAMES)[None]
    assert mel.shape == (1, 80, 3000)
    model = tw.ToyWhisper(tw.TINY_EN, device="cpu", seed=0, cache_backend="oracle", eot=ids.eot, eot_bias=30.0, eot_after=5)
    opt = od.DecodeOptions(greedy=True, sample_len=10)
    out = od.decode_task_reference(model, ids, opt, mel)
    again = od.decode_task_reference(model, ids, opt, mel)
    assert len(out) == 1 and out[0].tokens == again[0].tokens
    assert 0 < len(out[0].tokens) <= 10 and ids.eot not in out[0].tokens
    assert timestamp_grammar_ok(out[0].tokens, ids)
    assert 0.0 <= out[0].no_speech_prob <= 1.0 and out[0].avg_logprob <= 0.0
    assert out[0].n_steps < 10                                         # the EOT bias ended the loop


# ------------------------------------------------------------------------------------------------- GPU
def run_both(dims, ids, n_audio, *, beam=5, greedy=False, group=1, patience=1.0, prompt=None, sample_len=10, fused=True,
             eot_bias=12.0, eot_after=4, seed=3, mel=None, len_penalty=None, flag_cross=False):
    from whisper_rs_b200 import decode as wd
    from whisper_rs_b200.decode.task import DecodeTask
    if mel is None:
        g = torch.Generator().manual_seed(seed)
        mel = torch.randn(n_audio, 80, 3000, generator=g).cuda()
    mk = lambda backend: tw.ToyWhisper(dims, "cuda", seed, backend, eot=ids.eot, eot_bias=eot_bias, eot_after=eot_after,
                                       flag_cross_invariant=flag_cross and backend == "wrs")
    mode = wd.TokenExtractMode.Greedy(group) if greedy else wd.TokenExtractMode.BeamSearch(beam, patience)
    task = DecodeTask(mk("wrs"), ids, wd.DecodeOptions(token_extract_mode=mode, sample_len=sample_len, len_penalty=len_penalty),
                      prompt=prompt, max_audio=n_audio, fused=fused)
    got = task.run(mel)
    oopt = od.DecodeOptions(beam_size=beam, patience=patience, greedy=greedy, greedy_group=group, sample_len=sample_len, len_penalty=len_penalty)
    oprompt = None if prompt is None else torch.tensor(prompt, dtype=torch.int64)
    omodel = mk("oracle")
    want = od.decode_task_reference(omodel, ids, oopt, mel, oprompt)
    return got, want, task, omodel


def assert_same(got, want):
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g.tokens == w.tokens
        assert abs(g.avg_logprob - w.avg_logprob) <= 1e-5 * max(1.0, abs(w.avg_logprob))
        assert abs(g.no_speech_prob - w.no_speech_prob) <= 1e-5 * max(w.no_speech_prob, 1e-30) + 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("prompt", [None, [1212, 318, 257, 1332]])
def test_beam_task_matches_oracle(fused, prompt):
    ids = C.MULTILINGUAL
    got, want, task, omodel = run_both(tw.ModelDims(), ids, 4, prompt=prompt, fused=fused, sample_len=12)
    assert_same(got, want)
    assert task.model.reorders == omodel.reorders > 0                  # the KV reorder ran on the device every step
    for g in got:
        assert timestamp_grammar_ok(g.tokens, ids)


@pytest.mark.gpu
def test_beam_task_with_cross_caches_flagged_group_invariant():
    """The cross-attention K / V never change under the beam reorder (identical rows within a beam group): flagged, the
    library skips them -- same tokens, same scores as the oracle backend that gathers everything; the device-side
    verification (option kv_verify_group) finds nothing to object to."""
    from whisper_rs_b200 import _lib
    ids = C.MULTILINGUAL
    ctx = _lib.context(0)
    ctx.set_option("kv_verify_group", 5)
    try:
        got, want, task, omodel = run_both(tw.ModelDims(), ids, 4, sample_len=12, flag_cross=True)
        torch.cuda.synchronize()
        ctx.check()
    finally:
        ctx.set_option("kv_verify_group", 0)
    assert_same(got, want)
    assert task.model.reorders == omodel.reorders > 0


@pytest.mark.gpu
def test_beam_task_patience_and_length_penalty():
    ids = C.MULTILINGUAL
    got, want, _, _ = run_both(tw.ModelDims(), ids, 3, beam=4, patience=2.0, len_penalty=0.6, sample_len=14, eot_bias=9.0)
    assert_same(got, want)


@pytest.mark.gpu
def test_greedy_task_matches_oracle():
    ids = C.GPT2_EN
    dims = tw.ModelDims(n_vocab=ids.n_vocab)
    got, want, _, _ = run_both(dims, ids, 5, greedy=True, sample_len=9, eot_bias=25.0)
    assert_same(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("fused", [True, False])
def test_greedy_task_with_groups_of_three(fused):
    """TokenExtractMode::Greedy(3) (src/decode/extractor/mod.rs:14-18): three identical rows per audio; every audio must
    get the best of ITS OWN group (finalize groups rows by audio), i.e. what Greedy(1) gives it."""
    ids = C.GPT2_EN
    dims = tw.ModelDims(n_vocab=ids.n_vocab)
    got, want, _, _ = run_both(dims, ids, 4, greedy=True, group=3, sample_len=9, eot_bias=6.0, fused=fused)
    assert_same(got, want)
    single, _, _, _ = run_both(dims, ids, 4, greedy=True, group=1, sample_len=9, eot_bias=6.0, fused=fused)
    assert [g.tokens for g in got] == [g.tokens for g in single]
    assert [g.avg_logprob for g in got] == pytest.approx([g.avg_logprob for g in single], rel=1e-6)
    assert len({(tuple(g.tokens), round(g.avg_logprob, 6)) for g in got}) > 1   # (the audios do differ: a mix-up would show)


@pytest.mark.gpu
def test_config4_long_form_base_shaped():
    """BASELINE config 4: 1 h of AM-noise cut into 120 x 30 s segments (per-segment mode), front-end + base-shaped
    random-init decode in timestamp mode.  Parity on the first 6 segments (what the CPU oracle finishes in seconds),
    grammar + determinism properties over all 120."""
    from whisper_rs_b200 import audio as wa
    from whisper_rs_b200 import decode as wd
    from whisper_rs_b200.decode.task import DecodeTask
    ids = C.MULTILINGUAL
    n_seg = 120
    g = torch.Generator(device="cuda").manual_seed(0)
    t = torch.arange(n_seg * C.N_SAMPLES, device="cuda", dtype=torch.float32) / C.SAMPLE_RATE
    hour = 0.3 * torch.randn(n_seg * C.N_SAMPLES, device="cuda", generator=g) * (0.5 + 0.5 * torch.sin(2 * np.pi * 3 * t)) ** 4
    pcm = hour.view(n_seg, C.N_SAMPLES)
    mel = wa.log_mel_spectrogram_batch(pcm)
    want_mel = oa.log_mel_batch(pcm[:6].cpu(), oa.mel_filter())
    assert (mel[:6].cpu() - want_mel).abs().max().item() <= 1e-4
    got, want, _, _ = run_both(tw.BASE, ids, 6, sample_len=8, mel=mel[:6].contiguous(), seed=11)
    assert_same(got, want)
    # all 120 segments, 32 audios (160 rows) at a time
    model = tw.ToyWhisper(tw.BASE, "cuda", 11, "wrs", eot=ids.eot, eot_bias=12.0, eot_after=4)
    task = DecodeTask(model, ids, wd.DecodeOptions(sample_len=8), max_audio=32)
    outs = []
    for a in range(0, n_seg, 32):
        outs += task.run(mel[a:a + 32].contiguous())
    assert len(outs) == n_seg
    for o in outs:
        assert timestamp_grammar_ok(o.tokens, ids) and np.isfinite(o.avg_logprob) and 0.0 <= o.no_speech_prob <= 1.0

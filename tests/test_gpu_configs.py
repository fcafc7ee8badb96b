"""BASELINE.json configs at their stated sizes that the per-kernel test files do not already cover:

  config 3 variant -- "logits from a random-init decoder" (SURVEY.md §8d): x @ token_emb^T with tch's default N(0,1)
                      embedding init, 160 rows x 51865, against the oracle step;
  config 4 variant -- the hour of audio as ONE clip (src/transcribe.rs:111-112 semantics: reflect padding and max-8
                      floor over the whole clip), against the oracle at full size;
  config 5         -- 10 000 segments (19.2 GB of PCM) through the front-end and 50 000 rows through the beam step on one
                      GPU, checked through tiling: the bank the inputs are tiled from is checked against the oracle,
                      every tile must reproduce the bank's result bit for bit.
"""
import numpy as np
import pytest
import torch

from oracle import audio as oa
from oracle import decode as od
from whisper_rs_b200 import audio, synth
from whisper_rs_b200 import constants as C
from whisper_rs_b200 import decode as wd
from whisper_rs_b200 import shard
from test_gpu_decode import assert_beam_parity, close, fused_step

pytestmark = pytest.mark.gpu
IDS = C.MULTILINGUAL
TOL = 1e-4


def test_config3_logits_from_random_init_projection():
    n_audio, k, d = 32, 5, 512
    case = synth.decode_case(IDS, n_audio, k, "mixed", seed=77)
    g = torch.Generator().manual_seed(77)
    x = torch.nn.functional.layer_norm(torch.randn(n_audio * k, d, generator=g), (d,)).cuda()
    w = torch.randn(IDS.n_vocab, d, generator=g).cuda()                   # nn::embedding default init (model.rs:470)
    case["logits"] = (x @ w.T).cpu().numpy()                               # the projection of src/model.rs:521
    assert case["logits"].std() > 15.0                                     # sqrt(512): a very peaked softmax
    want = od.decode_step_reference(case, IDS)
    m = want["margins"][np.isfinite(want["margins"])]
    assert np.abs(m).min() > 1e-3                                          # no timestamp-vs-text decision on the fp edge
    got, _ = fused_step(IDS, case)
    assert_beam_parity(got, want, k)


def test_config4_hour_as_one_clip():
    n = 120 * C.N_SAMPLES
    g = torch.Generator(device="cuda").manual_seed(4)
    t = torch.arange(n, device="cuda", dtype=torch.float32) / C.SAMPLE_RATE
    clip = 0.3 * torch.randn(n, device="cuda", generator=g) * (0.5 + 0.5 * torch.sin(2 * np.pi * 3 * t)) ** 4
    del t
    got = audio.log_mel_spectrogram(clip, audio.mel_filter())
    assert got.shape == (80, 120 * C.N_FRAMES)
    want = oa.log_mel_spectrogram(clip.cpu(), oa.mel_filter())
    assert (got.cpu() - want).abs().max().item() <= TOL
    # the same samples cut into 120 independent segments: away from the cuts (2 frames) and above both floors the
    # values are the same arithmetic on the same samples
    per_seg = audio.log_mel_spectrogram_batch(clip.view(120, C.N_SAMPLES))
    whole = got.view(80, 120, C.N_FRAMES).permute(1, 0, 2)
    floor = torch.maximum(whole.amin(), per_seg.amin(dim=(1, 2), keepdim=True))
    inner = torch.zeros(C.N_FRAMES, dtype=torch.bool, device="cuda")
    inner[2:-2] = True
    sel = (whole > floor + 1e-6) & (per_seg > floor + 1e-6) & inner
    assert sel.float().mean().item() > 0.5
    assert (whole[sel] - per_seg[sel]).abs().max().item() <= 1e-6


def test_config5_ten_thousand_segments_front_end():
    n_seg = 10000
    assert shard.all_ranges(n_seg, 8)[-1][1] == n_seg and shard.segment_range(n_seg, 0, 1) == (0, n_seg)
    bank_np = synth.audio_batch(64, seed=5)
    bank = torch.from_numpy(bank_np).cuda()
    bank_out = audio.log_mel_spectrogram_batch(bank)
    for s in (0, 4, 8, 10, 11):                                            # one of each kind against the oracle
        want = oa.log_mel_spectrogram(torch.from_numpy(bank_np[s]), oa.mel_filter())
        assert (bank_out[s].cpu() - want).abs().max().item() <= TOL
    pcm = bank.repeat(n_seg // 64 + 1, 1)[:n_seg].contiguous()            # 19.2 GB resident
    out = audio.log_mel_spectrogram_batch(pcm)                             # 9.6 GB
    torch.cuda.synchronize()
    assert out.shape == (n_seg, 80, C.N_FRAMES)
    for a in range(0, n_seg, 64):
        b = min(a + 64, n_seg)
        assert torch.equal(out[a:b], bank_out[:b - a]), a


def test_config5_fifty_thousand_rows_beam_step():
    k, n_audio_bank = 5, 32
    case = synth.decode_case(IDS, n_audio_bank, k, "mixed", seed=55)
    want = od.decode_step_reference(case, IDS)
    got_bank, _ = fused_step(IDS, case)
    assert_beam_parity(got_bank, want, k)
    rows_bank = n_audio_bank * k
    n_rows = 25000                                                         # two such batches = 50 000 rows (10 000 x 5)
    reps = n_rows // rows_bank + 1
    prompt = wd.DecodePrompt(None, IDS)
    prompt.sample_begin = case["sample_begin"]
    step = wd.FusedDecodeStep(wd.DecodeOptions(), IDS, [prompt], max_rows=n_rows)
    logits = torch.from_numpy(case["logits"]).cuda().repeat(reps, 1)[:n_rows].contiguous()        # 5.2 GB
    tokens = torch.from_numpy(case["tokens"]).cuda().repeat(reps, 1)[:n_rows].contiguous()
    sums = torch.from_numpy(case["sum_logprobs"]).cuda().repeat(reps)[:n_rows].contiguous()
    for batch in range(2):
        step.reset()
        out = step.step(tokens, logits, sums)
        torch.cuda.synchronize()
        step.decoder.ctx.check()
        nt, src, so = out["next_tokens"].cpu().numpy(), out["source_idx"].cpu().numpy(), out["sum_logprobs"].cpu().numpy()
        for a in range(0, n_rows, rows_bank):
            b = min(a + rows_bank, n_rows)
            assert np.array_equal(nt[a:b], got_bank["next_tokens"][:b - a]), (batch, a)
            assert np.array_equal(src[a:b] - a, got_bank["source_idx"][:b - a]), (batch, a)
            # (a row's units are dealt to different warps at different row positions: same sums, another order)
            assert close(so[a:b], got_bank["sum_logprobs"][:b - a]), (batch, a)

"""Parity of the CUDA log-mel front-end (through the C ABI) with the oracle -- reference src/audio.rs:230-259.

Tolerance (north_star): 1e-4 absolute after normalisation.  The oracle issues the reference's ATen op
sequence on the CPU; both sides see the same seeded synthetic PCM bytes."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import audio as oa
from whisper_rs_b200 import _lib, audio, synth
from whisper_rs_b200 import constants as C

pytestmark = pytest.mark.gpu
TOL = 1e-4
VEC = np.load(os.path.join(GOLDEN, "oracle_vectors.npz"))


def oracle_batch(pcm: np.ndarray) -> np.ndarray:
    filt = oa.mel_filter()
    return np.stack([oa.log_mel_spectrogram(torch.from_numpy(row.copy()), filt).numpy() for row in pcm])


def run_cuda(pcm: np.ndarray, **kw) -> np.ndarray:
    out = audio.log_mel_spectrogram_batch(torch.from_numpy(pcm).cuda(), **kw)
    torch.cuda.synchronize()
    _lib.context(0).check()
    return out.cpu().numpy()


@pytest.mark.parametrize("kind", synth.AUDIO_KINDS)
def test_golden_vectors(kind):
    seed, n = [int(v) for v in VEC[f"logmel_{kind}_in_seed"]]
    got = run_cuda(synth.audio_segment(kind, seed, n)[None])[0]
    assert got.shape == (80, n // 160)
    assert np.abs(got - VEC[f"logmel_{kind}_out"]).max() <= TOL


@pytest.mark.parametrize("kind", synth.AUDIO_KINDS)
def test_full_segment_each_kind(kind):
    pcm = synth.audio_segment(kind, 21, C.N_SAMPLES)[None]
    got, want = run_cuda(pcm), oracle_batch(pcm)
    assert got.shape == (1, 80, 3000)
    assert np.abs(got - want).max() <= TOL
    if kind == "zeros":
        assert np.all(got == -1.5)      # log10(1e-10) = -10 everywhere -> (-10 + 4) / 4, exactly


@pytest.mark.parametrize("fused", [1, 0])
def test_batch_mixture_both_modes(fused):
    ctx = _lib.context(0)
    pcm = synth.audio_batch(12, seed=100)
    want = oracle_batch(pcm)
    ctx.set_option("logmel_fused", fused)
    try:
        got = run_cuda(pcm)
        assert ctx.get_stat("logmel_mode") == fused
    finally:
        ctx.set_option("logmel_fused", -1)
    assert np.abs(got - want).max() <= TOL
    # each segment has its own max-8 floor: a loud neighbour must not raise a quiet one's floor
    assert abs((got[11].max() - got[11].min()) - 2.0) < 1e-5


@pytest.mark.parametrize("n", [400, 401, 479, 480, 1000, 16000 + 37, 20480, 160 * 129, 160 * 257 + 159])
def test_ragged_lengths(n):
    pcm = np.stack([synth.audio_segment("am_noise", 3, n), synth.audio_segment("chirp", 4, n)])
    got, want = run_cuda(pcm), oracle_batch(pcm)
    assert got.shape == want.shape == (2, 80, n // 160)
    assert np.abs(got - want).max() <= TOL


def test_strided_and_unaligned_input():
    # seg_stride > n_samples, base pointer only 4-byte aligned: the scalar load path
    n = 160 * 300
    base = torch.from_numpy(synth.audio_batch(3, seed=5, n_samples=n + 7)).cuda()
    view = base[:, 3:3 + n]
    assert view.data_ptr() % 16 != 0 and view.stride(0) == n + 7
    out = audio.log_mel_spectrogram_batch(view)
    torch.cuda.synchronize()
    want = oracle_batch(view.cpu().numpy())
    assert np.abs(out.cpu().numpy() - want).max() <= TOL


def test_reference_signature_whole_clip_and_custom_filter():
    # log_mel_spectrogram(audio [N], mel_filter [80,201]) with a 3-minute clip: one "segment", global max (audio.rs:256)
    x = synth.audio_segment("am_noise", 9, 180 * 16000)
    filt = audio.mel_filter()
    got = audio.log_mel_spectrogram(torch.from_numpy(x).cuda(), filt)
    want = oa.log_mel_spectrogram(torch.from_numpy(x), oa.mel_filter())
    assert got.shape == (80, 18000)
    assert (got.cpu() - want).abs().max().item() <= TOL
    # a caller-supplied filterbank (the `mel_filter: &Tensor` argument is honoured, not assumed)
    rng = np.random.default_rng(0)
    custom = filt.clone() * torch.from_numpy(rng.uniform(0.5, 1.5, size=(80, 1)).astype(np.float32))
    seg = x[:C.N_SAMPLES]
    got = audio.log_mel_spectrogram(torch.from_numpy(seg).cuda(), custom)
    want = oa.log_mel_spectrogram(torch.from_numpy(seg), custom)
    assert (got.cpu() - want).abs().max().item() <= TOL
    got = audio.log_mel_spectrogram(torch.from_numpy(seg).cuda(), filt)   # and back to the built-in values
    want = oa.log_mel_spectrogram(torch.from_numpy(seg), oa.mel_filter())
    assert (got.cpu() - want).abs().max().item() <= TOL


def test_custom_filterbanks_of_every_band_class():
    """K1 projects a mel row over 4, 8 or 16 bins (band zero-padded) or, wider than that, over its own length: banded
    filterbanks that put rows into every class, including bands that end at bin 200."""
    pcm = synth.audio_batch(4, seed=77, n_samples=160 * 700)
    rng = np.random.default_rng(5)
    for widths in ([6] * 80, [3, 8, 16, 20, 1, 12] * 13 + [2, 9]):
        dense = np.zeros((80, 201), dtype=np.float32)
        for m, wd in enumerate(widths):
            lo = min(2 * m + (m % 3), 201 - wd)
            dense[m, lo:lo + wd] = rng.uniform(0.002, 0.02, size=wd).astype(np.float32)
        filt = torch.from_numpy(dense)
        for s_ in (1, 3):
            x = torch.from_numpy(pcm[s_])
            got = audio.log_mel_spectrogram(x.cuda(), filt)
            assert (got.cpu() - oa.log_mel_spectrogram(x, filt)).abs().max().item() <= TOL
    x = torch.from_numpy(pcm[0])
    got = audio.log_mel_spectrogram(x.cuda(), audio.mel_filter())          # back to the built-in filterbank
    assert (got.cpu() - oa.log_mel_spectrogram(x, oa.mel_filter())).abs().max().item() <= TOL


def test_long_clip_takes_the_two_launch_form():
    """a whole clip as one segment (src/transcribe.rs:112, maximum over the clip).  More slices than
    the grid holds in one round cannot use the in-kernel hand-over (waits could become circular): two launches."""
    from whisper_rs_b200 import _lib
    x = synth.audio_segment("am_noise", 11, 12 * 60 * 16000)                # 72 000 frames, 1125 slices
    got = audio.log_mel_spectrogram(torch.from_numpy(x).cuda(), audio.mel_filter())
    assert _lib.context(0).get_stat("logmel_mode") == 0
    want = oa.log_mel_spectrogram(torch.from_numpy(x), oa.mel_filter())
    assert got.shape == (80, 72000)
    assert (got.cpu() - want).abs().max().item() <= TOL
    # seek windowing of the clip's mel (src/transcribe.rs:150-152): narrow to a window, pad_or_trim to 3000 frames
    for seek in (0, 40000, 71000):
        win = got[:, seek:seek + C.N_FRAMES]
        w = audio.pad_or_trim(win, C.N_FRAMES)
        ref = oa.pad_or_trim(want[:, seek:seek + C.N_FRAMES], C.N_FRAMES)
        assert w.shape == (80, C.N_FRAMES) and torch.equal(w.cpu(), torch.where(ref == 0, ref, w.cpu()))
        assert (w.cpu() - ref).abs().max().item() <= TOL


def test_host_entry_point_matches_device_entry_point():
    pcm = synth.audio_batch(5, seed=40, n_samples=160 * 500)
    host = audio.log_mel_spectrogram_batch(torch.from_numpy(pcm))       # wrs_logmel_batch_host
    assert not host.is_cuda
    assert np.array_equal(host.numpy(), run_cuda(pcm))


def test_pad_or_trim_device():
    a = torch.arange(80 * 37, dtype=torch.float32).reshape(80, 37).cuda()
    assert torch.equal(audio.pad_or_trim(a, 20).cpu(), oa.pad_or_trim(a.cpu(), 20))
    assert torch.equal(audio.pad_or_trim(a, 3000).cpu(), oa.pad_or_trim(a.cpu(), 3000))
    assert torch.equal(audio.pad_or_trim(a, 37).cpu(), a.cpu())


def test_bad_arguments_raise():
    with pytest.raises(ValueError):
        audio.log_mel_spectrogram_batch(torch.zeros(2, 399).cuda())
    with pytest.raises(ValueError):
        audio.log_mel_spectrogram(torch.zeros(2, 4000).cuda(), audio.mel_filter())
    with pytest.raises(_lib.WrsError):
        _lib.check(_lib.lib().wrs_logmel_batch(_lib.context(0).handle, None, 1, 480000, 480000, None, None),
                   _lib.context(0).handle)


def test_config2_batch64_properties():
    """BASELINE config 2 at full size: 64 x 30 s.  Oracle on a sample of segments; size-independent
    properties on all of them (batch invariance, per-segment range, determinism)."""
    pcm = synth.audio_batch(64, seed=0)
    x = torch.from_numpy(pcm).cuda()
    got = audio.log_mel_spectrogram_batch(x)
    again = audio.log_mel_spectrogram_batch(x)
    torch.cuda.synchronize()
    assert torch.equal(got, again)                                   # bit-stable run to run
    for s in (0, 5, 9, 10, 11, 63):
        want = oa.log_mel_spectrogram(torch.from_numpy(pcm[s]), oa.mel_filter())
        assert (got[s].cpu() - want).abs().max().item() <= TOL, s
    # a segment's result does not depend on what else is in the batch, nor on its position
    perm = torch.randperm(64, generator=torch.Generator().manual_seed(1))
    shuffled = audio.log_mel_spectrogram_batch(x[perm.cuda()].contiguous())
    assert torch.equal(shuffled, got[perm.cuda()])
    alone = audio.log_mel_spectrogram_batch(x[17:18].contiguous())
    assert torch.equal(alone[0], got[17])
    rng_ = got.amax(dim=(1, 2)) - got.amin(dim=(1, 2))
    assert torch.all(rng_ <= 2.0 + 1e-6)                             # max-8 floor then /4


def test_fused_floor_skipping_is_bit_identical_to_the_unconditional_floor():
    """The fused kernel leaves a slice alone when its minimum is already at or above max - 8 (src/audio.rs:256 would
    not change it); the two-launch form floors every element.  Same bits, on a batch deep enough (110 x 47 slices over
    148 CTAs = 35 per CTA) to wrap the per-CTA ring of slice minima, with kinds that skip (gaussian, zeros) and kinds
    that do not (tone, burst, chirp) interleaved."""
    ctx = _lib.context(0)
    pcm = synth.audio_batch(110, seed=7)
    x = torch.from_numpy(pcm).cuda()
    fused = audio.log_mel_spectrogram_batch(x)
    assert ctx.get_stat("logmel_mode") == 1
    ctx.set_option("logmel_fused", 0)
    try:
        plain = audio.log_mel_spectrogram_batch(x)
        assert ctx.get_stat("logmel_mode") == 0
    finally:
        ctx.set_option("logmel_fused", 1)
    torch.cuda.synchronize()
    ctx.check()
    assert torch.equal(fused, plain)
    for s in (0, 1, 2, 3, 4, 5, 109):
        want = oa.log_mel_spectrogram(torch.from_numpy(pcm[s]), oa.mel_filter())
        assert (fused[s].cpu() - want).abs().max().item() <= TOL, s


@pytest.mark.parametrize("channels", [1, 2, 3, 6])
def test_pcm_i16_conversion_and_downmix_bit_exact(channels):
    """src/audio.rs:116-131: s / 32768 and the per-frame mean summed left to right in f32 -- integer input, bit-exact."""
    rng = np.random.default_rng(channels)
    for n in (1, 7, 4096 + 3, 160000):
        pcm = rng.integers(-32768, 32768, size=n * channels, dtype=np.int16)
        pcm[:channels] = -32768
        pcm[-channels:] = 32767
        got = audio.pcm_i16_to_f32(torch.from_numpy(pcm).cuda(), channels).cpu().numpy()
        assert np.array_equal(got, oa.pcm_i16_to_mono_f32(pcm, channels))
    odd = torch.from_numpy(rng.integers(-3000, 3000, size=1001, dtype=np.int16)).cuda()[1:]       # 2-byte aligned only
    assert np.array_equal(audio.pcm_i16_to_f32(odd, 1).cpu().numpy(), oa.pcm_i16_to_mono_f32(odd.cpu().numpy(), 1))


def test_host_i16_entry_point_matches_f32_entry_point():
    pcm = synth.audio_batch(6, seed=41, n_samples=160 * 500 + 80)
    i16 = np.round(np.clip(pcm, -1, 1) * 32767.0).astype(np.int16)
    got = audio.log_mel_spectrogram_batch_i16(torch.from_numpy(i16))
    conv = np.stack([oa.pcm_i16_to_mono_f32(r) for r in i16])
    assert np.array_equal(got.numpy(), run_cuda(conv))                      # same samples -> same bits
    assert np.abs(got.numpy() - oracle_batch(conv)).max() <= TOL

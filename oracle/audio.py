"""Oracle for the log-mel front-end -- restates reference src/audio.rs:175-259 (TEST INFRASTRUCTURE).

Every function issues the same ATen operators, in the same order, as the reference does through tch.
"""
from __future__ import annotations

import numpy as np
import torch

SAMPLE_RATE = 16000   # src/audio.rs:12
N_FFT = 400           # src/audio.rs:13
N_MELS = 80           # src/audio.rs:14
HOP_LENGTH = 160      # src/audio.rs:15
N_SAMPLES = 480000    # src/audio.rs:17
N_FRAMES = 3000       # src/audio.rs:18


def _hz_to_mel(f):
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, mels)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * m
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), freqs)


def mel_filter_np(sr: int = SAMPLE_RATE, n_fft: int = N_FFT, n_mels: int = N_MELS) -> np.ndarray:
    """src/audio.rs:175-188: `mel_filter::mel::<f32>(16000, 400, Some(80), None, None, false,
    NormalizationFactor::One)` -- crate mel_filter 0.1.1 (Cargo.lock:567-570, not vendored), a port of
    librosa.filters.mel(htk=False, norm="slaney").  Restated from the published algorithm in float64,
    rounded to f32; pinned by the reference KAT (tests/golden/reference_kat.json)."""
    fmin, fmax = 0.0, sr / 2.0
    fft_freqs = np.linspace(0.0, sr / 2.0, n_fft // 2 + 1)
    mel_pts = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_pts)
    ramps = mel_pts[:, None] - fft_freqs[None, :]
    w = np.zeros((n_mels, n_fft // 2 + 1), dtype=np.float64)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0.0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_pts[2:n_mels + 2] - mel_pts[:n_mels])   # slaney area normalisation
    w *= enorm[:, None]
    return w.astype(np.float32)


def mel_filter() -> torch.Tensor:
    """src/audio.rs:175-188 -> [80, 201] f32 tensor (`Tensor::of_slice2`)."""
    return torch.from_numpy(mel_filter_np())


def pad_or_trim(audio: torch.Tensor, length: int) -> torch.Tensor:
    """src/audio.rs:213-228: slice the last dim to `length`, or right-pad it with zeros."""
    n = audio.shape[-1]
    if n > length:
        return audio[..., :length]
    pad = [0] * (audio.dim() * 2)
    pad[1] = length - n
    return torch.nn.functional.pad(audio, pad, mode="constant", value=0.0)


def log_mel_spectrogram(audio: torch.Tensor, mel_filter: torch.Tensor) -> torch.Tensor:
    """src/audio.rs:230-259.  `audio` must be 1-D f32 (the reference's slice(1, ..) drops the last
    *frame* only for 1-D input).  Returns [80, len // 160] f32."""
    assert audio.dim() == 1 and audio.dtype == torch.float32
    window = torch.hann_window(N_FFT, dtype=torch.float32)                       # :233
    stft = torch.stft(audio, N_FFT, HOP_LENGTH, None, window, center=True,       # :236-246
                      pad_mode="reflect", normalized=False, onesided=True, return_complex=True)
    magnitudes = stft[:, :-1].abs().pow(2)                                       # :247
    mel_spec = mel_filter.matmul(magnitudes)                                     # :252
    log_spec = mel_spec.clamp_min(1e-10).log10()                                 # :254
    log_spec = log_spec.maximum(log_spec.max() - 8.0)                            # :256
    return (log_spec + 4.0) / 4.0                                                # :258


def log_mel_batch(pcm: torch.Tensor, mel_filter: torch.Tensor) -> torch.Tensor:
    """Per-segment semantics of the batched kernel: the reference function
    applied to every row of `pcm` [S, n] independently -> [S, 80, n // 160]."""
    return torch.stack([log_mel_spectrogram(row, mel_filter) for row in pcm])


def log_mel_spectrogram_f64(audio: np.ndarray, mel_filter: np.ndarray) -> np.ndarray:
    """Same sequence in float64 (numpy) -- used only to size the fp32 error budget in tests."""
    x = audio.astype(np.float64)
    n = len(x)
    xp = np.concatenate([x[1:201][::-1], x, x[-201:-1][::-1]])
    w = 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(N_FFT) / N_FFT)
    n_frames = n // HOP_LENGTH
    idx = np.arange(N_FFT)[None, :] + HOP_LENGTH * np.arange(n_frames)[:, None]
    spec = np.fft.rfft(xp[idx] * w[None, :], axis=1)
    power = (spec.real ** 2 + spec.imag ** 2).T
    mel = mel_filter.astype(np.float64) @ power
    ls = np.log10(np.maximum(mel, 1e-10))
    ls = np.maximum(ls, ls.max() - 8.0)
    return (ls + 4.0) / 4.0


def pcm_i16_to_mono_f32(pcm: np.ndarray, channels: int = 1) -> np.ndarray:
    """reference src/audio.rs:116-131 between the codec and the tensor: symphonia's SampleBuffer<f32> conversion of
    16-bit samples (s / 32768, exact in f32) and the downmix `c.iter().sum::<f32>() / c.len() as f32` of every
    interleaved frame -- summed left to right in f32.  pcm: int16 [n_frames * channels] interleaved."""
    x = pcm.astype(np.float32) * np.float32(1.0 / 32768.0)
    if channels == 1:
        return x
    fr = x.reshape(-1, channels)
    acc = np.zeros(fr.shape[0], dtype=np.float32)
    for c in range(channels):
        acc = (acc + fr[:, c]).astype(np.float32)
    return (acc / np.float32(channels)).astype(np.float32)

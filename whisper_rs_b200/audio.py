"""Host-side mirror of reference src/audio.rs (the front-end part): same names, same argument meaning,
same error behaviour (exceptions where the reference panics), computed by the CUDA library.

    mel_filter()                         src/audio.rs:175-188
    pad_or_trim(audio, len)              src/audio.rs:213-228
    log_mel_spectrogram(audio, filter)   src/audio.rs:230-259   (1-D audio, whole clip)
    log_mel_spectrogram_batch(pcm, filter)  -- the batched per-segment surface
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .constants import HOP_LENGTH, N_BINS, N_FFT, N_MELS, N_SAMPLES, N_FRAMES, SAMPLE_RATE, CHUNK_LENGTH  # noqa: F401

_filter_cache = {}   # device -> "builtin" or (the caller's filter tensor, its version) currently loaded in that device's context


def mel_filter() -> torch.Tensor:
    """[80, 201] f32 slaney filterbank on the CPU (reference returns a CPU tensor too)."""
    out = np.empty((N_MELS, N_BINS), dtype=np.float32)
    _lib.check(_lib.lib().wrs_mel_filter_host(out.ctypes.data_as(C.c_void_p)))
    return torch.from_numpy(out)


def _stream_of(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _ensure_filter(ctx: _lib.Context, filt: torch.Tensor | None):
    if filt is None:
        key = "builtin"
        if _filter_cache.get(ctx.device) != key:
            _lib.check(_lib.lib().wrs_logmel_set_filter(ctx.handle, None), ctx.handle)
            _filter_cache[ctx.device] = key
        return
    if tuple(filt.shape) != (N_MELS, N_BINS):
        raise ValueError(f"mel_filter must be [{N_MELS}, {N_BINS}], got {tuple(filt.shape)}")
    # the entry holds the tensor itself: its address cannot be reused by another filter while it is cached
    cur = _filter_cache.get(ctx.device)
    if isinstance(cur, tuple) and cur[0] is filt and cur[1] == filt._version:
        return
    host = filt.detach().to("cpu", torch.float32).contiguous().numpy()
    _lib.check(_lib.lib().wrs_logmel_set_filter(ctx.handle, host.ctypes.data_as(C.c_void_p)), ctx.handle)
    _filter_cache[ctx.device] = (filt, filt._version)


def pad_or_trim(audio: torch.Tensor, length: int) -> torch.Tensor:
    """Slice the last dimension to `length`, or right-pad it with zeros.  CUDA tensors go through
    wrs_pad_or_trim; CPU tensors are views/pads done by the caller's own framework in the reference as well,
    so they are handled with a device round trip only when they are CUDA."""
    n = audio.shape[-1]
    if audio.is_cuda:
        if audio.dtype != torch.float32:
            raise TypeError("pad_or_trim: f32 only")
        rows = audio.numel() // max(n, 1) if n else int(np.prod(audio.shape[:-1]))
        src = audio.contiguous()
        out = torch.empty(*audio.shape[:-1], length, dtype=torch.float32, device=audio.device)
        ctx = _lib.context(audio.device.index or 0)
        _lib.check(_lib.lib().wrs_pad_or_trim(ctx.handle, src.data_ptr(), rows, n, n, out.data_ptr(), length,
                                              _stream_of(audio)), ctx.handle)
        return out
    if n > length:
        return audio[..., :length]
    return torch.nn.functional.pad(audio, (0, length - n))


def log_mel_spectrogram_batch(pcm: torch.Tensor, mel_filter: torch.Tensor | None = None, device: int = 0) -> torch.Tensor:
    """pcm [S, n] f32 -> [S, 80, n // 160] f32; every row is treated as the reference treats one clip.
    CUDA input: result on the same device, enqueued on the current stream, no synchronisation (device-side
    failures -- an expired bounded wait -- surface through `_lib.context(dev).check()`, which synchronises).
    CPU input: host buffers in, host tensor out (copies inside the call)."""
    if pcm.dim() != 2 or pcm.dtype != torch.float32:
        raise TypeError("log_mel_spectrogram_batch: expected a 2-D f32 tensor")
    S, n = pcm.shape
    if n < N_FFT:
        raise ValueError("log_mel_spectrogram_batch: need at least 400 samples (reflect padding)")
    if pcm.stride(1) != 1:
        pcm = pcm.contiguous()
    n_frames = n // HOP_LENGTH
    if pcm.is_cuda:
        ctx = _lib.context(pcm.device.index or 0)
        _ensure_filter(ctx, mel_filter)
        out = torch.empty(S, N_MELS, n_frames, dtype=torch.float32, device=pcm.device)
        _lib.check(_lib.lib().wrs_logmel_batch(ctx.handle, pcm.data_ptr(), S, n, pcm.stride(0) if S > 1 else n,
                                               out.data_ptr(), _stream_of(pcm)), ctx.handle)
        return out
    ctx = _lib.context(device)
    _ensure_filter(ctx, mel_filter)
    out = torch.empty(S, N_MELS, n_frames, dtype=torch.float32)
    _lib.check(_lib.lib().wrs_logmel_batch_host(ctx.handle, pcm.data_ptr(), S, n, pcm.stride(0) if S > 1 else n,
                                                out.data_ptr()), ctx.handle)
    return out


def log_mel_spectrogram(audio: torch.Tensor, mel_filter: torch.Tensor) -> torch.Tensor:
    """Reference signature: 1-D f32 audio (any length >= 400), [80, 201] filter -> [80, len // 160]."""
    if audio.dim() != 1:
        raise ValueError("log_mel_spectrogram: 1-D audio only (the reference's frame slice assumes it, SURVEY.md §0.6)")
    return log_mel_spectrogram_batch(audio.unsqueeze(0), mel_filter)[0]


def pcm_i16_to_f32(pcm: torch.Tensor, channels: int = 1) -> torch.Tensor:
    """16-bit interleaved PCM (CUDA int16 tensor, n_frames * channels samples) -> mono f32 [n_frames]: the conversion and
    downmix of reference src/audio.rs:116-131 on the device (wrs_pcm_i16_to_f32)."""
    if pcm.dtype != torch.int16 or not pcm.is_cuda or pcm.dim() != 1:
        raise TypeError("pcm_i16_to_f32: expected a 1-D CUDA int16 tensor")
    if channels < 1 or pcm.numel() % channels:
        raise ValueError("pcm_i16_to_f32: the sample count is not a multiple of the channel count")
    pcm = pcm.contiguous()
    n = pcm.numel() // channels
    out = torch.empty(n, dtype=torch.float32, device=pcm.device)
    ctx = _lib.context(pcm.device.index or 0)
    _lib.check(_lib.lib().wrs_pcm_i16_to_f32(ctx.handle, pcm.data_ptr(), n, channels, out.data_ptr(), _stream_of(pcm)), ctx.handle)
    return out


def log_mel_spectrogram_batch_i16(pcm: torch.Tensor, mel_filter: torch.Tensor | None = None, device: int = 0) -> torch.Tensor:
    """Host int16 mono PCM [S, n] -> host log-mel [S, 80, n // 160] (wrs_logmel_batch_host_i16: half the bytes on the way in)."""
    if pcm.dim() != 2 or pcm.dtype != torch.int16 or pcm.is_cuda:
        raise TypeError("log_mel_spectrogram_batch_i16: expected a 2-D CPU int16 tensor")
    S, n = pcm.shape
    if n < N_FFT:
        raise ValueError("log_mel_spectrogram_batch_i16: need at least 400 samples (reflect padding)")
    pcm = pcm.contiguous()
    ctx = _lib.context(device)
    _ensure_filter(ctx, mel_filter)
    out = torch.empty(S, N_MELS, n // HOP_LENGTH, dtype=torch.float32)
    _lib.check(_lib.lib().wrs_logmel_batch_host_i16(ctx.handle, pcm.data_ptr(), S, n, n, out.data_ptr()), ctx.handle)
    return out

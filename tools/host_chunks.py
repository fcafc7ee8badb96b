#!/usr/bin/env python
"""Sweep the chunk size of the *_host log-mel pipelines (option host_chunk_segments): segments/s from pinned host buffers."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from whisper_rs_b200 import _lib, synth
L, ctx = _lib.lib(), _lib.context(0)
S, N = 64, 480000
pcm = torch.from_numpy(synth.audio_batch(16, seed=1)).repeat(4, 1).contiguous().pin_memory()
i16 = (pcm.clamp(-1, 1) * 32767).round().to(torch.int16).contiguous().pin_memory()
out = torch.empty(S, 80, 3000).pin_memory()
for ch in (2, 4, 8, 16, 32):
    ctx.set_option("host_chunk_segments", ch)
    res = []
    for fn, buf in ((L.wrs_logmel_batch_host, pcm), (L.wrs_logmel_batch_host_i16, i16)):
        for _ in range(3): _lib.check(fn(ctx.handle, buf.data_ptr(), S, N, N, out.data_ptr()), ctx.handle)
        t0 = time.perf_counter()
        for _ in range(20): fn(ctx.handle, buf.data_ptr(), S, N, N, out.data_ptr())
        res.append(S * 20 / (time.perf_counter() - t0))
    print(f"chunk {ch:3d} segments: f32 {res[0]:9.0f} seg/s   i16 {res[1]:9.0f} seg/s", flush=True)

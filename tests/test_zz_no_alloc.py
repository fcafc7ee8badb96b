"""wrs_logmel_batch_noalloc (include/wrs.h): the log-mel front-end with its scratch in a fixed static device array, for
callers that cannot call cudaMalloc.  Checked through the C ABI against the existing entry point (wrs_logmel_batch) on
two calls in a row -- the array is shared and keeps the previous call's words -- and against the oracle.  Collected last
(the file name sorts behind every other test file); passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import audio as oa
from whisper_rs_b200 import _lib, synth

pytestmark = pytest.mark.gpu
TOL = 1e-4                     # the tolerance of tests/test_gpu_logmel.py


def _call(fn_name, ctx, pcm_dev, out_dev):
    S, n = pcm_dev.shape
    stream = torch.cuda.current_stream().cuda_stream
    fn = getattr(_lib.lib(), fn_name)
    _lib.check(fn(ctx.handle, pcm_dev.data_ptr(), S, n, pcm_dev.stride(0), out_dev.data_ptr(), stream), ctx.handle)


def test_noalloc_matches_existing_entry_point_two_calls_in_a_row():
    ctx = _lib.context(0)
    _lib.check(_lib.lib().wrs_logmel_set_filter(ctx.handle, None), ctx.handle)
    # two different batches back to back: 5 x 30 s (47 slices each), then 3 ragged short clips (fewer words, new contents)
    batches = [synth.audio_batch(5, seed=7), synth.audio_batch(3, seed=8, n_samples=160 * 333 + 17)]
    # inputs and outputs are allocated up front: nothing is allocated between or inside the two calls under test
    dev = [torch.from_numpy(b).cuda() for b in batches]
    outs = [torch.empty(b.shape[0], 80, b.shape[1] // 160, dtype=torch.float32, device="cuda") for b in batches]
    refs = [torch.empty_like(o) for o in outs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for x, o in zip(dev, outs):
        _call("wrs_logmel_batch_noalloc", ctx, x, o)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    ctx.check()
    for x, r in zip(dev, refs):
        _call("wrs_logmel_batch", ctx, x, r)
    torch.cuda.synchronize()
    ctx.check()
    for b, o, r in zip(batches, outs, refs):
        assert (o - r).abs().max().item() <= TOL
        assert torch.equal(o, r)                     # same kernels, same arithmetic: in fact bit-equal
        want = oa.log_mel_spectrogram(torch.from_numpy(b[0]), oa.mel_filter())
        assert (o[0].cpu() - want).abs().max().item() <= TOL


def test_noalloc_rejects_more_than_the_fixed_scratch():
    ctx = _lib.context(0)
    # 65 536 words: one per 64-frame slice; 70 000 one-slice clips need 70 000 words.  Rejected before anything is
    # read, so the buffers only have to be non-null.
    dummy = torch.zeros(1024, dtype=torch.float32, device="cuda")
    rc = _lib.lib().wrs_logmel_batch_noalloc(ctx.handle, dummy.data_ptr(), 70000, 640, 640, dummy.data_ptr(), None)
    assert rc == _lib.WRS_ERR_UNSUPPORTED
    assert b"scratch" in _lib.lib().wrs_last_error(ctx.handle)

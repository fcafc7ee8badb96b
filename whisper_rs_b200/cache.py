"""Host mirror of the reference's cached keys / values and their beam reorder.

Reference: `Cached<T>` (src/model.rs:36-68) keeps one tensor per attention projection; during beam search
`Whisper::update_cache(source_indices)` (src/model.rs:591-594 -> :524-528 -> :321-327 -> :208-211 -> :53-62) replaces
every one of them by `cache.index(&[Some(sources_indices)])`.  Here all of a decoder's caches are reordered by ONE
launch (`wrs_kv_reorder`, csrc/kv_reorder.cu); the caches are double-buffered because the gather cannot run in place.
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import torch

from . import _lib


def _stream_of(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def update_cache(caches: Sequence[torch.Tensor], source_indices: torch.Tensor,
                 out: Sequence[torch.Tensor] | None = None,
                 group_invariant: Sequence[bool] | None = None) -> list[torch.Tensor]:
    """`[c.index_select(0, source_indices) for c in caches]` in one launch.

    group_invariant[i] marks a cache whose rows are identical within each beam group (the cross-attention K / V:
    projections of `audio_features.repeat_interleave(k)`, src/decode/mod.rs:311-313): the gather cannot change it, it is
    skipped and returned as it is (`wrs_kv_tensor.group_invariant`, include/wrs.h).

    caches: CUDA tensors whose dim 0 is the row (n_audio * beam) dimension, each contiguous; source_indices: CUDA
    int32/int64 `[rows]` (src/decode/extractor/beam_search.rs:111-114).  `out` (same shapes, different storage) is
    reused when given.  Raises WrsError when the extension is missing -- there is no CPU fallback.
    """
    if not caches:
        return []
    dev = caches[0].device
    if dev.type != "cuda":
        raise _lib.WrsError(_lib.WRS_ERR_INVALID, "update_cache: caches must be CUDA tensors (no CPU fallback)")
    rows = caches[0].shape[0]
    idx = source_indices.to(device=dev, dtype=torch.int32).contiguous()
    if idx.numel() != rows:
        raise _lib.WrsError(_lib.WRS_ERR_INVALID, "update_cache: one source index per row")
    inv = [bool(f) for f in group_invariant] if group_invariant is not None else [False] * len(caches)
    if out is None:
        out = [c if f else torch.empty_like(c) for c, f in zip(caches, inv)]
    else:
        out = [c if f else o for c, o, f in zip(caches, out, inv)]
    descs = (_lib.KvTensor * len(caches))()
    for i, (c, o) in enumerate(zip(caches, out)):
        if c.shape[0] != rows or not c.is_contiguous() or not o.is_contiguous() or o.shape != c.shape or o.dtype != c.dtype:
            raise _lib.WrsError(_lib.WRS_ERR_INVALID, "update_cache: caches must be contiguous [rows, ...] with matching outputs")
        descs[i].src, descs[i].dst = c.data_ptr(), o.data_ptr()
        descs[i].row_bytes = (c.numel() // rows) * c.element_size() if rows else 0
        descs[i].group_invariant = 1 if inv[i] else 0
    ctx = _lib.context(dev.index or 0)
    _lib.check(_lib.lib().wrs_kv_reorder(ctx.handle, descs, len(caches), rows, idx.data_ptr(), _stream_of(caches[0])),
               ctx.handle)
    return list(out)


class KVCache:
    """All cached tensors of one decoder with the reference's `update_cache` / `clear_cache` surface."""

    def __init__(self):
        self.tensors: list[torch.Tensor] = []
        self._spare: list[torch.Tensor] | None = None

    def set(self, tensors: Sequence[torch.Tensor], group_invariant: Sequence[bool] | None = None):
        self.tensors = [t.contiguous() for t in tensors]
        self.group_invariant = list(group_invariant) if group_invariant is not None else None
        self._spare = None

    def update_cache(self, source_indices: torch.Tensor):
        """Cached::update_cache (src/model.rs:53-62) for every tensor; ping-pongs two sets of buffers."""
        if not self.tensors:
            return
        spare = self._spare
        if spare is None or any(s.shape != t.shape or s.dtype != t.dtype for s, t in zip(spare, self.tensors)):
            spare = None
        new = update_cache(self.tensors, source_indices, spare, getattr(self, "group_invariant", None))
        self._spare, self.tensors = self.tensors, new

    def clear_cache(self):
        """Cached::clear_cache (src/model.rs:64-66)."""
        self.tensors, self._spare = [], None

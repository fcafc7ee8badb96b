"""ORACLE (test infrastructure only -- tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import it).

CPU restatement of the reference's KV-cache beam reorder: `Cached::update_cache` (src/model.rs:53-62) is
`cache.index(&[Some(sources_indices)])`, i.e. an index_select along dim 0, applied to every cached tensor of the decoder
(src/model.rs:208-211, 321-327, 524-528, 591-594).  Pinning: the reference has no test for this function; the op is a
pure gather, so the restatement is pinned by ATen's own `Tensor.index` (the call the reference makes), bit-exact.
"""
from __future__ import annotations

from typing import Sequence

import torch


def update_cache(caches: Sequence[torch.Tensor], source_indices: torch.Tensor) -> list[torch.Tensor]:
    """[cache.index([source_indices]) for cache in caches]  (src/model.rs:53-62)."""
    idx = source_indices.to(torch.int64)
    return [c[idx] for c in caches]

"""A random-init Whisper-shaped model in plain torch: the stand-in for the reference's tch forward in the tests.

TEST INFRASTRUCTURE.  It plays the part of `Whisper` (src/model.rs:538-600) at the boundary the hot path sees:
`encoder(mel)`, `decoder.forward_ext(tokens, audio_features, offset)` -> logits [rows, T, V], and the cached keys /
values of every attention (`Cached`, src/model.rs:36-68) that `update_cache(source_indices)` reorders after each
beam step.  `cache_backend` picks who performs that reorder: "wrs" = the CUDA library (wrs_kv_reorder, the product),
"oracle" = oracle/cache.py (index_select) -- so two instances with the same seed produce bit-identical logits as
long as both reorders agree.

`eot_bias`/`eot_after` add a bias to the EOT logit from decode step `eot_after` on, so that beams finish and the
finished-list / patience / finalize paths are exercised with model-produced logits.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn.functional as F


@dataclass(frozen=True)
class ModelDims:                      # src/model.rs:548-557
    n_vocab: int = 51865
    n_audio_ctxs: int = 1500
    n_audio_states: int = 64
    n_audio_heads: int = 2
    n_audio_layers: int = 1
    n_text_ctxs: int = 448
    n_text_states: int = 64
    n_text_heads: int = 2
    n_text_layers: int = 2
    n_mels: int = 80


TINY_EN = ModelDims(51864, 1500, 384, 6, 4, 448, 384, 6, 4)      # BASELINE config 1
BASE = ModelDims(51865, 1500, 512, 8, 6, 448, 512, 8, 6)         # BASELINE config 4 (examples/transcribe.rs:25-35 shape)


def _sinusoids(length: int, channels: int) -> torch.Tensor:
    inc = math.log(10000.0) / (channels // 2 - 1)
    inv = torch.exp(-inc * torch.arange(channels // 2, dtype=torch.float32))
    t = torch.arange(length, dtype=torch.float32)[:, None] * inv[None, :]
    return torch.cat([t.sin(), t.cos()], dim=1)


class _Attention:
    def __init__(self, gen, d: int, heads: int, device):
        s = d ** -0.5
        mk = lambda: (torch.randn(d, d, generator=gen) * s).to(device)
        self.wq, self.wk, self.wv, self.wo = mk(), mk(), mk(), mk()
        self.bq, self.bv, self.bo = (torch.zeros(d, device=device) for _ in range(3))
        self.heads = heads
        self.k_cache = None           # Cached<Linear> key / value (src/model.rs:36-68)
        self.v_cache = None

    def caches(self):
        return [c for c in (self.k_cache, self.v_cache) if c is not None]

    def set_caches(self, it):
        if self.k_cache is not None:
            self.k_cache = next(it)
        if self.v_cache is not None:
            self.v_cache = next(it)

    def clear(self):
        self.k_cache = self.v_cache = None

    def __call__(self, x, xa=None, mask=None, use_cache=False):
        q = x @ self.wq.T + self.bq
        if xa is None:                                   # self-attention: the cache grows by the new positions
            k, v = x @ self.wk.T, x @ self.wv.T + self.bv
            if use_cache:
                if self.k_cache is not None:
                    k, v = torch.cat([self.k_cache, k], 1), torch.cat([self.v_cache, v], 1)
                self.k_cache, self.v_cache = k.contiguous(), v.contiguous()
        else:                                            # cross-attention: computed once per window
            if use_cache and self.k_cache is not None:
                k, v = self.k_cache, self.v_cache
            else:
                k, v = xa @ self.wk.T, xa @ self.wv.T + self.bv
                if use_cache:
                    self.k_cache, self.v_cache = k.contiguous(), v.contiguous()
        B, T, D = q.shape
        h, dh = self.heads, D // self.heads
        scale = dh ** -0.25
        qh = (q.view(B, T, h, dh) * scale).permute(0, 2, 1, 3)
        kh = (k.view(B, -1, h, dh) * scale).permute(0, 2, 3, 1)
        vh = v.view(B, -1, h, dh).permute(0, 2, 1, 3)
        qk = qh @ kh
        if mask is not None:
            S = qk.shape[-1]
            qk = qk + mask[S - T:S, :S]
        w = qk.softmax(-1)
        return (w @ vh).permute(0, 2, 1, 3).flatten(2) @ self.wo.T + self.bo


class _Block:
    def __init__(self, gen, d, heads, cross, device):
        self.attn = _Attention(gen, d, heads, device)
        self.cross = _Attention(gen, d, heads, device) if cross else None
        self.w1 = (torch.randn(4 * d, d, generator=gen) * d ** -0.5).to(device)
        self.w2 = (torch.randn(d, 4 * d, generator=gen) * (4 * d) ** -0.5).to(device)
        self.d = d

    def __call__(self, x, xa=None, mask=None, use_cache=False):
        ln = lambda t: F.layer_norm(t, (self.d,))
        x = x + self.attn(ln(x), None, mask, use_cache)
        if self.cross is not None:
            x = x + self.cross(ln(x), xa, None, use_cache)
        return x + F.gelu(ln(x) @ self.w1.T) @ self.w2.T


class _Encoder:
    def __init__(self, gen, dims: ModelDims, device):
        d = dims.n_audio_states
        self.c1 = (torch.randn(d, dims.n_mels, 3, generator=gen) * (3 * dims.n_mels) ** -0.5).to(device)
        self.c2 = (torch.randn(d, d, 3, generator=gen) * (3 * d) ** -0.5).to(device)
        self.pos = _sinusoids(dims.n_audio_ctxs, d).to(device)
        self.blocks = [_Block(gen, d, dims.n_audio_heads, False, device) for _ in range(dims.n_audio_layers)]
        self.d = d

    def __call__(self, mel):
        x = F.gelu(F.conv1d(mel, self.c1, padding=1))
        x = F.gelu(F.conv1d(x, self.c2, stride=2, padding=1)).permute(0, 2, 1)
        x = x + self.pos
        for b in self.blocks:
            x = b(x)
        return F.layer_norm(x, (self.d,))


class _Decoder:
    def __init__(self, gen, dims: ModelDims, device, logit_scale, eot, eot_bias, eot_after):
        d = dims.n_text_states
        self.token_emb = torch.randn(dims.n_vocab, d, generator=gen).to(device)
        self.pos = (torch.randn(dims.n_text_ctxs, d, generator=gen) * 0.1).to(device)
        self.blocks = [_Block(gen, d, dims.n_text_heads, True, device) for _ in range(dims.n_text_layers)]
        self.mask = torch.full((dims.n_text_ctxs, dims.n_text_ctxs), float("-inf"), device=device).triu_(1)
        self.d, self.scale = d, logit_scale / math.sqrt(d)
        self.eot, self.eot_bias, self.eot_after = eot, eot_bias, eot_after

    def forward_hidden(self, tokens, xa, offset):                   # src/model.rs:488-519: everything before the projection
        x = self.token_emb[tokens] + self.pos[offset:offset + tokens.shape[-1]]
        for b in self.blocks:
            x = b(x, xa, self.mask, use_cache=True)
        return F.layer_norm(x, (self.d,))

    def forward_ext(self, tokens, xa, offset):                      # src/model.rs:488-521
        x = self.forward_hidden(tokens, xa, offset)
        logits = (x @ self.token_emb.T) * self.scale
        if self.eot_bias and offset >= self.eot_after:
            logits[:, :, self.eot] += self.eot_bias
        return logits


class ToyWhisper:
    def __init__(self, dims: ModelDims, device="cuda", seed=0, cache_backend="wrs", logit_scale=3.0, eot=50257,
                 eot_bias=0.0, eot_after=4, flag_cross_invariant=False):
        gen = torch.Generator().manual_seed(seed)
        self.dims, self.device, self.cache_backend = dims, torch.device(device), cache_backend
        self.encoder = _Encoder(gen, dims, self.device)
        self.decoder = _Decoder(gen, dims, self.device, logit_scale, eot, eot_bias, eot_after)
        self.reorders = 0
        # cross-attention K / V are projections of audio_features.repeat_interleave(k): identical within a beam group, so
        # their gather is a no-op that the library can skip (wrs_kv_tensor.group_invariant)
        self.flag_cross_invariant = flag_cross_invariant

    def _attentions(self):
        for b in self.decoder.blocks:
            yield b.attn
            yield b.cross

    def update_cache(self, source_indices):                         # src/model.rs:591-594
        caches = [c for a in self._attentions() for c in a.caches()]
        if not caches:
            return
        if self.cache_backend == "wrs":
            from whisper_rs_b200.cache import update_cache
            flags = None
            if self.flag_cross_invariant:
                flags = [a is b.cross for b in self.decoder.blocks for a in (b.attn, b.cross) for _ in a.caches()]
            new = update_cache(caches, source_indices, group_invariant=flags)
        else:
            from oracle.cache import update_cache
            new = update_cache(caches, source_indices.to(caches[0].device))
        it = iter(new)
        for a in self._attentions():
            a.set_caches(it)
        self.reorders += 1

    def clear_cache(self):                                          # src/model.rs:596-599
        for a in self._attentions():
            a.clear()

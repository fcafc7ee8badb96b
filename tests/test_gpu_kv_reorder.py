"""K3 parity: wrs_kv_reorder (through the C ABI) vs the oracle's `cache.index([source_indices])`, bit-exact."""
import numpy as np
import pytest
import torch

from oracle import cache as oc

pytestmark = pytest.mark.gpu


def _caches(rows, shapes, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    out = []
    for sh in shapes:
        if dtype.is_floating_point:
            out.append(torch.randn((rows, *sh), generator=g, dtype=torch.float32).to(dtype))
        else:
            out.append(torch.randint(-100, 100, (rows, *sh), generator=g, dtype=dtype))
    return out


def _beam_sources(n_audio, beam, seed):
    """source rows as beam search produces them: every row takes a row of its own audio, duplicates allowed."""
    rng = np.random.default_rng(seed)
    src = np.concatenate([a * beam + rng.integers(0, beam, size=beam) for a in range(n_audio)])
    return torch.from_numpy(src.astype(np.int64))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.int8])
def test_reorder_matches_index(dtype):
    from whisper_rs_b200 import cache
    rows = 15
    shapes = [(7, 64), (1500, 8), (3,), (5, 33)]          # row sizes: multiples of 16 B, tiny, odd byte counts
    cpu = _caches(rows, shapes, dtype, seed=1)
    src = _beam_sources(3, 5, seed=2)
    want = oc.update_cache(cpu, src)
    got = cache.update_cache([c.cuda() for c in cpu], src.cuda())
    for w, g in zip(want, got):
        assert torch.equal(w, g.cpu())


def test_identity_and_permutation_and_broadcast():
    from whisper_rs_b200 import cache
    rows = 10
    cpu = _caches(rows, [(9, 40)], torch.float32, seed=3)
    for src in (torch.arange(rows), torch.arange(rows).flip(0), torch.zeros(rows, dtype=torch.int64)):
        want = oc.update_cache(cpu, src)
        got = cache.update_cache([c.cuda() for c in cpu], src.cuda().to(torch.int32))
        assert torch.equal(want[0], got[0].cpu())


def test_unaligned_views_take_the_byte_path():
    from whisper_rs_b200 import cache
    rows = 6
    base = torch.randn(rows * 37 + 1, dtype=torch.float32)
    cpu = [base[1:].view(rows, 37)]                        # 4-byte aligned rows of 148 bytes
    src = torch.tensor([5, 5, 0, 1, 2, 2])
    dev = base.cuda()[1:].view(rows, 37)
    got = cache.update_cache([dev], src.cuda())
    assert torch.equal(oc.update_cache(cpu, src)[0], got[0].cpu())


def test_many_tensors_and_kvcache_object():
    """More descriptors than one launch carries (96), through the reference-shaped KVCache surface, two steps."""
    from whisper_rs_b200 import cache
    rows, n = 20, 130
    cpu = _caches(rows, [(4, 16)] * n, torch.float32, seed=4)
    kv = cache.KVCache()
    kv.set([c.cuda() for c in cpu])
    ref = cpu
    for step in range(2):
        src = _beam_sources(4, 5, seed=10 + step)
        ref = oc.update_cache(ref, src)
        kv.update_cache(src.cuda())
    for w, g in zip(ref, kv.tensors):
        assert torch.equal(w, g.cpu())
    kv.clear_cache()
    assert kv.tensors == []


def test_base_model_shape_one_layer():
    """BASELINE config 3 rows (160) with base.en-shaped caches of one decoder block: self [160, 32, 512], cross [160, 1500, 512]."""
    from whisper_rs_b200 import cache
    rows = 160
    g = torch.Generator(device="cuda").manual_seed(5)
    dev = [torch.randn((rows, 32, 512), generator=g, device="cuda"), torch.randn((rows, 1500, 512), generator=g, device="cuda")]
    src = _beam_sources(32, 5, seed=6).cuda()
    got = cache.update_cache(dev, src)
    for d, o in zip(dev, got):
        assert torch.equal(d[src], o)                      # ATen's own index on the device: same gather, bit-exact
    # size-independent property: reordering by a permutation and then by its inverse is the identity
    perm = torch.randperm(rows, device="cuda")
    inv = torch.empty_like(perm); inv[perm] = torch.arange(rows, device="cuda")
    back = cache.update_cache(cache.update_cache(dev, perm), inv)
    for d, o in zip(dev, back):
        assert torch.equal(d, o)


def test_bad_index_is_reported():
    from whisper_rs_b200 import _lib, cache
    dev = [torch.zeros((4, 8), device="cuda")]
    cache.update_cache(dev, torch.tensor([0, 1, 7, 3], device="cuda"))
    ctx = _lib.context(0)
    with pytest.raises(_lib.WrsError):
        ctx.check(torch.cuda.current_stream().cuda_stream)


def test_group_invariant_tensors_are_skipped_and_verified():
    """wrs_kv_tensor.group_invariant: the tensor is returned as it is (no traffic); option kv_verify_group checks the
    claim on the device and reports a violation -- differing rows inside a group, or an index that leaves its group."""
    from whisper_rs_b200 import _lib, cache
    n_audio, beam = 4, 5
    rows = n_audio * beam
    g = torch.Generator().manual_seed(9)
    cross = torch.randn((n_audio, 30, 16), generator=g).repeat_interleave(beam, 0).contiguous()     # identical within groups
    own = torch.randn((rows, 7, 16), generator=g)
    src = _beam_sources(n_audio, beam, seed=4)
    ctx = _lib.context(0)
    ctx.set_option("kv_verify_group", beam)
    try:
        dc, do = cross.cuda(), own.cuda()
        got = cache.update_cache([dc, do], src.cuda(), group_invariant=[True, False])
        torch.cuda.synchronize()
        ctx.check()
        assert got[0] is dc                                                  # skipped: the very same buffer
        assert torch.equal(got[0].cpu(), oc.update_cache([cross], src)[0])    # ... which is what the gather would give
        assert torch.equal(got[1].cpu(), oc.update_cache([own], src)[0])
        # a tensor that is not group-invariant, flagged: reported
        cache.update_cache([do], src.cuda(), group_invariant=[True])
        torch.cuda.synchronize()
        with pytest.raises(_lib.WrsError) as e:
            ctx.check()
        assert e.value.code == _lib.WRS_ERR_INVALID
        # an index that leaves its group: reported
        bad = src.clone(); bad[3] = 7
        cache.update_cache([dc], bad.cuda(), group_invariant=[True])
        torch.cuda.synchronize()
        with pytest.raises(_lib.WrsError):
            ctx.check()
        ctx.check()                                                           # the status word was re-armed
    finally:
        ctx.set_option("kv_verify_group", 0)

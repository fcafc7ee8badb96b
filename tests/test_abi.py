"""The C-ABI library: builds, loads, exports every symbol include/wrs.h declares, and its host-only entry
points agree with the oracle.  No compute call is made (there is no GPU here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from oracle import audio as oa
from oracle import decode as od


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "wrs.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(wrs_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(built_lib):
    handle = C.CDLL(built_lib)
    syms = declared_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(handle, s), f"{s} declared in wrs.h but not exported"
    from whisper_rs_b200 import _lib
    assert sorted(_lib.SIGNATURES) == syms          # the ctypes binding covers the whole header
    assert _lib.lib().wrs_abi_version() == 2


def test_cubin_is_sm_100a(built_lib):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", built_lib], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_mel_filter_host_matches_oracle(built_lib):
    from whisper_rs_b200 import audio
    f = audio.mel_filter().numpy()
    assert np.array_equal(f, oa.mel_filter_np())


def test_no_cpu_fallback(built_lib):
    from whisper_rs_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.WrsError) as e:
        _lib.Context(0)
    assert e.value.code == _lib.WRS_ERR_CUDA and "no CPU fallback" in str(e.value)


@pytest.mark.parametrize("penalty", [None, 0.6, 1.0])
def test_ranker_host_matches_oracle(built_lib, penalty):
    from whisper_rs_b200.decode import MaximumLikelihoodRanker
    rng = np.random.default_rng(0)
    tokens = [[list(range(int(rng.integers(1, 30)))) for _ in range(5)] for _ in range(7)]
    scores = -rng.random((7, 5)) * 20
    scores[3, 1] = scores[3, 4] = scores[3].max() + 0.0      # exact tie -> first index
    tokens[3][1] = tokens[3][4] = list(range(9))
    want = od.MaximumLikelihoodRanker(penalty).rank(tokens, torch.from_numpy(scores))
    got = MaximumLikelihoodRanker(penalty).rank(tokens, scores)
    assert got == want


def test_ranker_zero_length_division(built_lib):
    # len 0 => logprob / 0 (reference quirk): -inf for negative logprob
    from whisper_rs_b200.decode import MaximumLikelihoodRanker
    tokens = [[[], [1, 2]]]
    scores = np.array([[-1.0, -4.0]])
    assert MaximumLikelihoodRanker(None).rank(tokens, scores) == od.MaximumLikelihoodRanker(None).rank(tokens, torch.from_numpy(scores)) == [1]

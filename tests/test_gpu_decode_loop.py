"""Trait-level surface (in-place LogitFilter::apply, TokenExtractor::update/finalize, SequenceRanker::rank),
greedy, a multi-step decode loop, BASELINE config 3 at full size, and error behaviour -- CUDA vs oracle."""
import numpy as np
import pytest
import torch

from oracle import decode as od
from whisper_rs_b200 import _lib, synth
from whisper_rs_b200 import constants as C
from whisper_rs_b200 import decode as wd
from test_gpu_decode import assert_beam_parity, close, fused_step

pytestmark = pytest.mark.gpu
IDS = C.MULTILINGUAL


def make_prompt(sample_begin):
    p = wd.DecodePrompt(None, IDS)
    p.sample_begin = sample_begin
    return p


@pytest.mark.parametrize("step_kind,prompt_len", [("first", 0), ("second", 0), ("mixed", 0), ("mixed", 9)])
def test_inplace_filters_match_oracle_masks(step_kind, prompt_len):
    """LogitFilter::apply mutates logits: the -inf pattern must be bit-exact and finite values untouched."""
    case = synth.decode_case(IDS, 3, 5, step_kind, seed=8, prompt_len=prompt_len)
    want = od.decode_step_reference(case, IDS)
    filters = wd.build_filters(wd.DecodeOptions(), IDS, [make_prompt(case["sample_begin"])], max_rows=15)
    assert [type(f).__name__ for f in filters] == ["SuppressBlank", "SuppressTokens", "TimestampTokens"]
    logits = torch.from_numpy(case["logits"]).cuda()
    tokens = torch.from_numpy(case["tokens"]).cuda()
    for f in filters:
        f.apply(logits, tokens)
    torch.cuda.synchronize()
    got = logits.cpu().numpy()
    assert np.array_equal(np.isneginf(got), np.isneginf(want["masked_logits"]))
    assert np.array_equal(got, want["masked_logits"])            # survivors keep their exact bits
    assert np.array_equal(filters[-1].last_row_flags.cpu().numpy(), want["row_flags"])


def test_trait_level_path_equals_fused_path():
    """filters in place + BeamSearchTokenExtractor::update == the fused single pass == the oracle."""
    case = synth.decode_case(IDS, 4, 5, "mixed", seed=19)
    want = od.decode_step_reference(case, IDS)
    fused, _ = fused_step(IDS, case)
    logits = torch.from_numpy(case["logits"]).cuda()
    tokens = torch.from_numpy(case["tokens"]).cuda()
    sums = torch.from_numpy(case["sum_logprobs"]).cuda()
    for f in wd.build_filters(wd.DecodeOptions(), IDS, [make_prompt(case["sample_begin"])], max_rows=20):
        f.apply(logits, tokens)

    class Model:                      # stands in for Whisper::update_cache (beam_search.rs:114)
        seen = None

        def update_cache(self, idx):
            self.seen = idx.cpu().numpy()

    ex = wd.BeamSearchTokenExtractor(IDS, 5, 1.0, max_rows=20)
    ex.reset()
    model = Model()
    new_tokens, completed = ex.update(model, tokens, logits, sums, debug=True)
    assert np.array_equal(new_tokens.cpu().numpy(), want["next_tokens"])
    assert np.array_equal(new_tokens.cpu().numpy(), fused["next_tokens"])
    assert np.array_equal(model.seen, want["source_idx"])
    assert close(sums.cpu().numpy(), want["sum_logprobs"])      # &mut sum_logprobs
    assert np.array_equal(ex.last["cand_token"].cpu().numpy(), fused["cand_token"])
    assert completed == bool(want["completed"])


@pytest.mark.parametrize("step_kind", ["first", "mixed"])
def test_greedy_step_matches_oracle(step_kind):
    case = synth.decode_case(IDS, 6, 1, step_kind, seed=23)
    case["tokens"][2, -1] = IDS.eot if step_kind == "mixed" else case["tokens"][2, -1]    # a finished row stays EOT
    want = od.decode_step_reference(case, IDS, od.DecodeOptions(beam_size=1, greedy=True))
    got, _ = fused_step(IDS, case, greedy=True)
    assert np.array_equal(got["next_tokens"], want["next_tokens"])
    assert close(got["sum_logprobs"], want["sum_logprobs"])
    assert int(got["completed"][0]) == int(want["completed"])
    assert np.array_equal(got["row_flags"], want["row_flags"])


def step_logits(seed, step, R, V, eot, boost_eot):
    rng = np.random.default_rng([seed, step])
    x = (3.0 * rng.standard_normal((R, V))).astype(np.float32)
    x[:, IDS.ts_begin:] += np.float32(rng.choice([-4.0, 0.0, 3.0], size=(R, 1)))
    if boost_eot:
        x[rng.random(R) < 0.5, eot] = 25.0          # some beams finish
    return x


@pytest.mark.parametrize("patience,len_penalty", [(1.0, None), (2.0, 0.6)])
def test_multi_step_loop_finalize_and_rank(patience, len_penalty):
    """main_loop (decode/mod.rs:220-282) for 14 steps on seeded logits, then finalize + rank: device-resident
    tokens on one side, the oracle's host bookkeeping on the other."""
    n_audio, k, V = 3, 5, IDS.n_vocab
    R = n_audio * k
    opt = wd.DecodeOptions(token_extract_mode=wd.TokenExtractMode.BeamSearch(k, patience), len_penalty=len_penalty)
    prompt = wd.DecodePrompt(None, IDS)
    step = wd.FusedDecodeStep(opt, IDS, [prompt], max_rows=R)
    step.reset()
    oopt = od.DecodeOptions(beam_size=k, patience=patience)
    oprompt = [od.DecodePrompt(prompt.sample_begin, 0, torch.tensor(prompt.initial_tokens))]
    ofilters = od.build_filters(oopt, IDS, oprompt)
    oex = od.BeamSearchTokenExtractor(k, IDS.eot, patience)
    oex.reset()
    tok_o = torch.tensor(prompt.initial_tokens, dtype=torch.int64).repeat(R, 1)
    sum_o = torch.zeros(R)
    tok_g, sum_g = tok_o.clone().cuda(), sum_o.clone().cuda()
    for i in range(14):
        x = step_logits(1234, i, R, V, IDS.eot, boost_eot=(i >= 3))
        lo = torch.from_numpy(x).clone()
        for f in ofilters:
            f.apply(lo, tok_o)
        tok_o, done_o = oex.update(tok_o, lo, sum_o)
        out = step.step(tok_g, torch.from_numpy(x).cuda(), sum_g)
        tok_g, sum_g = out["next_tokens"], out["sum_logprobs"]
        assert np.array_equal(tok_g.cpu().numpy(), tok_o.numpy()), f"step {i}"
        assert close(sum_g.cpu().numpy(), sum_o.numpy()), f"step {i}"
        assert bool(out["completed"].item()) == done_o, f"step {i}"
        if done_o:
            break
    L = tok_o.shape[1]
    cand_o, score_o = oex.finalize(tok_o.reshape(n_audio, k, L), sum_o.reshape(n_audio, k))
    cand_g, score_g = step.finalize(tok_g.reshape(n_audio, k, L), sum_g.reshape(n_audio, k))
    assert cand_g == cand_o
    assert close(score_g, score_o.numpy())
    sl_o = od.slice_candidates(cand_o, prompt.sample_begin, IDS.eot)
    sel_o = od.MaximumLikelihoodRanker(len_penalty).rank(sl_o, score_o)
    sel_g = wd.MaximumLikelihoodRanker(len_penalty).rank(od.slice_candidates(cand_g, prompt.sample_begin, IDS.eot), score_g)
    assert sel_g == sel_o


def test_config3_full_size():
    """BASELINE config 3: beam 5 x batch 32 = 160 rows, V = 51865, all filters, each step kind."""
    for kind, seed in (("first", 1), ("second", 2), ("mixed", 3)):
        case = synth.decode_case(IDS, 32, 5, kind, seed=seed)
        want = od.decode_step_reference(case, IDS)
        got, _ = fused_step(IDS, case)
        assert_beam_parity(got, want, 5)


def test_nan_is_reported_not_hidden():
    # the reference panics on NaN (beam_search.rs:86 partial_cmp().unwrap()); the library raises WRS_ERR_NUMERIC
    case = synth.decode_case(IDS, 1, 5, "mixed", seed=4)
    case["logits"][0, 1000] = np.nan      # row 0: plain text history, id 1000 survives every mask
    with pytest.raises(_lib.WrsError) as e:
        fused_step(IDS, case)
    assert e.value.code == _lib.WRS_ERR_NUMERIC


def test_bad_arguments_raise():
    case = synth.decode_case(IDS, 1, 5, "mixed", seed=4)
    p = make_prompt(case["sample_begin"])
    step = wd.FusedDecodeStep(wd.DecodeOptions(), IDS, [p], max_rows=5)
    tokens = torch.from_numpy(case["tokens"]).cuda()
    sums = torch.zeros(5, device="cuda")
    with pytest.raises(_lib.WrsError):      # rows not a multiple of beam (beam_search.rs:29 assert)
        step.step(tokens[:4], torch.zeros(4, IDS.n_vocab, device="cuda"), sums[:4])
    with pytest.raises(TypeError):
        step.step(tokens, torch.zeros(5, IDS.n_vocab, device="cuda", dtype=torch.float16), sums)
    with pytest.raises(_lib.WrsError):      # vocab narrower than the decoder was built for
        step.step(tokens, torch.zeros(5, 1000, device="cuda"), sums)

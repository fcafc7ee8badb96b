"""the logits projection (reference src/model.rs:521, `x.matmul(token_emb.T)`) on the
tensor cores with the decode step (src/decode/mod.rs:257-270) as its epilogue -- through the C ABI.

Two separate questions, two kinds of test:
  * the EPILOGUE (masks, softmax statistics, exact top-K, tail): wrs_project_beam_step must equal wrs_beam_step run on
    the logits the same tensor-core arithmetic materialises (wrs_project_logits) -- ids, sources, flags bit-exact,
    log-probs 1e-5 relative -- and wrs_beam_step is what tests/test_gpu_decode.py pins to the oracle;
  * the ARITHMETIC: f16 hi/lo splits (22 bits per operand), three products, accumulation in the tensor core's f32
    accumulators -- whose additions are not IEEE roundings: measured |logit - f64 product| up to 1.2e-6 of
    sum_i |x_i w_i| (4e-5 absolute on logits of sigma 3 at n_state 512, 8e-5 at 1280), four to seven times the error of
    numpy's f32 matmul and growing like sqrt(n_state); with bf16 splits it was 1.8e-6.  NOT f32-faithful.  The bar written
    here is LOGIT_TOL of that sum, and against the oracle (numpy f32 matmul -> oracle decode step) ids must agree wherever
    the oracle's own candidate gaps exceed 2 LOGPROB_TOL, log-probs within LOGPROB_TOL."""
import numpy as np
import pytest
import torch

from oracle import decode as od
from whisper_rs_b200 import synth
from whisper_rs_b200 import constants as C
from whisper_rs_b200 import decode as wd
from whisper_rs_b200.decode.core import Projection

pytestmark = pytest.mark.gpu
LOGIT_TOL = 2.5e-6       # of sum |x_i w_i| per logit (measured up to 1.2e-6; numpy f32 matmul sits near 2e-7)
LOGPROB_TOL = 5e-4       # absolute, log-probs of the fused step against the oracle on f32-matmul logits (measured 3e-5)
RTOL = 1e-5


def make_model(ids, n_state, R, seed, dup=0, quant=0):
    """x [R, n_state], token_emb [V, n_state] with logits of sigma 3.  dup > 0: that many ids get an embedding copied
    from another id, and so does a twin of every row's best text id (exact logit ties at the top); quant > 0: every embedding is one of `quant` vectors (ties by the hundred)."""
    rng = np.random.default_rng([seed, n_state, R])
    x = rng.standard_normal((R, n_state)).astype(np.float32)
    if quant:
        base = (rng.standard_normal((quant, n_state)) * (3.0 / np.sqrt(n_state))).astype(np.float32)
        w = base[rng.integers(0, quant, size=ids.n_vocab)]
    else:
        w = (rng.standard_normal((ids.n_vocab, n_state)) * (3.0 / np.sqrt(n_state))).astype(np.float32)
    if dup:
        dst = rng.choice(ids.n_vocab, size=dup, replace=False)
        src = rng.choice(ids.n_vocab, size=dup, replace=False)
        w[dst] = w[src]
        for r in range(R):          # and every row's best plain-text id gets a twin 1000 ids further on
            a = 300 + int(np.argmax(x[r] @ w[300:49000].T))
            w[a + 1000] = w[a]
    return x, np.ascontiguousarray(w)


def build(ids, case, beam):
    opt = wd.DecodeOptions(token_extract_mode=wd.TokenExtractMode.BeamSearch(beam, 1.0))
    prompt = wd.DecodePrompt(None, ids)
    prompt.sample_begin = case["sample_begin"]
    step = wd.FusedDecodeStep(opt, ids, [prompt], max_rows=case["tokens"].shape[0])
    step.reset()
    return step


def to_np(out):
    return {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in out.items()}


@pytest.mark.parametrize("n_state,R", [(512, 160), (384, 5), (1280, 33)])
def test_projection_logits_error_bound(n_state, R, record_property):
    ids = C.MULTILINGUAL
    x, w = make_model(ids, n_state, R, seed=1)
    case = synth.decode_case(ids, 1, 1, "first", seed=0)
    step = wd.FusedDecodeStep(wd.DecodeOptions(token_extract_mode=wd.TokenExtractMode.BeamSearch(1, 1.0)), ids,
                              [wd.DecodePrompt(None, ids)], max_rows=max(R, 1))
    proj = Projection(step.decoder, torch.from_numpy(w).cuda())
    got = proj.logits(torch.from_numpy(x).cuda()).cpu().numpy()
    step.decoder.ctx.check()
    want64 = x.astype(np.float64) @ w.astype(np.float64).T
    scale = np.abs(x).astype(np.float64) @ np.abs(w).astype(np.float64).T
    rel = np.abs(got - want64) / scale
    f32 = x @ w.T                                            # what the reference computes (f32 matmul)
    record_property("proj_max_err_of_abs_sum", float(rel.max()))
    record_property("proj_max_abs_err", float(np.abs(got - want64).max()))
    record_property("f32_matmul_max_abs_err", float(np.abs(f32 - want64).max()))
    print(f"projection n_state={n_state} R={R}: max |err| {np.abs(got - want64).max():.3e} "
          f"({rel.max():.3e} of sum|x w|); numpy f32 matmul: {np.abs(f32 - want64).max():.3e}")
    assert got.shape == (R, ids.n_vocab)
    assert rel.max() <= LOGIT_TOL


@pytest.mark.parametrize("step_kind,beam,n_audio,dup,quant", [
    ("first", 5, 4, 0, 0), ("mixed", 5, 32, 0, 0), ("second", 1, 3, 0, 0), ("mixed", 8, 2, 0, 0), ("mixed", 12, 1, 0, 0),
    ("mixed", 5, 3, 4000, 0), ("mixed", 5, 2, 0, 300), ("second", 2, 2, 0, 7)])
def test_fused_step_equals_beam_step_on_projected_logits(step_kind, beam, n_audio, dup, quant):
    ids = C.MULTILINGUAL
    case = synth.decode_case(ids, n_audio, beam, step_kind, seed=41)
    R = n_audio * beam
    x, w = make_model(ids, 512, R, seed=beam + n_audio, dup=dup, quant=quant)
    step = build(ids, case, beam)
    proj = Projection(step.decoder, torch.from_numpy(w).cuda())
    xd = torch.from_numpy(x).cuda()
    tokens, sums = torch.from_numpy(case["tokens"]).cuda(), torch.from_numpy(case["sum_logprobs"]).cuda()
    logits = proj.logits(xd)
    want = to_np(step.decoder.beam_step(logits, tokens, case["sample_begin"], sums, debug=True))
    step.reset()
    got = to_np(proj.beam_step(xd, tokens, case["sample_begin"], sums, debug=True))
    torch.cuda.synchronize()
    step.decoder.ctx.check()
    for key in ("row_flags", "cand_token", "next_tokens", "source_idx", "fin_count"):
        assert np.array_equal(got[key], want[key]), key
    assert np.allclose(got["cand_logprob"], want["cand_logprob"], rtol=RTOL, atol=1e-5)
    assert np.allclose(got["sum_logprobs"], want["sum_logprobs"], rtol=RTOL, atol=1e-5)
    assert int(got["completed"][0]) == int(want["completed"][0])
    if dup or quant:
        assert (got["cand_logprob"][:, 0] == got["cand_logprob"][:, 1]).any() or step_kind == "first"


@pytest.mark.parametrize("ids,beam,n_audio,n_state,timestamps", [
    (C.MULTILINGUAL, 5, 60, 384, True),        # 300 rows: three M tiles, the last one 44 rows (packed compactly)
    (C.GPT2_EN, 5, 26, 512, True),             # 130 rows: a second M tile of 2 rows; V = 51864
    (C.MULTILINGUAL, 4, 8, 768, False),        # no timestamp rules: one partition, 295 tiles
    (C.MULTILINGUAL, 15, 2, 512, True),        # the largest beam: K = 16 entries per list
    (C.MULTILINGUAL, 1, 128, 512, True),       # beam 1 (K = 2), exactly one full M tile
    (C.MULTILINGUAL, 5, 400, 384, True)])      # 2 000 rows: sixteen M tiles per vocab tile (32 units per CTA)
def test_fused_step_geometries(ids, beam, n_audio, n_state, timestamps):
    case = synth.decode_case(ids, n_audio, beam, "mixed", seed=23)
    R = n_audio * beam
    x, w = make_model(ids, n_state, R, seed=beam, dup=500)
    opt = wd.DecodeOptions(token_extract_mode=wd.TokenExtractMode.BeamSearch(beam, 1.0), timestamps=timestamps)
    prompt = wd.DecodePrompt(None, ids)
    prompt.sample_begin = case["sample_begin"]
    step = wd.FusedDecodeStep(opt, ids, [prompt], max_rows=R)
    step.reset()
    proj = Projection(step.decoder, torch.from_numpy(w).cuda())
    xd = torch.from_numpy(x).cuda()
    tokens, sums = torch.from_numpy(case["tokens"]).cuda(), torch.from_numpy(case["sum_logprobs"]).cuda()
    logits = proj.logits(xd)
    want = to_np(step.decoder.beam_step(logits, tokens, case["sample_begin"], sums, debug=True))
    step.reset()
    got = to_np(proj.beam_step(xd, tokens, case["sample_begin"], sums, debug=True))
    torch.cuda.synchronize()
    step.decoder.ctx.check()
    for key in ("row_flags", "cand_token", "next_tokens", "source_idx", "fin_count"):
        assert np.array_equal(got[key], want[key]), key
    assert np.allclose(got["cand_logprob"], want["cand_logprob"], rtol=RTOL, atol=1e-5)
    assert np.allclose(got["sum_logprobs"], want["sum_logprobs"], rtol=RTOL, atol=1e-5)
    # and the projection itself on this geometry, against the f64 product
    want64 = x.astype(np.float64) @ w.astype(np.float64).T
    scale = np.abs(x).astype(np.float64) @ np.abs(w).astype(np.float64).T
    assert (np.abs(logits.cpu().numpy() - want64) / scale).max() <= LOGIT_TOL


@pytest.mark.parametrize("R_audio,beam", [(32, 5), (26, 5), (51, 5)])     # 160, 130 and 255 rows: two M tiles
def test_cluster_launch_is_bit_identical_to_the_single_cta_launch(R_audio, beam):
    """129 .. 256 rows run as clusters of two CTAs that share the weights by multicast (option proj_cluster, default 1): the
    same arithmetic per accumulator, so the same bits -- logits, records, everything the step returns."""
    from whisper_rs_b200 import _lib
    ids = C.MULTILINGUAL
    case = synth.decode_case(ids, R_audio, beam, "mixed", seed=77)
    R = R_audio * beam
    x, w = make_model(ids, 512, R, seed=R, dup=300)
    step = build(ids, case, beam)
    proj = Projection(step.decoder, torch.from_numpy(w).cuda())
    xd = torch.from_numpy(x).cuda()
    tokens, sums = torch.from_numpy(case["tokens"]).cuda(), torch.from_numpy(case["sum_logprobs"]).cuda()
    ctx = _lib.context(0)
    res = {}
    for mode in (1, 0):
        ctx.set_option("proj_cluster", mode)
        try:
            step.reset()
            logits = proj.logits(xd)
            out = to_np(proj.beam_step(xd, tokens, case["sample_begin"], sums, debug=True))
            torch.cuda.synchronize()
            ctx.check()
            res[mode] = (logits.cpu().numpy(), out)
        finally:
            ctx.set_option("proj_cluster", 1)
    assert np.array_equal(res[1][0], res[0][0])
    for key in ("row_flags", "cand_token", "cand_logprob", "next_tokens", "source_idx", "sum_logprobs", "fin_count"):
        assert np.array_equal(res[1][1][key], res[0][1][key]), key


@pytest.mark.parametrize("step_kind,R,dup", [("first", 7, 0), ("mixed", 160, 0), ("mixed", 37, 2000), ("second", 300, 0)])
def test_fused_greedy_step_equals_greedy_step_on_projected_logits(step_kind, R, dup):
    """TokenExtractMode::Greedy (src/decode/extractor/greedy.rs:14-51): argmax with the lowest id on ties, log-softmax
    gather, EOT propagation -- through the projection, against wrs_greedy_step on the logits it materialises."""
    ids = C.MULTILINGUAL
    case = synth.decode_case(ids, R, 1, step_kind, seed=19)
    x, w = make_model(ids, 512, R, seed=R, dup=dup)
    opt = wd.DecodeOptions(token_extract_mode=wd.TokenExtractMode.Greedy(1))
    prompt = wd.DecodePrompt(None, ids)
    prompt.sample_begin = case["sample_begin"]
    step = wd.FusedDecodeStep(opt, ids, [prompt], max_rows=R)
    step.reset()
    proj = Projection(step.decoder, torch.from_numpy(w).cuda())
    xd = torch.from_numpy(x).cuda()
    tokens, sums = torch.from_numpy(case["tokens"]).cuda(), torch.from_numpy(case["sum_logprobs"]).cuda()
    want = to_np(step.decoder.greedy_step(proj.logits(xd), tokens, case["sample_begin"], sums, debug=True))
    got = to_np(proj.greedy_step(xd, tokens, case["sample_begin"], sums, debug=True))
    torch.cuda.synchronize()
    step.decoder.ctx.check()
    for key in ("row_flags", "cand_token", "next_tokens"):
        assert np.array_equal(got[key], want[key]), key
    assert np.allclose(got["cand_logprob"], want["cand_logprob"], rtol=RTOL, atol=1e-5)
    assert np.allclose(got["sum_logprobs"], want["sum_logprobs"], rtol=RTOL, atol=1e-5)
    assert int(got["completed"][0]) == int(want["completed"][0])


@pytest.mark.parametrize("step_kind", ["first", "mixed"])
def test_fused_step_against_the_oracle_on_f32_logits(step_kind):
    ids = C.MULTILINGUAL
    beam, n_audio = 5, 8
    case = synth.decode_case(ids, n_audio, beam, step_kind, seed=9)
    R = n_audio * beam
    x, w = make_model(ids, 512, R, seed=3)
    case = dict(case)
    case["logits"] = (x @ w.T).astype(np.float32)            # src/model.rs:521 in f32, as the reference runs it
    want = od.decode_step_reference(case, ids)
    step = build(ids, case, beam)
    proj = Projection(step.decoder, torch.from_numpy(w).cuda())
    got = to_np(proj.beam_step(torch.from_numpy(x).cuda(), torch.from_numpy(case["tokens"]).cuda(), case["sample_begin"],
                               torch.from_numpy(case["sum_logprobs"]).cuda(), debug=True))
    torch.cuda.synchronize()
    step.decoder.ctx.check()
    assert np.array_equal(got["row_flags"], want["row_flags"])
    # rows whose k+1 candidates are separated by more than the tolerance must agree id for id
    gaps = -np.diff(want["cand_logprob"], axis=1)
    clear = np.all(gaps > 2 * LOGPROB_TOL, axis=1)
    assert clear.sum() >= R // 2
    assert np.array_equal(got["cand_token"][clear], want["cand_token"][clear])
    assert np.abs(got["cand_logprob"][clear] - want["cand_logprob"][clear]).max() <= LOGPROB_TOL


def test_nan_in_x_is_reported():
    ids = C.MULTILINGUAL
    case = synth.decode_case(ids, 2, 5, "mixed", seed=2)
    x, w = make_model(ids, 512, 10, seed=5)
    x[3, 17] = np.nan
    step = build(ids, case, 5)
    proj = Projection(step.decoder, torch.from_numpy(w).cuda())
    proj.beam_step(torch.from_numpy(x).cuda(), torch.from_numpy(case["tokens"]).cuda(), case["sample_begin"],
                   torch.from_numpy(case["sum_logprobs"]).cuda())
    torch.cuda.synchronize()
    with pytest.raises(Exception):
        step.decoder.ctx.check()


def test_bad_arguments_raise():
    ids = C.MULTILINGUAL
    case = synth.decode_case(ids, 1, 5, "first", seed=2)
    step = build(ids, case, 5)
    with pytest.raises(Exception):
        Projection(step.decoder, torch.zeros(ids.n_vocab, 500, device="cuda"))      # n_state not a multiple of 32
    with pytest.raises(ValueError):
        Projection(step.decoder, torch.zeros(ids.n_vocab - 1, 512, device="cuda"))


def test_decode_loop_through_the_projection_equals_the_loop_through_its_logits():
    """Ten steps of the reference's loop (src/decode/mod.rs:236-276) on a toy decoder, the KV caches reordered by every
    step's source rows: once with wrs_project_beam_step on the hidden state, once with wrs_beam_step on the logits
    wrs_project_logits materialises from the same hidden state.  Same tokens, sums and finished sequences, step by step."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "host"))
    import toy_whisper as tw
    ids = C.MULTILINGUAL
    n_audio, beam, n_steps = 3, 5, 10
    mel = torch.randn(n_audio, 80, 3000, generator=torch.Generator().manual_seed(4)).cuda()

    def loop(fused):
        model = tw.ToyWhisper(tw.ModelDims(), "cuda", 5, "wrs", eot=ids.eot, eot_bias=0.0)
        opt = wd.DecodeOptions(token_extract_mode=wd.TokenExtractMode.BeamSearch(beam, 1.0), sample_len=n_steps)
        prompt = wd.DecodePrompt(None, ids)
        step = wd.FusedDecodeStep(opt, ids, [prompt], max_rows=n_audio * beam)
        step.reset()
        proj = Projection(step.decoder, model.decoder.token_emb)
        with torch.no_grad():
            feats = model.encoder(mel).repeat_interleave(beam, 0)
            tokens = torch.tensor(prompt.initial_tokens, dtype=torch.int64, device="cuda").repeat(n_audio * beam, 1)
            sums = torch.zeros(n_audio * beam, dtype=torch.float32, device="cuda")
            trace = []
            for i in range(n_steps):
                t_in = tokens[:, -1:] if tokens.shape[-1] > prompt.sample_begin else tokens
                h = (model.decoder.forward_hidden(t_in, feats, i)[:, -1] * model.decoder.scale).contiguous()
                if fused:
                    out = proj.beam_step(h, tokens, prompt.sample_begin, sums)
                else:
                    out = step.decoder.beam_step(proj.logits(h), tokens, prompt.sample_begin, sums)
                tokens, sums = out["next_tokens"], out["sum_logprobs"]
                model.update_cache(out["source_idx"])
                torch.cuda.synchronize()
                step.decoder.ctx.check()
                trace.append((tokens.cpu().numpy().copy(), sums.cpu().numpy().copy(), out["source_idx"].cpu().numpy().copy(),
                              int(out["completed"].item())))
            cands, scores = step.decoder.finalize(tokens, sums)
        assert model.reorders == n_steps
        return trace, cands, scores

    a, ca, sa = loop(True)
    b, cb, sb = loop(False)
    for i, (x, y) in enumerate(zip(a, b)):
        assert np.array_equal(x[0], y[0]) and np.array_equal(x[2], y[2]) and x[3] == y[3], i
        assert np.allclose(x[1], y[1], rtol=RTOL, atol=1e-5), i
    assert ca == cb and np.allclose(sa, sb, rtol=RTOL, atol=1e-5)
    assert len({tuple(r) for r in a[-1][0].tolist()}) > n_audio         # beams diverged: the source rows were not trivial

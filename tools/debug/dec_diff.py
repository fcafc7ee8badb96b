import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))
import numpy as np, torch
from oracle import decode as od
from whisper_rs_b200 import constants as C, synth
from test_gpu_decode import fused_step
ids = C.MULTILINGUAL
for kind, pl in (("second", 0), ("mixed", 0), ("first", 0)):
    case = synth.decode_case(ids, 4, 5, kind, seed=31, prompt_len=pl)
    want = od.decode_step_reference(case, ids)
    got, _ = fused_step(ids, case)
    d = np.abs(got["cand_logprob"] - want["cand_logprob"])
    print(kind, "max abs diff", np.nanmax(d), "argmax", np.unravel_index(np.nanargmax(d), d.shape))
    r = np.unravel_index(np.nanargmax(d), d.shape)[0]
    print(" got ", got["cand_logprob"][r], got["cand_token"][r], got["row_flags"][r])
    print(" want", want["cand_logprob"][r], want["cand_token"][r], want["row_flags"][r])
    print(" per-row max diff", d.max(axis=1))

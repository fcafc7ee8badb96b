import ctypes as C, os, sys
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from whisper_rs_b200 import _lib, constants as K, synth
from whisper_rs_b200 import decode as wd
ids = K.MULTILINGUAL
R = 160
case = synth.decode_case(ids, 32, 5, "mixed", seed=3)
prompt = wd.DecodePrompt(None, ids); prompt.sample_begin = case["sample_begin"]
step = wd.FusedDecodeStep(wd.DecodeOptions(), ids, [prompt], max_rows=R)
step.reset()
lg = [torch.from_numpy(case["logits"]).cuda().roll(7 * b, 0).contiguous() for b in range(6)]
tok = torch.from_numpy(case["tokens"]).cuda(); sums = torch.from_numpy(case["sum_logprobs"]).cuda()
for b in range(6):
    step.step(tok, lg[b], sums)
torch.cuda.synchronize()
n = 148 * 16 * 8
buf = np.zeros(n, dtype=np.uint64)
h = _lib.lib()
h.wrs_debug_trace_read.argtypes = [C.c_void_p, C.c_int]
h.wrs_debug_trace_read(buf.ctypes.data_as(C.c_void_p), n)
t = buf.reshape(-1, 8).astype(np.int64)[160:]
t0 = t[:, 0].min()
dur = (t[:, 3] - t[:, 0]) / 1e3
wait = t[:, 4] / 1e3; red = t[:, 5] / 1e3
full = red > np.percentile(red, 60)
print("warps", len(t), "full-row warps", full.sum())
for name, v in (("duration", dur), ("wait on copies (sum)", wait), ("fetch+reduce (sum)", red), ("first landed", (t[:,1]-t[:,0])/1e3)):
    print(f"{name:24s} all: med {np.median(v):6.2f} p90 {np.percentile(v,90):6.2f} max {v.max():6.2f} | heavy: med {np.median(v[full]):6.2f} max {v[full].max():6.2f}")
print("other (duration - wait - reduce) heavy med", np.median((dur - wait - red)[full]))

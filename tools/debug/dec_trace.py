"""Per-warp timeline of one decode step (library built with WRS_NVCC_EXTRA=-DWRS_TRACE)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from whisper_rs_b200 import _lib, constants as K, synth
from whisper_rs_b200 import decode as wd
ids = K.MULTILINGUAL
R = 160
case = synth.decode_case(ids, 32, 5, sys.argv[1] if len(sys.argv) > 1 else "mixed", seed=3)
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
print("rc", h.wrs_debug_trace_read(buf.ctypes.data_as(C.c_void_p), n))
t = buf.reshape(-1, 8).astype(np.int64)
act = t[:, 0] > 0
t = t[act]
t0 = t[:, 0].min()
rel = lambda c: (t[:, c] - t0) / 1e3
print("warps", len(t))
for name, c in (("start", 0), ("first unit landed", 1), ("last unit reduced", 2), ("warp end", 3)):
    v = rel(c); print(f"{name:20s} min {v.min():7.2f} med {np.median(v):7.2f} p90 {np.percentile(v,90):7.2f} max {v.max():7.2f} us")
mg = t[:, 5] > 0
print("row merges", int(t[:, 5].sum()), "avg us", (t[mg, 4] / t[mg, 5]).mean() / 1e3, "max us", (t[mg, 4] / t[mg, 5]).max() / 1e3)
tl = t[:, 7] > 0
print("beam tails", int(t[:, 7].sum()), "avg us", (t[tl, 6] / t[tl, 7]).mean() / 1e3, "max us", (t[tl, 6] / t[tl, 7]).max() / 1e3)
end = rel(3); order = np.argsort(-end)[:8]
for i in order:
    print("late warp", i, "start %.2f first %.2f last %.2f end %.2f merges %d (%.2f us) tails %d (%.2f us)" % (
        rel(0)[i], rel(1)[i], rel(2)[i], end[i], t[i, 5], t[i, 4] / 1e3, t[i, 7], t[i, 6] / 1e3))
print("hist of last-unit-reduced (us):", np.histogram(rel(2), bins=10)[0], np.histogram(rel(2), bins=10)[1].round(1))

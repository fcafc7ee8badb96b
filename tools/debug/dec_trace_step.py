"""Timeline of one decode step (library built with -DWRS_TRACE): scan warps and tail rows on the global timer."""
import ctypes as C, sys
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
lg = [torch.from_numpy(case["logits"]).cuda().roll(7 * b, 0).contiguous() for b in range(8)]
tok = torch.from_numpy(case["tokens"]).cuda(); sums = torch.from_numpy(case["sum_logprobs"]).cuda()
for b in range(8):
    step.step(tok, lg[b], sums)
torch.cuda.synchronize()
NW = 148 * 24
n = (NW + 1024) * 8
buf = np.zeros(n, dtype=np.uint64)
h = _lib.lib()
h.wrs_debug_trace_read.argtypes = [C.c_void_p, C.c_int]
h.wrs_debug_trace_read(buf.ctypes.data_as(C.c_void_p), n)
t = buf.reshape(-1, 8).astype(np.int64)
sc = t[:NW]; sc = sc[sc[:, 0] > 0]
tl = t[NW:NW + R]
t0 = sc[:, 0].min()
pc = lambda v: [round(float(np.percentile(v, q)), 2) for q in (0, 10, 50, 90, 99, 100)]
print("scan warps", len(sc))
print(" start        ", pc((sc[:, 0] - t0) / 1e3))
print(" first landed ", pc((sc[:, 1] - t0) / 1e3))
print(" end          ", pc((sc[:, 3] - t0) / 1e3))
print(" wait sum     ", pc(sc[:, 4] / 1e3))
print(" fold sum     ", pc(sc[:, 5] / 1e3))
print("tail rows", R)
print(" start        ", pc((tl[:, 0] - t0) / 1e3))
print(" merged       ", pc((tl[:, 0] - t0 + tl[:, 4]) / 1e3))
print(" merge time   ", pc(tl[:, 4] / 1e3))
w0 = tl[::5]
print(" after barrier", pc((w0[:, 1] - t0) / 1e3))
print(" end (warp 0) ", pc((w0[:, 3] - t0) / 1e3))
print(" merge: sums done / cells chosen / cells loaded / merged (us from tail warp start)")
print("   ", pc(tl[:, 5] / 1e3), pc(tl[:, 6] / 1e3), pc(tl[:, 7] / 1e3), pc(tl[:, 4] / 1e3))

"""Which scan warps are slow?  (library built with -DWRS_TRACE)  Prints the slowest runs with the rows / history cases they cover."""
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
lg = [torch.from_numpy(case["logits"]).cuda().contiguous() for b in range(6)]
tok = torch.from_numpy(case["tokens"]).cuda(); sums = torch.from_numpy(case["sum_logprobs"]).cuda()
for b in range(6):
    step.step(tok, lg[b], sums)
torch.cuda.synchronize()
n = 148 * 16 * 8
buf = np.zeros(n, dtype=np.uint64)
h = _lib.lib()
h.wrs_debug_trace_read.argtypes = [C.c_void_p, C.c_int]
h.wrs_debug_trace_read(buf.ctypes.data_as(C.c_void_p), n)
t = buf.reshape(-1, 8).astype(np.int64)
rpr = 14
red = t[:, 5] / 1e3; wait = t[:, 4] / 1e3; dur = (t[:, 3] - t[:, 0]) / 1e3
n_runs = rpr * R
order = np.argsort(-red[:n_runs])
order = order[order >= 160][:25]
print("run   row:i  case  reduce  wait  dur  first-landed  end-rel")
t0 = t[160:n_runs, 0].min()
for g in order:
    print(f"{g:5d} {g//rpr:4d}:{g%rpr:2d}  {(g//rpr)%4}   {red[g]:6.2f} {wait[g]:5.2f} {dur[g]:6.2f} {(t[g,1]-t[g,0])/1e3:6.2f} {(t[g,3]-t0)/1e3:6.2f}")
v = red[160:n_runs]; e = (t[160:n_runs, 3] - t0) / 1e3; st = (t[160:n_runs, 0] - t0) / 1e3
print("reduce percentiles", [round(float(np.percentile(v, p)), 2) for p in (10, 50, 75, 90, 95, 99, 100)])
print("start  percentiles", [round(float(np.percentile(st, p)), 2) for p in (10, 50, 75, 90, 95, 99, 100)])
print("end    percentiles", [round(float(np.percentile(e, p)), 2) for p in (10, 50, 75, 90, 95, 99, 100)])
print("first-landed percentiles", [round(float(np.percentile((t[160:n_runs,1]-t[160:n_runs,0])/1e3, p)), 2) for p in (10, 50, 90, 100)])
for i in range(rpr):
    m = np.array([g % rpr == i for g in range(160, n_runs)])
    print("run-in-row", i, "median reduce", round(float(np.median(v[m])), 2), "p90", round(float(np.percentile(v[m], 90)), 2), "max", round(float(v[m].max()), 2))
# ---- per-CTA view (interleaved dealing: CTA of run g = g % 148)
G = 148
ends = np.zeros(G); reds = np.zeros(G); fl = np.zeros(G); units = np.zeros(G)
tt0 = t[:n_runs, 0][t[:n_runs, 0] > 0].min()
for g in range(n_runs):
    if t[g, 3] == 0: continue
    b = g % G
    ends[b] = max(ends[b], (t[g, 3] - tt0) / 1e3); reds[b] += red[g]; fl[b] = max(fl[b], (t[g, 1] - tt0) / 1e3)
    units[b] += 5 if g % rpr < 5 else (4 if g % rpr < 13 else 2)
o = np.argsort(-ends)
print("CTA  end   sum-reduce  last-first-landed  units   (slowest 12, fastest 5)")
for b in list(o[:12]) + list(o[-5:]):
    print(f"{b:4d} {ends[b]:6.2f} {reds[b]:8.2f} {fl[b]:8.2f} {int(units[b]):5d}")
print("corr(end, units)", np.corrcoef(ends, units)[0, 1], "corr(end, sum-reduce)", np.corrcoef(ends, reds)[0, 1])
# within the slowest CTA: all its warps
b = o[0]
print("warps of CTA", b)
for g in range(b, n_runs, G):
    print(f"  gen {t[g,6]:3d} merges {t[g,7]:8d}", end="")
    print(f"  run {g:5d} row {g//rpr:3d} i {g%rpr:2d} case {(g//rpr)%4} start {(t[g,0]-tt0)/1e3:5.2f} landed {(t[g,1]-tt0)/1e3:5.2f} end {(t[g,3]-tt0)/1e3:6.2f} reduce {red[g]:6.2f} wait {wait[g]:5.2f}")

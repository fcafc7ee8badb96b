"""Per-phase cycles of the log-mel slice loop (library built with WRS_NVCC_EXTRA=-DWRS_TRACE)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
from whisper_rs_b200 import _lib, audio, synth
if len(sys.argv) > 1:
    _lib.context(0).set_option("logmel_fused", int(sys.argv[1]))
pcm = torch.from_numpy(synth.audio_batch(16, seed=1)).repeat(4, 1).contiguous().cuda()
for _ in range(3):
    audio.log_mel_spectrogram_batch(pcm)
torch.cuda.synchronize()
buf = np.zeros(148 * 12, dtype=np.uint64)
h = _lib.lib()
h.wrs_debug_logmel_trace_read.argtypes = [C.c_void_p, C.c_int]
print("rc", h.wrs_debug_logmel_trace_read(buf.ctypes.data_as(C.c_void_p), 148 * 12))
t = buf.reshape(148, 12).astype(np.float64)
slices = 3008 / 148.0
names = ["stage A", "stage B (+copy issue, poll)", "C: raw loads / edge staging (warp 0)", "C: mel (warp 0)", "C: finalize + max (warp 0)", "C: barrier wait (warp 0)", "raw store + queue", "wait for PCM tile", "B: arrival of warp 15 (count/poll) after barrier A", "B: arrival of warp 5 (q = 5)", "loop top before the tile wait", "B: arrival of warp 13 (copy issue)"]
for k, n in enumerate(names[:12]):
    print(f"{n:34s} mean {t[:, k].mean() / slices:9.0f} cycles/slice   (min CTA {t[:, k].min() / slices:.0f}, max {t[:, k].max() / slices:.0f})")
print("sum per slice", t[:, [0,1,2,3,4,5,6,7,10]].sum(1).mean() / slices)

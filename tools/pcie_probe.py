#!/usr/bin/env python
"""Host <-> device copy rates on this box: each direction alone and both at once, 61 MB and 123 MB pinned buffers."""
import time, torch
dev = torch.device("cuda", 0)
for mb in (61.44, 122.88):
    n = int(mb * 1e6 / 4)
    h_in, h_out = torch.empty(n).pin_memory(), torch.empty(n).pin_memory()
    d_in, d_out = torch.empty(n, device=dev), torch.empty(n, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    def run(h2d, d2h, reps=20):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            if h2d:
                with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps
    for _ in range(2): run(True, True, 3)
    a, b, c = run(True, False), run(False, True), run(True, True)
    print(f"{mb:7.2f} MB: H2D alone {mb/a/1e3:6.1f} GB/s, D2H alone {mb/b/1e3:6.1f} GB/s, both at once {mb/c/1e3:6.1f} GB/s each ({1e3*c:.2f} ms)")

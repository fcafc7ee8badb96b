#!/usr/bin/env python
"""Timing-only experiment: build libwrs variants with -DWRS_EXP_* (parts of K1 switched off: results are wrong) and time
wrs_logmel_batch on 64 x 30 s.  usage: k1_variants.py build | run"""
import ctypes as C, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VDIR = os.path.join(ROOT, "whisper_rs_b200", "lib", "variants")
VARIANTS = {"base": [], "noF": ["-DWRS_EXP_NO_F"], "noC_noF": ["-DWRS_EXP_NO_C", "-DWRS_EXP_NO_F"],
            "noLog_noF": ["-DWRS_EXP_NO_LOG", "-DWRS_EXP_NO_F"], "noCstore_noF": ["-DWRS_EXP_NO_CSTORE", "-DWRS_EXP_NO_F"],
            "noLog_noCstore_noF": ["-DWRS_EXP_NO_LOG", "-DWRS_EXP_NO_CSTORE", "-DWRS_EXP_NO_F"],
            "noAmath_noF": ["-DWRS_EXP_NO_AMATH", "-DWRS_EXP_NO_F"], "noBmath_noF": ["-DWRS_EXP_NO_BMATH", "-DWRS_EXP_NO_F"],
            "noAmath_noBmath_noC_noF": ["-DWRS_EXP_NO_AMATH", "-DWRS_EXP_NO_BMATH", "-DWRS_EXP_NO_C", "-DWRS_EXP_NO_F"]}
if sys.argv[1] == "build":
    os.makedirs(VDIR, exist_ok=True)
    csrc = os.path.join(ROOT, "whisper_rs_b200", "csrc")
    for name, defs in VARIANTS.items():
        cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-shared", *defs,
               os.path.join(csrc, "ctx.cu"), os.path.join(csrc, "logmel.cu"), "-o", os.path.join(VDIR, f"lib_{name}.so"), "-lcudart"]
        subprocess.run(cmd, check=True)
    print("built", list(VARIANTS))
else:
    import torch
    x = (0.1 * torch.randn(64, 480000, device="cuda")).contiguous()
    out = torch.empty(64, 80, 3000, device="cuda")
    for name in VARIANTS:
        lib = C.CDLL(os.path.join(VDIR, f"lib_{name}.so"))
        ctx = C.c_void_p()
        assert lib.wrs_ctx_create(0, C.byref(ctx)) == 0
        f = lib.wrs_logmel_batch
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(5): assert f(ctx, x.data_ptr(), 64, 480000, 480000, out.data_ptr(), st) == 0
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): f(ctx, x.data_ptr(), 64, 480000, 480000, out.data_ptr(), st)
        e1.record(); torch.cuda.synchronize()
        print(f"{name:20s} {e0.elapsed_time(e1) / 20 * 1000:8.1f} us", flush=True)

#!/usr/bin/env python
"""Timing-only experiment: libwrs variants with -DWRS_EXP_PJ_* (parts of the projection kernel switched off: results are
wrong) timed on config 3's shape.  usage: proj_variants.py build | run"""
import ctypes as C, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VDIR = os.path.join(ROOT, "whisper_rs_b200", "lib", "variants")
VARIANTS = {"base": [], "noepi": ["-DWRS_EXP_PJ_NOEPI"], "nomma_noepi": ["-DWRS_EXP_PJ_NOMMA", "-DWRS_EXP_PJ_NOEPI"],
            "oneprod_noepi": ["-DWRS_EXP_PJ_ONEPROD", "-DWRS_EXP_PJ_NOEPI"], "oneprod": ["-DWRS_EXP_PJ_ONEPROD"], "noins": ["-DWRS_EXP_PJ_NOINS"], "nosum": ["-DWRS_EXP_PJ_NOSUM"],
            "noins_nosum": ["-DWRS_EXP_PJ_NOINS", "-DWRS_EXP_PJ_NOSUM"], "times": ["-DWRS_EXP_PJ_TIMES"]}
if sys.argv[1] == "build":
    os.makedirs(VDIR, exist_ok=True)
    csrc = os.path.join(ROOT, "whisper_rs_b200", "csrc")
    srcs = [os.path.join(csrc, f) for f in ("ctx.cu", "decode.cu", "decode_host.cu", "logmel.cu", "kv_reorder.cu", "no_speech.cu")]
    procs = []
    for name, defs in VARIANTS.items():
        cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-shared", *defs,
               *srcs, "-o", os.path.join(VDIR, f"libpj_{name}.so"), "-lcudart"]
        procs.append(subprocess.Popen(cmd))
    assert all(p.wait() == 0 for p in procs)
    print("built", list(VARIANTS))
else:
    import numpy as np, torch
    sys.path.insert(0, ROOT)
    from whisper_rs_b200 import _lib, constants as Cn, synth
    from whisper_rs_b200 import decode as wd
    ids = Cn.MULTILINGUAL
    R, D, V = 160, 512, ids.n_vocab
    case = synth.decode_case(ids, 32, 5, "mixed", seed=3)
    cfg = wd.build_decoder_cfg(wd.DecodeOptions(), ids, filters=True)
    from whisper_rs_b200.decode.core import suppress_list
    sup = np.asarray(list(suppress_list(wd.DecodeOptions(), ids)), dtype=np.int32)
    emb = torch.randn(V, D, device="cuda") * (3.0 / D ** 0.5)
    xs = [torch.randn(R, D, device="cuda") for _ in range(8)]
    tok = torch.from_numpy(case["tokens"]).cuda(); sums = torch.from_numpy(case["sum_logprobs"]).cuda()
    Lc = tok.shape[1]
    nxt = torch.empty(R, Lc + 1, dtype=torch.int64, device="cuda"); src = torch.empty(R, dtype=torch.int32, device="cuda")
    out = torch.empty(R, device="cuda"); done = torch.empty(1, dtype=torch.int32, device="cuda")
    logits = torch.empty(R, V, device="cuda")
    P = C.c_void_p
    for name in (sys.argv[2:] or VARIANTS):
        lib = C.CDLL(os.path.join(VDIR, f"libpj_{name}.so"))
        ctx, dec = P(), P()
        assert lib.wrs_ctx_create(0, C.byref(ctx)) == 0
        lib.wrs_decoder_create.argtypes = [P, P, P, C.c_int32, C.c_int32, P]
        assert lib.wrs_decoder_create(ctx, C.byref(cfg), sup.ctypes.data_as(P), len(sup), R, C.byref(dec)) == 0
        st = torch.cuda.current_stream().cuda_stream
        lib.wrs_projection_create.argtypes = [P, P, C.c_int64, C.c_int32, P, P]
        pjs = []
        for _ in range(2):
            pj = P()
            assert lib.wrs_projection_create(dec, emb.data_ptr(), emb.stride(0), D, st, C.byref(pj)) == 0
            pjs.append(pj)
        i64 = C.c_int64
        lib.wrs_project_logits.argtypes = [P, P, i64, i64, P, i64, P]
        lib.wrs_project_beam_step.argtypes = [P, P, i64, i64, P, i64, i64, i64, P, P, i64, P, P, P, P, P]
        lib.wrs_decoder_reset.argtypes = [P, P]

        def plain(i):
            assert lib.wrs_project_logits(pjs[i & 1], xs[i % 8].data_ptr(), D, R, logits.data_ptr(), V, st) == 0

        def fused(i):
            if i % 8 == 0: lib.wrs_decoder_reset(dec, st)
            assert lib.wrs_project_beam_step(pjs[i & 1], xs[i % 8].data_ptr(), D, R, tok.data_ptr(), tok.stride(0), Lc, case["sample_begin"],
                                             sums.data_ptr(), nxt.data_ptr(), nxt.stride(0), src.data_ptr(), out.data_ptr(), done.data_ptr(), None, st) == 0
        res = []
        for fn in (plain, fused):
            for i in range(8): fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(64): fn(i)
            e1.record(); torch.cuda.synchronize()
            res.append(e0.elapsed_time(e1) / 64 * 1000)
        print(f"{name:16s} logits {res[0]:8.1f} us   fused step {res[1]:8.1f} us", flush=True)
        if name == "times":
            buf = (C.c_ulonglong * (148 * 32))()
            lib.wrs_debug_pj_times(buf, 148 * 32)
            clk = (C.c_longlong * (148 * 32))()
            lib.wrs_debug_pj_clk(clk, 148 * 32)
            for cta in (1, 73):
                print(f"cta {cta} unit 2 select warp 0, per chunk [cycles to max, to staged, to loop end, iterations]: " +
                      " ".join(str([clk[cta * 32 + 4 * c + k] for k in range(4)]) for c in range(6)))
            for cta in (0, 73, 147):
                t = [buf[cta * 32 + k] for k in range(16)]
                t0 = min(x for x in t if x)
                print(f"cta {cta} u0 chunks (loaded, done): " + " ".join(f"{(buf[cta*32+16+2*c]-t0)/1e3:5.1f},{(buf[cta*32+17+2*c]-t0)/1e3:5.1f}" for c in range(6)))
                print(f"cta {cta}: " + " | ".join(f"u{u}: mma {(t[4*u]-t0)/1e3:5.1f}-{(t[4*u+1]-t0)/1e3:5.1f} epi {(t[4*u+2]-t0)/1e3:5.1f}-{(t[4*u+3]-t0)/1e3:5.1f}" for u in range(4)))

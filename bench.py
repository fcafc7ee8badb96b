#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on B200: 30 s-segment log-mels/s (headline, config 2: 64 x 30 s segments)
and beam decode row-steps/s (config 3: beam 5 x batch 32, V 51865), each with its HBM roofline fraction,
an end-to-end figure through the host-buffer C-ABI call, and the oracle (reference ATen op sequence on the
CPU) timed on this box's host cores.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path over one batch: one fused log-mel launch over 64 segments (headline
region), and -- timed separately, reported under "decode_step" -- one fused filter+log-softmax+top-k+beam-merge
pass over 160 rows.  Segments/rows shard across ranks with no collective (weak scaling: every rank has its own
batch); NCCL is used only for the barrier and the max-over-ranks of the timings.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

N_SEG, N_SAMPLES, N_MELS, N_FRAMES = 64, 480000, 80, 3000
N_AUDIO, BEAM, VOCAB = 32, 5, 51865
LOGMEL_BYTES_PER_SEG = N_SAMPLES * 4 + N_MELS * N_FRAMES * 4          # 2 880 000 B  
DECODE_BYTES_PER_ROW = VOCAB * 4                                      # 207 460 B
# dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu --set full captures of this same
# command (profiles/r1_v4_logmel_ncu.txt, profiles/r1_v8_decode_scan_ncu.txt, profiles/r1_v8_decode_tail_ncu.txt);
# re-measure when a kernel's traffic changes
LOGMEL_DRAM_TRAFFIC_PER_LAUNCH = 122.96e6 + 30.77e6                   # 64 segments: PCM read once, output half evicted
# 160 rows: logits read once by the scan launch (33.24 MB) + the tail launch (1.48 MB under ncu's cold caches: partials and
# the ~6 winning 28-id cells per row it reads again; in a running step most of that is still in L2)
DECODE_DRAM_TRAFFIC_PER_STEP = 33.24e6 + 1.48e6
TRAFFIC_SOURCE = ("ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch "
                  "(profiles/r1_v4_logmel_ncu.txt, profiles/r1_v8_decode_*_ncu.txt)")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled while the GPU works (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 7] or [r for _, r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons, "samples": len(rows),
                "power_w_max": max(float(r[2]) for r in rows if r[2].replace(".", "").isdigit())}


def cpu_logmel_time(n_segments, threads):
    """Oracle (reference src/audio.rs:230-259 op sequence) over `n_segments` 30 s segments, one call per
    segment as TranscribeTask would (it is not batch-safe)."""
    from oracle import audio as oa
    from whisper_rs_b200 import synth
    torch.set_num_threads(threads)
    bank = synth.audio_batch(4, seed=0)
    filt = oa.mel_filter()
    xs = [torch.from_numpy(bank[i % 4]) for i in range(n_segments)]
    oa.log_mel_spectrogram(xs[0], filt)
    t0 = time.perf_counter()
    for x in xs:
        oa.log_mel_spectrogram(x, filt)
    return time.perf_counter() - t0


def cpu_decode_time(case, ids, n_steps, threads):
    from oracle import decode as od
    torch.set_num_threads(threads)
    od.decode_step_reference(case, ids)
    t0 = time.perf_counter()
    for _ in range(n_steps):
        od.decode_step_reference(case, ids)
    return time.perf_counter() - t0


def run_reference(args, rank):
    """--impl reference: the reference's own CPU path for the headline metric (the oracle = the same ATen op
    sequence; the Rust crate cannot be built here), all host threads, bounded sample per step."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    for _ in range(max(args.warmup, 1)):
        t_seg = cpu_logmel_time(2, threads) / 2
    # segments per step: a bounded sample of the 64-segment batch, sized so that K steps take about a minute
    per_step = int(max(1, min(N_SEG, 60.0 / (args.steps * t_seg))))
    t = 0.0
    for _ in range(args.steps):
        t += cpu_logmel_time(per_step, threads)
    value = per_step * args.steps / t
    from whisper_rs_b200 import constants as C, synth
    case = synth.decode_case(C.MULTILINGUAL, N_AUDIO, BEAM, "mixed", seed=3)
    dsteps = max(1, min(args.steps, 5))
    dt = cpu_decode_time(case, C.MULTILINGUAL, dsteps, threads)
    line = {
        "impl": "reference", "metric": "log-mels/s (30 s segments)", "value": value, "unit": "segments/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "config 2: log-mel front-end, 64 x 30 s synthetic 16 kHz segments (bounded sample: " + str(per_step) + " segments/step)"},
        "cpu_baseline": {"value": value, "unit": "segments/s", "cores": threads, "kind": "port",
                         "sample": f"{per_step} segments x {args.steps} steps, torch CPU ATen op sequence of src/audio.rs:230-259"},
        "e2e": {"value": value, "unit": "segments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "decode_step": {"value": N_AUDIO * BEAM * dsteps / dt, "unit": "row-steps/s", "cores": threads, "kind": "port",
                        "sample": f"{dsteps} steps of 160 rows x V {VOCAB}, filters + beam update"},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def timed(stream_fn, steps, dist_barrier):
    """Time `steps` calls of stream_fn(i) with CUDA events on the current stream, barrier + sync both sides."""
    torch.cuda.synchronize()
    dist_barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        stream_fn(i)
    e1.record()
    torch.cuda.synchronize()
    dist_barrier()
    return e0.elapsed_time(e1) * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank)
    args.warmup = max(args.warmup, 3)

    import torch.distributed as dist
    from whisper_rs_b200 import _lib, audio, synth
    from whisper_rs_b200 import constants as C
    from whisper_rs_b200 import decode as wd
    from whisper_rs_b200 import build as wbuild
    if rank == 0:
        wbuild.build()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    barrier()
    if rank != 0:
        _lib.lib()
    dev = torch.device("cuda", local)
    ctx = _lib.context(local)
    L = _lib.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    hbm_peak, peak_src = peaks()
    sampler = ClockSampler(local) if rank == 0 else None
    t_gpu0 = time.time()

    # ---------------- log-mel, config 2: 64 x 30 s segments per GPU per step ----------------
    NBUF = 3   # rotate 3 in/out sets (553 MB > 126 MB L2): every step reads cold PCM
    bank = torch.from_numpy(synth.audio_batch(16, seed=1000 * rank))
    host_pcm = bank.repeat(N_SEG // 16, 1).contiguous().pin_memory()
    d_in = [host_pcm.to(dev, non_blocking=True).clone() for _ in range(NBUF)]
    d_out = [torch.empty(N_SEG, N_MELS, N_FRAMES, dtype=torch.float32, device=dev) for _ in range(NBUF)]
    host_out = torch.empty(N_SEG, N_MELS, N_FRAMES, dtype=torch.float32).pin_memory()

    def logmel_step(i):
        b = i % NBUF
        _lib.check(L.wrs_logmel_batch(ctx.handle, d_in[b].data_ptr(), N_SEG, N_SAMPLES, N_SAMPLES, d_out[b].data_ptr(), stream),
                   ctx.handle)

    for i in range(args.warmup):
        logmel_step(i)
    t_logmel = max_over_ranks(timed(logmel_step, args.steps, barrier))
    ctx.check(stream)
    logmel_mode = ctx.get_stat("logmel_mode")
    seg_per_s = world * N_SEG * args.steps / t_logmel
    ach = N_SEG * LOGMEL_BYTES_PER_SEG / (t_logmel / args.steps) / 1e9

    def logmel_e2e(i):
        _lib.check(L.wrs_logmel_batch_host(ctx.handle, host_pcm.data_ptr(), N_SEG, N_SAMPLES, N_SAMPLES, host_out.data_ptr()),
                   ctx.handle)

    e2e_steps = max(3, min(args.steps, 30))
    for i in range(3):
        logmel_e2e(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        logmel_e2e(i)          # synchronous: H2D + kernel + D2H + stream sync inside the call
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    barrier()

    # ---------------- decode step, config 3: 32 audios x beam 5 = 160 rows, V = 51865 ----------------
    ids = C.MULTILINGUAL
    R = N_AUDIO * BEAM
    DBUF = 8   # rotate 8 logits sets (266 MB > L2)
    case = synth.decode_case(ids, N_AUDIO, BEAM, "mixed", seed=3 + rank)
    prompt = wd.DecodePrompt(None, ids)
    prompt.sample_begin = case["sample_begin"]
    step = wd.FusedDecodeStep(wd.DecodeOptions(), ids, [prompt], max_rows=R, device=local)
    step.reset()
    dec = step.decoder
    host_logits = torch.from_numpy(case["logits"]).pin_memory()
    d_logits = [host_logits.to(dev).roll(7 * b, dims=0).contiguous() for b in range(DBUF)]
    d_tokens = torch.from_numpy(case["tokens"]).to(dev)
    d_sums = torch.from_numpy(case["sum_logprobs"]).to(dev)
    Lcur = d_tokens.shape[1]
    # what the host reads back every step (next tokens, sums, completed) sits in ONE device block: one D2H copy per step
    nb_next, nb_sum = R * (Lcur + 1) * 8, R * 4
    d_outs = torch.zeros(nb_next + nb_sum + 4, dtype=torch.uint8, device=dev)
    h_outs = torch.zeros(nb_next + nb_sum + 4, dtype=torch.uint8).pin_memory()
    d_next = d_outs[:nb_next].view(torch.int64).view(R, Lcur + 1)
    d_sum_out = d_outs[nb_next:nb_next + nb_sum].view(torch.float32)
    d_done = d_outs[nb_next + nb_sum:].view(torch.int32)
    h_done = h_outs[nb_next + nb_sum:].view(torch.int32)
    d_src = torch.empty(R, dtype=torch.int32, device=dev)

    def decode_call(logits):
        _lib.check(L.wrs_beam_step(dec.handle, logits.data_ptr(), logits.stride(0), R, d_tokens.data_ptr(), d_tokens.stride(0),
                                   Lcur, prompt.sample_begin, d_sums.data_ptr(), d_next.data_ptr(), d_next.stride(0),
                                   d_src.data_ptr(), d_sum_out.data_ptr(), d_done.data_ptr(), None, stream), ctx.handle)

    def decode_step(i):
        if i % 64 == 0:
            dec.reset()          # keep the finished store from saturating: same work every step
        decode_call(d_logits[i % DBUF])

    for i in range(args.warmup):
        decode_step(i)
    t_dec_eager = max_over_ranks(timed(decode_step, args.steps, barrier))
    # A step is a few microseconds of device time, less than one Python/ctypes launch: in the reference loop the
    # launch hides behind the decoder forward (src/decode/mod.rs:236-256); here DBUF consecutive steps (one per
    # logits buffer, reset first) are captured once in a CUDA graph and the graph is replayed.
    GSTEPS = DBUF
    cap_stream = torch.cuda.Stream(dev)
    graph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(graph, stream=cap_stream):
        cs = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(L.wrs_decoder_reset(dec.handle, cs), ctx.handle)
        for b in range(GSTEPS):
            lg = d_logits[b]
            _lib.check(L.wrs_beam_step(dec.handle, lg.data_ptr(), lg.stride(0), R, d_tokens.data_ptr(), d_tokens.stride(0),
                                       Lcur, prompt.sample_begin, d_sums.data_ptr(), d_next.data_ptr(), d_next.stride(0),
                                       d_src.data_ptr(), d_sum_out.data_ptr(), d_done.data_ptr(), None, cs), ctx.handle)
    n_replays = max(1, args.steps // GSTEPS)
    dec_steps = n_replays * GSTEPS
    for i in range(3):
        graph.replay()
    t_dec = max_over_ranks(timed(lambda i: graph.replay(), n_replays, barrier))
    rows_per_s = world * R * dec_steps / t_dec
    dec_ach = R * DECODE_BYTES_PER_ROW / (t_dec / dec_steps) / 1e9

    # step-0 no-speech probability (src/decode/mod.rs:250-255): one unfiltered softmax per row
    d_ns = torch.empty(R, dtype=torch.float32, device=dev)
    ns_graph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(ns_graph, stream=cap_stream):
        cs = torch.cuda.current_stream(dev).cuda_stream
        for b in range(GSTEPS):
            lg = d_logits[b]
            _lib.check(L.wrs_no_speech_probs(dec.handle, lg.data_ptr(), lg.stride(0), R, d_ns.data_ptr(), cs), ctx.handle)
    for i in range(3):
        ns_graph.replay()
    t_ns = max_over_ranks(timed(lambda i: ns_graph.replay(), n_replays, barrier))
    ns_ach = R * DECODE_BYTES_PER_ROW / (t_ns / dec_steps) / 1e9

    # end to end as the reference's caller sees it (decode/mod.rs:257-276): logits are already on the device
    # (north_star: raw device pointers from the tch forward), the host reads back tokens + scores + completed.
    def decode_e2e(i, h2d_logits):
        lg = d_logits[i % DBUF]
        if h2d_logits:
            lg.copy_(host_logits, non_blocking=True)
        decode_call(lg)
        h_outs.copy_(d_outs, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return bool(h_done[0])          # the `completed` read is the per-step sync in the reference loop too

    e2e_dec = {}
    for name, h2d in (("device_logits", False), ("host_logits", True)):
        for i in range(3):
            decode_e2e(i, h2d)
        n = max(3, min(args.steps, 200))
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for i in range(n):
            decode_e2e(i, h2d)
        torch.cuda.synchronize()
        e2e_dec[name] = world * R * n / max_over_ranks(time.perf_counter() - t0)
    ctx.check(stream)

    # ---------------- KV-cache beam reorder: base-shaped decoder, config 3 rows ----------------
    # Cached::update_cache (src/model.rs:53-62) on every cached K/V of a 6-block, 512-wide decoder (examples/transcribe.rs
    # dims): self-attention caches [160, 32, 512] and cross-attention caches [160, 1500, 512], f32, 24 tensors, one launch.
    KV_LAYERS, KV_STATE, KV_SELF, KV_CROSS = 6, 512, 32, 1500
    kv_src, kv_dst = [], []
    for _ in range(KV_LAYERS):
        for ctx_len in (KV_SELF, KV_SELF, KV_CROSS, KV_CROSS):       # self K, self V, cross K, cross V
            kv_src.append(torch.empty((R, ctx_len, KV_STATE), dtype=torch.float32, device=dev).normal_())
            kv_dst.append(torch.empty_like(kv_src[-1]))
    kv_bytes = 2 * sum(t.numel() * 4 for t in kv_src)                # every byte read once and written once
    rng = np.random.default_rng(7 + rank)
    kv_idx = torch.from_numpy(np.concatenate([a * BEAM + rng.integers(0, BEAM, size=BEAM) for a in range(N_AUDIO)])
                              .astype(np.int32)).to(dev)
    from whisper_rs_b200 import cache as wcache

    def kv_step(i):
        if i % 2 == 0:
            wcache.update_cache(kv_src, kv_idx, kv_dst)
        else:
            wcache.update_cache(kv_dst, kv_idx, kv_src)

    kv_steps = max(4, min(args.steps, 40))
    for i in range(4):
        kv_step(i)
    t_kv = max_over_ranks(timed(kv_step, kv_steps, barrier))
    kv_ach = kv_bytes / (t_kv / kv_steps) / 1e9
    kv_ok = bool(torch.equal(kv_src[2][kv_idx.long()], wcache.update_cache(kv_src[2:3], kv_idx)[0]))
    ctx.check(stream)
    del kv_src, kv_dst
    t_gpu1 = time.time()
    clocks = sampler.stop(t_gpu0, t_gpu1) if sampler else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---------------- CPU baseline: the oracle on this box's host cores (rank 0, N = 1 only) ----------------
    cpu = cpu_dec = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        t1 = cpu_logmel_time(4, 1) / 4
        n_all = 64
        t_all = cpu_logmel_time(n_all, threads) / n_all
        cpu = {"value": 1.0 / t_all, "unit": "segments/s", "cores": threads, "kind": "port",
               "sample": f"{n_all} of the 64 x 30 s segments, one call per segment, torch {torch.__version__} CPU ATen op "
                         f"sequence of src/audio.rs:230-259 (reference Rust crate cannot be built here)",
               "single_thread_value": 1.0 / t1}
        nd = 5
        td = cpu_decode_time({k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in case.items()}, ids, nd, threads)
        cpu_dec = {"value": R * nd / td, "unit": "row-steps/s", "cores": threads, "kind": "port",
                   "sample": f"{nd} steps of 160 rows x V {VOCAB}: 3 filters + BeamSearchTokenExtractor::update restatement"}

    cpu_kv = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import cache as ocache
        torch.set_num_threads(threads)
        blk = [torch.randn((R, n, KV_STATE)) for n in (KV_SELF, KV_SELF, KV_CROSS, KV_CROSS)]
        hidx = kv_idx.cpu().long()
        ocache.update_cache(blk, hidx)
        reps = 3
        t0 = time.perf_counter()
        for _ in range(reps):
            ocache.update_cache(blk, hidx)
        tk = (time.perf_counter() - t0) / reps * KV_LAYERS           # one block timed, scaled to the decoder's six
        cpu_kv = {"value": R / tk, "unit": "row-steps/s", "cores": threads, "kind": "port",
                  "sample": f"one of the {KV_LAYERS} decoder blocks (4 tensors, 1.0 GB read + 1.0 GB written), x{KV_LAYERS}: "
                            "torch CPU Tensor.index, the call src/model.rs:59 makes"}

    line = {
        "metric": "log-mels/s (30 s segments)", "value": seg_per_s, "unit": "segments/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_logmel / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "config 2: log-mel front-end, 64 x 30 s synthetic 16 kHz segments per GPU per step",
                   "segments_per_step_per_gpu": N_SEG, "samples_per_segment": N_SAMPLES,
                   "cache": f"inputs/outputs rotate over {NBUF} buffer sets (553 MB total > 126 MB L2)",
                   "logmel_mode": "fused single launch" if logmel_mode == 1 else "two launches",
                   "parallelism": f"segment-sharded x{world}, no collective"},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                     "traffic": LOGMEL_DRAM_TRAFFIC_PER_LAUNCH, "traffic_source": TRAFFIC_SOURCE,
                     "peak_source": peak_src, "frac_of_8TBps": ach / 8000.0,
                     "algorithmic_bytes_per_launch": N_SEG * LOGMEL_BYTES_PER_SEG,
                     "kernel": "wrs::logmel_kernel<true> (persistent, cooperative; one launch per batch)"},
        "cpu_baseline": cpu,
        "e2e": {"value": world * N_SEG * e2e_steps / t_e2e, "unit": "segments/s",
                "h2d_bytes_per_step": N_SEG * N_SAMPLES * 4, "d2h_bytes_per_step": N_SEG * N_MELS * N_FRAMES * 4,
                "api": "wrs_logmel_batch_host (pinned host PCM in, host log-mel out, synchronous)", "steps": e2e_steps},
        "gpu_launches": args.steps * (1 if logmel_mode == 1 else 2),
        "clocks": clocks,
        "decode_step": {
            "metric": "beam decode row-steps/s", "value": rows_per_s, "unit": "row-steps/s",
            "batch_steps_per_s": rows_per_s / R / world, "ms_per_step": 1e3 * t_dec / dec_steps, "steps": dec_steps,
            "launch": f"CUDA graph of {GSTEPS} steps (+1 reset) replayed {n_replays}x",
            "eager_ms_per_step": 1e3 * t_dec_eager / args.steps,
            "config": {"workload": "config 3: beam 5 x batch 32 = 160 rows, V 51865, SuppressBlank + SuppressTokens + "
                                   "TimestampTokens + log_softmax + top-6 + beam merge", "step_kind": "mixed histories",
                       "cache": f"logits rotate over {DBUF} buffers (266 MB > L2)"},
            "roofline": {"bound": "hbm", "achieved": dec_ach, "peak": hbm_peak, "unit": "GB/s", "frac": dec_ach / hbm_peak,
                         "traffic": DECODE_DRAM_TRAFFIC_PER_STEP, "traffic_source": TRAFFIC_SOURCE,
                         "frac_of_8TBps": dec_ach / 8000.0,
                         "algorithmic_bytes_per_launch": R * DECODE_BYTES_PER_ROW,
                         "kernel": "wrs::decode_scan_kernel (reads the logits once) + wrs::decode_tail_kernel<0> (row merge over cells, "
                                   "beam update); achieved = algorithmic bytes / time of the whole two-launch step"},
            "cpu_baseline": cpu_dec,
            "e2e": {"value": e2e_dec["device_logits"], "unit": "row-steps/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": R * (Lcur + 1) * 8 + R * 4 + 4,
                    "api": "wrs_beam_step on device logits + D2H of tokens, scores, completed each step",
                    "with_host_logits": {"value": e2e_dec["host_logits"], "h2d_bytes_per_step": R * VOCAB * 4}},
            "gpu_launches": 2 * dec_steps,
        },
        "no_speech_probs": {
            "metric": "step-0 no-speech probability", "value": world * R * dec_steps / t_ns, "unit": "row-steps/s",
            "ms_per_step": 1e3 * t_ns / dec_steps, "steps": dec_steps,
            "config": {"workload": f"softmax over the unfiltered row, P(<|nospeech|>): {R} rows x V {VOCAB} (src/decode/mod.rs:250-255)",
                       "launch": f"CUDA graph of {GSTEPS} calls replayed {n_replays}x"},
            "roofline": {"bound": "hbm", "achieved": ns_ach, "peak": hbm_peak, "unit": "GB/s", "frac": ns_ach / hbm_peak,
                         "traffic": None, "algorithmic_bytes_per_launch": R * DECODE_BYTES_PER_ROW,
                         "kernel": "wrs::no_speech_kernel"},
            "gpu_launches": dec_steps,
        },
        "kv_reorder": {
            "metric": "KV-cache beam reorder", "value": world * R * kv_steps / t_kv, "unit": "row-steps/s",
            "ms_per_step": 1e3 * t_kv / kv_steps, "steps": kv_steps,
            "config": {"workload": f"Cached::update_cache on a base-shaped decoder: {KV_LAYERS} blocks x (self K, V "
                                   f"[{R}, {KV_SELF}, {KV_STATE}] + cross K, V [{R}, {KV_CROSS}, {KV_STATE}]) f32, 24 tensors, "
                                   f"one launch, beam-shaped source indices",
                       "cache": "12 GB of caches per step, ping-pong between two sets (>> L2)"},
            "roofline": {"bound": "hbm", "achieved": kv_ach, "peak": hbm_peak, "unit": "GB/s", "frac": kv_ach / hbm_peak,
                         "traffic": None, "algorithmic_bytes_per_launch": kv_bytes, "kernel": "wrs::kv_reorder_kernel"},
            "cpu_baseline": cpu_kv, "parity_spot_check": kv_ok, "gpu_launches": kv_steps,
        },
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

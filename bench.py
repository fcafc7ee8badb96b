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
WORKLOAD = "config 2: log-mel front-end, 64 x 30 s synthetic 16 kHz segments per GPU per step"     # both arms, verbatim
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r2_traffic.json")


def traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, as tools/prof_all.py extracted it from the
    committed `ncu --set full` captures of this same command (profiles/r2_traffic.json); None when there is no capture."""
    try:
        rec = json.load(open(TRAFFIC_FILE))
        return float(rec["kernels"][kernel]["dram_bytes"]), rec.get("source", "profiles/r2_traffic.json")
    except Exception:
        return None, "no capture (profiles/r2_traffic.json missing or without this kernel)"


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
        "config": {"workload": WORKLOAD, "bounded_sample": f"{per_step} of the 64 segments per step"},
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


class Env:
    """What every section needs: rank plumbing, the library, the context, the roofline denominator."""

    def __init__(self, args):
        import torch.distributed as dist
        from whisper_rs_b200 import _lib
        from whisper_rs_b200 import build as wbuild
        self.args, self.dist = args, dist
        self.rank, self.world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
        self.local = int(os.environ.get("LOCAL_RANK", 0))
        if self.rank == 0:
            wbuild.build()
        pin_to_gpu_node(self.local)                      # before the first pinned allocation
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.barrier()
        self.lib = _lib.lib()
        self._lib = _lib
        self.dev = torch.device("cuda", self.local)
        self.ctx = _lib.context(self.local)
        self.stream = torch.cuda.current_stream(self.dev).cuda_stream
        self.hbm_peak, self.peak_src = peaks()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, x):
        if self.world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def check(self, rc):
        self._lib.check(rc, self.ctx.handle)

    def roofline(self, bytes_per_launch, seconds_per_launch, kernel, kernel_key=None):
        ach = bytes_per_launch / seconds_per_launch / 1e9
        tr, src = traffic(kernel_key) if kernel_key else (None, None)
        r = {"bound": "hbm", "achieved": ach, "peak": self.hbm_peak, "unit": "GB/s", "frac": ach / self.hbm_peak,
             "traffic": tr, "peak_source": self.peak_src, "frac_of_8TBps": ach / 8000.0,
             "algorithmic_bytes_per_launch": bytes_per_launch, "kernel": kernel}
        if kernel_key:
            r["traffic_source"] = src
        return r


def pin_to_gpu_node(local):
    """Bind this rank to the cores (and, through first touch, the memory) next to its GPU: nvidia-smi topo's CPU affinity,
    split evenly between the ranks that share it.  Pinned host buffers allocated afterwards land on that node."""
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", 1)))
        for line in out.splitlines():
            cols = line.split()
            if cols and cols[0] == f"GPU{local}":
                aff = next((c for c in cols[1:] if c[0].isdigit() and ("-" in c or "," in c) and not c.startswith("0x")), None)
                if not aff:
                    return None
                cores = []
                for part in aff.split(","):
                    lo, _, hi = part.partition("-")
                    cores += list(range(int(lo), int(hi or lo) + 1))
                allowed = sorted(set(cores) & os.sched_getaffinity(0))
                if not allowed:
                    return None
                share = max(1, len(allowed) // max(world, 1))
                mine = allowed[(local % max(world, 1)) * share:][:share] or allowed
                os.sched_setaffinity(0, mine)
                return {"cores": f"{mine[0]}-{mine[-1]}", "n": len(mine)}
    except Exception:
        return None
    return None


def bench_logmel(env):
    """config 2: 64 x 30 s segments per GPU per step -- the headline."""
    from whisper_rs_b200 import synth
    args, L, ctx, dev = env.args, env.lib, env.ctx, env.dev
    NBUF = 3   # rotate 3 in/out sets (553 MB > 126 MB L2): every step reads cold PCM
    bank = torch.from_numpy(synth.audio_batch(16, seed=1000 * env.rank))
    host_pcm = bank.repeat(N_SEG // 16, 1).contiguous().pin_memory()
    d_in = [host_pcm.to(dev, non_blocking=True).clone() for _ in range(NBUF)]
    d_out = [torch.empty(N_SEG, N_MELS, N_FRAMES, dtype=torch.float32, device=dev) for _ in range(NBUF)]
    host_out = torch.empty(N_SEG, N_MELS, N_FRAMES, dtype=torch.float32).pin_memory()

    def logmel_step(i):
        b = i % NBUF
        env.check(L.wrs_logmel_batch(ctx.handle, d_in[b].data_ptr(), N_SEG, N_SAMPLES, N_SAMPLES, d_out[b].data_ptr(), env.stream))

    for i in range(args.warmup):
        logmel_step(i)
    t_logmel = env.max_over_ranks(timed(logmel_step, args.steps, env.barrier))
    ctx.check(env.stream)
    logmel_mode = ctx.get_stat("logmel_mode")

    # ---- end to end: host PCM in, host log-mel out, through the call a CPU-resident caller makes
    def logmel_e2e(i):
        env.check(L.wrs_logmel_batch_host(ctx.handle, host_pcm.data_ptr(), N_SEG, N_SAMPLES, N_SAMPLES, host_out.data_ptr()))

    def wall(fn, n):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        env.barrier()
        t0 = time.perf_counter()
        for i in range(n):
            fn(i)
        torch.cuda.synchronize()
        t = env.max_over_ranks(time.perf_counter() - t0)
        env.barrier()
        return t

    e2e_steps = max(3, min(args.steps, 30))
    t_e2e = wall(logmel_e2e, e2e_steps)
    # ---- the copy-only floor of that call: the same pinned bytes in the same chunks over the same two copy streams
    #      (wrs_logmel_batch_host: chunks of 4 segments, H2D and D2H concurrently), no kernel in between
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    CH = 4

    def copy_only(i):
        evs = []
        with torch.cuda.stream(s_in):
            for c in range(0, N_SEG, CH):
                d_in[0][c:c + CH].copy_(host_pcm[c:c + CH], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(s_in)
                evs.append(ev)
        with torch.cuda.stream(s_out):
            for j, c in enumerate(range(0, N_SEG, CH)):
                s_out.wait_event(evs[j])
                host_out[c:c + CH].copy_(d_out[0][c:c + CH], non_blocking=True)
        s_in.synchronize()
        s_out.synchronize()

    t_floor = wall(copy_only, e2e_steps)

    # ---- 16-bit PCM in (what the reference's decoders produce before the f32 conversion, src/audio.rs:116-152): half the H2D bytes
    host_i16 = (bank.repeat(N_SEG // 16, 1).clamp(-1, 1) * 32767.0).round().to(torch.int16).contiguous().pin_memory()
    e2e_i16 = None
    if hasattr(L, "wrs_logmel_batch_host_i16"):
        def logmel_e2e_i16(i):
            env.check(L.wrs_logmel_batch_host_i16(ctx.handle, host_i16.data_ptr(), N_SEG, N_SAMPLES, N_SAMPLES, host_out.data_ptr()))
        t_i16 = wall(logmel_e2e_i16, e2e_steps)
        e2e_i16 = {"value": env.world * N_SEG * e2e_steps / t_i16, "unit": "segments/s", "h2d_bytes_per_step": N_SEG * N_SAMPLES * 2,
                   "d2h_bytes_per_step": N_SEG * N_MELS * N_FRAMES * 4, "api": "wrs_logmel_batch_host_i16 (pinned 16-bit PCM in)"}
    ctx.check(env.stream)
    per_launch = t_logmel / args.steps
    return {
        "value": env.world * N_SEG * args.steps / t_logmel, "ms_per_step": 1e3 * per_launch, "mode": logmel_mode, "nbuf": NBUF,
        "roofline": env.roofline(N_SEG * LOGMEL_BYTES_PER_SEG, per_launch,
                                 "wrs::logmel_kernel<true> (persistent, warp-specialised, cooperative; one launch per batch)", "logmel_kernel"),
        "e2e": {"value": env.world * N_SEG * e2e_steps / t_e2e, "unit": "segments/s",
                "h2d_bytes_per_step": N_SEG * N_SAMPLES * 4, "d2h_bytes_per_step": N_SEG * N_MELS * N_FRAMES * 4,
                "api": "wrs_logmel_batch_host (pinned host PCM in, host log-mel out, synchronous)", "steps": e2e_steps,
                "copy_floor": {"value": env.world * N_SEG * e2e_steps / t_floor, "unit": "segments/s",
                               "what": "the same pinned H2D + D2H bytes in the same 4-segment chunks on two copy streams, no kernel"},
                "frac_of_copy_floor": t_floor / t_e2e, "pcm_i16": e2e_i16},
        "gpu_launches": args.steps * (1 if logmel_mode == 1 else 2),
    }


def bench_decode(env):
    """config 3: 32 audios x beam 5 = 160 rows, V = 51865 -- plus the step-0 no-speech probability, the trait-level
    in-place filters, and the same step at 25 000 rows (config 5's shape on one GPU: the scan's streaming rate)."""
    from whisper_rs_b200 import constants as C, synth
    from whisper_rs_b200 import decode as wd
    args, L, ctx, dev = env.args, env.lib, env.ctx, env.dev
    ids = C.MULTILINGUAL
    R = N_AUDIO * BEAM
    DBUF = 8   # rotate 8 logits sets (266 MB > L2)
    case = synth.decode_case(ids, N_AUDIO, BEAM, "mixed", seed=3 + env.rank)
    prompt = wd.DecodePrompt(None, ids)
    prompt.sample_begin = case["sample_begin"]
    step = wd.FusedDecodeStep(wd.DecodeOptions(), ids, [prompt], max_rows=R, device=env.local)
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

    def beam_call(logits, s):
        env.check(L.wrs_beam_step(dec.handle, logits.data_ptr(), logits.stride(0), R, d_tokens.data_ptr(), d_tokens.stride(0),
                                  Lcur, prompt.sample_begin, d_sums.data_ptr(), d_next.data_ptr(), d_next.stride(0),
                                  d_src.data_ptr(), d_sum_out.data_ptr(), d_done.data_ptr(), None, s))

    def decode_step(i):
        if i % 64 == 0:
            dec.reset()          # keep the finished store from saturating: same work every step
        beam_call(d_logits[i % DBUF], env.stream)

    for i in range(args.warmup):
        decode_step(i)
    eager_steps = max(200, min(args.steps, 2000))
    t_dec_eager = env.max_over_ranks(timed(decode_step, eager_steps, env.barrier))
    # A step is a few microseconds of device time, less than one Python/ctypes launch: in the reference loop the
    # launch hides behind the decoder forward (src/decode/mod.rs:236-256); here DBUF consecutive steps (one per
    # logits buffer, reset first) are captured once in a CUDA graph and the graph is replayed -- at least 50 times,
    # whatever --steps says (a sub-benchmark keeps its own minimum sample).
    GSTEPS = DBUF
    cap_stream = torch.cuda.Stream(dev)
    n_replays = max(50, args.steps // GSTEPS)
    dec_steps = n_replays * GSTEPS

    def graph_of(body):
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=cap_stream):
            cs = torch.cuda.current_stream(dev).cuda_stream
            body(cs)
        for _ in range(3):
            g.replay()
        return env.max_over_ranks(timed(lambda i: g.replay(), n_replays, env.barrier))

    def beam_body(cs):
        env.check(L.wrs_decoder_reset(dec.handle, cs))
        for b in range(GSTEPS):
            beam_call(d_logits[b], cs)
    t_dec = graph_of(beam_body)

    # step-0 no-speech probability (src/decode/mod.rs:250-255): one unfiltered softmax per row
    d_ns = torch.empty(R, dtype=torch.float32, device=dev)
    t_ns = graph_of(lambda cs: [env.check(L.wrs_no_speech_probs(dec.handle, d_logits[b].data_ptr(), d_logits[b].stride(0), R,
                                                                d_ns.data_ptr(), cs)) for b in range(GSTEPS)])
    # the trait-level path: LogitFilter::apply x 3 in place (src/decode/filter/*.rs), what a caller that keeps the
    # reference's own extractor would run (the logits are rewritten with the same masks every replay: same work)
    d_flags = torch.empty(R, dtype=torch.uint8, device=dev)
    t_flt = graph_of(lambda cs: [env.check(L.wrs_filter_apply(dec.handle, env._lib.FILTER_ALL, d_logits[b].data_ptr(), d_logits[b].stride(0), R,
                                                              d_tokens.data_ptr(), d_tokens.stride(0), Lcur, prompt.sample_begin,
                                                              d_flags.data_ptr(), cs)) for b in range(GSTEPS)])
    d_logits = [host_logits.to(dev).roll(7 * b, dims=0).contiguous() for b in range(DBUF)]       # (undo the in-place masks)

    # end to end as the reference's caller sees it (decode/mod.rs:257-276): logits are already on the device
    # (north_star: raw device pointers from the tch forward), the host reads back tokens + scores + completed.
    def decode_e2e(i, h2d_logits):
        lg = d_logits[i % DBUF]
        if h2d_logits:
            lg.copy_(host_logits, non_blocking=True)
        beam_call(lg, env.stream)
        h_outs.copy_(d_outs, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return bool(h_done[0])          # the `completed` read is the per-step sync in the reference loop too

    e2e_dec = {}
    for name, h2d in (("device_logits", False), ("host_logits", True)):
        for i in range(3):
            decode_e2e(i, h2d)
        n = max(50, min(args.steps, 200))
        torch.cuda.synchronize()
        env.barrier()
        t0 = time.perf_counter()
        for i in range(n):
            decode_e2e(i, h2d)
        torch.cuda.synchronize()
        e2e_dec[name] = env.world * R * n / env.max_over_ranks(time.perf_counter() - t0)
    ctx.check(env.stream)
    # one debug step of the benchmark case, kept for the parity figures of the CPU leg
    dbg = step.step(d_tokens, d_logits[0], d_sums, debug=True)
    torch.cuda.synchronize()
    got = {k: v.cpu().numpy() for k, v in dbg.items() if isinstance(v, torch.Tensor)}
    del d_logits, step

    # ---- the same fused step at config 5's row count on one GPU: 5 000 audios x beam 5 = 25 000 rows, 5.2 GB of logits
    RL = 25000
    g = torch.Generator(device=dev).manual_seed(11 + env.rank)
    big = torch.empty((RL, VOCAB), dtype=torch.float32, device=dev).normal_(0.0, 3.0, generator=g)
    ltok = torch.from_numpy(np.tile(case["tokens"], (RL // R + 1, 1))[:RL]).to(dev)
    lsum = torch.zeros(RL, dtype=torch.float32, device=dev)
    lnext = torch.empty((RL, Lcur + 1), dtype=torch.int64, device=dev)
    lsrc = torch.empty(RL, dtype=torch.int32, device=dev)
    lout = torch.empty(RL, dtype=torch.float32, device=dev)
    ldone = torch.empty(1, dtype=torch.int32, device=dev)
    lstep = wd.FusedDecodeStep(wd.DecodeOptions(), ids, [prompt], max_rows=RL, device=env.local)
    lstep.reset()

    def large_step(i):
        if i % 4 == 0:
            lstep.decoder.reset()
        env.check(L.wrs_beam_step(lstep.decoder.handle, big.data_ptr(), big.stride(0), RL, ltok.data_ptr(), ltok.stride(0), Lcur,
                                  prompt.sample_begin, lsum.data_ptr(), lnext.data_ptr(), lnext.stride(0), lsrc.data_ptr(),
                                  lout.data_ptr(), ldone.data_ptr(), None, env.stream))
    for i in range(3):
        large_step(i)
    large_steps = max(20, min(args.steps, 100))
    t_large = env.max_over_ranks(timed(large_step, large_steps, env.barrier))
    ctx.check(env.stream)
    del big, lstep

    per_step, bytes_step = t_dec / dec_steps, R * DECODE_BYTES_PER_ROW
    roof = env.roofline(bytes_step, per_step, "wrs::decode_scan_kernel (reads the logits once) + wrs::decode_tail_kernel<0> (row merge over "
                        "cells, beam update); achieved = algorithmic bytes / time of the whole two-launch step", "decode_scan_kernel")
    tail_tr, _ = traffic("decode_tail_kernel")
    if roof["traffic"] is not None and tail_tr is not None:
        roof["traffic"] += tail_tr
    return {"case": case, "ids": ids, "got": got, "R": R, "Lcur": Lcur, "line": {
        "decode_step": {
            "metric": "beam decode row-steps/s", "value": env.world * R * dec_steps / t_dec, "unit": "row-steps/s",
            "batch_steps_per_s": dec_steps / t_dec, "ms_per_step": 1e3 * per_step, "steps": dec_steps,
            "launch": f"CUDA graph of {GSTEPS} steps (+1 reset) replayed {n_replays}x",
            "eager_ms_per_step": 1e3 * t_dec_eager / eager_steps,
            "config": {"workload": "config 3: beam 5 x batch 32 = 160 rows, V 51865, SuppressBlank + SuppressTokens + "
                                   "TimestampTokens + log_softmax + top-6 + beam merge", "step_kind": "mixed histories",
                       "cache": f"logits rotate over {DBUF} buffers (266 MB > L2)"},
            "roofline": roof,
            "e2e": {"value": e2e_dec["device_logits"], "unit": "row-steps/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": R * (Lcur + 1) * 8 + R * 4 + 4,
                    "api": "wrs_beam_step on device logits + D2H of tokens, scores, completed each step",
                    "with_host_logits": {"value": e2e_dec["host_logits"], "h2d_bytes_per_step": R * VOCAB * 4}},
            "gpu_launches": 2 * dec_steps,
        },
        "decode_step_large": {
            "metric": "beam decode row-steps/s", "value": env.world * RL * large_steps / t_large, "unit": "row-steps/s",
            "ms_per_step": 1e3 * t_large / large_steps, "steps": large_steps,
            "config": {"workload": f"config 5's decode shape on one GPU: {RL} rows (5 000 audios x beam 5), V {VOCAB}, the same fused step",
                       "cache": "5.2 GB of logits per step (>> L2)", "launch": "eager (a step is ~1 ms)"},
            "roofline": env.roofline(RL * DECODE_BYTES_PER_ROW, t_large / large_steps, "wrs::decode_scan_kernel + wrs::decode_tail_kernel<0>"),
            "gpu_launches": 2 * large_steps,
        },
        "no_speech_probs": {
            "metric": "step-0 no-speech probability", "value": env.world * R * dec_steps / t_ns, "unit": "row-steps/s",
            "ms_per_step": 1e3 * t_ns / dec_steps, "steps": dec_steps,
            "config": {"workload": f"softmax over the unfiltered row, P(<|nospeech|>): {R} rows x V {VOCAB} (src/decode/mod.rs:250-255)",
                       "launch": f"CUDA graph of {GSTEPS} calls replayed {n_replays}x"},
            "roofline": env.roofline(bytes_step, t_ns / dec_steps, "wrs::no_speech_kernel", "no_speech_kernel"),
            "gpu_launches": dec_steps,
        },
        "filter_apply": {
            "metric": "in-place LogitFilter::apply x 3", "value": env.world * R * dec_steps / t_flt, "unit": "row-steps/s",
            "ms_per_step": 1e3 * t_flt / dec_steps, "steps": dec_steps,
            "config": {"workload": f"SuppressBlank + SuppressTokens + TimestampTokens in place on {R} rows x V {VOCAB} (wrs_filter_apply, "
                                   "the trait-level path)", "launch": f"CUDA graph of {GSTEPS} calls replayed {n_replays}x"},
            "roofline": env.roofline(bytes_step, t_flt / dec_steps, "filter_index_fill_kernel x 2 + decode_scan_kernel + decode_tail_kernel<2>"),
            "gpu_launches": 4 * dec_steps,
        }}}


def bench_kv(env):
    """KV-cache beam reorder: Cached::update_cache (src/model.rs:53-62) on every cached K/V of a 6-block,
    512-wide decoder (examples/transcribe.rs dims): self-attention caches [160, 32, 512] and cross-attention caches
    [160, 1500, 512], f32, 24 tensors, one launch -- as the reference does it (everything gathered), and with the
    cross-attention caches flagged group-invariant (their gather is a bit-exact no-op: skipped)."""
    from whisper_rs_b200 import cache as wcache
    args, dev, R = env.args, env.dev, N_AUDIO * BEAM
    KV_LAYERS, KV_STATE, KV_SELF, KV_CROSS = 6, 512, 32, 1500
    kv_src, kv_dst, flags = [], [], []
    for _ in range(KV_LAYERS):
        for ctx_len, cross in ((KV_SELF, False), (KV_SELF, False), (KV_CROSS, True), (KV_CROSS, True)):     # self K, V, cross K, V
            if cross:    # what the reference holds: projections of audio_features.repeat_interleave(k) (src/decode/mod.rs:311-313)
                t = torch.empty((N_AUDIO, ctx_len, KV_STATE), dtype=torch.float32, device=dev).normal_().repeat_interleave(BEAM, 0).contiguous()
            else:
                t = torch.empty((R, ctx_len, KV_STATE), dtype=torch.float32, device=dev).normal_()
            kv_src.append(t)
            kv_dst.append(torch.empty_like(t))
            flags.append(cross)
    kv_bytes = 2 * sum(t.numel() * 4 for t in kv_src)                # every byte read once and written once
    moving_bytes = 2 * sum(t.numel() * 4 for t, f in zip(kv_src, flags) if not f)
    rng = np.random.default_rng(7 + env.rank)
    kv_idx = torch.from_numpy(np.concatenate([a * BEAM + rng.integers(0, BEAM, size=BEAM) for a in range(N_AUDIO)])
                              .astype(np.int32)).to(dev)

    def kv_step(i, fl=None):
        a, b = (kv_src, kv_dst) if i % 2 == 0 else (kv_dst, kv_src)
        wcache.update_cache(a, kv_idx, b, fl)

    kv_steps = max(4, min(args.steps, 40))
    for i in range(4):
        kv_step(i)
    t_kv = env.max_over_ranks(timed(kv_step, kv_steps, env.barrier))
    # flagged: the self-attention caches ping-pong, the cross-attention caches stay where they are
    self_src = [t for t, f in zip(kv_src, flags) if not f]
    self_dst = [t for t, f in zip(kv_dst, flags) if not f]
    cross = [t for t, f in zip(kv_src, flags) if f]
    fl = [False] * len(self_src) + [True] * len(cross)

    def kv_step_flagged(i):
        a, b = (self_src, self_dst) if i % 2 == 0 else (self_dst, self_src)
        wcache.update_cache(a + cross, kv_idx, b + cross, fl)

    fsteps = max(50, min(args.steps, 400))
    for i in range(4):
        kv_step_flagged(i)
    t_kvf = env.max_over_ranks(timed(kv_step_flagged, fsteps, env.barrier))
    ok = bool(torch.equal(kv_src[2][kv_idx.long()], wcache.update_cache(kv_src[2:3], kv_idx)[0])) and \
        bool(torch.equal(kv_src[2][kv_idx.long()], kv_src[2]))        # the flagged tensors' gather is the identity
    env.ctx.check(env.stream)
    return {"kv_idx": kv_idx.cpu().long(), "dims": (KV_LAYERS, KV_STATE, KV_SELF, KV_CROSS), "line": {
        "metric": "KV-cache beam reorder", "value": env.world * R * kv_steps / t_kv, "unit": "row-steps/s",
        "ms_per_step": 1e3 * t_kv / kv_steps, "steps": kv_steps,
        "config": {"workload": f"Cached::update_cache on a base-shaped decoder: {KV_LAYERS} blocks x (self K, V "
                               f"[{R}, {KV_SELF}, {KV_STATE}] + cross K, V [{R}, {KV_CROSS}, {KV_STATE}]) f32, 24 tensors, "
                               f"one launch, beam-shaped source indices",
                   "cache": "12 GB of caches per step, ping-pong between two sets (>> L2)"},
        "roofline": env.roofline(kv_bytes, t_kv / kv_steps, "wrs::kv_reorder_kernel", "kv_reorder_kernel"),
        "cross_flagged_group_invariant": {
            "ms_per_step": 1e3 * t_kvf / fsteps, "steps": fsteps, "value": env.world * R * fsteps / t_kvf, "unit": "row-steps/s",
            "what": "the 12 cross-attention caches flagged wrs_kv_tensor.group_invariant (identical rows within a beam group: "
                    "their gather is a bit-exact no-op) -- only the 12 self-attention caches move",
            "roofline": env.roofline(moving_bytes, t_kvf / fsteps, "wrs::kv_reorder_kernel (moving tensors only)")},
        "parity_spot_check": ok, "gpu_launches": kv_steps + fsteps}}


def bench_configs(env):
    """BASELINE configs 4 and 5.  config 4: 1 h = 120 x 30 s segments on every rank (front-end per segment, the whole clip
    as one segment -- the two-launch form -- and timestamp-mode beam steps for its 600 rows).  config 5: 10 000 segments /
    50 000 rows split over the ranks by shard.segment_range (strong scaling), front-end + beam=5 decode-step filtering."""
    from whisper_rs_b200 import constants as C, shard, synth
    from whisper_rs_b200 import decode as wd
    L, ctx, dev = env.lib, env.ctx, env.dev
    ids = C.MULTILINGUAL
    bank = torch.from_numpy(synth.audio_batch(16, seed=5)).to(dev)
    out = {}

    def logmel(pcm, n_seg, n_samples, dst):
        env.check(L.wrs_logmel_batch(ctx.handle, pcm.data_ptr(), n_seg, n_samples, n_samples, dst.data_ptr(), env.stream))

    def beam_setup(rows, seed):
        case = synth.decode_case(ids, 32, BEAM, "mixed", seed=seed)
        prompt = wd.DecodePrompt(None, ids)
        prompt.sample_begin = case["sample_begin"]
        st = wd.FusedDecodeStep(wd.DecodeOptions(), ids, [prompt], max_rows=rows, device=env.local)
        st.reset()
        g = torch.Generator(device=dev).manual_seed(seed)
        lg = torch.empty((rows, VOCAB), dtype=torch.float32, device=dev).normal_(0.0, 3.0, generator=g)
        tok = torch.from_numpy(np.tile(case["tokens"], (rows // 160 + 1, 1))[:rows]).to(dev)
        Lc = tok.shape[1]
        bufs = (torch.zeros(rows, dtype=torch.float32, device=dev), torch.empty((rows, Lc + 1), dtype=torch.int64, device=dev),
                torch.empty(rows, dtype=torch.int32, device=dev), torch.empty(rows, dtype=torch.float32, device=dev),
                torch.empty(1, dtype=torch.int32, device=dev))

        def call(i):
            if i % 4 == 0:
                st.decoder.reset()
            env.check(L.wrs_beam_step(st.decoder.handle, lg.data_ptr(), lg.stride(0), rows, tok.data_ptr(), tok.stride(0), Lc,
                                      prompt.sample_begin, bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[1].stride(0),
                                      bufs[2].data_ptr(), bufs[3].data_ptr(), bufs[4].data_ptr(), None, env.stream))
        return call, (st, lg, tok, bufs)

    # ---- config 4
    n4 = 120
    pcm4 = bank.repeat(n4 // 16 + 1, 1)[:n4].contiguous()
    mel4 = torch.empty(n4, N_MELS, N_FRAMES, dtype=torch.float32, device=dev)
    for _ in range(3):
        logmel(pcm4, n4, N_SAMPLES, mel4)
    reps = 20
    t_seg = env.max_over_ranks(timed(lambda i: logmel(pcm4, n4, N_SAMPLES, mel4), reps, env.barrier)) / reps
    clip = pcm4.view(1, n4 * N_SAMPLES)                              # the hour as ONE clip (src/transcribe.rs:112): global max
    for _ in range(2):
        logmel(clip, 1, n4 * N_SAMPLES, mel4)
    mode_clip = ctx.get_stat("logmel_mode")
    t_clip = env.max_over_ranks(timed(lambda i: logmel(clip, 1, n4 * N_SAMPLES, mel4), reps, env.barrier)) / reps
    call4, keep4 = beam_setup(n4 * BEAM, 41)
    for i in range(3):
        call4(i)
    t_b4 = env.max_over_ranks(timed(call4, 40, env.barrier)) / 40
    ctx.check(env.stream)
    out["config4"] = {
        "workload": "config 4: 1 h synthetic audio = 120 x 30 s segments per GPU: front-end + timestamp-mode beam=5 decode steps (600 rows, V 51865)",
        "logmel_per_segment": {"ms": 1e3 * t_seg, "segments_per_s": env.world * n4 / t_seg,
                               "roofline_frac": n4 * LOGMEL_BYTES_PER_SEG / t_seg / 1e9 / env.hbm_peak, "mode": "fused single launch"},
        "logmel_whole_clip": {"ms": 1e3 * t_clip, "segments_per_s": env.world * n4 / t_clip,
                              "roofline_frac": n4 * LOGMEL_BYTES_PER_SEG / t_clip / 1e9 / env.hbm_peak,
                              "mode": "two launches (global maximum over the clip)" if mode_clip == 0 else "fused"},
        "decode_step": {"rows": n4 * BEAM, "ms_per_step": 1e3 * t_b4, "row_steps_per_s": env.world * n4 * BEAM / t_b4,
                        "roofline_frac": n4 * BEAM * DECODE_BYTES_PER_ROW / t_b4 / 1e9 / env.hbm_peak, "launch": "eager"},
        "scaling": "weak (every rank has its own hour)"}
    del keep4, call4, pcm4, mel4, clip

    # ---- config 5: this rank's share of the 10 000 segments (contiguous block), then of the 50 000 rows
    n5 = 10000
    lo, hi = shard.segment_range(n5, env.rank, env.world)
    mine = hi - lo
    pcm5 = torch.empty((mine, N_SAMPLES), dtype=torch.float32, device=dev)
    for s0 in range(0, mine, 16):
        n = min(16, mine - s0)
        pcm5[s0:s0 + n] = bank[:n] * (0.5 + 0.5 * ((lo + s0) % 7) / 7.0)      # distinct levels per tile of the bank
    mel5 = torch.empty((mine, N_MELS, N_FRAMES), dtype=torch.float32, device=dev)
    logmel(pcm5, mine, N_SAMPLES, mel5)
    t_l5 = env.max_over_ranks(timed(lambda i: logmel(pcm5, mine, N_SAMPLES, mel5), 3, env.barrier)) / 3
    del pcm5, mel5
    rows5 = mine * BEAM
    chunk = 25000                                                        # a decoder holds fewer than 32 768 rows
    calls, keep = [], []
    for r0 in range(0, rows5, chunk):
        c, k = beam_setup(min(chunk, rows5 - r0), 51 + r0)
        calls.append(c)
        keep.append(k)

    def step5(i):
        for c in calls:
            c(i)
    for i in range(2):
        step5(i)
    t_d5 = env.max_over_ranks(timed(step5, 10, env.barrier)) / 10
    ctx.check(env.stream)
    out["config5"] = {
        "workload": "config 5: 10 000 synthetic 30 s segments, front-end + beam=5 decode-step filtering (50 000 rows, V 51865), "
                    "segments split over the ranks in contiguous blocks (shard.segment_range), no collective",
        "segments_this_rank": mine, "logmel_ms": 1e3 * t_l5, "segments_per_s": n5 / t_l5,
        "logmel_roofline_frac_per_gpu": mine * LOGMEL_BYTES_PER_SEG / t_l5 / 1e9 / env.hbm_peak,
        "rows_this_rank": rows5, "decode_ms_per_step": 1e3 * t_d5, "row_steps_per_s": n5 * BEAM / t_d5,
        "decode_roofline_frac_per_gpu": rows5 * DECODE_BYTES_PER_ROW / t_d5 / 1e9 / env.hbm_peak,
        "scaling": "strong (total work fixed: 10 000 segments / 50 000 rows over all ranks; times are the max over ranks)"}
    return out


def cpu_legs(env, dec, kv):
    """The oracle on this box's host cores (rank 0, N = 1 only): the reported CPU baseline of every sub-benchmark, and --
    the oracle being here as the checker -- the measured parity of the benchmark's own decode step against it."""
    threads = os.cpu_count() or 1
    t1 = cpu_logmel_time(4, 1) / 4
    n_all = 64
    t_all = cpu_logmel_time(n_all, threads) / n_all
    cpu = {"value": 1.0 / t_all, "unit": "segments/s", "cores": threads, "kind": "port",
           "sample": f"{n_all} of the 64 x 30 s segments, one call per segment, torch {torch.__version__} CPU ATen op "
                     f"sequence of src/audio.rs:230-259 (reference Rust crate cannot be built here)",
           "single_thread_value": 1.0 / t1}
    from oracle import decode as od
    case, ids, R = dec["case"], dec["ids"], dec["R"]
    nd = 5
    td = cpu_decode_time({k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in case.items()}, ids, nd, threads)
    cpu_dec = {"value": R * nd / td, "unit": "row-steps/s", "cores": threads, "kind": "port",
               "sample": f"{nd} steps of 160 rows x V {VOCAB}: 3 filters + BeamSearchTokenExtractor::update restatement"}
    want, got = od.decode_step_reference(case, ids), dec["got"]
    lp64 = torch.from_numpy(want["masked_logits"]).double().log_softmax(-1).numpy()
    want64 = np.take_along_axis(lp64, want["cand_token"].astype(np.int64), 1)

    def rel(a, b):
        a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
        ok = np.isfinite(a) & np.isfinite(b) & (np.abs(b) >= 0.1)
        return float((np.abs(a - b)[ok] / np.abs(b)[ok]).max()) if ok.any() else 0.0
    parity = {"ids_bit_exact": bool(np.array_equal(got["cand_token"], want["cand_token"]) and np.array_equal(got["next_tokens"], want["next_tokens"])
                                    and np.array_equal(got["source_idx"], want["source_idx"])),
              "row_flags_bit_exact": bool(np.array_equal(got["row_flags"], want["row_flags"])),
              "cand_logprob_max_rel_err_vs_oracle": rel(got["cand_logprob"], want["cand_logprob"]),
              "cand_logprob_max_rel_err_vs_f64": rel(got["cand_logprob"], want64),
              "oracle_max_rel_err_vs_f64": rel(want["cand_logprob"], want64),
              "sum_logprobs_max_rel_err_vs_oracle": rel(got["sum_logprobs"], want["sum_logprobs"]),
              "note": "relative errors over |log-prob| >= 0.1 (north_star: 1e-5); the oracle is ATen's f32 log_softmax"}
    from oracle import cache as ocache
    torch.set_num_threads(threads)
    KV_LAYERS, KV_STATE, KV_SELF, KV_CROSS = kv["dims"]
    blk = [torch.randn((R, n, KV_STATE)) for n in (KV_SELF, KV_SELF, KV_CROSS, KV_CROSS)]
    ocache.update_cache(blk, kv["kv_idx"])
    reps = 3
    t0 = time.perf_counter()
    for _ in range(reps):
        ocache.update_cache(blk, kv["kv_idx"])
    tk = (time.perf_counter() - t0) / reps * KV_LAYERS           # one block timed, scaled to the decoder's six
    cpu_kv = {"value": R / tk, "unit": "row-steps/s", "cores": threads, "kind": "port",
              "sample": f"one of the {KV_LAYERS} decoder blocks (4 tensors, 1.0 GB read + 1.0 GB written), x{KV_LAYERS}: "
                        "torch CPU Tensor.index, the call src/model.rs:59 makes"}
    return cpu, cpu_dec, parity, cpu_kv


def bench_project(env):
    """`x.matmul(token_emb.T)` (reference src/model.rs:521) + the decode step at config 3's shape
    (160 rows, n_state 512 = Whisper base, V 51865), three ways: the fused tcgen05 kernel (logits never leave the SM),
    our projection kernel writing logits + wrs_beam_step, and the library f32 GEMM (what tch runs) + wrs_beam_step."""
    from whisper_rs_b200 import constants as C, synth
    from whisper_rs_b200 import decode as wd
    from whisper_rs_b200.decode.core import Projection
    args, L, ctx, dev = env.args, env.lib, env.ctx, env.dev
    ids = C.MULTILINGUAL
    R, D, NB = N_AUDIO * BEAM, 512, 8
    case = synth.decode_case(ids, N_AUDIO, BEAM, "mixed", seed=3 + env.rank)
    prompt = wd.DecodePrompt(None, ids)
    prompt.sample_begin = case["sample_begin"]
    step = wd.FusedDecodeStep(wd.DecodeOptions(), ids, [prompt], max_rows=R, device=env.local)
    step.reset()
    dec = step.decoder
    g = torch.Generator(device=dev).manual_seed(5 + env.rank)
    emb = torch.empty((VOCAB, D), dtype=torch.float32, device=dev).normal_(0.0, 3.0 / D ** 0.5, generator=g)
    # two packed copies of the embedding, alternated step by step: 2 x 107 MB > L2, so no step finds its weights cached
    # (in the decode loop a whole decoder forward runs between two projections)
    projs = [Projection(dec, emb), Projection(dec, emb)]
    embs = [emb, emb.clone()]
    xs = [torch.empty((R, D), dtype=torch.float32, device=dev).normal_(generator=g) for _ in range(NB)]
    d_tokens = torch.from_numpy(case["tokens"]).to(dev)
    d_sums = torch.from_numpy(case["sum_logprobs"]).to(dev)
    Lcur = d_tokens.shape[1]
    d_next = torch.empty((R, Lcur + 1), dtype=torch.int64, device=dev)
    d_src = torch.empty(R, dtype=torch.int32, device=dev)
    d_out = torch.empty(R, dtype=torch.float32, device=dev)
    d_done = torch.empty(1, dtype=torch.int32, device=dev)
    logits = [torch.empty((R, VOCAB), dtype=torch.float32, device=dev) for _ in range(NB)]
    n_replays = max(50, args.steps // NB)
    n_steps = n_replays * NB
    cap_stream = torch.cuda.Stream(dev)

    def graph_of(body):
        gr = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        with torch.cuda.graph(gr, stream=cap_stream):
            body(torch.cuda.current_stream(dev).cuda_stream)
        for _ in range(3):
            gr.replay()
        return env.max_over_ranks(timed(lambda i: gr.replay(), n_replays, env.barrier)) / n_steps

    def beam(lg, cs):
        env.check(L.wrs_beam_step(dec.handle, lg.data_ptr(), lg.stride(0), R, d_tokens.data_ptr(), d_tokens.stride(0), Lcur,
                                  prompt.sample_begin, d_sums.data_ptr(), d_next.data_ptr(), d_next.stride(0), d_src.data_ptr(),
                                  d_out.data_ptr(), d_done.data_ptr(), None, cs))

    def fused(cs):
        env.check(L.wrs_decoder_reset(dec.handle, cs))
        for b in range(NB):
            env.check(L.wrs_project_beam_step(projs[b & 1].handle, xs[b].data_ptr(), xs[b].stride(0), R, d_tokens.data_ptr(),
                                              d_tokens.stride(0), Lcur, prompt.sample_begin, d_sums.data_ptr(), d_next.data_ptr(),
                                              d_next.stride(0), d_src.data_ptr(), d_out.data_ptr(), d_done.data_ptr(), None, cs))

    def unfused(cs):
        env.check(L.wrs_decoder_reset(dec.handle, cs))
        for b in range(NB):
            env.check(L.wrs_project_logits(projs[b & 1].handle, xs[b].data_ptr(), xs[b].stride(0), R, logits[b].data_ptr(),
                                           logits[b].stride(0), cs))
            beam(logits[b], cs)

    def proj_only(cs):
        for b in range(NB):
            env.check(L.wrs_project_logits(projs[b & 1].handle, xs[b].data_ptr(), xs[b].stride(0), R, logits[b].data_ptr(),
                                           logits[b].stride(0), cs))

    def library(cs):
        env.check(L.wrs_decoder_reset(dec.handle, cs))
        for b in range(NB):
            torch.matmul(xs[b], embs[b & 1].t(), out=logits[b])
            beam(logits[b], cs)

    def library_gemm(cs):
        for b in range(NB):
            torch.matmul(xs[b], embs[b & 1].t(), out=logits[b])

    t_fused = graph_of(fused)
    env.check(L.wrs_ctx_set_option(ctx.handle, b"proj_cluster", 0))
    t_fused_single = graph_of(fused)          # A/B: every CTA on its own (two M tiles per vocab tile in one CTA)
    env.check(L.wrs_ctx_set_option(ctx.handle, b"proj_cluster", 1))
    t_unfused = graph_of(unfused)
    t_proj = graph_of(proj_only)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.matmul(xs[0], embs[0].t(), out=logits[0])          # (cuBLAS workspace / heuristics outside the capture)
    t_lib, t_lib_gemm = graph_of(library), graph_of(library_gemm)
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.matmul(xs[0], embs[0].t(), out=logits[0])
    t_lib_tf32 = graph_of(library_gemm)
    torch.backends.cuda.matmul.allow_tf32 = prev
    ctx.check(env.stream)
    mp = os.path.join(ROOT, "MEASURED_PEAKS.json")
    tc_peak = float(json.load(open(mp)).get("bf16_tflops", 0.0)) if os.path.exists(mp) else 0.0
    emb_bytes = VOCAB * D * 4 + R * D * 4                    # the f32 embedding and x, read once
    flops = 3 * 2.0 * R * D * VOCAB                          # three f16 products per f32 product
    roof = env.roofline(emb_bytes, t_fused, "wrs::proj_kernel<1> (tcgen05 projection + row scan; reads the packed f16 hi/lo "
                        "embedding, 4 B per weight) + wrs::proj_pack_kernel (x) + wrs::decode_tail_kernel<0>; achieved = bytes of the f32 "
                        "embedding and x / time of the whole three-launch step", "proj_kernel")
    roof["tensor"] = {"achieved": flops / t_fused / 1e12, "peak": tc_peak or None, "unit": "TFLOP/s",
                      "frac": flops / t_fused / 1e12 / tc_peak if tc_peak else None,
                      "note": "3 f16 products per f32 product on the 160 real rows; the second M tile of 128 is 3/4 padding"}
    return {"project_beam_step": {
        "metric": "projection + beam decode row-steps/s", "value": env.world * R / t_fused, "unit": "row-steps/s",
        "ms_per_step": 1e3 * t_fused, "steps": n_steps, "launch": f"CUDA graph of {NB} steps (+1 reset) replayed {n_replays}x",
        "config": {"workload": f"x [160, {D}] f32 @ token_emb [{VOCAB}, {D}] f32 (src/model.rs:521, Whisper base) + config 3's decode step; "
                               "logits never written", "arithmetic": "f16 hi/lo x 3 products, f32 accumulate in TMEM (about 4e-5 abs on "
                               "logits of sigma 3; an f32 matmul is near 9e-6)",
                   "cache": "two packed embedding copies alternate (2 x 107 MB > L2); x rotates over 8 buffers"},
        "roofline": roof,
        "alternatives_ms_per_step": {
            "wrs_project_beam_step without two-CTA clusters (option proj_cluster = 0)": 1e3 * t_fused_single,
            "wrs_project_logits + wrs_beam_step": 1e3 * t_unfused, "wrs_project_logits alone": 1e3 * t_proj,
            "torch.matmul f32 (cuBLAS, TF32 off: what tch runs) + wrs_beam_step": 1e3 * t_lib,
            "torch.matmul f32 alone": 1e3 * t_lib_gemm, "torch.matmul TF32 alone": 1e3 * t_lib_tf32},
        "gpu_launches": 3 * n_steps}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--only", default="", help="comma list of sections to run besides the headline: decode,kv,configs,project (default: all)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args, int(os.environ.get("RANK", 0)))
    args.warmup = max(args.warmup, 3)
    want = set(args.only.split(",")) if args.only else {"decode", "kv", "configs", "project"}
    env = Env(args)
    sampler = ClockSampler(env.local) if env.rank == 0 else None
    t_gpu0 = time.time()
    lm = bench_logmel(env)
    dec = bench_decode(env) if "decode" in want else None
    kv = bench_kv(env) if "kv" in want else None
    cfg = bench_configs(env) if "configs" in want else {}
    if "project" in want:
        cfg.update(bench_project(env))
    t_gpu1 = time.time()
    clocks = sampler.stop(t_gpu0, t_gpu1) if sampler else None
    if env.rank != 0:
        if env.world > 1:
            env.dist.destroy_process_group()
        return
    cpu = cpu_dec = parity = cpu_kv = None
    if env.world == 1 and not args.no_cpu_baseline and dec and kv:
        cpu, cpu_dec, parity, cpu_kv = cpu_legs(env, dec, kv)
    line = {
        "metric": "log-mels/s (30 s segments)", "value": lm["value"], "unit": "segments/s", "n_gpus": env.world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": lm["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "segments_per_step_per_gpu": N_SEG, "samples_per_segment": N_SAMPLES,
                   "cache": f"inputs/outputs rotate over {lm['nbuf']} buffer sets (553 MB total > 126 MB L2)",
                   "logmel_mode": "fused single launch" if lm["mode"] == 1 else "two launches",
                   "parallelism": f"segment-sharded x{env.world}, no collective (NCCL / gloo only for the barrier and the max of the timings)"},
        "roofline": lm["roofline"], "cpu_baseline": cpu, "e2e": lm["e2e"], "gpu_launches": lm["gpu_launches"], "clocks": clocks,
    }
    if dec:
        line.update(dec["line"])
        line["decode_step"]["cpu_baseline"] = cpu_dec
        line["decode_step"]["parity"] = parity
    if kv:
        line["kv_reorder"] = kv["line"]
        line["kv_reorder"]["cpu_baseline"] = cpu_kv
    line.update(cfg)
    print(json.dumps(line), flush=True)
    if env.world > 1:
        env.dist.destroy_process_group()


if __name__ == "__main__":
    main()

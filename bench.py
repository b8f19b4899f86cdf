"""Headline benchmark: audio-seconds decoded per second (RTF^-1) on BASELINE.json config 2 --
batch 64 x 10 s utterances (125 GLM tokens each) with a 37-token / 148-frame speech prompt and speaker embedding,
CFG on, 10 Euler steps, bf16 storage / fp32 accumulation, through AudioDecoder's batched decode (flow front end ->
libmossdec Euler solve -> libmossdec HiFT vocoder), random-init weights of the reference architecture, synthetic tokens.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-extras] [--no-cpu-baseline]

One JSON line.  `value`: inputs resident in HBM.  `e2e` (the headline against the reference arm): the public batched API
`AudioDecoder.decode_batch(tokens, prompt_tokens, prompt_feats, embeddings)` with per-utterance HOST tensors, H2D copies,
list handling and the D2H copy of the waveform inside the timed region.  Next to them (N = 1, unless --no-extras):
`latency_b1_ms` (one 10 s utterance through token2wav), `config3` (256 / 16 / 1 concurrent streams, p50 / p99 per chunk
step), `config4` (32 x 60 s), `gpu_eager_baseline` (the oracle restatement = plain eager PyTorch fp32 on this GPU, batch 1),
`cpu_baseline` (the oracle on the host cores); at every N: `config5` (512 mixed 2-30 s utterances, LPT-sharded over the
ranks, no collective on the data path).

N > 1 is launched by torchrun (one rank per GPU); utterances are independent, so every rank decodes its own batch of 64
(weak scaling); NCCL only carries the barriers and the max-over-ranks times.  `--impl reference` times the reference's CPU
path (the pinned oracle restatement, oracle/port.py -- the reference is pure Python and /root/reference does not travel to
the GPU box) on the host cores for the same workload shape, same --steps / --warmup.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH, N_TOK, N_PROMPT = 64, 125, 37
EST_MFLOP_PER_FRAME = 132.162e6      # estimator GEMM-class FLOPs per frame per CFG branch per step
ATTN_MFLOP_PER_FRAME_T = 0.114688e6  # attention FLOPs per frame per branch per step, times T
HIFT_FLOP_PER_FRAME = 612.296e6      # vocoder GEMM-class FLOPs per generated frame
# 56 transformer blocks x (attention out-projection 512 -> 256, fused into the kernel as its prologue GEMM, + FF1 + FF2),
# per frame per branch per step
FFN_FLOP_PER_FRAME = 56 * (2 * 256 * 512 + 2 * 2 * 256 * 1024)
WORKLOAD = "config2: batch 64 x 10 s utterances (125 tokens), 37-token/148-frame prompt + speaker embedding, CFG 0.7, 10 Euler steps, flow + HiFT"


def prompt(g):
    return (torch.randint(0, 16384, (1, N_PROMPT), generator=g).to(torch.int32),
            torch.clamp(torch.randn(1, 4 * N_PROMPT, 80, generator=g) * 2 - 5, min=-11.5129))


def workload(batch, seed0=2000):
    """BASELINE/SURVEY 8(d) config 2 inputs (host tensors, one set per utterance)."""
    toks, ptoks, pfeats = [], [], []
    for i in range(batch):
        g = torch.Generator().manual_seed(seed0 + i)
        toks.append(torch.randint(0, 16384, (1, N_TOK), generator=g))
        pt, pf = prompt(g)
        ptoks.append(pt)
        pfeats.append(pf)
    emb = torch.randn(batch, 192, generator=torch.Generator().manual_seed(seed0 - 1))
    return toks, ptoks, pfeats, emb


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"tflops": float(p.get("bf16_tflops_sustained", p.get("bf16_tflops", 1400.0))), "hbm": float(p["hbm_gbs"]),
                "src": "measured (MEASURED_PEAKS.json: sustained bf16, copy bandwidth)"}
    return {"tflops": 1400.0, "hbm": 6650.0, "src": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path).read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 8:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------------------------------- reference / baselines
def port_decoder(device="cpu"):
    from moss_speech_decoder_cosy_b200 import synth
    from oracle import port
    return port.PortDecoder(synth.synth_flow_state_dict(0), synth.synth_hift_state_dict(0), device=device)


def run_port(dec, n_utts, toks, ptoks, pfeats, emb):
    """The reference path restated (oracle/port.py, fp32, batch 1 like the reference) on `n_utts` utterances."""
    for i in range(n_utts):
        dec.token2wav(toks[i], ptoks[i], pfeats[i], emb[i:i + 1])


def run_reference(args, rank):
    """`--impl reference`: the reference's CPU implementation of the path (oracle port), a bounded sample per step, the
    same --steps / --warmup as the repo's arm."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    n = max(2, args.ref_utts)
    toks, ptoks, pfeats, emb = workload(n)
    dec = port_decoder("cpu")
    for _ in range(args.warmup):
        run_port(dec, n, toks, ptoks, pfeats, emb)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run_port(dec, n, toks, ptoks, pfeats, emb)
    dt = time.perf_counter() - t0
    val = args.steps * n * N_TOK * 0.08 / dt
    sample = f"{n} utterances of the config-2 batch per step (10 s + 37-token prompt each), fp32, batch 1 like the reference"
    print(json.dumps({
        "impl": "reference", "metric": "audio_seconds_per_second", "value": val, "unit": "audio-s/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": val, "unit": "audio-s/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": val, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def median(v):
    v = sorted(v)
    return v[len(v) // 2]


# ---------------------------------------------------------------------------------------------------- extra configurations
def event_ms(fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), out


def bench_latency_b1(dec):
    """One 10 s utterance (config-2 shape: 125 tokens + 37-token prompt) through token2wav, enqueue -> waveform on host."""
    toks, ptoks, pfeats, emb = workload(1, seed0=7000)
    lat = []
    for i in range(8):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        wav, _ = dec.token2wav(toks[0], f"b1-{i}", prompt_token=ptoks[0], prompt_feat=pfeats[0], embedding=emb)
        wav.cpu()
        lat.append((time.perf_counter() - t0) * 1e3)
    return {"p50_ms": median(lat[3:]), "min_ms": min(lat[3:]), "audio_s": N_TOK * 0.08, "rtf_inv": N_TOK * 0.08 / (median(lat[3:]) / 1e3),
            "graph_replays": dec.engine.graph_replays()}


def bench_config3(dec, n_streams, n_tokens=80, prefix_cache=True):
    """`n_streams` concurrent streams, block 5, look-ahead 3, window 40, mel cache 8, 37-token prompt: latency of one chunk
    step (enqueue -> all waveform chunks on the host) when all streams step together.  A first pass warms the allocator
    and captures the CUDA graphs of the chunk geometries; the second pass is measured.  prefix_cache: the frozen-prefix
    cache of StreamScheduler (exact; the first 100 frames of every stream are recorded by its first chunk)."""
    from moss_speech_decoder_cosy_b200.streaming import StreamScheduler
    out = None
    for rep in range(2):
        g = torch.Generator().manual_seed(3000)
        sched = StreamScheduler(dec, block_size=5, max_token_len=40, prefix_cache=prefix_cache)
        for _ in range(n_streams):
            pt, pf = prompt(g)
            sched.add(torch.randint(0, 16384, (n_tokens,), generator=g).tolist(), pt, pf, torch.randn(1, 192, generator=g))
        lat = []
        while any(not s.done for s in sched.streams):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            em = sched.step()
            _ = [o.cpu() for o in em.values()]
            lat.append((time.perf_counter() - t0) * 1e3)
        steady = sorted(lat[8:-1]) or sorted(lat)     # windows are full (40 tokens, T = 296 frames) from the 8th chunk on
        p50 = steady[len(steady) // 2]
        out = {"streams": n_streams, "chunks": len(lat), "first_chunk_ms": lat[0], "p50_ms": p50,
               "p99_ms": steady[min(len(steady) - 1, int(0.99 * len(steady)))], "chunk_audio_s": 0.4,
               "audio_s_per_s": n_streams * 0.4 / (p50 / 1e3), "realtime_margin": 400.0 / p50,
               "prefix_cache": bool(prefix_cache and sched._prefix_frames), "prefix_frames": sched._prefix_frames}
    return out


def bench_config4(dec):
    g = torch.Generator().manual_seed(4000)
    B, N = 32, 750
    toks = [torch.randint(0, 16384, (1, N), generator=g) for _ in range(B)]
    pr = [prompt(g) for _ in range(B)]
    emb = torch.randn(B, 192, generator=g)
    run = lambda: dec.decode_batch(toks, [p[0] for p in pr], [p[1] for p in pr], emb)["wav"].cpu()
    run()
    run()
    ms, _ = event_ms(run)
    return {"batch": B, "audio_s": B * 60.0, "ms": ms, "audio_s_per_s": B * 60.0 / (ms / 1e3), "frames_per_utterance": 4 * (N + N_PROMPT)}


def bench_config5(dec, rank, world, dist, dev):
    """512 utterances of 2-30 s (uniform, seed 5000) with a shared 37-token prompt, LPT-sharded over the ranks by the
    path's cost model (sharding.balanced_partition, identical on every rank: no communication), length-bucketed inside
    a rank.  Time = max over ranks between two barriers."""
    from moss_speech_decoder_cosy_b200 import sharding
    rs = np.random.RandomState(5000)
    lens = [int(round(12.5 * s)) for s in rs.uniform(2, 30, size=512)]
    g = torch.Generator().manual_seed(5000)
    toks = [torch.randint(0, 16384, (1, n), generator=g) for n in lens]
    pt, pf = prompt(g)
    emb = torch.randn(512, 192, generator=g)
    shares = sharding.balanced_partition([sharding.utterance_cost(n, N_PROMPT) for n in lens], world)
    mine = shares[rank]
    buckets = sharding.length_buckets(mine, lens, max_frames=48000)
    wav_h = torch.empty(sum(lens[i] for i in mine) * 1920, dtype=torch.float32).pin_memory()

    def run():
        off = 0
        for b in buckets:
            r = dec.decode_batch([toks[i] for i in b], [pt] * len(b), [pf] * len(b), emb[b])
            n = r["wav"].numel()
            wav_h[off:off + n].copy_(r["wav"], non_blocking=True)
            off += n
    run()
    run()     # second pass: the graphs of the bucket geometries are captured
    if dist is not None:
        dist.barrier()
    ms, _ = event_ms(run)
    ms_max, ms_mean = ms, ms
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_max = float(t)
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms_mean = float(t) / world
    audio = sum(lens) * 0.08
    return {"utterances": 512, "audio_s": audio, "ranks": world, "buckets_rank0": len(buckets), "ms": ms_max,
            "audio_s_per_s": audio / (ms_max / 1e3), "imbalance_max_over_mean": ms_max / ms_mean,
            "partition": "LPT on T * F_est(T) + HiFT cost (sharding.balanced_partition)"}


def bench_gpu_eager(n_utts=2):
    """'the existing implementation on the same box' -- the reference path in eager PyTorch fp32
    on this GPU, batch 1 looped like the reference (oracle/port.py with its weights and inputs on cuda; cuBLAS / cuDNN /
    ATen kernels).  Context for the speed-up, not the ratio's denominator."""
    toks, ptoks, pfeats, emb = workload(n_utts, seed0=2000)
    dec = port_decoder("cuda")
    run_port(dec, 1, toks, ptoks, pfeats, emb)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run_port(dec, n_utts, toks, ptoks, pfeats, emb)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"value": n_utts * N_TOK * 0.08 / dt, "unit": "audio-s/s", "kind": "oracle/port.py on cuda, fp32, batch 1 looped",
            "utterances": n_utts, "s": dt}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--ref-utts", type=int, default=2, help="utterances per CPU-reference step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only (skip configs 3 / 4 / 5, latency, eager baseline)")
    ap.add_argument("--streams", type=int, default=256)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU baseline")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner to stdout when the first communicator comes up: keep stdout to the ONE JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    dec = AudioDecoder(seed=0)
    dev = dec.device
    B = args.batch
    toks, ptoks, pfeats, emb = workload(B, seed0=2000 + 1000 * rank)
    lens_tok = [N_PROMPT + N_TOK] * B
    F_ = [4 * N_PROMPT] * B
    T = 4 * (N_PROMPT + N_TOK)
    dec.engine.reserve(flow_rows=B * T, flow_seqs=B, hift_frames=8192)
    # device-resident copies for the kernel-side leg; the end-to-end leg goes through decode_batch with the host lists
    tok_d = torch.cat([torch.cat([ptoks[b].reshape(-1).long(), toks[b].reshape(-1)]) for b in range(B)]).to(dev)
    feat_d = torch.cat([pf.reshape(-1, 80) for pf in pfeats], 0).to(dev)
    emb_d = emb.to(dev)
    wav_h = torch.empty(B * N_TOK * 1920, dtype=torch.float32).pin_memory()
    audio_s = B * N_TOK * 0.08

    def step_resident():
        r = dec.flow_device(tok_d, lens_tok, feat_d, F_, emb_d, streaming=False, finalize=True)
        return dec.vocode_generated(r)["wav"]

    def step_e2e():   # the call a user makes: per-utterance host tensors in, waveform on the host out
        r = dec.decode_batch(toks, ptoks, pfeats, emb)
        wav_h.copy_(r["wav"], non_blocking=True)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms

    for _ in range(args.warmup):
        step_resident()
    step_e2e()
    torch.cuda.synchronize()
    sampler = ClockSampler(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else local)
    dec.engine.reset_launch_count()
    sampler.start()
    ms = timed(step_resident, args.steps)
    clocks = sampler.stop()
    launches = dec.engine.launch_count()
    # Per-kernel device times for the roofline: a separate pass of the same K steps with CUDA event pairs around every
    # launch.  With profiling on the engine runs the two CFG halves serially and without graph replay (in the timed pass
    # above they overlap on two streams, which would stretch the per-launch durations of co-running kernels).
    dec.engine.profile(True)
    dec.engine.profile_read()
    dec.engine.profile_bandwidth()
    ms_prof = timed(step_resident, args.steps)
    prof = dec.engine.profile_read()
    bw = dec.engine.profile_bandwidth()
    dec.engine.profile(False)
    ms_e2e = timed(step_e2e, args.steps)
    # encoder + 10-step solve alone in the headline mode (two streams, graph replay): what the tensor kernels deliver when
    # the CFG halves fill each other's partial waves, which the serial per-kernel pass above cannot show
    ms_flow = timed(lambda: dec.flow_device(tok_d, lens_tok, feat_d, F_, emb_d, streaming=False, finalize=True), args.steps)
    wav_ok = bool(torch.isfinite(wav_h[:4096]).all()) and float(wav_h.abs().max()) > 0

    value = world * audio_s * args.steps / (ms / 1e3)
    e2e = world * audio_s * args.steps / (ms_e2e / 1e3)
    pk = peaks()
    # Algorithmic FLOPs per step of the three tensor-core kernels (SURVEY 8d figures):
    frames_steps = B * 10 * 2 * T                                    # frame . CFG branch . Euler step
    f0_flop = B * 4 * N_TOK * 6.54e6                                 # the f0 predictor runs in fp32 on the CUDA cores
    flops = {
        "attention_ws_kernel": frames_steps * ATTN_MFLOP_PER_FRAME_T * T,
        "ffn_tc_kernel": frames_steps * FFN_FLOP_PER_FRAME,
        "gemm_tc_kernel": frames_steps * (EST_MFLOP_PER_FRAME - FFN_FLOP_PER_FRAME) + B * 4 * N_TOK * HIFT_FLOP_PER_FRAME - f0_flop,
    }
    times = {"attention_ws_kernel": (prof["attn_ms"], prof["attn_launches"]),
             "ffn_tc_kernel": (prof["ffn_ms"], prof["ffn_launches"]),
             "gemm_tc_kernel": (prof["gemm_ms"], prof["gemm_launches"])}
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "ncu_dram_traffic.json")   # bytes per launch from the committed ncu --set full capture
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    desc = {"attention_ws_kernel": "warp-specialised tcgen05 flash attention (QK^T and PV on the tensor cores, double-buffered S and P in TMEM)",
            "ffn_tc_kernel": "fused transformer tail (attention out-projection + residual + LayerNorm + FF1 + GELU + FF2 + residual + LayerNorm, tcgen05, hidden on chip)",
            "gemm_tc_kernel": "tcgen05 implicit GEMM family (convs, projections, vocoder) with fused epilogues"}
    kernels = {}
    for k, (kms, kn) in times.items():
        sec = kms / 1e3 / args.steps
        ach = flops[k] / sec / 1e12 if sec > 0 else 0.0
        kernels[k] = {"ms_per_step": kms / args.steps, "launches_per_step": kn / args.steps, "share_of_step": kms / ms_prof if ms_prof > 0 else None,
                      "achieved": ach, "frac": ach / pk["tflops"], "algorithmic_tflop_per_step": flops[k] / 1e12,
                      "traffic": traffic.get(k)}
    # bandwidth-class kernels: achieved HBM GB/s = algorithmic bytes / device time (CUDA events on the launch stream)
    bw_kernels = {}
    for k, v in bw.items():
        if v["launches"] and v["ms"] > 0:
            gbs = v["bytes"] / (v["ms"] / 1e3) / 1e9
            bw_kernels[k] = {"achieved": gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                             "us_per_launch": v["ms"] / v["launches"] * 1e3, "launches_per_step": v["launches"] / args.steps,
                             "algorithmic_mb_per_launch": v["bytes"] / v["launches"] / 1e6, "traffic": traffic.get(k)}
    flow_flop = frames_steps * (EST_MFLOP_PER_FRAME + ATTN_MFLOP_PER_FRAME_T * T)   # estimator only (the encoder's share is < 1 %)
    dom = "attention_ws_kernel"   # the dominant SINGLE kernel of the step (gemm_tc_kernel is a family of five epilogue modes)
    line = {
        "metric": "audio_seconds_per_second", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD + f" (batch {B} per GPU)", "global_batch": B * world,
                   "l2": "activation working set per step (> 2 GB) exceeds the 126 MB L2; no explicit flush",
                   "parallelism": f"dp{world} (independent utterances, no data-path collective)",
                   "headline": "e2e (public API, host tensors in, waveform on host out); value = same step with inputs resident in HBM"},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "audio-s/s", "h2d_bytes_per_step": int(dec.last_h2d_bytes),
                "d2h_bytes_per_step": int(wav_h.nbytes), "ms_per_step": ms_e2e / args.steps, "output_ok": wav_ok,
                "api": "AudioDecoder.decode_batch(tokens, prompt_tokens, prompt_feats, embeddings) with per-utterance host tensors"},
        "gpu_launches": int(launches),
        "graph_replays": dec.engine.graph_replays(),
        "roofline": {"bound": "tensor", "kernel": f"{dom} ({desc[dom]})", "achieved": kernels[dom]["achieved"],
                     "peak": pk["tflops"], "unit": "TFLOP/s", "frac": kernels[dom]["frac"],
                     "traffic": kernels[dom]["traffic"], "peak_source": pk["src"],
                     "launches_per_step": kernels[dom]["launches_per_step"],
                     "kernel_ms_per_step": kernels[dom]["ms_per_step"], "share_of_step": kernels[dom]["share_of_step"],
                     "timed_in": f"a serial profiling pass of the same {args.steps} steps ({ms_prof / args.steps:.1f} ms/step; the headline pass overlaps the two CFG halves on two streams and replays a CUDA graph)",
                     "all_tensor_kernels": kernels,
                     "tensor_kernels_total": {"ms_per_step": sum(k["ms_per_step"] for k in kernels.values()),
                                              "achieved": sum(flops.values()) / (sum(t[0] for t in times.values()) / 1e3 / args.steps) / 1e12,
                                              "frac": sum(flops.values()) / (sum(t[0] for t in times.values()) / 1e3 / args.steps) / 1e12 / pk["tflops"]},
                     "flow_two_stream": {"ms_per_step": ms_flow / args.steps,
                                         "achieved": flow_flop / (ms_flow / 1e3 / args.steps) / 1e12,
                                         "frac": flow_flop / (ms_flow / 1e3 / args.steps) / 1e12 / pk["tflops"],
                                         "what": "estimator FLOPs (attention + tail + other estimator GEMMs) / device time of encoder + 10-step solve "
                                                 "as the headline runs it (CFG halves on two streams, CUDA graph); the per-kernel figures above come from "
                                                 "the serial pass and carry each launch's partial last wave (324 row tiles on 148 SMs)"},
                     "bandwidth_kernels": bw_kernels},
    }
    extras = not args.no_extras
    if extras:
        # config 5 at every N (sharded); the single-GPU extras only at N = 1
        line["config5"] = bench_config5(dec, rank, world, dist, dev)
    if extras and world == 1:
        line["latency_b1_ms"] = bench_latency_b1(dec)
        line["config4"] = bench_config4(dec)
        c3 = bench_config3(dec, args.streams)
        c3["p50_ms_16_streams"] = bench_config3(dec, 16, n_tokens=60)["p50_ms"]
        c3["p50_ms_1_stream"] = bench_config3(dec, 1, n_tokens=60)["p50_ms"]
        # the reference's semantics taken literally (full recompute of prompt + window per chunk, no frozen-prefix cache)
        c3["p50_ms_recompute"] = bench_config3(dec, args.streams, prefix_cache=False)["p50_ms"]
        line["config3"] = c3
        try:
            line["gpu_eager_baseline"] = bench_gpu_eager()
        except Exception as e:  # context only: never fail the bench line on it
            line["gpu_eager_baseline"] = {"unavailable": repr(e)[:200]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        cdec = port_decoder("cpu")
        run_port(cdec, 1, toks, ptoks, pfeats, emb)           # warm-up
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            run_port(cdec, 1, toks, ptoks, pfeats, emb)
            ts.append(time.perf_counter() - t0)
        dt = median(ts)
        line["cpu_baseline"] = {"value": N_TOK * 0.08 / dt, "unit": "audio-s/s", "cores": torch.get_num_threads(), "kind": "port",
                                "sample": f"1 utterance of the batch (10 s + 37-token prompt), fp32, warm-up 1, median of 3 ({dt:.1f} s each)"}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Headline benchmark: audio-seconds decoded per second (RTF^-1) on BASELINE.json config 2 --
batch 64 x 10 s utterances (125 GLM tokens each) with a 37-token / 148-frame speech prompt and speaker embedding,
CFG on, 10 Euler steps, bf16 storage / fp32 accumulation, through AudioDecoder's batched decode (flow front end ->
libmossdec Euler solve -> libmossdec HiFT vocoder), random-init weights of the reference architecture, synthetic tokens.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N > 1 is launched by torchrun (one rank per GPU); utterances are independent, so every rank decodes its own batch of 64
(weak scaling) with no collective on the data path; NCCL only carries the barrier and the max-over-ranks time.
`--impl reference` times the reference's CPU path (the pinned oracle restatement, oracle/port.py -- the reference is
pure Python and /root/reference does not travel to the GPU box) on the host cores for the same workload shape.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH, N_TOK, N_PROMPT = 64, 125, 37
EST_MFLOP_PER_FRAME = 132.162e6      # estimator GEMM-class FLOPs per frame per CFG branch per step
ATTN_MFLOP_PER_FRAME_T = 0.114688e6  # attention FLOPs per frame per branch per step, times T
HIFT_FLOP_PER_FRAME = 612.296e6      # vocoder GEMM-class FLOPs per generated frame
FFN_FLOP_PER_FRAME = 56 * 2 * 2 * 256 * 1024   # 56 transformer blocks x (FF1 + FF2), per frame per branch per step


def workload(batch, seed0=2000):
    """BASELINE/SURVEY 8(d) config 2 inputs (host tensors)."""
    toks, ptoks, pfeats = [], [], []
    for i in range(batch):
        g = torch.Generator().manual_seed(seed0 + i)
        toks.append(torch.randint(0, 16384, (1, N_TOK), generator=g))
        ptoks.append(torch.randint(0, 16384, (1, N_PROMPT), generator=g).to(torch.int32))
        pfeats.append(torch.clamp(torch.randn(1, 4 * N_PROMPT, 80, generator=g) * 2 - 5, min=-11.5129))
    emb = torch.randn(batch, 192, generator=torch.Generator().manual_seed(seed0 - 1))
    return toks, ptoks, pfeats, emb


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"tflops": float(p.get("bf16_tflops_sustained", p.get("bf16_tflops", 1400.0))), "hbm": float(p["hbm_gbs"]),
                "src": "measured (MEASURED_PEAKS.json, sustained bf16)"}
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


def cpu_port_decoder():
    from moss_speech_decoder_cosy_b200 import synth
    from oracle import port
    return port.PortDecoder(synth.synth_flow_state_dict(0), synth.synth_hift_state_dict(0))


def time_cpu_port(dec, n_utts, toks, ptoks, pfeats, emb):
    """Reference CPU path (oracle restatement, fp32, all host threads) on `n_utts` utterances of the workload."""
    t0 = time.perf_counter()
    for i in range(n_utts):
        dec.token2wav(toks[i], ptoks[i], pfeats[i], emb[i:i + 1])
    dt = time.perf_counter() - t0
    return n_utts * N_TOK * 0.08 / dt, dt


def run_reference(args, rank):
    """`--impl reference`: the reference's CPU implementation of the path (oracle port), bounded sample per step."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    toks, ptoks, pfeats, emb = workload(max(1, args.ref_utts))
    dec = cpu_port_decoder()
    for _ in range(args.warmup_ref):
        time_cpu_port(dec, 1, toks, ptoks, pfeats, emb)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        time_cpu_port(dec, args.ref_utts, toks, ptoks, pfeats, emb)
    dt = time.perf_counter() - t0
    val = args.steps * args.ref_utts * N_TOK * 0.08 / dt
    sample = f"{args.ref_utts} utterance(s) of the config-2 batch per step (10 s + 37-token prompt each), fp32"
    print(json.dumps({
        "impl": "reference", "metric": "audio_seconds_per_second", "value": val, "unit": "audio-s/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup_ref, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "config2: batch 64 x 10 s utterances, 37-token prompt, CFG, 10 Euler steps"},
        "cpu_baseline": {"value": val, "unit": "audio-s/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": val, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--ref-utts", type=int, default=1, help="utterances per CPU-reference step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    args.warmup_ref = 1
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU baseline")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    dec = AudioDecoder(seed=0)
    dev = dec.device
    B = args.batch
    toks, ptoks, pfeats, emb = workload(B, seed0=2000 + 1000 * rank)
    lens_tok = [N_PROMPT + N_TOK] * B
    F_ = [4 * N_PROMPT] * B
    # host (pinned) copies for the end-to-end leg, device-resident copies for the kernel-side leg
    tok_h = torch.stack([torch.cat([ptoks[b].reshape(-1).long(), toks[b].reshape(-1)]) for b in range(B)]).pin_memory()
    feat_h = torch.cat([pf.reshape(-1, 80) for pf in pfeats], 0).pin_memory()
    emb_h = emb.pin_memory()
    wav_h = torch.empty(B * N_TOK * 1920, dtype=torch.float32).pin_memory()
    tok_d, feat_d, emb_d = tok_h.to(dev), feat_h.to(dev), emb_h.to(dev)
    audio_s = B * N_TOK * 0.08

    def step_resident():
        r = dec.flow_device(tok_d, lens_tok, feat_d, F_, emb_d, streaming=False, finalize=True)
        return dec.vocode_generated(r)["wav"]

    def step_e2e():
        r = dec.flow_device(tok_h.to(dev, non_blocking=True), lens_tok, feat_h.to(dev, non_blocking=True), F_,
                            emb_h.to(dev, non_blocking=True), streaming=False, finalize=True)
        wav_h.copy_(dec.vocode_generated(r)["wav"], non_blocking=True)

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
    # launch.  With profiling on the engine runs the two CFG halves serially (in the timed pass above they overlap on two
    # streams, which would stretch the per-launch durations of co-running kernels).
    dec.engine.profile(True)
    dec.engine.profile_read()
    ms_prof = timed(step_resident, args.steps)
    prof = dec.engine.profile_read()
    dec.engine.profile(False)
    ms_e2e = timed(step_e2e, args.steps)
    wav_ok = bool(torch.isfinite(wav_h[:4096]).all()) and float(wav_h.abs().max()) > 0

    value = world * audio_s * args.steps / (ms / 1e3)
    e2e = world * audio_s * args.steps / (ms_e2e / 1e3)
    T = 4 * (N_PROMPT + N_TOK)
    pk = peaks()
    # Algorithmic FLOPs per step of the three tensor-core kernels (SURVEY 8d figures):
    frames_steps = B * 10 * 2 * T                                    # frame . CFG branch . Euler step
    flops = {
        "attention_ws_kernel": frames_steps * ATTN_MFLOP_PER_FRAME_T * T,
        "ffn_tc_kernel": frames_steps * FFN_FLOP_PER_FRAME,
        "gemm_tc_kernel": frames_steps * (EST_MFLOP_PER_FRAME - FFN_FLOP_PER_FRAME) + B * 4 * N_TOK * HIFT_FLOP_PER_FRAME,
    }
    times = {"attention_ws_kernel": (prof["attn_ms"], prof["attn_launches"]),
             "ffn_tc_kernel": (prof["ffn_ms"], prof["ffn_launches"]),
             "gemm_tc_kernel": (prof["gemm_ms"], prof["gemm_launches"])}
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "ncu_dram_traffic.json")   # bytes per launch from the committed ncu --set full capture
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    desc = {"attention_ws_kernel": "warp-specialised tcgen05 flash attention (QK^T and PV on the tensor cores, double-buffered S and P in TMEM)",
            "ffn_tc_kernel": "fused feed-forward (FF1 + GELU + FF2 + residual + LayerNorm, tcgen05, hidden on chip)",
            "gemm_tc_kernel": "tcgen05 implicit GEMM (convs, projections, vocoder) with fused epilogues"}
    kernels = {}
    for k, (kms, kn) in times.items():
        sec = kms / 1e3 / args.steps
        ach = flops[k] / sec / 1e12 if sec > 0 else 0.0
        kernels[k] = {"ms_per_step": kms / args.steps, "launches_per_step": kn / args.steps, "share_of_step": kms / ms_prof if ms_prof > 0 else None,
                      "achieved": ach, "frac": ach / pk["tflops"], "algorithmic_tflop_per_step": flops[k] / 1e12,
                      "traffic": traffic.get(k)}
    dom = max(kernels, key=lambda k: kernels[k]["ms_per_step"])       # the dominant kernel of the step
    line = {
        "metric": "audio_seconds_per_second", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"config2: batch {B} x 10 s utterances (125 tokens) per GPU, 37-token/148-frame prompt + "
                               "speaker embedding, CFG 0.7, 10 Euler steps, flow + HiFT", "global_batch": B * world,
                   "l2": "activation working set per step (> 2 GB) exceeds the 126 MB L2; no explicit flush",
                   "parallelism": f"dp{world} (independent utterances, no data-path collective)"},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "audio-s/s", "h2d_bytes_per_step": int(tok_h.nbytes + feat_h.nbytes + emb_h.nbytes),
                "d2h_bytes_per_step": int(wav_h.nbytes), "ms_per_step": ms_e2e / args.steps, "output_ok": wav_ok},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": f"{dom} ({desc[dom]})", "achieved": kernels[dom]["achieved"],
                     "peak": pk["tflops"], "unit": "TFLOP/s", "frac": kernels[dom]["frac"],
                     "traffic": kernels[dom]["traffic"], "peak_source": pk["src"],
                     "launches_per_step": kernels[dom]["launches_per_step"],
                     "kernel_ms_per_step": kernels[dom]["ms_per_step"], "share_of_step": kernels[dom]["share_of_step"],
                     "timed_in": f"a serial profiling pass of the same {args.steps} steps ({ms_prof / args.steps:.1f} ms/step; the headline pass overlaps the two CFG halves on two streams)",
                     "all_tensor_kernels": kernels,
                     "tensor_kernels_total": {"ms_per_step": sum(k["ms_per_step"] for k in kernels.values()),
                                              "achieved": sum(flops.values()) / (sum(t[0] for t in times.values()) / 1e3 / args.steps) / 1e12}},
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        cdec = cpu_port_decoder()
        v, dt = time_cpu_port(cdec, 1, toks, ptoks, pfeats, emb)
        line["cpu_baseline"] = {"value": v, "unit": "audio-s/s", "cores": torch.get_num_threads(), "kind": "port",
                                "sample": f"1 utterance of the batch (10 s + 37-token prompt), fp32, {dt:.1f} s of CPU work"}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

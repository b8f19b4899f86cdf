"""Secondary measurements (BASELINE.json configs 3, 4, 5) on one GPU.  Prints one JSON line per config.

    python scripts/bench_configs.py [--configs 3,4,5] [--streams 256]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moss_speech_decoder_cosy_b200 import sharding  # noqa: E402
from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder  # noqa: E402
from moss_speech_decoder_cosy_b200.streaming import StreamScheduler  # noqa: E402

P = 37


def prompt(g):
    return (torch.randint(0, 16384, (1, P), generator=g).int(),
            torch.clamp(torch.randn(1, 4 * P, 80, generator=g) * 2 - 5, min=-11.5129))


def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), out


def config3(dec, n_streams):
    """256 concurrent streams, block 5, look-ahead 3, window 40, mel cache 8: per-chunk latency when all step together."""
    g = torch.Generator().manual_seed(3000)
    sched = StreamScheduler(dec, block_size=5, max_token_len=40)
    for _ in range(n_streams):
        pt, pf = prompt(g)
        sched.add(torch.randint(0, 16384, (80,), generator=g).tolist(), pt, pf, torch.randn(1, 192, generator=g))
    lat = []
    while any(not s.done for s in sched.streams):
        t0 = time.perf_counter()
        out = sched.step()
        host = [o.cpu() for o in out.values()]       # chunk latency = enqueue -> waveform on the host
        lat.append((time.perf_counter() - t0) * 1e3)
    steady = sorted(lat[8:-1]) or sorted(lat)        # windows are full (40 tokens) from the 8th chunk on
    audio = n_streams * 0.4
    p50 = steady[len(steady) // 2]
    return {"config": 3, "streams": n_streams, "chunks": len(lat), "first_chunk_ms": lat[0], "p50_ms": p50,
            "p99_ms": steady[min(len(steady) - 1, int(0.99 * len(steady)))], "audio_s_per_s": audio / (p50 / 1e3),
            "realtime_margin": 400.0 / p50}


def config4(dec):
    g = torch.Generator().manual_seed(4000)
    B, N = 32, 750
    toks = [torch.randint(0, 16384, (1, N), generator=g) for _ in range(B)]
    pr = [prompt(g) for _ in range(B)]
    emb = torch.randn(B, 192, generator=g)
    run = lambda: dec.decode_batch(toks, [p[0] for p in pr], [p[1] for p in pr], emb)
    run()
    ms, _ = timed(run)
    return {"config": 4, "batch": B, "audio_s": B * 60.0, "ms": ms, "audio_s_per_s": B * 60.0 / (ms / 1e3)}


def config5(dec):
    rs = np.random.RandomState(5000)
    lens = [int(round(12.5 * s)) for s in rs.uniform(2, 30, size=512)]
    g = torch.Generator().manual_seed(5000)
    toks = [torch.randint(0, 16384, (1, n), generator=g) for n in lens]
    pt, pf = prompt(g)
    emb = torch.randn(512, 192, generator=g)
    buckets = sharding.length_buckets(list(range(512)), lens, max_frames=48000)

    def run():
        for b in buckets:
            dec.decode_batch([toks[i] for i in b], [pt] * len(b), [pf] * len(b), emb[b])
    run()
    ms, _ = timed(run)
    audio = sum(lens) * 0.08
    return {"config": 5, "utterances": 512, "buckets": len(buckets), "audio_s": audio, "ms": ms,
            "audio_s_per_s": audio / (ms / 1e3)}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="3,4,5")
    ap.add_argument("--streams", type=int, default=256)
    a = ap.parse_args()
    dec = AudioDecoder(seed=0)
    for c in a.configs.split(","):
        r = {"3": lambda: config3(dec, a.streams), "4": lambda: config4(dec), "5": lambda: config5(dec)}[c]()
        print(json.dumps(r), flush=True)

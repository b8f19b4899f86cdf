"""Device timing of the row-A9 building blocks (cosyvoice1 / Matcha ResnetBlock1D, Downsample1D, Upsample1D) at the
config-2 row count (B2 = 128 CFG-doubled utterances x T = 648 frames): CUDA events on the launch stream, 3 warm-up + N timed
calls per op, inputs larger than L2 in aggregate (x alone is 106 MB fp32).  One JSON line per op.

    python scripts/bench_cv1.py [--iters 10] [--B 128] [--T 648]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moss_speech_decoder_cosy_b200 import runtime, synth, weights  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--B", type=int, default=128)
    ap.add_argument("--T", type=int, default=648)
    a = ap.parse_args()
    eng = runtime.build_engine(seed=0)
    sd = synth.synth_cv1_blocks_state_dict(0)
    t = {}
    t.update(weights.pack_cv1_resnet(sd, "res_in", "cv1.res_in"))
    t.update(weights.pack_cv1_resnet(sd, "res_mid", "cv1.res_mid"))
    t.update(weights.pack_cv1_downsample(sd, "down", "cv1.down"))
    t.update(weights.pack_cv1_upsample(sd, "up", "cv1.up"))
    t.update(weights.pack_cv1_block1d(sd, "final", "cv1.final"))
    eng.load_tensors(t)
    B, T = a.B, a.T
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(B, 320, T, device="cuda", generator=g)
    h = torch.randn(B, 256, T, device="cuda", generator=g)
    temb = torch.randn(B, 1024, device="cuda", generator=g)
    mask = torch.ones(B, 1, T, device="cuda")
    rows = B * T
    ops = {
        # algorithmic flops: two k3 convs + 1x1 residual; HBM floor: read x fp32, write out fp32
        "resnet_block_320": (lambda: eng.cv1_resnet_block("cv1.res_in", x, mask, temb),
                             2.0 * rows * 256 * (3 * 320 + 3 * 256 + 320), rows * (320 + 256) * 4),
        "resnet_block_256": (lambda: eng.cv1_resnet_block("cv1.res_mid", h, mask, temb),
                             2.0 * rows * 256 * (3 * 256 + 3 * 256 + 256), rows * (256 + 256) * 4),
        "block1d": (lambda: eng.cv1_block1d("cv1.final", h, mask), 2.0 * rows * 256 * 768, rows * 512 * 4),
        "downsample": (lambda: eng.cv1_downsample("cv1.down", h), 2.0 * (rows / 2) * 256 * 768, rows * 256 * 4 * 1.5),
        "upsample": (lambda: eng.cv1_upsample("cv1.up", h), 2.0 * rows * 256 * 4 * 256, rows * 256 * 4 * 3),
    }
    for name, (fn, flops, io_bytes) in ops.items():
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = eng.launch_count()
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        print(json.dumps({"op": name, "B": B, "T": T, "ms": round(ms, 4), "launches_per_call": (eng.launch_count() - n0) // a.iters,
                          "algorithmic_tflops": round(flops / ms / 1e9, 2), "boundary_io_gbps": round(io_bytes / ms / 1e6, 1),
                          "frames_per_s": round(rows / ms * 1e3)}))


if __name__ == "__main__":
    main()

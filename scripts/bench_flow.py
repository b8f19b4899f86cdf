"""Developer probe: times mossdec_flow_solve alone (config-2 shape by default).  Not the judged benchmark."""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from moss_speech_decoder_cosy_b200 import runtime  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--T", type=int, default=648)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--chunk", type=int, default=0)
a = ap.parse_args()

eng = runtime.build_engine(seed=0, with_hift=False)
dev = "cuda"
g = torch.Generator().manual_seed(1)
mu = torch.randn(a.B * 80 * a.T, generator=g).to(dev)
cond = torch.zeros_like(mu)
spks = torch.randn(a.B, 80, generator=g).to(dev)
lens = [a.T] * a.B
for _ in range(2):
    eng.flow_solve(lens, mu, spks, cond, n_steps=a.steps, chunk=a.chunk)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
eng.reset_launch_count()
e0.record()
for _ in range(a.reps):
    mel = eng.flow_solve(lens, mu, spks, cond, n_steps=a.steps, chunk=a.chunk)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.reps
flops = a.steps * 2 * a.B * a.T * (132.162e6 + 0.114688e6 * a.T)
eng.profile(True)
for _ in range(a.reps):
    eng.flow_solve(lens, mu, spks, cond, n_steps=a.steps, chunk=a.chunk)
pr = eng.profile_read()
eng.profile(False)
print(f"  serial per solve: attention {pr['attn_ms'] / a.reps:.2f} ms  ffn {pr['ffn_ms'] / a.reps:.2f} ms  other gemm {pr['gemm_ms'] / a.reps:.2f} ms")
print(f"flow_solve B={a.B} T={a.T} steps={a.steps}: {ms:.2f} ms  {flops / ms / 1e9:.1f} TFLOP/s  "
      f"launches/solve={eng.launch_count() / a.reps:.0f}  finite={bool(torch.isfinite(mel).all())}")

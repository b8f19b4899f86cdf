"""Developer probe for A/B library builds (gpu_ab.sh): one 10-step flow_solve on fixed inputs, saved to argv[1];
with two more arguments, compares two saved results."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if len(sys.argv) == 3:
    a, b = torch.load(sys.argv[1]), torch.load(sys.argv[2])
    print(f"A/B mel rel_l2 {float((a - b).norm() / b.norm()):.3e}  max abs {float((a - b).abs().max()):.3e}  finite {bool(torch.isfinite(a).all())}")
else:
    from moss_speech_decoder_cosy_b200 import runtime
    eng = runtime.build_engine(seed=0, with_hift=False)
    g = torch.Generator().manual_seed(1)
    B, T = 8, 648
    mu = torch.randn(B * 80 * T, generator=g).cuda()
    spks = torch.randn(B, 80, generator=g).cuda()
    mel = eng.flow_solve([T] * B, mu, spks, torch.zeros_like(mu), n_steps=10, chunk=int(os.environ.get("AB_CHUNK", "0")))
    torch.cuda.synchronize()
    torch.save(mel.float().cpu(), sys.argv[1])

"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference on CPU (fp32).

    python -m oracle.make_golden [estimator] [solve] [hift] [e2e] [stream]

The reference has no tests or golden vectors of its own, so parity is pinned on outputs of the
reference code itself, produced here with the deterministic random-init checkpoint of
moss_speech_decoder_cosy_b200/synth.py (seed 0) and seeded inputs.  Fixtures hold inputs + outputs only; the weights are
regenerated from the seed on the GPU box.
"""
import os
import sys

import numpy as np
import torch

from . import build_ref

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _save(name, **arrays):
    os.makedirs(GOLD, exist_ok=True)
    path = os.path.join(GOLD, name)
    np.savez_compressed(path, **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v))
                                 for k, v in arrays.items()})
    print("wrote", path, {k: tuple(np.asarray(v).shape) for k, v in arrays.items()})


def _logmel_like(g, n):
    return torch.clamp(torch.randn(1, n, 80, generator=g) * 2 - 5, min=float(np.log(1e-5)))


def gen_estimator(flow):
    est = flow.decoder.estimator
    for T in (36, 150):
        g = torch.Generator().manual_seed(100 + T)
        x = torch.randn(2, 80, T, generator=g)
        mu = torch.randn(2, 80, T, generator=g)
        cond = torch.randn(2, 80, T, generator=g) * 0.5
        spks = torch.randn(2, 80, generator=g)
        t = torch.tensor([0.3, 0.71])
        mask = torch.ones(2, 1, T)
        with torch.inference_mode():
            out_full = est(x, mask, mu, t, spks, cond, streaming=False)
            out_chunk = est(x, mask, mu, t, spks, cond, streaming=True)
        _save(f"estimator_T{T}.npz", x=x, mu=mu, cond=cond, spks=spks, t=t, out_full=out_full, out_chunk=out_chunk,
              chunk=np.int32(est.static_chunk_size))


def gen_solve(flow):
    """CausalConditionalCFM.forward (flow_matching.py:206-230): 10 Euler steps, CFG 0.7, fixed noise bank."""
    for T, P in ((120, 40), (64, 0)):
        g = torch.Generator().manual_seed(200 + T)
        mu = torch.randn(1, 80, T, generator=g)
        spks = torch.randn(1, 80, generator=g)
        cond = torch.zeros(1, 80, T)
        if P:
            cond[:, :, :P] = _logmel_like(g, P).transpose(1, 2)
        mask = torch.ones(1, 1, T)
        with torch.inference_mode():
            mel, _ = flow.decoder(mu=mu.clone(), mask=mask, spks=spks, cond=cond, n_timesteps=10, streaming=False)
            mel_s, _ = flow.decoder(mu=mu.clone(), mask=mask, spks=spks, cond=cond, n_timesteps=10, streaming=True)
        _save(f"solve_T{T}.npz", mu=mu, spks=spks, cond=cond, noise=flow.decoder.rand_noise[0, :, :T], mel=mel,
              mel_stream=mel_s, chunk=np.int32(flow.decoder.estimator.static_chunk_size))


def main(argv):
    what = set(argv) or {"estimator", "solve", "hift", "e2e", "stream"}
    flow = build_ref.build_flow() if what & {"estimator", "solve", "e2e", "stream"} else None
    if "estimator" in what:
        gen_estimator(flow)
    if "solve" in what:
        gen_solve(flow)
    from . import make_golden_hift
    if what & {"hift", "e2e", "stream"}:
        make_golden_hift.run(what, flow)


if __name__ == "__main__":
    main(sys.argv[1:])

"""TEST INFRASTRUCTURE ONLY -- HiFT / end-to-end / streaming golden vectors from the UNMODIFIED reference (CPU, fp32).

HiFT draws its NSF randomness from torch's global generator (generator.py:270,334,388).  The fixtures pin it by
seeding the generator right before `HiFTGenerator.inference` and re-drawing the same stream afterwards
(rand(B, 9) -> randn(B, L, 9)), so the injected-noise input of the CUDA path is exactly what the reference consumed.
"""
import numpy as np
import torch

from . import build_ref


def nsf_draws(seed, L):
    """The draws SineGen2 consumes for one utterance of L samples, in the reference's order.

    `sine_waves` is a transposed view ([1, L, 9] with strides (9L, 1, L), generator.py:281), so
    `torch.randn_like(sine_waves)` (generator.py:334) fills a non-contiguous tensor: torch's CPU generator then takes
    its scalar path, whose stream differs from `torch.randn(1, L, 9)`.  Reproduce it with the same strides."""
    torch.manual_seed(seed)
    rand_ini = torch.rand(1, 9)          # generator.py:270 (provably unused: only sample 0 of the phase is touched)
    noise = torch.randn_like(torch.empty(1, 9, L).transpose(1, 2))
    return rand_ini, noise.contiguous()


def hift_case(hift, name, T, seed, cache_len=0):
    from .make_golden import _save
    g = torch.Generator().manual_seed(seed)
    mel = torch.randn(1, 80, T, generator=g)
    cache = torch.zeros(1, 1, 0)
    if cache_len:
        cache = torch.tanh(torch.randn(1, 1, cache_len, generator=g) * 0.2)
    torch.manual_seed(seed + 1)
    with torch.inference_mode():
        wav, s = hift.inference(speech_feat=mel, cache_source=cache)
        f0 = hift.f0_predictor(mel)
    _, noise = nsf_draws(seed + 1, 480 * T)
    _save(name, mel=mel[0], f0=f0[0], noise=noise[0], wav=wav[0],
          source=s[0, 0], cache=cache[0, 0], seed=np.int64(seed + 1))


def run(what, flow):
    if "hift" in what:
        hift = build_ref.build_hift()
        hift_case(hift, "hift_T40.npz", 40, 300)
        hift_case(hift, "hift_T28_cache.npz", 28, 301, cache_len=3840)

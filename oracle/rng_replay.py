"""TEST INFRASTRUCTURE ONLY -- replays the draws HiFT's NSF source takes from torch's global CPU generator.

Per `HiFTGenerator.inference` call on L samples the reference draws, in order (generator.py:270,334,388):
  torch.rand(1, 9); torch.randn_like(sine_waves)  -- a transposed view [1, L, 9] with strides (9L, 1, L), filled through
  torch's scalar (non-contiguous) path; torch.randn_like(uv) [1, L, 1] (contiguous, vectorised path, result unused).
Seeding the generator once and replaying this sequence reproduces the noise of every call of a session."""
import torch


def replay(seed, sample_lens):
    """-> list of [L, 9] fp32 noise tensors, one per vocoder call."""
    torch.manual_seed(int(seed))
    out = []
    for L in sample_lens:
        torch.rand(1, 9)
        n = torch.randn_like(torch.empty(1, 9, int(L)).transpose(1, 2))
        torch.randn(1, int(L), 1)
        out.append(n[0].contiguous())
    return out


def fingerprint(noise):
    """Small digest stored in fixtures to detect a torch build whose CPU generator differs."""
    n = noise.double()
    return [float(n.sum()), float(n.abs().sum()), float(n[0, 0]), float(n[-1, -1])]

"""TEST INFRASTRUCTURE ONLY -- seeded input builders shared by the golden generators (run in the build container
against the unmodified reference) and the tests (run anywhere).  Fixtures then only need to hold OUTPUTS plus a small
fingerprint of the regenerated inputs, which detects a torch build whose CPU generator differs.
"""
import numpy as np
import torch


def logmel_like(g, n):
    """Prompt-mel-shaped values: clamp(N(-5, 2), min=log 1e-5) (floor of matcha/utils/audio.py:23-24)."""
    return torch.clamp(torch.randn(1, n, 80, generator=g) * 2 - 5, min=float(np.log(1e-5)))


def estimator_inputs(T, seed, B2=2, lens=None):
    """Inputs of one CausalConditionalDecoder.forward call (decoder.py:405): x, mask, mu, t, spks, cond."""
    g = torch.Generator().manual_seed(int(seed))
    x = torch.randn(B2, 80, T, generator=g)
    mu = torch.randn(B2, 80, T, generator=g)
    cond = torch.randn(B2, 80, T, generator=g) * 0.5
    spks = torch.randn(B2, 80, generator=g)
    t = torch.linspace(0.07, 0.93, B2) if B2 > 2 else torch.tensor([0.3, 0.71])[:B2]
    mask = torch.ones(B2, 1, T)
    if lens is not None:
        for b, n in enumerate(lens):
            mask[b, :, n:] = 0
    return dict(x=x, mask=mask, mu=mu, t=t, spks=spks, cond=cond)


def solve_inputs(lens, seed, prompt_frames=None):
    """Inputs of CausalConditionalCFM.forward (flow_matching.py:206) for a list of utterances: per utterance
    mu [80, T], cond [80, T] (prompt mel in the first P frames), spks [80]."""
    g = torch.Generator().manual_seed(int(seed))
    out = []
    for i, T in enumerate(lens):
        mu = torch.randn(80, T, generator=g)
        spks = torch.randn(80, generator=g)
        cond = torch.zeros(80, T)
        P = 0 if prompt_frames is None else min(int(prompt_frames[i]), T)
        if P:
            cond[:, :P] = logmel_like(g, P)[0].t()
        out.append(dict(mu=mu, spks=spks, cond=cond))
    return out


def utterance(seed, n_tokens, n_prompt, vocab=16384):
    """token [1, N], prompt_token int32 [1, P], prompt_feat [1, 4P, 80], embedding [1, 192]."""
    g = torch.Generator().manual_seed(int(seed))
    token = torch.randint(0, vocab, (1, n_tokens), generator=g)
    prompt_token = torch.randint(0, vocab, (1, n_prompt), generator=g).to(torch.int32)
    prompt_feat = logmel_like(g, 4 * n_prompt)
    emb = torch.randn(1, 192, generator=g)
    return token, prompt_token, prompt_feat, emb


def fingerprint(*tensors):
    """Digest of regenerated inputs stored next to the golden outputs."""
    out = []
    for t in tensors:
        d = t.double().reshape(-1)
        out += [float(d.sum()), float(d.abs().sum()), float(d[0]), float(d[-1])]
    return np.asarray(out)


def prompt_waveform(seed, n_samples, sr=24000):
    """A speech-like test waveform in [-1, 1]: a few harmonics of a gliding f0 under a slow amplitude envelope with a
    near-silent stretch, plus low-level noise -- so the log-mel covers its whole dynamic range down to the 1e-5 clamp."""
    g = torch.Generator().manual_seed(int(seed))
    t = torch.arange(n_samples, dtype=torch.float64) / sr
    f0 = 110.0 + 60.0 * torch.sin(2 * np.pi * 0.7 * t)
    phase = 2 * np.pi * torch.cumsum(f0, 0) / sr
    y = sum((0.5 / k) * torch.sin(k * phase + k) for k in range(1, 12))
    env = (0.55 + 0.45 * torch.sin(2 * np.pi * 1.3 * t)).clamp_min(0.0) ** 2
    env[int(0.45 * n_samples):int(0.6 * n_samples)] *= 1e-4
    y = 0.6 * env * y + 2e-4 * torch.randn(n_samples, generator=g, dtype=torch.float64)
    return y.clamp(-1, 1).float().unsqueeze(0)

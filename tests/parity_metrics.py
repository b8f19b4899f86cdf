"""Distance measures shared by the GPU parity tests."""
import numpy as np
import torch


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float(torch.linalg.norm(a - b) / torch.linalg.norm(b))


def snr_db(x, ref):
    x, ref = x.reshape(-1).double().cpu(), ref.reshape(-1).double().cpu()
    return float(10 * torch.log10(ref.pow(2).sum() / (x - ref).pow(2).sum().clamp_min(1e-30)))


def windowed_snr_db(x, ref, win):
    """SNR per consecutive window of `win` samples (how a phase drift grows along an utterance)."""
    x, ref = x.reshape(-1).double().cpu(), ref.reshape(-1).double().cpu()
    n = (ref.numel() // win) * win
    e = (x[:n] - ref[:n]).view(-1, win).pow(2).sum(1).clamp_min(1e-30)
    s = ref[:n].view(-1, win).pow(2).sum(1).clamp_min(1e-30)
    return (10 * torch.log10(s / e)).tolist()


def band_db(x, ref, n_fft=1024, hop=256, nb=40, floor_db=50.0):
    """Mean absolute difference (dB) between log band energies of two waveforms: 40 log-spaced bands of a 1024-point
    STFT, floored 50 dB under the reference's peak band energy (a mel-cepstral-distance-like, phase-blind metric)."""
    x, ref = x.reshape(-1).float().cpu(), ref.reshape(-1).float().cpu()
    w = torch.hann_window(n_fft)
    X = torch.stft(x, n_fft, hop, window=w, return_complex=True).abs() ** 2
    R = torch.stft(ref, n_fft, hop, window=w, return_complex=True).abs() ** 2
    edges = torch.unique(torch.logspace(np.log10(2), np.log10(n_fft // 2 + 1), nb + 1).long().clamp(max=n_fft // 2 + 1))
    bx = torch.stack([X[edges[i]:max(edges[i + 1], edges[i] + 1)].mean(0) for i in range(len(edges) - 1)])
    br = torch.stack([R[edges[i]:max(edges[i + 1], edges[i] + 1)].mean(0) for i in range(len(edges) - 1)])
    fl = br.max() * 10 ** (-floor_db / 10)
    return float((10 * torch.log10(bx.clamp_min(fl)) - 10 * torch.log10(br.clamp_min(fl))).abs().mean())

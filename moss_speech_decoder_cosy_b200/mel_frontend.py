"""Prompt mel front end behind `AudioDecoder.feat_extractor`.

The reference's `feat_extractor` is matcha `mel_spectrogram` (third_party/Matcha-TTS/matcha/utils/audio.py:45-82) bound by
the hyperpyyaml config to n_fft 1920, 80 mels, 24 kHz, hop 480, win 1920, fmin 0, fmax 8000, center False; callers do
`decoder.feat_extractor(speech).squeeze(dim=0).transpose(0, 1)` with `speech` a [1, L] waveform
(whisper_encoder_decoder.py:191-195).  Here the transform runs on the device (csrc/mel_kernels.cu: reflect pad, windowed
DFT as an fp32 GEMM, magnitude, filterbank GEMM, log-clamp); this module builds the two constant operand tables on the
host -- the hann-windowed DFT basis and the Slaney mel filterbank that `librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax)`
(htk=False, norm='slaney'; librosa pinned 0.10.2 / 0.11.0 by the reference's requirements) produces.
"""
import math

import numpy as np
import torch

N_FFT, HOP, WIN, N_MELS, SR, FMIN, FMAX = 1920, 480, 1920, 80, 24000, 0.0, 8000.0
N_BINS = N_FFT // 2 + 1           # 961
LD_HALF = 1024                    # padded bin count: the spectrum is stored as [re (1024) | im (1024)]


def _hz_to_mel(f):
    """Slaney / Auditory-Toolbox mel scale (librosa.hz_to_mel, htk=False): linear below 1 kHz, logarithmic above."""
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, math.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-10) / min_log_hz) / logstep, f / f_sp)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, math.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def slaney_mel_filterbank(sr=SR, n_fft=N_FFT, n_mels=N_MELS, fmin=FMIN, fmax=FMAX):
    """[n_mels, 1 + n_fft // 2] float32: triangular filters with centres equally spaced on the Slaney mel scale between
    fmin and fmax, each normalised to unit area in Hz (norm='slaney': 2 / (f[i + 2] - f[i]))."""
    fftfreqs = np.linspace(0.0, sr / 2.0, 1 + n_fft // 2)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    lower = -ramps[:-2] / fdiff[:-1, None]
    upper = ramps[2:] / fdiff[1:, None]
    w = np.maximum(0.0, np.minimum(lower, upper))
    w *= (2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels]))[:, None]
    return w.astype(np.float32)


def pack_mel_tensors():
    """Engine tensors: "mel.dft" [1920, 2048] = [hann * cos | -hann * sin] (torch.stft with the periodic hann window of
    torch.hann_window(1920), audio.py:57,65-71), "mel.fb" [1024, 128] = filterbank^T zero-padded."""
    n = np.arange(N_FFT, dtype=np.float64)
    hann = 0.5 - 0.5 * np.cos(2.0 * np.pi * n / WIN)                       # periodic window
    k = np.arange(N_BINS, dtype=np.float64)
    ang = 2.0 * np.pi * np.outer(n, k) / N_FFT
    dft = np.zeros((N_FFT, 2 * LD_HALF), dtype=np.float32)
    dft[:, :N_BINS] = (hann[:, None] * np.cos(ang)).astype(np.float32)
    dft[:, LD_HALF:LD_HALF + N_BINS] = (-hann[:, None] * np.sin(ang)).astype(np.float32)
    fb = np.zeros((LD_HALF, 128), dtype=np.float32)
    fb[:N_BINS, :N_MELS] = slaney_mel_filterbank().T
    return {"mel.dft": torch.from_numpy(dft), "mel.fb": torch.from_numpy(fb)}


class MelExtractor:
    """`feat_extractor(speech)`: speech float [1, L] (or [L]) in [-1, 1] at 24 kHz -> log-mel [1, 80, L // 480], on the
    device of the input (the reference's callers pass CPU tensors and move the result later)."""

    def __init__(self, engine, device):
        self.engine, self.device = engine, device

    @torch.inference_mode()
    def __call__(self, speech):
        y = speech.reshape(1, -1) if speech.dim() == 1 else speech
        if y.dim() != 2 or y.shape[0] != 1:
            raise ValueError("feat_extractor expects one waveform [1, L]")
        if float(y.min()) < -1.0 or float(y.max()) > 1.0:                  # audio.py:46-49 prints; keep the hint
            print(f"feat_extractor: waveform outside [-1, 1] (min {float(y.min()):.3f}, max {float(y.max()):.3f})")
        mel = self.engine.mel_spectrogram(y[0].to(self.device, torch.float32))
        return mel.unsqueeze(0).to(speech.device)

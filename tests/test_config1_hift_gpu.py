"""GPU parity at the lengths that are benchmarked, for the vocoder and the whole `token2wav` call.

  * BASELINE.json config 1 (one 10 s utterance, 125 tokens, no prompt) through `AudioDecoder.token2wav` against the golden
    of the unmodified reference's `AudioDecoder.token2wav` (CPU fp32): mel at T = 500 and the phase-blind waveform distance;
  * HiFT at T' = 500 frames on the REFERENCE's mel (not randn) with the reference's NSF noise: f0 error, source SNR,
    waveform SNR (whole utterance and per second), band distance -- the length-dependent bounds are stated below;
  * HiFT at T' = 3000 frames (config 4, 60 s) against the oracle run on this box's host cores.

Length dependence: the NSF source integrates the predicted f0 into the sine phase (generator.py:275-283), so an f0 BIAS
of df Hz over a voiced stretch of D seconds shifts the k-th harmonic by 2 pi k df D radians.  The tensor cores accumulate
with truncation, which biased a split-bf16 f0 predictor by -1.9e-5 relative and let the waveform SNR fall from 30 dB in
the first second to 10 dB in the tenth (profiles/r02_f0_bias.txt).  The predictor therefore runs in fp32 with
round-to-nearest FMAs (csrc/f0_fp32.cu): f0 is unbiased at the 1e-6 level and the SNR no longer depends on length --
measured 53 / 54 dB (source) and 41 / 41 dB (waveform) at 10 s / 60 s, flat per second.  Bounds, for ANY length up to the
60 s of config 4: f0 rtol 5e-5 (atol 5e-4 Hz), source SNR >= 45 dB, waveform SNR >= 36 dB whole and >= 34 dB in every
window, band distance <= 0.3 dB.
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from parity_metrics import band_db, rel_l2, snr_db, windowed_snr_db

pytestmark = pytest.mark.gpu

MEL_REL_L2 = 1.5e-2
MEL_RTOL, MEL_ATOL = 5e-2, 0.15
BAND_DB = 3.0            # end to end through a bf16 mel (HiFT is ill-conditioned in its mel input, see test_e2e_gpu.py)
F0_RTOL, F0_ATOL = 5e-5, 5e-4


@pytest.fixture(scope="module")
def decoder():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    return AudioDecoder(seed=0)


def _config1():
    from oracle import cases, rng_replay
    z = np.load(os.path.join(GOLDEN, "config1_N125.npz"))
    token, _, _, emb = cases.utterance(int(z["seed"]), 125, 0)
    noise = rng_replay.replay(int(z["noise_seed"]), [480 * 500])[0]
    if not np.allclose(cases.fingerprint(token.float(), emb), z["fp"], rtol=1e-6, atol=1e-6) or \
            not np.allclose(rng_replay.fingerprint(noise), z["noise_fp"], rtol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    return z, token, emb, noise


def test_token2wav_config1(decoder):
    z, token, emb, noise = _config1()
    decoder.nsf_noise_hook = lambda lens: noise.cuda()
    try:
        wav, mel = decoder.token2wav(token, "config1", embedding=emb)
    finally:
        decoder.nsf_noise_hook = None
    ref_mel, ref_wav = torch.from_numpy(z["mel"]), torch.from_numpy(z["wav"])
    assert mel.shape == (1, 80, 500) and wav.shape == (1, 240000)
    err = (mel[0].cpu() - ref_mel).abs()
    bound = MEL_ATOL + MEL_RTOL * ref_mel.abs()
    e, d = rel_l2(mel[0], ref_mel), band_db(wav, ref_wav)
    print(f"config 1 token2wav: mel rel_l2 {e:.3e}, max abs err {float(err.max()):.3e} (worst err/bound "
          f"{float((err / bound).max()):.2f}), wav band distance {d:.2f} dB")
    assert e < MEL_REL_L2
    assert bool((err <= bound).all())
    assert d < BAND_DB


def test_hift_500_frames_on_reference_mel(decoder):
    z, token, emb, noise = _config1()
    mel = torch.from_numpy(z["mel"]).cuda()
    wav, src = decoder.engine.hift([500], mel.reshape(-1), nsf_noise=noise.cuda())
    torch.cuda.synchronize()
    f0 = decoder.engine.debug_read("hift.f0", (500 + 8) * 4).reshape(-1, 4)[4:504, 0]
    f0_ref = torch.from_numpy(z["f0"])
    voiced = float((f0_ref > 10).float().mean())
    f0_err = float(((f0 - f0_ref).abs() / f0_ref.abs().clamp_min(1.0)).max())
    ref_wav, ref_src = torch.from_numpy(z["wav"]), torch.from_numpy(z["source"])
    s_src, s_wav, d = snr_db(src, ref_src), snr_db(wav, ref_wav), band_db(wav, ref_wav)
    per_s = windowed_snr_db(wav, ref_wav, 24000)
    print(f"HiFT T'=500 on the reference mel: voiced {voiced:.0%}, f0 max rel err {f0_err:.2e}, source SNR {s_src:.1f} dB, "
          f"wav SNR {s_wav:.1f} dB (per second: {', '.join(f'{v:.0f}' for v in per_s)}), band distance {d:.2f} dB")
    assert torch.allclose(f0, f0_ref, rtol=F0_RTOL, atol=F0_ATOL)
    assert voiced > 0.05, "fixture must exercise the sine branch of SineGen2"
    assert s_src > 45.0
    assert s_wav > 36.0
    assert min(per_s) > 34.0
    assert d < 0.3


def test_hift_3000_frames_vs_oracle(decoder):
    """Config 4 vocodes 3000 frames per utterance: compared with the oracle on a mel made of the reference's config-1 mel
    tiled with small seeded perturbations (log-mel statistics of the path, not randn)."""
    from moss_speech_decoder_cosy_b200 import synth
    from oracle import port
    torch.set_num_threads(os.cpu_count() or 1)
    z, token, emb, _ = _config1()
    g = torch.Generator().manual_seed(3000)
    base = torch.from_numpy(z["mel"])
    mel = torch.cat([base + 0.05 * torch.randn(80, 500, generator=g) for _ in range(6)], 1)
    noise = torch.randn(480 * 3000, 9, generator=g)
    hsd = {k: v.float() for k, v in synth.synth_hift_state_dict(0).items()}
    with torch.inference_mode():
        ref_wav, ref_src = port.hift_inference(hsd, mel, noise)
    wav, src = decoder.engine.hift([3000], mel.cuda().reshape(-1), nsf_noise=noise.cuda())
    torch.cuda.synchronize()
    s_src, s_wav, d = snr_db(src, ref_src), snr_db(wav, ref_wav), band_db(wav, ref_wav)
    per10 = windowed_snr_db(wav, ref_wav, 240000)
    print(f"HiFT T'=3000 vs oracle: source SNR {s_src:.1f} dB, wav SNR {s_wav:.1f} dB (per 10 s: "
          f"{', '.join(f'{v:.0f}' for v in per10)}), band distance {d:.2f} dB")
    assert torch.isfinite(wav).all() and wav.shape == ref_wav.shape
    assert s_src > 45.0
    assert s_wav > 36.0
    assert min(per10) > 34.0
    assert d < 0.3

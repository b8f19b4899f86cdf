"""GPU parity of the AudioDecoder mirror against the reference's REAL AudioDecoder (CPU fp32 golden, made by
oracle/make_golden_hift.py): token2wav with prompt, stream_inference with a sliding window, and the batched paths.

Tolerances: mel rel-L2 <= 1.5e-2 (bf16 storage / fp32 accumulation; PyTorch's own all-bf16 run: 4.8e-2).  The NSF phase integrates f0 over the
utterance, so a waveform compared through a bf16 mel is only phase-coherent over short spans: end-to-end waveforms are
held to a log-spectral (mel-cepstral-like) distance bound, the strict waveform SNR bound lives in test_hift_gpu.py where
both sides get the same mel.
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

MEL_REL_L2 = 1.5e-2
# End-to-end waveform bound.  HiFT is ill-conditioned in its mel input: the NSF phase integrates the predicted f0, so
# the REFERENCE ITSELF (CPU fp32), fed its own mel perturbed by 0.8 % relative noise -- the size of our bf16 mel
# error -- moves by band_db = 2.2 dB and drops to 2.5 dB waveform SNR (0.1 %: 1.3 dB / 9.7 dB; another NSF noise seed
# alone: 0.9 dB).  Measured with oracle/ on this fixture; our end-to-end output sits at 2.2 dB, i.e. at that floor.
BAND_DB = 3.0
REFMEL_SNR_DB = 36.0   # same mel + same NSF noise on both sides (fp32 f0 predictor: no phase drift with length)


from parity_metrics import band_db, rel_l2, snr_db  # noqa: E402


@pytest.fixture(scope="module")
def decoder():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    return AudioDecoder(seed=0)


def _noise_hook(seed, frames_expected):
    from oracle import rng_replay
    calls = []

    def hook(lens):
        calls.append(list(lens))
        k = len(calls) - 1
        assert list(lens) == [int(frames_expected[k])], (lens, frames_expected, k)
        # replay from the start: the generator state after call k depends on all earlier calls
        n = rng_replay.replay(seed, [480 * int(t) for t in frames_expected[:k + 1]])[k]
        return n.cuda()
    return hook


def test_token2wav_matches_reference(decoder):
    z = np.load(os.path.join(GOLDEN, "e2e_N30_P10.npz"))
    from oracle import rng_replay
    n0 = rng_replay.replay(int(z["seed"]), [480 * int(z["hift_frames"][0])])[0]
    if not np.allclose(rng_replay.fingerprint(n0), z["noise_fp"], rtol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    decoder.nsf_noise_hook = _noise_hook(int(z["seed"]), z["hift_frames"])
    try:
        wav, mel = decoder.token2wav(torch.from_numpy(z["token"]), "t", prompt_token=torch.from_numpy(z["prompt_token"]),
                                     prompt_feat=torch.from_numpy(z["prompt_feat"]),
                                     embedding=torch.from_numpy(z["embedding"]))
    finally:
        decoder.nsf_noise_hook = None
    ref_mel, ref_wav = torch.from_numpy(z["mel"]).cuda(), torch.from_numpy(z["wav"]).cuda()
    assert mel.shape == (1, 80, 120) and wav.shape == (1, 57600)
    e, d = rel_l2(mel[0], ref_mel), band_db(wav, ref_wav)
    print(f"token2wav: mel rel_l2 {e:.3e}, wav band distance {d:.2f} dB")
    assert e < MEL_REL_L2
    assert d < BAND_DB
    # vocoder alone on the reference's mel with the same injected noise: phase-coherent waveform match
    decoder.nsf_noise_hook = _noise_hook(int(z["seed"]), z["hift_frames"])
    try:
        w2, _ = decoder.hift.inference(speech_feat=torch.from_numpy(z["mel"]).unsqueeze(0))
    finally:
        decoder.nsf_noise_hook = None
    s2 = snr_db(w2, ref_wav)
    print(f"token2wav: vocoder on reference mel: wav SNR {s2:.1f} dB, band distance {band_db(w2, ref_wav):.2f} dB")
    assert s2 > REFMEL_SNR_DB


def test_stream_inference_matches_reference(decoder):
    z = np.load(os.path.join(GOLDEN, "stream_N33_P7.npz"))
    decoder.nsf_noise_hook = _noise_hook(int(z["seed"]), z["hift_frames"])
    try:
        wav = decoder.stream_inference(torch.from_numpy(z["token"]), prompt_token=torch.from_numpy(z["prompt_token"]),
                                       prompt_feat=torch.from_numpy(z["prompt_feat"]),
                                       embedding=torch.from_numpy(z["embedding"]), block_size=int(z["block_size"]),
                                       max_token_len=int(z["max_token_len"]))
    finally:
        decoder.nsf_noise_hook = None
    ref = torch.from_numpy(z["wav"])
    assert wav.shape == (1, ref.numel())
    d = band_db(wav, ref)
    print(f"stream_inference: {len(z['hift_frames'])} chunks, wav band distance {d:.2f} dB")
    assert d < BAND_DB


def test_batch_equals_single(decoder):
    """decode_batch over ragged utterances with prompts == token2wav per utterance (same device RNG seed)."""
    g = torch.Generator().manual_seed(3)
    toks = [torch.randint(0, 16384, (1, n), generator=g) for n in (12, 25, 7)]
    ptoks = [torch.randint(0, 16384, (1, p), generator=g).int() for p in (5, 0, 9)]
    pfeats = [torch.randn(1, 4 * p.shape[1], 80, generator=g) - 4 for p in ptoks]
    emb = torch.randn(3, 192, generator=g)
    res = decoder.token2wav_batch(toks, ptoks, pfeats, emb, seed=5)
    for b in range(3):
        r = decoder.flow_batch([toks[b]], [ptoks[b]], [pfeats[b]], emb[b:b + 1], streaming=False, finalize=True)
        mel1 = r["mel"][r["prompt_frames"][0]:]
        mel_b = res[b][1][0].t()
        assert rel_l2(mel_b, mel1) < 2e-2, rel_l2(mel_b, mel1)   # bf16 encoder: padded vs unpadded GEMM shapes
        assert res[b][0].shape == (1, 1920 * toks[b].shape[1])


def test_stream_scheduler_matches_single_stream(decoder):
    from moss_speech_decoder_cosy_b200.streaming import StreamScheduler
    g = torch.Generator().manual_seed(4)
    sched = StreamScheduler(decoder, block_size=5, max_token_len=20)
    items = []
    for n, p in ((23, 7), (31, 4)):
        tok = torch.randint(0, 16384, (1, n), generator=g)
        pt = torch.randint(0, 16384, (1, p), generator=g).int()
        pf = torch.randn(1, 4 * p, 80, generator=g) - 4
        e = torch.randn(1, 192, generator=g)
        items.append((tok, pt, pf, e))
        sched.add(tok.reshape(-1).tolist(), pt, pf, e)
    outs = sched.run()
    for (tok, pt, pf, e), o in zip(items, outs):
        single = decoder.stream_inference(tok, pt, pf, e, block_size=5, max_token_len=20)
        assert o.shape == single.shape == (1, 1920 * tok.shape[1])
        assert band_db(o.cpu(), single) < BAND_DB

"""CPU: the oracle restatement (oracle/port.py) against golden outputs of the UNMODIFIED reference (tests/golden,
produced from /root/reference by oracle/make_golden*.py).  This pins the oracle that the GPU box uses as checker and as
timed CPU baseline.  fp32 on both sides: tolerances are round-off level."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_speech_decoder_cosy_b200 import synth
from oracle import port, rng_replay


@pytest.fixture(scope="module")
def sds():
    return synth.synth_flow_state_dict(0), synth.synth_hift_state_dict(0)


def _rel(a, b):
    return float(torch.linalg.norm(a.double() - b.double()) / torch.linalg.norm(b.double()))


@pytest.mark.parametrize("mode", ["full", "chunk"])
def test_estimator_restatement(sds, mode):
    z = np.load(os.path.join(GOLDEN, "estimator_T36.npz"))
    t = {k: torch.from_numpy(z[k]) for k in ("x", "mu", "cond", "spks", "t")}
    with torch.inference_mode():
        out = port.estimator(sds[0], t["x"], torch.ones(2, 1, 36), t["mu"], t["t"], t["spks"], t["cond"],
                             chunk=int(z["chunk"]) if mode == "chunk" else 0)
    ref = torch.from_numpy(z["out_full" if mode == "full" else "out_chunk"])
    torch.testing.assert_close(out, ref, rtol=1e-2, atol=1e-4)   # the reference's own bar, export_onnx.py:105
    assert _rel(out, ref) < 1e-5


def test_euler_solve_restatement(sds):
    z = np.load(os.path.join(GOLDEN, "solve_T64.npz"))
    with torch.inference_mode():
        mel = port.solve_euler(sds[0], torch.from_numpy(z["mu"]), torch.from_numpy(z["spks"]),
                               torch.from_numpy(z["cond"]), torch.from_numpy(z["noise"]))
    assert _rel(mel, torch.from_numpy(z["mel"])) < 1e-4
    assert torch.equal(port.noise_bank()[:, :64], torch.from_numpy(z["noise"]))


@pytest.mark.parametrize("name", ["hift_T40", "hift_T28_cache"])
def test_hift_restatement(sds, name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    mel, noise = torch.from_numpy(z["mel"]), torch.from_numpy(z["noise"])
    with torch.inference_mode():
        f0 = port.f0_predictor(sds[1], mel.unsqueeze(0))[0]
        wav, src = port.hift_inference(sds[1], mel, noise, torch.from_numpy(z["cache"]))
    torch.testing.assert_close(f0, torch.from_numpy(z["f0"]), rtol=1e-5, atol=1e-4)
    ref_w, ref_s = torch.from_numpy(z["wav"]), torch.from_numpy(z["source"])
    snr = lambda x, r: float(10 * torch.log10(r.pow(2).sum() / (x - r).pow(2).sum().clamp_min(1e-30)))
    assert snr(src, ref_s) > 80.0
    assert snr(wav, ref_w) > 60.0


def test_token2wav_restatement(sds):
    z = np.load(os.path.join(GOLDEN, "e2e_N30_P10.npz"))
    noise = rng_replay.replay(int(z["seed"]), [480 * int(z["hift_frames"][0])])[0]
    if not np.allclose(rng_replay.fingerprint(noise), z["noise_fp"], rtol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    dec = port.PortDecoder(*sds)
    wav, mel = dec.token2wav(torch.from_numpy(z["token"]), torch.from_numpy(z["prompt_token"]),
                             torch.from_numpy(z["prompt_feat"]), torch.from_numpy(z["embedding"]), noise=noise)
    assert _rel(mel[0], torch.from_numpy(z["mel"])) < 1e-4
    ref = torch.from_numpy(z["wav"])
    assert float(10 * torch.log10(ref.pow(2).sum() / (wav[0] - ref).pow(2).sum())) > 40.0


@pytest.mark.parametrize("tag", ["a", "b"])
def test_cv1_blocks_restatement(tag):
    """Row A9: the restated Matcha Block1D / ResnetBlock1D / Downsample1D / Upsample1D against the reference's modules."""
    z = np.load(os.path.join(GOLDEN, "cv1_blocks.npz"))
    sd = synth.synth_cv1_blocks_state_dict(0)
    t = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith(tag + "_")}
    with torch.inference_mode():
        h = port.cv1_resnet_block(sd, "res_in.", t["x"], t["mask"], t["temb"])
        h2 = port.cv1_resnet_block(sd, "res_mid.", h, t["mask"], t["temb"])
        d = port.cv1_downsample(sd, "down.", h2 * t["mask"])
        u = port.cv1_upsample(sd, "up.", d)
        fb = port.cv1_block1d(h2, t["mask"], sd, "final.")
    for got, key in ((h, "res_in"), (h2, "res_mid"), (d, "down"), (u, "up"), (fb, "final")):
        assert got.shape == t[key].shape
        assert _rel(got, t[key]) < 1e-5, key


def _check_fp(z, *tensors):
    from oracle import cases
    if not np.allclose(cases.fingerprint(*tensors), z["fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")


@pytest.mark.parametrize("T", [150, 296])
@pytest.mark.parametrize("mode", ["full", "chunk"])
def test_estimator_restatement_chunk20(sds, T, mode):
    """static_chunk_size = 5 tokens = 20 mel frames (every shipped MOSS script streams with block 5): the visibility edge
    of the block-causal mask (decoder.py:439-440, mask.py:127-158) falls inside key and query tiles."""
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, f"estimator_c20_T{T}.npz"))
    c = cases.estimator_inputs(T, int(z["seed"]))
    _check_fp(z, c["x"], c["cond"], c["spks"])
    with torch.inference_mode():
        out = port.estimator(sds[0], c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"],
                             chunk=int(z["chunk"]) if mode == "chunk" else 0)
    ref = torch.from_numpy(z["out_full" if mode == "full" else "out_chunk"])
    torch.testing.assert_close(out, ref, rtol=1e-2, atol=1e-4)
    assert _rel(out, ref) < 1e-5


@pytest.mark.parametrize("mode", ["full", "chunk"])
def test_estimator_restatement_padding_mask(sds, mode):
    """Ragged key-padding mask (lens 150 / 97) through the reference's mask path (decoder.py:433-494)."""
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, "estimator_c20_mask.npz"))
    c = cases.estimator_inputs(150, int(z["seed"]), lens=[int(v) for v in z["lens"]])
    _check_fp(z, c["x"], c["cond"], c["spks"])
    with torch.inference_mode():
        out = port.estimator(sds[0], c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"],
                             chunk=int(z["chunk"]) if mode == "chunk" else 0)
    ref = torch.from_numpy(z["out_full" if mode == "full" else "out_chunk"])
    # the reference leaves garbage-free zeros on masked frames (output * mask)
    assert float(ref[1, :, 97:].abs().max()) == 0.0
    torch.testing.assert_close(out, ref, rtol=1e-2, atol=1e-4)
    assert _rel(out, ref) < 1e-5


@pytest.mark.parametrize("tag", ["a", "b"])
def test_encoder_restatement_taps(sds, tag):
    """oracle/port.py:encoder_forward against taps of the reference's UpsampleConformerEncoder: non-streaming (a) and
    streaming with a 3-token look-ahead context and 5-token chunk masks (b)."""
    z = np.load(os.path.join(GOLDEN, "encoder_taps.npz"))
    fsd = {k: v.float() for k, v in sds[0].items()}
    taps = {}
    with torch.inference_mode():
        mu = port.encoder_forward(fsd, torch.from_numpy(z[f"{tag}_tokens"]), int(z[f"{tag}_ctx"]),
                                  int(z[f"{tag}_chunk_tokens"]), taps=taps)
    for k in ("pre_lookahead", "encoders", "up_layer", "final"):
        ref = torch.from_numpy(z[f"{tag}_{k}"])
        assert taps[k].shape == ref.shape, (k, taps[k].shape, ref.shape)
        assert _rel(taps[k], ref) < 3e-6, (k, _rel(taps[k], ref))
    assert _rel(mu, torch.from_numpy(z[f"{tag}_mu"])) < 3e-6


def test_slaney_filterbank_matches_torchaudio():
    """librosa is absent offline: the restated filterbank (oracle) and the product's table builder are held against
    torchaudio's independent Slaney implementation (same published algorithm as librosa.filters.mel)."""
    import torchaudio
    from moss_speech_decoder_cosy_b200 import mel_frontend
    ref = torchaudio.functional.melscale_fbanks(961, 0.0, 8000.0, 80, 24000, norm="slaney", mel_scale="slaney").t().double()
    ours = torch.from_numpy(port.slaney_mel_filterbank(24000, 1920, 80, 0, 8000)).double()
    prod = torch.from_numpy(mel_frontend.slaney_mel_filterbank()).double()
    assert ours.shape == ref.shape == prod.shape == (80, 961)
    assert float((ours - ref).abs().max()) < 1e-7 and float((prod - ref).abs().max()) < 1e-7
    assert float(ours.sum(1).min()) > 0        # no empty filter at n_fft = 1920


@pytest.mark.parametrize("tag", ["a", "b"])
def test_mel_spectrogram_restatement(tag):
    """oracle/port.py:mel_spectrogram against the UNMODIFIED matcha mel_spectrogram (golden), incl. a length that is not a
    multiple of the hop."""
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, "mel_frontend.npz"))
    n = int(z[f"{tag}_n"])
    y = cases.prompt_waveform(700 + n, n)
    if not np.allclose(cases.fingerprint(y), z[f"{tag}_fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    with torch.inference_mode():
        mel = port.mel_spectrogram(y)[0]
    ref = torch.from_numpy(z[f"{tag}_mel"])
    assert mel.shape == ref.shape == (80, n // 480)
    assert float(ref.min()) < -8.5 and float(ref.max()) > 1.0       # the fixture spans 10 nepers: near-silence to loud frames
    torch.testing.assert_close(mel, ref, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_cv1_decoder_restatement(tag):
    """Row A9: oracle/port.py:cv1_decoder against the UNMODIFIED non-causal ConditionalDecoder (channels [256, 256])."""
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, "cv1_decoder.npz"))
    nb, nm = int(z["n_blocks"]), int(z["n_mid"])
    T, lens = int(z[f"{tag}_T"]), [int(v) for v in z[f"{tag}_lens"]]
    c = cases.estimator_inputs(T, 8800 + T, lens=lens)
    if not np.allclose(cases.fingerprint(c["x"], c["cond"], c["spks"]), z[f"{tag}_fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    sd = synth.synth_cv1_decoder_state_dict(0, nb, nm)
    with torch.inference_mode():
        out = port.cv1_decoder(sd, c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"], nb, nm)
    ref = torch.from_numpy(z[f"{tag}_out"])
    assert out.shape == ref.shape and float(ref[1, :, lens[1]:].abs().max()) == 0.0
    assert _rel(out, ref) < 1e-5

"""GPU: the caller-facing boundary of the mirror beyond token2wav -- what `GLM4Encoder.decode` touches before it reaches
the decode path (whisper_encoder_decoder.py:191-227): `feat_extractor` (matcha mel_spectrogram on the device, row N2),
`campplus_session`, `sample_rate`, and the constructor's error behaviour."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def decoder():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    return AudioDecoder(seed=0)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_feat_extractor_matches_matcha_mel_spectrogram(decoder, tag):
    """Golden = the UNMODIFIED matcha mel_spectrogram (audio.py:45-82).  fp32 DFT-as-GEMM vs torch's FFT: the log-mel is
    held to atol 2e-3 everywhere (a 0.2 % energy error) down to the quietest bins, 10 nepers below the loudest."""
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, "mel_frontend.npz"))
    n = int(z[f"{tag}_n"])
    y = cases.prompt_waveform(700 + n, n)
    if not np.allclose(cases.fingerprint(y), z[f"{tag}_fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    mel = decoder.feat_extractor(y)
    ref = torch.from_numpy(z[f"{tag}_mel"])
    assert mel.shape == (1, 80, n // 480) and mel.device == y.device
    err = (mel[0] - ref).abs()
    print(f"feat_extractor {tag}: max abs err {float(err.max()):.2e}, mean {float(err.mean()):.2e} (log-mel range "
          f"{float(ref.min()):.1f} .. {float(ref.max()):.1f})")
    torch.testing.assert_close(mel[0], ref, rtol=0, atol=2e-3)


def test_extract_speech_feat_body_runs_against_the_mirror(decoder):
    """The body of GLM4Encoder._extract_speech_feat and the 24 kHz trim of GLM4Encoder.decode
    (whisper_encoder_decoder.py:191-195, 224-227), verbatim, with the mirror as `self.audio_decoder`; then token2wav."""
    from oracle import cases, port
    speech = cases.prompt_waveform(801, 24000 * 2 + 777)
    audio_decoder = decoder
    # --- whisper_encoder_decoder.py:192-195
    speech_feat = audio_decoder.feat_extractor(speech).squeeze(dim=0).transpose(0, 1)
    speech_feat = speech_feat.unsqueeze(dim=0)
    speech_feat_len = torch.tensor([speech_feat.shape[1]], dtype=torch.int32)
    # --- whisper_encoder_decoder.py:224-227 (speech tokens come from the tokenizer side: synthetic here)
    speech_token = torch.randint(0, 16384, (1, 27), generator=torch.Generator().manual_seed(1))
    speech_token_len = torch.tensor([[speech_token.shape[-1]]])
    assert audio_decoder.sample_rate == 24000
    token_len = min(int(speech_feat.shape[1] / 4), speech_token.shape[1])
    speech_feat, speech_feat_len[:] = speech_feat[:, :4 * token_len], 4 * token_len
    speech_token, speech_token_len[:] = speech_token[:, :token_len], token_len
    assert speech_feat.shape == (1, 4 * token_len, 80)
    with torch.inference_mode():
        ref = port.mel_spectrogram(speech)[0].t()[:4 * token_len]
    torch.testing.assert_close(speech_feat[0], ref, rtol=0, atol=2e-3)
    codes = torch.randint(0, 16384, (1, 12), generator=torch.Generator().manual_seed(2))
    wav, mel = audio_decoder.token2wav(codes, "boundary", prompt_token=speech_token.to(torch.int32), prompt_feat=speech_feat,
                                       embedding=torch.randn(1, 192, generator=torch.Generator().manual_seed(3)))
    assert wav.shape == (1, 12 * 1920) and mel.shape == (1, 80, 48) and torch.isfinite(wav).all()


def test_campplus_session_and_constructor_errors(decoder, tmp_path):
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    with pytest.raises(RuntimeError, match="campplus_session is unavailable"):
        decoder.campplus_session.run(None, {})
    with pytest.raises(FileNotFoundError):
        AudioDecoder(flow_ckpt_path=str(tmp_path / "missing_flow.pt"))
    with pytest.raises(FileNotFoundError):
        AudioDecoder(config_path=str(tmp_path / "missing.yaml"))
    cfg = tmp_path / "cfg.yaml"
    cfg.write_text("chunk_size: 5\nflow:\n    static_chunk_size: !ref <chunk_size>\n    inference_cfg_rate: 0.5\n")
    d = AudioDecoder(config_path=str(cfg))
    assert d.static_chunk_size == 5 and d.estimator_chunk_frames == 20 and d.inference_cfg_rate == 0.5


def test_two_devices_in_one_process():
    """Per-device kernel setup (opt-in shared memory, constant STFT tables): an engine on a second GPU of the same
    process must produce the same waveform, also when called while another device is current."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    g = torch.Generator().manual_seed(5)
    tok = torch.randint(0, 16384, (1, 10), generator=g)
    emb = torch.randn(1, 192, generator=g)
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        d = AudioDecoder(seed=0, device=dev)
        d._seed = 0
        torch.cuda.set_device(0)
        wav, mel = d.token2wav(tok, "x", embedding=emb)
        outs.append((wav.cpu(), mel.cpu()))
    assert torch.allclose(outs[0][1], outs[1][1], atol=1e-5)
    assert torch.allclose(outs[0][0], outs[1][0], atol=1e-4)

"""GPU parity of the streaming path (rows A7 / A8) chunk by chunk, against goldens of the unmodified reference's
`stream_inference` with a spy on every `hift.inference` call (oracle/make_golden_r2.py).

A phase-blind distance on the concatenated waveform cannot see a wrong `token_offset` crop, a missing cross-fade, a stale
mel cache or a one-chunk slip.  These tests can:
  * the mel handed to the vocoder in EVERY chunk (cached 8 frames + cropped new frames) is held to the mel bound;
  * "teacher forced": with the reference's per-chunk mel and the reference's NSF noise fed to our vocoder + cross-fade +
    cache logic, every returned speech chunk is held to the waveform SNR bound, and the cached source / mel tensors are
    compared after every chunk;
  * `speed != 1` on the final call (flow_inference.py:158-160; the reference itself raises NameError there because `F` is
    never imported -- the intended F.interpolate(..., mode='linear') is checked against torch on the CPU).
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from parity_metrics import band_db, rel_l2, snr_db

pytestmark = pytest.mark.gpu

MEL_REL_L2 = 1.5e-2        # per chunk, bf16 storage / fp32 accumulation
CHUNK_SNR_DB = 36.0        # same mel + same NSF noise (measured 41 dB)
BAND_DB = 3.0

CASES = ["stream_chunks_N33_P7", "stream_chunks_c5_N33_P7", "stream_chunks_N60_P38"]


def _decoder(static_chunk_tokens):
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    dec = AudioDecoder(seed=0)
    dec.static_chunk_size = static_chunk_tokens
    dec.estimator_chunk_frames = static_chunk_tokens * 4
    return dec


@pytest.fixture(scope="module")
def decoders():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    cache = {}

    def get(static_chunk_tokens):
        if static_chunk_tokens not in cache:
            cache[static_chunk_tokens] = _decoder(static_chunk_tokens)
        return cache[static_chunk_tokens]
    return get


def _load(name):
    from oracle import cases, rng_replay
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    token, ptok, pfeat, emb = cases.utterance(int(z["seed"]), int(z["n_tokens"]), int(z["n_prompt"]))
    if not np.allclose(cases.fingerprint(token.float(), pfeat, emb), z["fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    frames = [int(v) for v in z["hift_frames"]]
    noise = rng_replay.replay(int(z["noise_seed"]), [480 * t for t in frames])
    if not np.allclose([rng_replay.fingerprint(n) for n in noise], z["noise_fp"], rtol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    return z, (token, ptok, pfeat, emb), frames, noise


@pytest.mark.parametrize("name", CASES)
def test_stream_per_chunk_mel(decoders, name):
    z, (token, ptok, pfeat, emb), frames, noise = _load(name)
    dec = decoders(int(z["static_chunk_tokens"]))
    seen = []
    orig_hift = dec.engine.hift

    def spy(lens, mel, **kw):
        seen.append((list(lens), mel.view(-1, 80).clone()))
        return orig_hift(lens, mel, **kw)
    k = [0]

    def hook(lens):
        assert list(lens) == [frames[k[0]]], (lens, frames, k[0])
        k[0] += 1
        return noise[k[0] - 1].cuda()
    dec.engine.hift, dec.nsf_noise_hook = spy, hook
    try:
        wav = dec.stream_inference(token, prompt_token=ptok, prompt_feat=pfeat, embedding=emb,
                                   block_size=int(z["block_size"]), max_token_len=int(z["max_token_len"]))
    finally:
        dec.engine.hift, dec.nsf_noise_hook = orig_hift, None
    assert len(seen) == len(frames)
    worst = 0.0
    for i, (lens, mel) in enumerate(seen):
        ref = torch.from_numpy(z[f"mel_{i}"]).t()
        assert lens == [ref.shape[0]], (i, lens, ref.shape)
        e = rel_l2(mel, ref)
        worst = max(worst, e)
        assert e < MEL_REL_L2, (name, i, e)
    ref_wav = torch.from_numpy(z["wav"])
    assert wav.shape == (1, ref_wav.numel())
    d = band_db(wav, ref_wav)
    print(f"{name}: {len(frames)} chunks, worst per-chunk mel rel_l2 {worst:.3e}, whole-waveform band distance {d:.2f} dB")
    assert d < BAND_DB


@pytest.mark.parametrize("name", CASES)
def test_stream_teacher_forced_chunks(decoders, name):
    """Reference mel + reference noise into OUR vocoder, cross-fade, tail hold-back and cache updates."""
    from moss_speech_decoder_cosy_b200 import streaming
    z, (token, ptok, pfeat, emb), frames, noise = _load(name)
    dec = decoders(int(z["static_chunk_tokens"]))
    plan = streaming.chunk_plan(int(z["n_tokens"]), int(z["n_prompt"]), int(z["block_size"]), int(z["max_token_len"]))
    assert len(plan) == len(frames)
    uuid, S = "teacher", dec.source_cache_len
    dec.hift_cache_dict[uuid] = None
    k = [0]

    def fake_flow_batch(tokens, prompt_tokens, prompt_feats, embeddings, streaming, finalize):
        i = k[0]
        mel_in = torch.from_numpy(z[f"mel_{i}"]).t()
        new = mel_in if i == 0 else mel_in[dec.mel_cache_len:]          # strip the cached frames the reference prepended
        off = plan[i][2] * 4
        assert streaming == (not plan[i][3]) and finalize == plan[i][3]
        assert tokens[0].shape[1] == plan[i][1] - plan[i][0]
        mel = torch.cat([torch.zeros(off, 80), new], 0).cuda()
        return {"mel": mel, "frames": [int(mel.shape[0])], "prompt_frames": [0]}

    def hook(lens):
        assert list(lens) == [frames[k[0]]]
        return noise[k[0]].cuda()
    orig = dec.flow_batch
    dec.flow_batch, dec.nsf_noise_hook = fake_flow_batch, hook
    toks = token.reshape(-1).tolist()
    worst = 1e9
    try:
        for i, (start, end, actual, fin) in enumerate(plan):
            sp = dec.token2wav_streaming(torch.tensor(toks[start:end]).unsqueeze(0), ptok, pfeat, actual, uuid, emb,
                                         finalize=fin, stream=not fin)
            ref = torch.from_numpy(z[f"speech_{i}"])
            assert sp.shape == (1, ref.numel()), (i, sp.shape, ref.shape)
            s = snr_db(sp, ref)
            worst = min(worst, s)
            assert s > CHUNK_SNR_DB, (name, i, s)
            if not fin:
                c = dec.hift_cache_dict[uuid]
                assert rel_l2(c["mel"][0], torch.from_numpy(z[f"mel_{i}"])[:, -dec.mel_cache_len:]) < 1e-6
                assert snr_db(c["source"], torch.from_numpy(z[f"hsrc_{i}"])[-S:]) > 45.0
                assert torch.allclose(c["source"].reshape(-1).cpu(), torch.from_numpy(z[f"cache_{i + 1}"]), atol=5e-2)
            k[0] += 1
    finally:
        dec.flow_batch, dec.nsf_noise_hook = orig, None
        del dec.hift_cache_dict[uuid]
    print(f"{name}: teacher-forced, worst per-chunk speech SNR {worst:.1f} dB over {len(plan)} chunks")


def test_finalize_with_speed(decoders):
    """speed = 1.2 on a non-streamed final call: the mel is linearly resampled to int(T / speed) frames before the vocoder."""
    from oracle import cases
    dec = decoders(25)
    token, ptok, pfeat, emb = cases.utterance(55, 24, 6)
    seen = []
    orig_hift = dec.engine.hift

    def spy(lens, mel, **kw):
        seen.append((list(lens), mel.view(-1, 80).clone()))
        return orig_hift(lens, mel, **kw)
    dec.engine.hift = spy
    try:
        dec.hift_cache_dict["sp"] = None
        base = dec.token2wav_streaming(token, ptok, pfeat, 0, "sp", emb, finalize=True, speed=1.0)
        fast = dec.token2wav_streaming(token, ptok, pfeat, 0, "sp", emb, finalize=True, speed=1.2)
    finally:
        dec.engine.hift = orig_hift
    T = 4 * 24
    assert base.shape == (1, 480 * T) and fast.shape == (1, 480 * int(T / 1.2))
    (l0, m0), (l1, m1) = seen
    assert l0 == [T] and l1 == [int(T / 1.2)]
    ref = torch.nn.functional.interpolate(m0.cpu().t().unsqueeze(0), size=int(T / 1.2), mode="linear")[0].t()
    assert torch.allclose(m1.cpu(), ref, atol=1e-4)
    with pytest.raises(AssertionError):            # flow_inference.py:159: speed change only without a streaming cache
        dec.hift_cache_dict["sp2"] = {"mel": torch.zeros(1, 80, 8), "source": torch.zeros(1, 1, 3840),
                                      "speech": torch.zeros(1, 3840)}
        dec.token2wav_streaming(token, ptok, pfeat, 0, "sp2", emb, finalize=True, speed=1.2)
    del dec.hift_cache_dict["sp2"]

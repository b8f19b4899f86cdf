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
    hift = build_ref.build_hift()
    if what & {"e2e", "stream"}:
        e2e_cases(flow, hift)
    if "hift" in what:
        hift_case(hift, "hift_T40.npz", 40, 300)
        hift_case(hift, "hift_T28_cache.npz", 28, 301, cache_len=3840)


def _spy_hift_lengths(hift):
    """Wrap hift.inference to record the number of mel frames of every vocoder call."""
    calls = []
    orig = hift.inference

    def wrapped(speech_feat, cache_source=torch.zeros(1, 1, 0)):
        calls.append(int(speech_feat.shape[2]))
        return orig(speech_feat=speech_feat, cache_source=cache_source)
    hift.inference = wrapped
    return calls


def e2e_cases(flow, hift):
    """token2wav (flow_inference.py:94-123) and stream_inference (:166-243) through the reference's REAL AudioDecoder."""
    from .make_golden import _save, _logmel_like
    from . import rng_replay
    dec = build_ref.make_audio_decoder(flow, hift)
    calls = _spy_hift_lengths(dec.hift)
    # ---- non-streaming, with prompt
    g = torch.Generator().manual_seed(400)
    N, P = 30, 10
    token = torch.randint(0, 16384, (1, N), generator=g)
    prompt_token = torch.randint(0, 16384, (1, P), generator=g).to(torch.int32)
    prompt_feat = _logmel_like(g, 4 * P)
    emb = torch.randn(1, 192, generator=g)
    seed = 401
    torch.manual_seed(seed)
    calls.clear()
    wav, mel = dec.token2wav(token, "golden-e2e", prompt_token=prompt_token, prompt_feat=prompt_feat, embedding=emb)
    noise = rng_replay.replay(seed, [480 * t for t in calls])
    _save("e2e_N30_P10.npz", token=token, prompt_token=prompt_token, prompt_feat=prompt_feat, embedding=emb, wav=wav[0],
          mel=mel[0], seed=np.int64(seed), hift_frames=np.array(calls), noise_fp=np.array(rng_replay.fingerprint(noise[0])))
    # ---- streaming: 33 tokens, P = 7, block 5, window 20 (the window slides), chunked flow masks of 25 tokens
    g = torch.Generator().manual_seed(410)
    N, P = 33, 7
    token = torch.randint(0, 16384, (1, N), generator=g)
    prompt_token = torch.randint(0, 16384, (1, P), generator=g).to(torch.int32)
    prompt_feat = _logmel_like(g, 4 * P)
    emb = torch.randn(1, 192, generator=g)
    seed = 411
    torch.manual_seed(seed)
    calls.clear()
    wav = dec.stream_inference(token, prompt_token=prompt_token, prompt_feat=prompt_feat, embedding=emb, block_size=5,
                               max_token_len=20)
    noise = rng_replay.replay(seed, [480 * t for t in calls])
    _save("stream_N33_P7.npz", token=token, prompt_token=prompt_token, prompt_feat=prompt_feat, embedding=emb,
          wav=wav[0], seed=np.int64(seed), hift_frames=np.array(calls), block_size=np.int64(5), max_token_len=np.int64(20),
          noise_fp=np.array([rng_replay.fingerprint(n) for n in noise]))

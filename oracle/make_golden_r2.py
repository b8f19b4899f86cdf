"""TEST INFRASTRUCTURE ONLY -- second set of golden vectors from the UNMODIFIED reference (CPU, fp32):

    python -m oracle.make_golden_r2 [est_c20] [stream] [config1] [encoder]

  est_c20 : estimator with static_chunk_size = 5 tokens = 20 mel frames (the mask edge falls inside a 32-key tile and
            inside a 128-row query tile; cosyvoice/flow/decoder.py:439-440, utils/mask.py:127-158) at T = 150 and 296, and a
            ragged key-padding mask (lens 150 / 97) with and without chunking.
  stream  : `stream_inference` (flow_inference.py:166-243) with a spy on every `hift.inference` call: per chunk the mel
            handed to the vocoder (cached 8 frames + new frames), its waveform / source outputs, and the speech chunk
            `token2wav_streaming` returns after the cross-fade -- the three cache tensors are slices of these.
  encoder : taps of the reference's UpsampleConformerEncoder (pins oracle/port.py:encoder_forward)
  config1 : BASELINE.json config 1 -- one 10 s utterance (125 tokens, no prompt) through `AudioDecoder.token2wav`.

Inputs are regenerated from seeds by oracle/cases.py; fixtures hold outputs + an input fingerprint.
"""
import sys

import numpy as np
import torch

from . import build_ref, cases, rng_replay
from .make_golden import _save


def gen_est_c20():
    flow = build_ref.build_flow(static_chunk_tokens=5)
    est = flow.decoder.estimator
    assert est.static_chunk_size == 20
    for T in (150, 296):
        c = cases.estimator_inputs(T, 500 + T)
        with torch.inference_mode():
            out_full = est(c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"], streaming=False)
            out_chunk = est(c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"], streaming=True)
        _save(f"estimator_c20_T{T}.npz", out_full=out_full, out_chunk=out_chunk, chunk=np.int32(20), seed=np.int64(500 + T),
              fp=cases.fingerprint(c["x"], c["cond"], c["spks"]))
    lens = [150, 97]
    c = cases.estimator_inputs(150, 777, lens=lens)
    with torch.inference_mode():
        out_full = est(c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"], streaming=False)
        out_chunk = est(c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"], streaming=True)
    _save("estimator_c20_mask.npz", out_full=out_full, out_chunk=out_chunk, chunk=np.int32(20), seed=np.int64(777),
          lens=np.asarray(lens), fp=cases.fingerprint(c["x"], c["cond"], c["spks"]))


def _spy_hift(hift):
    calls = []
    orig = hift.inference

    def wrapped(speech_feat, cache_source=torch.zeros(1, 1, 0)):
        wav, src = orig(speech_feat=speech_feat, cache_source=cache_source)
        calls.append(dict(mel=speech_feat[0].clone(), cache=cache_source.reshape(-1).clone(), wav=wav[0].clone(),
                          source=src[0, 0].clone()))
        return wav, src
    hift.inference = wrapped
    return calls


def _stream_case(name, static_chunk_tokens, N, P, block, window, seed):
    flow = build_ref.build_flow(static_chunk_tokens=static_chunk_tokens)
    hift = build_ref.build_hift()
    dec = build_ref.make_audio_decoder(flow, hift)
    calls = _spy_hift(dec.hift)
    chunks = []
    orig = dec.token2wav_streaming

    def spy_stream(*a, **kw):
        out = orig(*a, **kw)
        chunks.append(out[0].clone())
        return out
    dec.token2wav_streaming = spy_stream
    token, prompt_token, prompt_feat, emb = cases.utterance(seed, N, P)
    torch.manual_seed(seed + 1)
    wav = dec.stream_inference(token, prompt_token=prompt_token, prompt_feat=prompt_feat, embedding=emb,
                               block_size=block, max_token_len=window)
    frames = [int(c["mel"].shape[1]) for c in calls]
    noise = rng_replay.replay(seed + 1, [480 * t for t in frames])
    arrays = dict(wav=wav[0], seed=np.int64(seed), noise_seed=np.int64(seed + 1), hift_frames=np.array(frames),
                  block_size=np.int64(block), max_token_len=np.int64(window), n_tokens=np.int64(N), n_prompt=np.int64(P),
                  static_chunk_tokens=np.int64(static_chunk_tokens),
                  noise_fp=np.array([rng_replay.fingerprint(n) for n in noise]),
                  fp=cases.fingerprint(token.float(), prompt_feat, emb))
    for k, (c, sp) in enumerate(zip(calls, chunks)):
        arrays[f"mel_{k}"] = c["mel"]
        arrays[f"cache_{k}"] = c["cache"]
        arrays[f"hwav_{k}"] = c["wav"]
        arrays[f"hsrc_{k}"] = c["source"]
        arrays[f"speech_{k}"] = sp
    _save(name, **arrays)


def gen_stream():
    _stream_case("stream_chunks_N33_P7.npz", 25, 33, 7, 5, 20, 420)
    _stream_case("stream_chunks_c5_N33_P7.npz", 5, 33, 7, 5, 20, 430)
    _stream_case("stream_chunks_N60_P38.npz", 25, 60, 38, 5, 40, 440)   # BASELINE config 3 geometry: T = 296 + per chunk


def gen_config1():
    flow, hift = build_ref.build_flow(), build_ref.build_hift()
    dec = build_ref.make_audio_decoder(flow, hift)
    calls = _spy_hift(dec.hift)
    seed = 1001
    token, _, _, emb = cases.utterance(seed, 125, 0)
    torch.manual_seed(seed + 1)
    wav, mel = dec.token2wav(token, "golden-config1", embedding=emb)
    with torch.inference_mode():
        f0 = hift.f0_predictor(mel)
    noise = rng_replay.replay(seed + 1, [480 * int(mel.shape[2])])
    _save("config1_N125.npz", mel=mel[0], wav=wav[0], source=calls[0]["source"], f0=f0[0], seed=np.int64(seed),
          noise_seed=np.int64(seed + 1), noise_fp=np.array(rng_replay.fingerprint(noise[0])),
          fp=cases.fingerprint(token.float(), emb))


def gen_encoder_taps():
    """Taps of the reference's UpsampleConformerEncoder (upsample_encoder.py:244-307) behind input_embedding: after the
    pre-lookahead layer, after the 6 token-rate layers, after the up-conv, after after_norm, and mu = encoder_proj(h).
    Case a: 20 tokens, non-streaming, finalize (no context).  Case b: 23 tokens of which 3 are look-ahead context,
    streaming with static_chunk_size = 5 tokens (chunk masks at token and mel rate)."""
    arrays = {}
    for tag, L, ctx, chunk_tokens, streaming, seed in (("a", 20, 0, 25, False, 610), ("b", 23, 3, 5, True, 611)):
        flow = build_ref.build_flow(static_chunk_tokens=chunk_tokens)
        enc = flow.encoder
        taps = {}
        hooks = [enc.pre_lookahead_layer.register_forward_hook(lambda m, i, o: taps.__setitem__("pre_lookahead", o[0].clone())),
                 enc.encoders[-1].register_forward_hook(lambda m, i, o: taps.__setitem__("encoders", o[0][0].clone())),
                 enc.up_layer.register_forward_hook(lambda m, i, o: taps.__setitem__("up_layer", o[0][0].t().clone())),
                 enc.after_norm.register_forward_hook(lambda m, i, o: taps.__setitem__("final", o[0].clone()))]
        g = torch.Generator().manual_seed(seed)
        tokens = torch.randint(0, 16384, (1, L), generator=g)
        with torch.inference_mode():
            emb = flow.input_embedding(tokens)
            if ctx:
                h, _ = enc(emb[:, :-ctx], torch.tensor([L]), context=emb[:, -ctx:], streaming=streaming)
            else:
                h, _ = enc(emb, torch.tensor([L]), streaming=streaming)
            mu = flow.encoder_proj(h)
        for hk in hooks:
            hk.remove()
        arrays[f"{tag}_tokens"] = tokens[0]
        arrays[f"{tag}_ctx"] = np.int64(ctx)
        arrays[f"{tag}_chunk_tokens"] = np.int64(chunk_tokens if streaming else 0)
        arrays[f"{tag}_mu"] = mu[0]
        for k, v in taps.items():
            arrays[f"{tag}_{k}"] = v
    _save("encoder_taps.npz", **arrays)


def gen_mel_frontend():
    """The reference's feat_extractor = matcha mel_spectrogram (matcha/utils/audio.py:45-82, the UNMODIFIED function, with
    librosa.filters.mel supplied by the restated Slaney filterbank) on a seeded 2.0 s waveform and on one whose length
    is not a multiple of the hop."""
    from . import ref_shims
    ref_shims.install()
    from matcha.utils.audio import mel_spectrogram
    arrays = {}
    for tag, n in (("a", 48000), ("b", 30000 + 123)):
        y = cases.prompt_waveform(700 + n, n)
        with torch.inference_mode():
            mel = mel_spectrogram(y, 1920, 80, 24000, 480, 1920, 0, 8000, center=False)
        arrays[f"{tag}_n"] = np.int64(n)
        arrays[f"{tag}_mel"] = mel[0]
        arrays[f"{tag}_fp"] = cases.fingerprint(y)
    _save("mel_frontend.npz", **arrays)


def main(argv):
    what = set(argv) or {"est_c20", "stream", "config1", "encoder", "mel"}
    if "mel" in what:
        gen_mel_frontend()
    if "encoder" in what:
        gen_encoder_taps()
    if "est_c20" in what:
        gen_est_c20()
    if "stream" in what:
        gen_stream()
    if "config1" in what:
        gen_config1()


if __name__ == "__main__":
    main(sys.argv[1:])

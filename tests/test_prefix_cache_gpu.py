"""GPU: the frozen-prefix cache of streaming chunk calls (include/mossdec.h mossdec_token2mel_prefix).

The reference recomputes the flow over prompt + window for every streamed chunk (flow_inference.py:188-218).  Here the
first whole attention chunks inside the prompt are recorded once per stream and reused.  Checked:
  * reuse is EXACT: a reuse call on a slid window equals, bit for bit on the live frames, a recording call (full solve)
    of the same inputs in another slot;
  * a recording call equals the plain recompute path at the level of the bf16 key tiling (the plain kernel's key tiles
    straddle the prefix boundary), and the scheduler with the cache matches the REFERENCE's per-chunk mel goldens;
  * misuse fails loudly: another stream's tag, a prompt that does not cover the prefix, a non-streaming call."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from parity_metrics import band_db, rel_l2

pytestmark = pytest.mark.gpu

PREFIX = 100


@pytest.fixture(scope="module")
def decoder():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    return AudioDecoder(seed=0)


def _inputs(seed, windows, P=37):
    """One stream per entry of `windows` (window token counts incl. 3 look-ahead tokens): shared prompt, own tokens."""
    g = torch.Generator().manual_seed(seed)
    pt = torch.randint(0, 16384, (1, P), generator=g).int()
    pf = torch.clamp(torch.randn(1, 4 * P, 80, generator=g) * 2 - 5, min=-11.5129)
    toks = [torch.randint(0, 16384, (1, n), generator=g) for n in windows]
    emb = torch.randn(len(windows), 192, generator=g)
    return toks, [pt] * len(windows), [pf] * len(windows), emb


def _live(r, b):
    off = sum(r["frames"][:b])
    return r["mel"][off + PREFIX: off + r["frames"][b]]


def test_reuse_is_bit_exact_on_a_slid_window(decoder):
    dec = decoder
    dec.engine.prefix_reserve(8, PREFIX, dec.n_timesteps)
    toksA, pts, pfs, emb = _inputs(1, [43, 28])                     # windows at chunk k
    g = torch.Generator().manual_seed(2)
    toksB = [torch.cat([t[:, 5:], torch.randint(0, 16384, (1, 5), generator=g)], 1) for t in toksA]   # slid by one hop
    kw = dict(streaming=True, finalize=False)
    dec.flow_batch(toksA, pts, pfs, emb, prefix={"mode": 1, "slots": [0, 1], "tags": [11, 12]}, **kw)      # record
    rB_reuse = dec.flow_batch(toksB, pts, pfs, emb, prefix={"mode": 2, "slots": [0, 1], "tags": [11, 12]}, **kw)
    rB_record = dec.flow_batch(toksB, pts, pfs, emb, prefix={"mode": 1, "slots": [2, 3], "tags": [21, 22]}, **kw)
    rB_plain = dec.flow_batch(toksB, pts, pfs, emb, **kw)
    torch.cuda.synchronize()
    for b in range(2):
        a, c, p = _live(rB_reuse, b), _live(rB_record, b), _live(rB_plain, b)
        assert torch.isfinite(a).all()
        assert torch.equal(a, c), (b, rel_l2(a, c))                 # reuse == full recompute with the same key tiling
        e = rel_l2(c, p)
        print(f"stream {b}: prefix-tiled vs plain key tiling rel_l2 {e:.3e} over {a.shape[0]} live frames")
        assert e < 5e-3
    # a second reuse (graph replay of the geometry) stays exact
    rB2 = dec.flow_batch(toksB, pts, pfs, emb, prefix={"mode": 2, "slots": [0, 1], "tags": [11, 12]}, **kw)
    rB3 = dec.flow_batch(toksB, pts, pfs, emb, prefix={"mode": 2, "slots": [0, 1], "tags": [11, 12]}, **kw)
    for b in range(2):
        assert torch.equal(_live(rB2, b), _live(rB_record, b)) and torch.equal(_live(rB3, b), _live(rB_record, b))


def test_prefix_misuse_fails_loudly(decoder):
    dec = decoder
    dec.engine.prefix_reserve(4, PREFIX, dec.n_timesteps)
    toks, pts, pfs, emb = _inputs(3, [43])
    kw = dict(streaming=True, finalize=False)
    dec.flow_batch(toks, pts, pfs, emb, prefix={"mode": 1, "slots": [0], "tags": [5]}, **kw)
    with pytest.raises(RuntimeError, match="tag mismatch"):
        dec.flow_batch(toks, pts, pfs, emb, prefix={"mode": 2, "slots": [0], "tags": [6]}, **kw)
    with pytest.raises(RuntimeError, match="tag mismatch"):
        dec.flow_batch(toks, pts, pfs, emb, prefix={"mode": 2, "slots": [1], "tags": [5]}, **kw)      # empty slot
    short = _inputs(4, [43], P=20)                                  # 80 prompt frames < 100
    with pytest.raises(RuntimeError, match="does not cover the frozen prefix"):
        dec.flow_batch(*short, prefix={"mode": 1, "slots": [2], "tags": [7]}, **kw)
    with pytest.raises(RuntimeError, match="whole number of"):
        dec.flow_batch(toks, pts, pfs, emb, streaming=False, finalize=True, prefix={"mode": 1, "slots": [3], "tags": [8]})


def test_scheduler_with_prefix_cache_matches_reference_chunks(decoder):
    """stream_inference geometry of BASELINE config 3 (P = 38, block 5, window 40) through the batched scheduler WITH the
    prefix cache, per-chunk mel against the reference's golden (oracle/make_golden_r2.py)."""
    from oracle import cases, rng_replay
    from moss_speech_decoder_cosy_b200.streaming import StreamScheduler
    z = np.load(os.path.join(GOLDEN, "stream_chunks_N60_P38.npz"))
    token, ptok, pfeat, emb = cases.utterance(int(z["seed"]), int(z["n_tokens"]), int(z["n_prompt"]))
    if not np.allclose(cases.fingerprint(token.float(), pfeat, emb), z["fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    frames = [int(v) for v in z["hift_frames"]]
    noise = rng_replay.replay(int(z["noise_seed"]), [480 * t for t in frames])
    dec = decoder
    seen, k = [], [0]
    orig_hift = dec.engine.hift

    def spy(lens, mel, **kw):
        seen.append(mel.view(-1, 80).clone())
        return orig_hift(lens, mel, **kw)

    def hook(lens):
        k[0] += 1
        return noise[k[0] - 1].cuda()
    sched = StreamScheduler(dec, block_size=int(z["block_size"]), max_token_len=int(z["max_token_len"]), prefix_cache=True)
    sched.add(token.reshape(-1).tolist(), ptok, pfeat, emb)
    dec.engine.hift, dec.nsf_noise_hook = spy, hook
    try:
        wav = sched.run()[0]
    finally:
        dec.engine.hift, dec.nsf_noise_hook = orig_hift, None
    s = sched.streams[0]
    assert s.slot == 0 and s.recorded and sched._prefix_frames == PREFIX
    assert len(seen) == len(frames)
    worst = max(rel_l2(m, torch.from_numpy(z[f"mel_{i}"]).t()) for i, m in enumerate(seen))
    d = band_db(wav, torch.from_numpy(z["wav"]))
    print(f"scheduler + prefix cache: {len(frames)} chunks, worst per-chunk mel rel_l2 {worst:.3e} vs the reference, band distance {d:.2f} dB")
    assert worst < 1.5e-2 and d < 3.0

"""GPU parity of the native token -> mu front end (rows A2 / N1: mossdec_token2mel's encoder on gemm_tc_kernel,
attention_ws_kernel<bias> and the LayerNorm / gather kernels) against taps of the reference's UpsampleConformerEncoder
(golden) and against the oracle's own restatement at the benchmarked size.

Tolerance: mu rel-L2 <= 1.5e-2 (bf16 operands / fp32 accumulation and residual stream; 10 pre-norm layers)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from parity_metrics import rel_l2

pytestmark = pytest.mark.gpu

MU_REL_L2 = 1.5e-2


@pytest.fixture(scope="module")
def decoder():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    return AudioDecoder(seed=0)


def _run_encoder(dec, token_rows, ctx, chunk_tokens, prompt_feats=None, emb=None):
    """token_rows: list of int64 [L_b].  Runs mossdec_token2mel with 1 Euler step and reads the encoder taps back."""
    dev = dec.device
    B = len(token_rows)
    n_tok = [int(t.numel()) for t in token_rows]
    F_ = [0] * B if prompt_feats is None else [int(p.shape[0]) for p in prompt_feats]
    feat = torch.zeros(1, 80) if not sum(F_) else torch.cat([p for p in prompt_feats if p.shape[0]], 0)
    emb = torch.zeros(B, 192) if emb is None else emb
    dec.engine.token2mel(n_tok, torch.cat(token_rows).to(dev), F_, feat.to(dev), emb.to(dev), ctx, chunk_tokens,
                         chunk_tokens * 4, n_steps=1)
    torch.cuda.synchronize()
    frames = [4 * (n - ctx) for n in n_tok]
    total = sum(frames)
    mu = dec.engine.debug_read("enc.mu", total * 80).view(total, 80)
    cond = dec.engine.debug_read("enc.cond", total * 80).view(total, 80)
    spks = dec.engine.debug_read("enc.spks", B * 80).view(B, 80)
    return mu, cond, spks, frames


@pytest.mark.parametrize("tag", ["a", "b"])
def test_encoder_matches_reference_taps(decoder, tag):
    """a: 20 tokens, non-streaming; b: 20 tokens + 3 look-ahead context tokens, streaming with 5-token chunk masks."""
    z = np.load(os.path.join(GOLDEN, "encoder_taps.npz"))
    tokens = torch.from_numpy(z[f"{tag}_tokens"]).long()
    mu, _, _, frames = _run_encoder(decoder, [tokens], int(z[f"{tag}_ctx"]), int(z[f"{tag}_chunk_tokens"]))
    ref = torch.from_numpy(z[f"{tag}_mu"])
    assert mu.shape == ref.shape == (frames[0], 80)
    e = rel_l2(mu, ref)
    print(f"encoder {tag}: mu rel_l2 {e:.3e} vs the reference's UpsampleConformerEncoder + encoder_proj")
    assert e < MU_REL_L2


def test_encoder_ragged_batch_at_benchmark_size(decoder):
    """64 utterances of the config-2 shape (162 tokens -> T = 648) next to short ones in one ragged batch; checked against
    the oracle (fp32) for the first, a short and the last utterance; cond / spks exactly."""
    from moss_speech_decoder_cosy_b200 import synth
    from oracle import port
    torch.set_num_threads(os.cpu_count() or 1)
    g = torch.Generator().manual_seed(31)
    lens = [162] * 30 + [7, 1, 40] + [162] * 31
    rows = [torch.randint(0, 16384, (n,), generator=g) for n in lens]
    pf = [torch.randn(min(148, 4 * n), 80, generator=g) - 4 for n in lens]
    emb = torch.randn(len(lens), 192, generator=g)
    mu, cond, spks, frames = _run_encoder(decoder, rows, 0, 0, prompt_feats=pf, emb=emb)
    fsd = {k: v.float() for k, v in synth.synth_flow_state_dict(0).items()}
    off = np.concatenate([[0], np.cumsum(frames)])
    for b in (0, 30, 31, 32, len(lens) - 1):
        with torch.inference_mode():
            ref = port.encoder_forward(fsd, rows[b], 0, 0)
        e = rel_l2(mu[off[b]:off[b + 1]], ref)
        print(f"encoder batch utterance {b} ({lens[b]} tokens): mu rel_l2 {e:.3e}")
        assert e < MU_REL_L2, b
        c = cond[off[b]:off[b + 1]]
        assert torch.equal(c[:pf[b].shape[0]], pf[b]) and float(c[pf[b].shape[0]:].abs().sum()) == 0.0
    with torch.inference_mode():
        ref_spk = port.speaker_projection(fsd, emb)
    torch.testing.assert_close(spks, ref_spk, rtol=1e-5, atol=1e-5)


def test_encoder_streaming_context_batch(decoder):
    """Streaming chunks of several streams: 3 look-ahead context tokens each, 25-token chunk masks, ragged lengths."""
    from moss_speech_decoder_cosy_b200 import synth
    from oracle import port
    g = torch.Generator().manual_seed(32)
    lens = [77, 43, 80, 12]
    rows = [torch.randint(0, 16384, (n,), generator=g) for n in lens]
    mu, _, _, frames = _run_encoder(decoder, rows, 3, 25)
    fsd = {k: v.float() for k, v in synth.synth_flow_state_dict(0).items()}
    off = np.concatenate([[0], np.cumsum(frames)])
    for b in range(len(lens)):
        with torch.inference_mode():
            ref = port.encoder_forward(fsd, rows[b], 3, 25)
        e = rel_l2(mu[off[b]:off[b + 1]], ref)
        print(f"encoder streaming utterance {b} ({lens[b]} tokens incl. 3 context): mu rel_l2 {e:.3e}")
        assert e < MU_REL_L2, b

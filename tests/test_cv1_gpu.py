"""GPU parity, SURVEY section 8 row A9: the cosyvoice1 / Matcha non-causal building blocks (GroupNorm(8) + Mish Block1D /
ResnetBlock1D, stride-2 Downsample1D, ConvTranspose k4 s2 Upsample1D) through the C-ABI against golden outputs of the
unmodified reference modules (tests/golden/cv1_blocks.npz, CPU fp32) and, at a larger size, against the CPU oracle.

Tolerance: bf16 operands / fp32 accumulation.  One conv with bf16-rounded inputs and weights differs from fp32 by
rel-L2 ~3e-3 (2^-9 per operand, averaged over K); a resnet block chains two convs through a normalisation -> 1e-2.
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

CONV_REL_L2 = 5e-3
RESNET_REL_L2 = 1e-2


def rel_l2(a, b):
    return float(torch.linalg.norm(a.double() - b.double()) / torch.linalg.norm(b.double()))


@pytest.fixture(scope="module")
def cv1(engine):
    from moss_speech_decoder_cosy_b200 import synth, weights
    sd = synth.synth_cv1_blocks_state_dict(0)
    t = {}
    t.update(weights.pack_cv1_resnet(sd, "res_in", "cv1.res_in"))
    t.update(weights.pack_cv1_resnet(sd, "res_mid", "cv1.res_mid"))
    t.update(weights.pack_cv1_downsample(sd, "down", "cv1.down"))
    t.update(weights.pack_cv1_upsample(sd, "up", "cv1.up"))
    t.update(weights.pack_cv1_block1d(sd, "final", "cv1.final"))
    engine.load_tensors(t)
    return engine, sd


@pytest.mark.parametrize("tag", ["a", "b"])
def test_cv1_blocks_match_reference(cv1, tag):
    engine, _ = cv1
    z = np.load(os.path.join(GOLDEN, "cv1_blocks.npz"))
    t = {k[2:]: torch.from_numpy(z[k]).cuda() for k in z.files if k.startswith(tag + "_")}
    n0 = engine.launch_count()
    h = engine.cv1_resnet_block("cv1.res_in", t["x"], t["mask"], t["temb"])
    h2 = engine.cv1_resnet_block("cv1.res_mid", t["res_in"], t["mask"], t["temb"])
    d = engine.cv1_downsample("cv1.down", t["res_mid"] * t["mask"])
    u = engine.cv1_upsample("cv1.up", t["down"])
    fb = engine.cv1_block1d("cv1.final", t["res_mid"], t["mask"])
    torch.cuda.synchronize()
    assert engine.launch_count() > n0
    for got, key, tol in ((h, "res_in", RESNET_REL_L2), (h2, "res_mid", RESNET_REL_L2), (d, "down", CONV_REL_L2),
                          (u, "up", CONV_REL_L2), (fb, "final", CONV_REL_L2)):
        assert got.shape == t[key].shape, key
        assert torch.isfinite(got).all(), key
        err = rel_l2(got, t[key])
        print(f"cv1 {tag} {key}: rel_l2={err:.3e} max_abs={float((got - t[key]).abs().max()):.3e}")
        assert err < tol, key


def test_cv1_chain_matches_oracle_large(cv1):
    """down -> resnet -> up chain at config-2-like size (B2 = 16, T = 648, odd half length 324 -> 162) vs the CPU oracle."""
    from oracle import port
    engine, sd = cv1
    g = torch.Generator().manual_seed(7)
    B, T = 16, 648
    x = torch.randn(B, 320, T, generator=g)
    temb = torch.randn(B, 1024, generator=g)
    mask = torch.ones(B, 1, T)
    mask[3, :, 500:] = 0
    mask[9, :, 301:] = 0
    with torch.inference_mode():
        r1 = port.cv1_resnet_block(sd, "res_in.", x, mask, temb)
        rd = port.cv1_downsample(sd, "down.", r1 * mask)
        mask2 = mask[:, :, ::2]
        r2 = port.cv1_resnet_block(sd, "res_mid.", rd, mask2, temb)
        ru = port.cv1_upsample(sd, "up.", r2 * mask2)
    h1 = engine.cv1_resnet_block("cv1.res_in", x.cuda(), mask.cuda(), temb.cuda())
    hd = engine.cv1_downsample("cv1.down", h1 * mask.cuda())
    h2 = engine.cv1_resnet_block("cv1.res_mid", hd, mask2.contiguous().cuda(), temb.cuda())
    hu = engine.cv1_upsample("cv1.up", h2 * mask2.cuda())
    torch.cuda.synchronize()
    for got, ref, name, tol in ((h1, r1, "resnet", 1e-2), (hd, rd, "down", 1.5e-2), (h2, r2, "resnet2", 2.5e-2),
                                (hu, ru, "up", 3e-2)):
        err = rel_l2(got.cpu(), ref)
        print(f"cv1 chain {name}: rel_l2={err:.3e}")
        assert got.shape == ref.shape and err < tol, name


def test_cv1_missing_weights_fail_loudly(engine):
    x = torch.zeros(1, 256, 8, device="cuda")
    with pytest.raises(RuntimeError, match="missing"):
        engine.cv1_downsample("cv1.nope", x)


@pytest.mark.parametrize("B,T", [(1, 1), (3, 2), (2, 5), (1, 33), (5, 127)])
def test_cv1_edge_lengths_match_oracle(cv1, B, T):
    """Shortest and tile-straddling lengths (T = 1: a single frame per utterance; 127 / 33: partial 32-frame tiles; B * (T + 2)
    far below one 128-row GEMM tile) against the CPU oracle, with one fully masked tail."""
    from oracle import port
    engine, sd = cv1
    g = torch.Generator().manual_seed(1000 * B + T)
    x = torch.randn(B, 320, T, generator=g)
    temb = torch.randn(B, 1024, generator=g)
    mask = torch.ones(B, 1, T)
    if T > 2:
        mask[-1, :, T // 2:] = 0
    with torch.inference_mode():
        r = port.cv1_resnet_block(sd, "res_in.", x, mask, temb)
        f = port.cv1_block1d(r, mask, sd, "final.")
        d = port.cv1_downsample(sd, "down.", r)
        u = port.cv1_upsample(sd, "up.", r)
    xr = r.cuda()
    got = (engine.cv1_resnet_block("cv1.res_in", x.cuda(), mask.cuda(), temb.cuda()), engine.cv1_block1d("cv1.final", xr, mask.cuda()),
           engine.cv1_downsample("cv1.down", xr), engine.cv1_upsample("cv1.up", xr))
    torch.cuda.synchronize()
    for o, ref, name, tol in zip(got, (r, f, d, u), ("resnet", "block1d", "down", "up"), (1.5e-2, 1e-2, 5e-3, 5e-3)):
        assert o.shape == ref.shape, name
        assert torch.isfinite(o).all(), name
        err = rel_l2(o.cpu(), ref)
        print(f"cv1 edge B={B} T={T} {name}: rel_l2={err:.3e}")
        assert err < tol, name


# ------------------------------------------------------------------ the assembled two-level decoder (mossdec_cv1_estimator)
def _cv1_decoder_engine(n_blocks, n_mid):
    from moss_speech_decoder_cosy_b200 import lib, synth, weights
    eng = lib.Engine()
    t = weights.pack_estimator(synth.synth_flow_state_dict(0))         # the engine needs the live estimator to finalize
    t.update(weights.pack_cv1_decoder(synth.synth_cv1_decoder_state_dict(0, n_blocks, n_mid), "cv1d", n_blocks, n_mid))
    eng.load_tensors(t)
    eng.finalize()
    return eng


@pytest.mark.parametrize("tag", ["a", "b"])
def test_cv1_decoder_matches_reference(tag):
    """ConditionalDecoder.forward (non-causal, channels [256, 256], cosyvoice/flow/decoder.py:206-291): two resolutions, skip
    concatenation, GroupNorm resnets, key-padding attention, stride-2 down / ConvTranspose up -- vs the golden of the
    unmodified reference class; ragged masks, odd (77) and even (130) lengths."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, "cv1_decoder.npz"))
    nb, nm = int(z["n_blocks"]), int(z["n_mid"])
    T, lens = int(z[f"{tag}_T"]), [int(v) for v in z[f"{tag}_lens"]]
    c = cases.estimator_inputs(T, 8800 + T, lens=lens)
    if not np.allclose(cases.fingerprint(c["x"], c["cond"], c["spks"]), z[f"{tag}_fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    eng = _cv1_decoder_engine(nb, nm)
    d = {k: v.cuda() for k, v in c.items()}
    out = eng.cv1_estimator("cv1d", nb, nm, d["x"], d["mask"], d["mu"], d["t"], d["spks"], d["cond"])
    torch.cuda.synchronize()
    ref = torch.from_numpy(z[f"{tag}_out"]).cuda()
    e = float(torch.linalg.norm(out.double() - ref.double()) / torch.linalg.norm(ref.double()))
    print(f"cv1 decoder {tag} (T={T}, lens {lens}, {nb} blocks x {nm + 4} levels): rel_l2 {e:.3e}")
    assert torch.isfinite(out).all()
    assert float(out[1, :, lens[1]:].abs().max()) == 0.0
    assert e < 1.6e-2


def test_cv1_decoder_full_depth_vs_oracle():
    """The CosyVoice-300M depth (4 blocks per level, 12 mid blocks) at T = 500 against the oracle on the host cores."""
    import os
    from moss_speech_decoder_cosy_b200 import synth
    from oracle import cases, port
    torch.set_num_threads(os.cpu_count() or 1)
    nb, nm, T, lens = 4, 12, 500, [500, 333]
    c = cases.estimator_inputs(T, 8899, lens=lens)
    eng = _cv1_decoder_engine(nb, nm)
    d = {k: v.cuda() for k, v in c.items()}
    out = eng.cv1_estimator("cv1d", nb, nm, d["x"], d["mask"], d["mu"], d["t"], d["spks"], d["cond"])
    torch.cuda.synchronize()
    with torch.inference_mode():
        ref = port.cv1_decoder(synth.synth_cv1_decoder_state_dict(0, nb, nm), c["x"], c["mask"], c["mu"], c["t"], c["spks"],
                               c["cond"], nb, nm)
    e = float(torch.linalg.norm(out.double().cpu() - ref.double()) / torch.linalg.norm(ref.double()))
    print(f"cv1 decoder full depth T={T}: rel_l2 {e:.3e}")
    assert e < 1.6e-2

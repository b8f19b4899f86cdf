"""GPU parity at the sizes BASELINE.json benchmarks, through the C-ABI, against the CPU oracle (oracle/port.py, pinned to
the reference by tests/test_oracle_cpu.py) run on this box's host cores, and against goldens of the unmodified reference.

Covered code paths that the small goldens (T <= 150) never reach:
  * config 2: estimator at B2 = 128 x T = 648 (654 row tiles, six 128-row query tiles the last of which holds 8 rows,
    21 key tiles, A-stationary GEMM units) and the full 10-step CFG solve at B = 64 x T = 648 (two CFG halves on two
    streams at row0 = rows_half, CUDA-graph replay);
  * config 4: T = 3148 (99 key tiles: the lazy-rescale path of the attention kernel) with full and block-causal masks;
  * config 3: T = 296 with chunk 100 and chunk 20; chunk in {20, 100} x T in {150, 296, 648}: the visibility edge inside a
    32-key tile and inside a 128-row query tile;
  * edge lengths T in {4, 5, 127, 129, 641} (4 = the TensorRT profile minimum, cosyvoice/cli/model.py:96-101) in one
    ragged solve with padding tails; key-padding masks through mossdec_estimator.

Tolerances (bf16 storage / fp32 accumulation): estimator rel-L2 <= 1.6e-2 per call, 10-step mel rel-L2 <= 1.5e-2 and
element-wise |err| <= MEL_ATOL + MEL_RTOL * |ref|.  PyTorch's own all-bf16 run of the reference sits at 2.2e-2 / 4.8e-2.
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

EST_REL_L2 = 1.6e-2
MEL_REL_L2 = 1.5e-2
MEL_RTOL, MEL_ATOL = 5e-2, 0.15


def rel_l2(a, b):
    return float(torch.linalg.norm(a.double().cpu() - b.double().cpu()) / torch.linalg.norm(b.double().cpu()))


def assert_mel_close(got, ref, what):
    got, ref = got.float().cpu(), ref.float().cpu()
    err = (got - ref).abs()
    bound = MEL_ATOL + MEL_RTOL * ref.abs()
    r = rel_l2(got, ref)
    print(f"{what}: rel_l2 {r:.3e}, max abs err {float(err.max()):.3e}, worst err/bound {float((err / bound).max()):.2f}")
    assert r < MEL_REL_L2, what
    assert bool((err <= bound).all()), what


@pytest.fixture(scope="module")
def flow_sd():
    from moss_speech_decoder_cosy_b200 import synth
    return synth.synth_flow_state_dict(0)


def _oracle_estimator(flow_sd, c, idx, chunk):
    from oracle import port
    sel = lambda k: c[k][idx]
    with torch.inference_mode():
        return port.estimator(flow_sd, sel("x"), sel("mask"), sel("mu"), sel("t"), sel("spks"), sel("cond"), chunk=chunk)


def test_estimator_config2_size(engine, flow_sd):
    """B2 = 128 x T = 648: the shape of one estimator call of the headline benchmark (both CFG halves)."""
    from oracle import cases
    torch.set_num_threads(os.cpu_count() or 1)
    c = cases.estimator_inputs(648, 6480, B2=128)
    dev = {k: v.cuda() for k, v in c.items()}
    out = engine.estimator(dev["x"], dev["mask"], dev["mu"], dev["t"], dev["spks"], dev["cond"], chunk=0)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    idx = [0, 1, 63, 64, 100, 127]            # first / last sequences, both sides of the half boundary, one in between
    ref = _oracle_estimator(flow_sd, c, idx, 0)
    for k, b in enumerate(idx):
        e = rel_l2(out[b], ref[k])
        print(f"estimator 128 x 648, sequence {b}: rel_l2 {e:.3e}")
        assert e < EST_REL_L2, b


@pytest.mark.parametrize("chunk", [0, 100])
def test_estimator_config4_size(engine, flow_sd, chunk):
    """T = 3148 (60 s + prompt): 99 key tiles per query tile."""
    from oracle import cases
    torch.set_num_threads(os.cpu_count() or 1)
    c = cases.estimator_inputs(3148, 31480 + chunk, B2=2)
    dev = {k: v.cuda() for k, v in c.items()}
    out = engine.estimator(dev["x"], dev["mask"], dev["mu"], dev["t"], dev["spks"], dev["cond"], chunk=chunk)
    torch.cuda.synchronize()
    ref = _oracle_estimator(flow_sd, c, [0, 1], chunk)
    e = rel_l2(out, ref)
    print(f"estimator 2 x 3148 chunk {chunk}: rel_l2 {e:.3e}")
    assert torch.isfinite(out).all()
    assert e < EST_REL_L2


@pytest.mark.parametrize("T", [150, 296, 648])
@pytest.mark.parametrize("chunk", [20, 100])
def test_estimator_chunk_masks(engine, flow_sd, T, chunk):
    """Block-causal masks whose edge falls inside key tiles (32) and query tiles (128): chunk 20 = static_chunk_size 5
    tokens (the MOSS streaming scripts), chunk 100 = upstream CosyVoice2's 25 tokens."""
    from oracle import cases
    torch.set_num_threads(os.cpu_count() or 1)
    c = cases.estimator_inputs(T, 500 + T)
    dev = {k: v.cuda() for k, v in c.items()}
    out = engine.estimator(dev["x"], dev["mask"], dev["mu"], dev["t"], dev["spks"], dev["cond"], chunk=chunk)
    torch.cuda.synchronize()
    gold = os.path.join(GOLDEN, f"estimator_c20_T{T}.npz")
    if chunk == 20 and os.path.exists(gold):      # golden of the unmodified reference built with static_chunk_size = 5
        z = np.load(gold)
        if np.allclose(cases.fingerprint(c["x"], c["cond"], c["spks"]), z["fp"], rtol=1e-6, atol=1e-6):
            e = rel_l2(out, torch.from_numpy(z["out_chunk"]))
            print(f"estimator T={T} chunk 20 vs reference golden: rel_l2 {e:.3e}")
            assert e < EST_REL_L2
    ref = _oracle_estimator(flow_sd, c, [0, 1], chunk)
    e = rel_l2(out, ref)
    full = _oracle_estimator(flow_sd, c, [0, 1], 0)
    print(f"estimator T={T} chunk {chunk}: rel_l2 {e:.3e} (distance of the unmasked result: {rel_l2(full, ref):.3e})")
    assert e < EST_REL_L2
    if T > chunk:
        assert rel_l2(full, ref) > 2.5 * EST_REL_L2   # the mask matters at this size: a kernel that ignored it would fail


@pytest.mark.parametrize("chunk", [0, 20])
def test_estimator_padding_mask(engine, chunk):
    """Ragged key-padding mask through mossdec_estimator vs the golden of the unmodified reference; the output is zero
    on masked frames and may be written over x (the TensorRT contract aliases the output onto x)."""
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, "estimator_c20_mask.npz"))
    lens = [int(v) for v in z["lens"]]
    c = cases.estimator_inputs(150, int(z["seed"]), lens=lens)
    if not np.allclose(cases.fingerprint(c["x"], c["cond"], c["spks"]), z["fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    dev = {k: v.cuda() for k, v in c.items()}
    ref = torch.from_numpy(z["out_chunk" if chunk else "out_full"])
    out = engine.estimator(dev["x"], dev["mask"], dev["mu"], dev["t"], dev["spks"], dev["cond"], chunk=chunk)
    e = rel_l2(out, ref)
    print(f"estimator padding mask chunk {chunk}: rel_l2 {e:.3e}")
    assert e < EST_REL_L2
    assert float(out[1, :, lens[1]:].abs().max()) == 0.0
    x2 = dev["x"].clone()
    out2 = engine.estimator(x2, dev["mask"], dev["mu"], dev["t"], dev["spks"], dev["cond"], chunk=chunk, out_alias_x=True)
    assert out2.data_ptr() == x2.data_ptr() and torch.equal(out2, out)
    bad = dev["mask"].clone()
    bad[0, 0, 10] = 0                              # a hole: not a padding mask
    with pytest.raises(RuntimeError, match="prefix"):
        engine.estimator(dev["x"], bad, dev["mu"], dev["t"], dev["spks"], dev["cond"], chunk=chunk)


def _solve_vs_oracle(engine, flow_sd, lens, prompt_frames, chunk, seed, check):
    from oracle import cases, port
    torch.set_num_threads(os.cpu_count() or 1)
    items = cases.solve_inputs(lens, seed, prompt_frames)
    mu = torch.cat([it["mu"].reshape(-1) for it in items]).cuda()
    cond = torch.cat([it["cond"].reshape(-1) for it in items]).cuda()
    spks = torch.stack([it["spks"] for it in items]).cuda()
    mel = engine.flow_solve(lens, mu, spks, cond, n_steps=10, cfg_rate=0.7, chunk=chunk)
    torch.cuda.synchronize()
    assert torch.isfinite(mel).all()
    bank = port.noise_bank()
    off = [0]
    for T in lens:
        off.append(off[-1] + 80 * T)
    for b in check:
        it, T = items[b], lens[b]
        with torch.inference_mode():
            ref = port.solve_euler(flow_sd, it["mu"].unsqueeze(0), it["spks"].unsqueeze(0), it["cond"].unsqueeze(0), bank,
                                   chunk=chunk)[0]
        assert_mel_close(mel[off[b]:off[b + 1]].view(80, T), ref, f"solve lens={lens} chunk={chunk} utterance {b} (T={T})")
    return mel


def test_solve_edge_lengths_ragged(engine, flow_sd):
    """One ragged 10-step solve holding T = 648 next to the edge lengths (padding tails in every tile)."""
    lens = [648, 4, 641, 5, 129, 127]
    _solve_vs_oracle(engine, flow_sd, lens, [148, 0, 148, 0, 40, 0], 0, 901, check=range(len(lens)))


def test_solve_config2_batch(engine, flow_sd):
    """The headline configuration itself: B = 64 x T = 648 (148 prompt frames), CFG, 10 steps; three utterances
    (first, one from the middle, last) against the oracle.  Run twice: the second solve of a geometry replays the CUDA
    graph captured from it and must reproduce the first bit for bit."""
    lens = [648] * 64
    m1 = _solve_vs_oracle(engine, flow_sd, lens, [148] * 64, 0, 902, check=[0, 37, 63])
    r0 = engine.graph_replays()
    m2 = _solve_vs_oracle(engine, flow_sd, lens, [148] * 64, 0, 902, check=[])
    m3 = _solve_vs_oracle(engine, flow_sd, lens, [148] * 64, 0, 902, check=[])
    assert engine.graph_replays() >= r0 + 1, "the Euler loop was not replayed from a CUDA graph"
    assert torch.equal(m1, m2) and torch.equal(m1, m3)


def test_solve_config3_chunk(engine, flow_sd):
    """Streaming geometry of config 3: T = 296 with the block-causal masks (chunk 100 and chunk 20), 8 streams."""
    for chunk in (100, 20):
        _solve_vs_oracle(engine, flow_sd, [296] * 8, [148] * 8, chunk, 903 + chunk, check=[0, 7])


def test_solve_config4_length(engine, flow_sd):
    """One 60 s utterance (T = 3148) through the 10-step solve."""
    _solve_vs_oracle(engine, flow_sd, [3148, 3148], [148, 148], 0, 904, check=[1])

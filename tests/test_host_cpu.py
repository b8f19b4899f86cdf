"""CPU tests of the host logic: C-ABI surface, weight packer layouts, streaming chunk schedule, sharding (incl. a
world_size-2 gloo run)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import ROOT
from moss_speech_decoder_cosy_b200 import lib, sharding, streaming, synth, weights


@pytest.fixture(scope="module")
def shared_lib():
    if not os.path.exists(lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return ctypes.CDLL(lib.LIB_PATH)


def test_cabi_exports_every_declared_symbol(shared_lib):
    header = open(os.path.join(ROOT, "include", "mossdec.h")).read()
    declared = sorted(set(re.findall(r"\b(mossdec_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 16
    for name in declared:
        assert hasattr(shared_lib, name), name
    assert sorted(lib.EXPORTED_SYMBOLS) == declared


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError):
        lib.Engine()
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    with pytest.raises(RuntimeError):
        AudioDecoder()


def test_conv_taps_layout_is_a_shifted_gemm():
    g = torch.Generator().manual_seed(0)
    w, x = torch.randn(8, 80, 3, generator=g), torch.randn(1, 80, 20, generator=g)
    ref = F.conv1d(F.pad(x, (2, 0)), w)[0].t()                       # causal conv, [T, Cout]
    wt = weights.conv_weight_taps(w)                                  # [Cout, 3 * 128]
    rows = F.pad(x[0].t(), (0, 48, 2, 0))                             # channels-last, 128 wide, 2 zero guard rows
    a = torch.cat([rows[k:k + 20] for k in range(3)], 1)              # taps -2, -1, 0 as K-slabs
    torch.testing.assert_close(a @ wt.t(), ref, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("stride,k", weights.HIFT_UPS)
def test_convtranspose_polyphase(stride, k):
    g = torch.Generator().manual_seed(stride)
    cin, cout, T = 16, 8, 12
    w, x = torch.randn(cin, cout, k, generator=g), torch.randn(1, cin, T, generator=g)
    ref = F.conv_transpose1d(x, w, stride=stride, padding=(k - stride) // 2)[0].t()      # [stride * T, cout]
    wp = weights.convtranspose_polyphase(w, stride, (k - stride) // 2)                   # [stride * cout, 3 * cin]
    rows = F.pad(x[0].t(), (0, 0, 1, 1))
    a = torch.cat([rows[d:d + T] for d in range(3)], 1)                                  # input offsets -1, 0, +1
    out = (a @ wp.t()).reshape(T * stride, cout)
    assert ref.shape[0] == stride * T
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("k,stride,pad", [(30, 15, 7), (6, 3, 1), (1, 1, 0)])
def test_source_down_window_weight(k, stride, pad):
    g = torch.Generator().manual_seed(k)
    w, x = torch.randn(4, 18, k, generator=g), torch.randn(1, 18, 121, generator=g)
    ref = F.conv1d(x, w, stride=stride, padding=pad)[0].t()
    wp = weights.source_down_weight(w)
    rows = F.pad(x[0].t(), (0, weights.STFT_LD - 18, pad, pad + 64)).reshape(-1)        # 24-wide rows, flattened
    n = ref.shape[0]
    a = torch.stack([rows[q * stride * 24: q * stride * 24 + wp.shape[1]] for q in range(n)])
    torch.testing.assert_close(a @ wp.t(), ref, rtol=1e-4, atol=1e-4)


def test_weight_norm_fold_and_f0_fp32_layout():
    conv = torch.nn.utils.parametrizations.weight_norm(torch.nn.Conv1d(6, 5, 3))
    sd = {"c." + k: v for k, v in conv.state_dict().items()}
    torch.testing.assert_close(weights.fold_weight_norm(sd, "c"), conv.weight.detach())
    # f0 predictor weights (fp32, csrc/f0_fp32.cu): Wt [3 * Cin, Cout], k = tap * Cin + c, tap j reads row r + j - 1
    h = weights.pack_hift(synth.synth_hift_state_dict(0))
    wt = h["hift.f0.c1.wt"]
    assert wt.dtype == torch.float32 and wt.shape == (3 * 512, 512) and h["hift.f0.c0.wt"].shape == (240, 512)
    w = weights.fold_weight_norm(synth.synth_hift_state_dict(0), "f0_predictor.condnet.2").float()
    g = torch.Generator().manual_seed(0)
    x = torch.randn(1, 512, 9, generator=g)
    ref = F.conv1d(x, w, padding=1)[0].t()                                       # [T, Cout]
    rows = F.pad(x[0].t(), (0, 0, 1, 1))                                          # zero guard rows = conv padding
    a = torch.cat([rows[j:j + 9] for j in range(3)], 1)                           # taps -1, 0, +1 as K slabs
    torch.testing.assert_close(a @ wt, ref, rtol=1e-4, atol=1e-4)


def test_pack_shapes_and_schema():
    t = weights.pack_estimator(synth.synth_flow_state_dict(0))
    assert t["est.res0.c1.w"].shape == (256, 3 * 320) and t["est.res13.c1.w"].shape == (256, 3 * 512)
    assert t["est.tf0.0.qkv.w"].shape == (1536, 256) and t["est.tbias.w"].shape == (14 * 256, 1024)
    h = weights.pack_hift(synth.synth_hift_state_dict(0))
    assert h["hift.up0.w"].shape == (2048, 1536) and h["hift.sd0.w"].shape == (256, 768)
    assert h["hift.f0.cls.w"].shape == (512,) and h["hift.post.w"].shape == (18, 448)


def test_layernorm_affine_folded_into_consumer_weights():
    """pack_estimator folds norm1 into q|k|v and norm3 into ff1 (W (g n + b) = (W diag g) n + W b): with the folded
    tensors a plain (non-affine) LayerNorm reproduces the reference's affine LayerNorm + Linear."""
    sd = synth.synth_flow_state_dict(0)
    p = "decoder.estimator.mid_blocks.3.1.2."
    g = torch.Generator().manual_seed(5)
    sd = dict(sd)
    for n in ("norm1", "norm3"):   # random-init LayerNorms are the identity affine: perturb them so the fold is exercised
        sd[p + n + ".weight"] = 1 + 0.3 * torch.randn(256, generator=g)
        sd[p + n + ".bias"] = 0.2 * torch.randn(256, generator=g)
    t = weights.pack_estimator(sd)
    x = torch.randn(9, 256, generator=g) * 3 + 1
    plain = torch.nn.functional.layer_norm(x, (256,))
    ref1 = torch.nn.functional.layer_norm(x, (256,), sd[p + "norm1.weight"], sd[p + "norm1.bias"])
    want = torch.cat([ref1 @ sd[p + "attn1.to_q.weight"].t() / 8.0, ref1 @ sd[p + "attn1.to_k.weight"].t(),
                      ref1 @ sd[p + "attn1.to_v.weight"].t()], 1)
    got = plain @ t["est.tf4.2.qkv.w"].float().t() + t["est.tf4.2.qkv.b"]
    assert (got - want).abs().max() < 2e-2 * want.abs().max()            # bf16 weight rounding only
    ref3 = torch.nn.functional.layer_norm(x, (256,), sd[p + "norm3.weight"], sd[p + "norm3.bias"])
    want = ref3 @ sd[p + "ff.net.0.proj.weight"].t() + sd[p + "ff.net.0.proj.bias"]
    got = plain @ t["est.tf4.2.ff1.w"].float().t() + t["est.tf4.2.ff1.b"]
    assert (got - want).abs().max() < 2e-2 * want.abs().max()
    assert "est.tf4.2.n1.g" not in t and "est.tf4.2.n3.b" not in t


def test_tiled_f32_layout_is_a_permutation_with_coalesced_rows():
    """gemm.h tiled_f32_offset (mirrored here): the float4 of columns 4j..4j+3 of the 128 rows of a tile are contiguous."""
    def off(row, col):
        return (((row >> 7) * 64 + (col >> 2)) * 128 + (row & 127)) * 4 + (col & 3)
    M = 384
    idx = np.array([[off(r, c) for c in range(256)] for r in range(M)])
    assert sorted(idx.reshape(-1).tolist()) == list(range(M * 256))      # bijection onto [0, M * 256)
    assert (idx[1:128, 8] - idx[:127, 8] == 4).all()                      # consecutive rows: consecutive float4s
    assert idx[128, 0] == 128 * 256                                       # tiles are contiguous 128 KB blocks


def test_chunk_plan_matches_reference_schedule():
    # SURVEY 3.2 probe of the reference: 60 tokens, P = 38, block 5 -> windows of 10, 15, ..., 40, 40, ... tokens,
    # every call emits 9 600 samples, 12 calls in total
    plan = streaming.chunk_plan(60, 38, 5, 40)
    assert len(plan) == 12 and [p[3] for p in plan] == [False] * 11 + [True]
    sizes = [e - s for s, e, _, _ in plan]
    assert sizes[:7] == [10, 15, 20, 25, 30, 35, 40] and set(sizes[7:]) == {40}
    new = [(e - s) - (0 if fin else 3) - off for (s, e, off, fin) in plan]
    assert new == [7] + [5] * 10 + [3]
    # emitted samples: the first call holds back 3840, the last one releases them -> 9 600 per call
    emitted = [n * 1920 - (3840 if i == 0 else 0) + (3840 if i == len(new) - 1 else 0) for i, n in enumerate(new)]
    assert emitted == [9600] * 12
    assert sum(((e - s) - (0 if f else 3) - o) for s, e, o, f in plan) == 60
    # no window: every call re-runs the prefix (flow_inference.py:199-201)
    plan = streaming.chunk_plan(23, 0, 8, None)
    assert [p[:3] for p in plan] == [(0, 11, 0), (0, 19, 8), (0, 23, 16)]


def test_sharding_partitions():
    lens = list(np.random.RandomState(0).randint(25, 376, size=512))
    costs = [sharding.utterance_cost(n, 37) for n in lens]
    for world in (1, 2, 4, 8):
        shares = sharding.balanced_partition(costs, world)
        assert sorted(i for s in shares for i in s) == list(range(512))
        load = [sum(costs[i] for i in s) for s in shares]
        assert max(load) / (sum(load) / world) < 1.02
    b = sharding.length_buckets(shares[0], lens, max_frames=40000)
    assert sorted(i for g in b for i in g) == sorted(shares[0])
    assert all(sum(4 * lens[i] for i in g) <= 40000 or len(g) == 1 for g in b)
    assert sharding.stride_partition(10, 1, 4) == [1, 5, 9]


def test_sharded_decode_world2_gloo(tmp_path):
    """Two gloo ranks shard one utterance list with no data-path collective; the optional gather returns all shares."""
    script = tmp_path / "w.py"
    script.write_text(
        "import os, sys, torch, torch.distributed as dist\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "from moss_speech_decoder_cosy_b200 import sharding\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "lens = [30, 200, 75, 120, 41, 333, 90]\n"
        "mine = sharding.balanced_partition([sharding.utterance_cost(n) for n in lens], w)[r]\n"
        "done = [(i, lens[i] * 1920) for i in mine]          # stand-in for decoded waveform lengths\n"
        "out = [None] * w\n"
        "dist.all_gather_object(out, done)\n"
        "flat = sorted(x for s in out for x in s)\n"
        "assert [i for i, _ in flat] == list(range(len(lens))), flat\n"
        "t = torch.tensor([float(sum(n for _, n in done))]); dist.all_reduce(t, op=dist.ReduceOp.MAX)\n"
        "if r == 0: print('OK', t.item())\n"
        "dist.destroy_process_group()\n")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert res.returncode == 0 and "OK" in res.stdout, res.stdout + res.stderr


def test_mish_rational_form_matches_definition():
    """epilogue.cuh fast_mish: x * n / (n + 2) with n = u (u + 2), u = e^min(x, 20) is x * tanh(softplus(x)) (matcha
    decoder.py:56 nn.Mish); the kernel evaluates it with ex2.approx / rcp.approx, no branches."""
    x = np.linspace(-30.0, 30.0, 24001)
    u = np.exp(np.minimum(x, 20.0))
    n = u * (u + 2.0)
    fast = np.where(x > 20.0, x, x * (n / (n + 2.0)))
    want = x * np.tanh(np.log1p(np.exp(-np.abs(x))) + np.maximum(x, 0.0))
    assert np.max(np.abs(fast - want)) < 1e-9


def test_ffn_p_operand_columns_in_tensor_memory():
    """ffn_tc.cu: the epilogue part that drained accumulator columns [32 p, 32 p + 32) writes its 16 packed bf16x2 words
    over columns [32 p, 32 p + 16); K-step kk of FF2 (hidden [16 kk, 16 kk + 16)) must read exactly those words."""
    written = {}
    for part in range(4):
        for w in range(16):                       # word w holds hidden 32 part + 2 w, + 1
            written[32 * part + w] = (32 * part + 2 * w, 32 * part + 2 * w + 1)
    for kk in range(8):
        col = (kk >> 1) * 32 + (kk & 1) * 8
        hidden = [h for c in range(col, col + 8) for h in written[c]]
        assert hidden == list(range(16 * kk, 16 * kk + 16))


def test_gemm_work_units_cover_every_tile_once_and_share_the_a_tile():
    """gemm_tc.cu producer / MMA / epilogue loops: unit = `group` consecutive N tiles of one M tile (A fetched once)."""
    for tiles_m, tiles_n, group, grid in [(654, 6, 3, 148), (654, 6, 1, 148), (7, 4, 2, 5), (3, 6, 3, 148)]:
        num_tiles = tiles_m * tiles_n
        num_units = (num_tiles + group - 1) // group
        seen = []
        for cta in range(min(grid, num_units)):
            for unit in range(cta, num_units, grid):
                ms = set()
                for j in range(group):
                    tile = unit * group + j
                    if tile >= num_tiles:
                        break
                    seen.append(tile)
                    ms.add(tile // tiles_n)
                assert len(ms) == 1               # tiles_n % group == 0: one A tile per unit
        assert sorted(seen) == list(range(num_tiles))


# ---- row A9 layouts (csrc/cv1_ops.cu): the operand views the GEMM core is given, emulated in fp32 on the host -----------
def _rows_layout(x, S, g, rows_total):
    """[B, C, T] -> channels-last rows, row of (b, t) = b * S + g + t, zero elsewhere (cv1_pack_kernel)."""
    B, C, T = x.shape
    rows = torch.zeros(rows_total, C)
    for b in range(B):
        rows[b * S + g: b * S + g + T] = x[b].t()
    return rows


def _shifted(rows, shift):
    out = torch.zeros_like(rows)
    n = rows.shape[0]
    if shift >= 0:
        out[: n - shift] = rows[shift:]
    else:
        out[-shift:] = rows[: n + shift]
    return out


def test_cv1_stride2_window_view_equals_conv():
    """Downsample1D as ONE K = 3C segment over the overlapping-window view (row stride 2C, 3C columns)."""
    from moss_speech_decoder_cosy_b200 import weights
    g = torch.Generator().manual_seed(3)
    for T in (9, 12):
        B, C = 2, 64
        x, w, bias = torch.randn(B, C, T, generator=g), torch.randn(C, C, 3, generator=g), torch.randn(C, generator=g)
        ref = torch.nn.functional.conv1d(x, w, bias, stride=2, padding=1)
        T_out = (T - 1) // 2 + 1
        assert ref.shape[-1] == T_out
        S_out = T_out + 1
        S_in = 2 * S_out
        Mp = B * S_out
        flat = _rows_layout(x, S_in, 1, 2 * Mp + 2).reshape(-1)
        view = torch.as_strided(flat, (Mp, 3 * C), (2 * C, 1))
        y = view @ weights.conv_weight_taps(w).t() + bias
        for b in range(B):
            torch.testing.assert_close(y[b * S_out: b * S_out + T_out].t(), ref[b], rtol=1e-4, atol=1e-4)


def test_cv1_polyphase_convtranspose_and_symmetric_taps():
    from moss_speech_decoder_cosy_b200 import weights
    g = torch.Generator().manual_seed(4)
    B, C, T = 2, 64, 11
    x = torch.randn(B, C, T, generator=g)
    S, Mp = T + 2, B * (T + 2)
    rows = _rows_layout(x, S, 1, Mp)
    taps = torch.cat([_shifted(rows, -1), rows, _shifted(rows, 1)], dim=1)   # segments with row_shift -1 / 0 / +1
    # ConvTranspose1d(C, C, 4, 2, 1): N = 2C, output row q holds frames 2q | 2q + 1
    wt, bt = torch.randn(C, C, 4, generator=g), torch.randn(C, generator=g)
    ref = torch.nn.functional.conv_transpose1d(x, wt, bt, stride=2, padding=1)
    y = taps @ weights.convtranspose_polyphase(wt, 2, 1).t() + bt.repeat(2)
    for b in range(B):
        got = y[b * S + 1: b * S + 1 + T].reshape(T, 2, C).reshape(2 * T, C).t()
        torch.testing.assert_close(got, ref[b], rtol=1e-4, atol=1e-4)
    # symmetric k3 conv (Block1D)
    w3, b3 = torch.randn(C, C, 3, generator=g), torch.randn(C, generator=g)
    ref3 = torch.nn.functional.conv1d(x, w3, b3, padding=1)
    y3 = taps @ weights.conv_weight_taps(w3).t() + b3
    for b in range(B):
        torch.testing.assert_close(y3[b * S + 1: b * S + 1 + T].t(), ref3[b], rtol=1e-4, atol=1e-4)


def test_cv1_packers_shapes():
    from moss_speech_decoder_cosy_b200 import synth, weights
    sd = synth.synth_cv1_blocks_state_dict(0)
    t = weights.pack_cv1_resnet(sd, "res_in", "cv1.res_in")
    assert t["cv1.res_in.c1.w"].shape == (256, 3 * 320) and t["cv1.res_in.c1.w"].dtype == torch.bfloat16
    assert t["cv1.res_in.res.w"].shape == (256, 320) and t["cv1.res_in.mlp.w"].shape == (256, 1024)
    assert weights.pack_cv1_downsample(sd, "down", "d")["d.w"].shape == (256, 768)
    u = weights.pack_cv1_upsample(sd, "up", "u")
    assert u["u.w"].shape == (512, 768) and u["u.b"].shape == (512,)

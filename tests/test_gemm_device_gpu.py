"""GPU: the standalone device check of the tcgen05 implicit-GEMM kernel (csrc/tests/test_gemm.cu) -- every epilogue mode of
gemm_tc_kernel against a host fp64 reference on shapes the decode path uses (K-slab segments with row shifts, ragged M / N,
LayerNorm / Mish / Snake epilogues, the tiled fp32 stream, the vocoder residual epilogue with its TMA-staged fp32 tiles).
The binary is built in-tree by __graft_entry__.build() / `make all` and travels to the GPU box with the snapshot."""
import os
import subprocess

import pytest
import torch

pytestmark = pytest.mark.gpu

BINARY = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "moss_speech_decoder_cosy_b200", "lib", "test_gemm")


def test_gemm_kernel_matches_host_reference():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    assert os.path.exists(BINARY), "lib/test_gemm is missing: run __graft_entry__.build()"
    r = subprocess.run([BINARY], capture_output=True, text=True, timeout=600)
    tail = "\n".join(r.stdout.strip().splitlines()[-12:])
    print(tail)
    assert r.returncode == 0 and "ALL TESTS PASSED" in r.stdout, tail + r.stderr[-2000:]
    n_pass = r.stdout.count(" PASS")
    assert n_pass >= 18 and "FAIL" not in r.stdout, r.stdout[-3000:]
    for case in ("vocres_n64_snake", "vocres_n128_accum", "vocres_n64_f32only", "n64_snake_tma_store", "conv3_ln_mish_res_ln2_lnact"):
        assert f"[{case}]" in r.stdout, case


FFN_BINARY = os.path.join(os.path.dirname(BINARY), "test_ffn")


def test_fused_transformer_tail_matches_host_reference():
    """csrc/tests/test_ffn.cu: ffn_tc_kernel (LayerNorm + FF1 + exact-erf GELU + FF2 + residual, 2-CTA multicast clusters) against
    a host fp64 reference, single tiles, ragged tile counts and the config-2 size."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    assert os.path.exists(FFN_BINARY), "lib/test_ffn is missing: run __graft_entry__.build()"
    r = subprocess.run([FFN_BINARY], capture_output=True, text=True, timeout=600)
    tail = "\n".join(r.stdout.strip().splitlines()[-8:])
    print(tail)
    assert r.returncode == 0 and "ALL OK" in r.stdout and "FAILED" not in r.stdout, tail + r.stderr[-2000:]

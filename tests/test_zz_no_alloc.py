"""GPU: mossdec_mel_spectrogram_static, the allocation-free variant of the prompt mel front end (include/mossdec.h).

Not a reference feature: added for callers that cannot call cudaMalloc.  Its workspace
is a file-scope `static __device__` array in csrc/mel_kernels.cu (<= 512 frames, 7.2 MiB per GPU).  Checked through the C-ABI:
  * two calls in a row (a long input, then a shorter one over the stale scratch) equal mossdec_mel_spectrogram -- same kernels,
    so bit for bit, which is inside the atol = 2e-3 the front-end test uses against matcha (test_boundary_gpu.py);
  * the committed matcha golden is met by the static path itself;
  * free device memory does not move across the calls (nothing is allocated), the maximum size works, one frame more is
    rejected with a clear error and leaves the output untouched.
The file name sorts last on purpose: a failure here must not hide other tests."""
import ctypes
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

ATOL = 2e-3                      # log-mel tolerance of tests/test_boundary_gpu.py::test_feat_extractor_matches_matcha_mel_spectrogram
MAX_FRAMES = 512                 # MOSSDEC_MEL_STATIC_MAX_FRAMES


@pytest.fixture(scope="module")
def mel_engine():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200 import lib, mel_frontend
    eng = lib.Engine()
    eng.load_tensors(mel_frontend.pack_mel_tensors())
    return eng


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _both(eng, wav, out_static, out_plain):
    """The two entry points through the C-ABI on the same stream; outputs are preallocated by the caller."""
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    L = wav.numel()
    Ts = eng.lib.mossdec_mel_spectrogram_static(eng.h, _ptr(wav), L, _ptr(out_static), st)
    assert Ts == L // 480, eng.lib.mossdec_last_error(eng.h).decode()
    Tp = eng.lib.mossdec_mel_spectrogram(eng.h, _ptr(wav), L, _ptr(out_plain), st)
    assert Tp == Ts
    torch.cuda.synchronize()


def test_static_entry_point_is_exported_and_declared():
    from moss_speech_decoder_cosy_b200 import lib
    header = open(os.path.join(os.path.dirname(GOLDEN), "..", "include", "mossdec.h")).read()
    assert "mossdec_mel_spectrogram_static" in header and "MOSSDEC_MEL_STATIC_MAX_FRAMES 512" in header
    assert "mossdec_mel_spectrogram_static" in lib.EXPORTED_SYMBOLS
    assert hasattr(lib.load_library(), "mossdec_mel_spectrogram_static")


def test_static_matches_existing_entry_point_on_two_calls_in_a_row(mel_engine):
    eng = mel_engine
    g = torch.Generator().manual_seed(11)
    # call 1: 7.3 s; call 2: 2.1 s with a ragged tail (T = 210 is not a multiple of the 128-row GEMM tile) over stale scratch
    lengths = [350 * 480 + 17, 210 * 480 + 479]
    wavs = [(torch.rand(n, generator=g) * 2 - 1).mul_(0.8).cuda() for n in lengths]
    outs = [(torch.full((80, n // 480), 7.0, device="cuda"), torch.full((80, n // 480), -7.0, device="cuda")) for n in lengths]
    torch.cuda.synchronize()
    for wav, (o_s, o_p) in zip(wavs, outs):
        _both(eng, wav, o_s, o_p)
    for n, (o_s, o_p) in zip(lengths, outs):
        assert torch.isfinite(o_s).all()
        err = float((o_s - o_p).abs().max())
        print(f"static vs growable workspace, {n} samples ({n // 480} frames): max abs diff {err:.1e}")
        torch.testing.assert_close(o_s, o_p, rtol=0, atol=ATOL)
        assert torch.equal(o_s, o_p)          # same kernels in the same order: identical
    # the same two inputs again, in the other order: the result does not depend on what the scratch held
    again = [torch.empty_like(o[0]) for o in outs]
    for k in (1, 0):
        _both(eng, wavs[k], again[k], torch.empty_like(again[k]))
    assert torch.equal(again[0], outs[0][0]) and torch.equal(again[1], outs[1][0])


@pytest.mark.parametrize("tag", ["a", "b"])
def test_static_meets_the_matcha_golden(mel_engine, tag):
    """The golden of the UNMODIFIED matcha mel_spectrogram (tests/golden/mel_frontend.npz), same bound as the existing path."""
    from oracle import cases
    z = np.load(os.path.join(GOLDEN, "mel_frontend.npz"))
    n = int(z[f"{tag}_n"])
    y = cases.prompt_waveform(700 + n, n)
    if not np.allclose(cases.fingerprint(y), z[f"{tag}_fp"], rtol=1e-6, atol=1e-6):
        pytest.skip("torch CPU generator differs from the build that produced the fixture")
    y = y.float().reshape(-1).cuda().contiguous()
    ref = torch.from_numpy(z[f"{tag}_mel"]).cuda()
    out = torch.empty(80, n // 480, device="cuda")
    mel_engine.mel_spectrogram_static(y, out)
    torch.cuda.synchronize()
    torch.testing.assert_close(out, ref, rtol=0, atol=ATOL)


def test_static_allocates_nothing_and_enforces_its_limit(mel_engine):
    eng = mel_engine
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator().manual_seed(12)
    n_max = MAX_FRAMES * 480 + 479                                   # the largest accepted input
    big = (torch.rand(n_max + 1, generator=g) * 2 - 1).cuda()
    out = torch.empty(80, MAX_FRAMES + 1, device="cuda")
    ref = torch.empty(80, MAX_FRAMES, device="cuda")
    small = big[: 40 * 480].contiguous()
    # warm-up (module load, symbol address), then nothing may allocate: free device memory stays where it is
    assert eng.lib.mossdec_mel_spectrogram_static(eng.h, _ptr(small), small.numel(), _ptr(out), st) == 40
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for n in (n_max, 40 * 480, 300 * 480 + 5, n_max):
        assert eng.lib.mossdec_mel_spectrogram_static(eng.h, _ptr(big), n, _ptr(out), st) == n // 480
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the static entry point changed the amount of free device memory"
    got = out.reshape(-1)[: 80 * MAX_FRAMES].reshape(80, MAX_FRAMES).clone()
    assert eng.lib.mossdec_mel_spectrogram(eng.h, _ptr(big), n_max, _ptr(ref), st) == MAX_FRAMES
    torch.cuda.synchronize()
    assert torch.equal(got, ref)
    # one frame more: rejected, message names the limit, output untouched
    out.fill_(3.0)
    rc = eng.lib.mossdec_mel_spectrogram_static(eng.h, _ptr(big), n_max + 1, _ptr(out), st)
    msg = eng.lib.mossdec_last_error(eng.h).decode()
    torch.cuda.synchronize()
    assert rc < 0 and "MOSSDEC_MEL_STATIC_MAX_FRAMES" in msg and "513 frames" in msg, (rc, msg)
    assert bool((out == 3.0).all())
    with pytest.raises(RuntimeError, match="exceed the static scratch"):
        eng.mel_spectrogram_static(big, out)

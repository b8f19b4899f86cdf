"""ctypes binding of libmossdec.so (include/mossdec.h).  No fallback: a missing library or GPU raises."""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MOSSDEC_LIB") or os.path.join(_HERE, "lib", "libmossdec.so")   # MOSSDEC_LIB: A/B builds of the same library

F32, BF16, I32 = 0, 1, 2


class MossdecConfig(ctypes.Structure):
    _fields_ = [("n_feats", ctypes.c_int32), ("channels", ctypes.c_int32), ("heads", ctypes.c_int32),
                ("head_dim", ctypes.c_int32), ("n_blocks", ctypes.c_int32), ("n_mid", ctypes.c_int32),
                ("noise_len", ctypes.c_int32), ("hift_base", ctypes.c_int32), ("n_harmonics", ctypes.c_int32),
                ("sample_rate", ctypes.c_float), ("nsf_alpha", ctypes.c_float), ("nsf_sigma", ctypes.c_float),
                ("voiced_threshold", ctypes.c_float), ("lrelu_slope", ctypes.c_float),
                ("audio_limit", ctypes.c_float)]


_lib = None


def load_library():
    """dlopen libmossdec.so and declare every prototype of include/mossdec.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(the decode path has no CPU or PyTorch fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_float
    lib.mossdec_default_config.argtypes = [ctypes.POINTER(MossdecConfig)]
    lib.mossdec_default_config.restype = None
    lib.mossdec_create.argtypes = [ctypes.POINTER(MossdecConfig), ctypes.POINTER(vp)]
    lib.mossdec_create.restype = i32
    lib.mossdec_destroy.argtypes = [vp]
    lib.mossdec_destroy.restype = None
    lib.mossdec_last_error.argtypes = [vp]
    lib.mossdec_last_error.restype = ctypes.c_char_p
    lib.mossdec_set_tensor.argtypes = [vp, ctypes.c_char_p, vp, i32, i32, ctypes.POINTER(i64)]
    lib.mossdec_set_tensor.restype = i32
    lib.mossdec_finalize.argtypes = [vp]
    lib.mossdec_finalize.restype = i32
    lib.mossdec_estimator.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, i32, vp]
    lib.mossdec_estimator.restype = i32
    lib.mossdec_flow_solve.argtypes = [vp, i32, ctypes.POINTER(i32), vp, vp, vp, vp, i32, i32, f32, f32, i32, vp, i32, vp]
    lib.mossdec_flow_solve.restype = i32
    lib.mossdec_hift.argtypes = [vp, i32, ctypes.POINTER(i32), vp, vp, i32, ctypes.POINTER(i32), vp, vp, vp,
                                 ctypes.c_uint64, i32, vp]
    lib.mossdec_hift.restype = i32
    lib.mossdec_hift_skip.argtypes = [vp, i32, ctypes.POINTER(i32), ctypes.POINTER(i32), vp, vp, vp, vp, ctypes.c_uint64, vp]
    lib.mossdec_hift_skip.restype = i32
    lib.mossdec_token2mel.argtypes = [vp, i32, ctypes.POINTER(i32), vp, ctypes.POINTER(i32), vp, vp, i32, i32, i32, i32, f32, vp, vp]
    lib.mossdec_token2mel.restype = i32
    lib.mossdec_prefix_reserve.argtypes = [vp, i32, i32, i32]
    lib.mossdec_prefix_reserve.restype = i32
    lib.mossdec_token2mel_prefix.argtypes = [vp, i32, ctypes.POINTER(i32), vp, ctypes.POINTER(i32), vp, vp, i32, i32, i32, i32, f32,
                                             i32, ctypes.POINTER(i32), ctypes.POINTER(ctypes.c_uint64), vp, vp]
    lib.mossdec_token2mel_prefix.restype = i32
    lib.mossdec_fade_in_out.argtypes = [vp, vp, vp, vp, i32, vp]
    lib.mossdec_fade_in_out.restype = i32
    lib.mossdec_mel_spectrogram.argtypes = [vp, vp, i64, vp, vp]
    lib.mossdec_mel_spectrogram.restype = i32
    lib.mossdec_mel_spectrogram_static.argtypes = [vp, vp, i64, vp, vp]
    lib.mossdec_mel_spectrogram_static.restype = i32
    lib.mossdec_cv1_resnet_block.argtypes = [vp, ctypes.c_char_p, i32, i32, i32, vp, vp, vp, i32, vp, vp]
    lib.mossdec_cv1_resnet_block.restype = i32
    lib.mossdec_cv1_block1d.argtypes = [vp, ctypes.c_char_p, i32, i32, i32, vp, vp, vp, vp]
    lib.mossdec_cv1_block1d.restype = i32
    lib.mossdec_cv1_downsample.argtypes = [vp, ctypes.c_char_p, i32, i32, i32, vp, vp, vp]
    lib.mossdec_cv1_downsample.restype = i32
    lib.mossdec_cv1_upsample.argtypes = [vp, ctypes.c_char_p, i32, i32, i32, vp, vp, vp]
    lib.mossdec_cv1_upsample.restype = i32
    lib.mossdec_cv1_estimator.argtypes = [vp, ctypes.c_char_p, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.mossdec_cv1_estimator.restype = i32
    lib.mossdec_debug_read.argtypes = [vp, ctypes.c_char_p, vp, i64]
    lib.mossdec_debug_read.restype = i32
    lib.mossdec_profile.argtypes = [vp, i32]
    lib.mossdec_profile.restype = None
    lib.mossdec_profile_read.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    lib.mossdec_profile_read.restype = i32
    lib.mossdec_profile_named.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(ctypes.c_double)]
    lib.mossdec_profile_named.restype = i32
    lib.mossdec_set_hift_group_frames.argtypes = [vp, i64]
    lib.mossdec_set_hift_group_frames.restype = None
    lib.mossdec_reserve.argtypes = [vp, i64, i32, i64]
    lib.mossdec_reserve.restype = i32
    lib.mossdec_set_graph_mode.argtypes = [vp, i32]
    lib.mossdec_set_graph_mode.restype = None
    lib.mossdec_graph_replays.argtypes = [vp]
    lib.mossdec_graph_replays.restype = i64
    lib.mossdec_device.argtypes = [vp]
    lib.mossdec_device.restype = i32
    lib.mossdec_launch_count.argtypes = [vp]
    lib.mossdec_launch_count.restype = i64
    lib.mossdec_reset_launch_count.argtypes = [vp]
    lib.mossdec_reset_launch_count.restype = None
    _lib = lib
    return lib


EXPORTED_SYMBOLS = [
    "mossdec_default_config", "mossdec_create", "mossdec_destroy", "mossdec_last_error", "mossdec_set_tensor",
    "mossdec_finalize", "mossdec_estimator", "mossdec_flow_solve", "mossdec_hift", "mossdec_fade_in_out",
    "mossdec_launch_count", "mossdec_reset_launch_count", "mossdec_debug_read", "mossdec_profile",
    "mossdec_profile_read", "mossdec_set_hift_group_frames", "mossdec_cv1_resnet_block", "mossdec_cv1_downsample",
    "mossdec_cv1_upsample", "mossdec_cv1_block1d", "mossdec_reserve", "mossdec_set_graph_mode", "mossdec_graph_replays",
    "mossdec_device", "mossdec_profile_named", "mossdec_mel_spectrogram", "mossdec_mel_spectrogram_static", "mossdec_hift_skip", "mossdec_token2mel", "mossdec_prefix_reserve", "mossdec_token2mel_prefix", "mossdec_cv1_estimator",
]


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _i32_array(values):
    return (ctypes.c_int32 * len(values))(*[int(v) for v in values])


class Engine:
    """One decode engine on the current CUDA device (wraps a mossdec_handle)."""

    def __init__(self, config: MossdecConfig = None):
        if not torch.cuda.is_available():
            raise RuntimeError("mossdec needs a CUDA device (sm_100a); there is no CPU path")
        self.lib = load_library()
        self.cfg = config or self.default_config()
        h = ctypes.c_void_p()
        rc = self.lib.mossdec_create(ctypes.byref(self.cfg), ctypes.byref(h))
        if rc != 0:
            raise RuntimeError(f"mossdec_create failed ({rc})")
        self.h = h
        self._device_index = None

    @staticmethod
    def default_config():
        cfg = MossdecConfig()
        load_library().mossdec_default_config(ctypes.byref(cfg))
        return cfg

    def close(self):
        if getattr(self, "h", None):
            self.lib.mossdec_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed: {self.lib.mossdec_last_error(self.h).decode()}")

    def load_tensors(self, tensors):
        for name, t in tensors.items():
            t = t.contiguous().cpu()
            dt = {torch.float32: F32, torch.bfloat16: BF16, torch.int32: I32}[t.dtype]
            shape = (ctypes.c_int64 * t.ndim)(*t.shape)
            self._check(self.lib.mossdec_set_tensor(self.h, name.encode(), ctypes.c_void_p(t.data_ptr()), dt, t.ndim,
                                                    shape), f"set_tensor({name})")

    def finalize(self):
        self._check(self.lib.mossdec_finalize(self.h), "finalize")

    def _stream(self):
        # torch's current stream ON THE ENGINE'S DEVICE (the caller's current device may be another GPU)
        return ctypes.c_void_p(torch.cuda.current_stream(self.device_index).cuda_stream)

    def estimator(self, x, mask, mu, t, spks, cond, chunk=0, out_alias_x=False):
        """Same contract as ConditionalCFM.forward_estimator (flow_matching.py:129): fp32 CUDA tensors [B2,80,T]."""
        B2, _, T = x.shape
        x, mu = x.float().contiguous(), mu.float().contiguous()
        t = t.float().contiguous()
        spks = None if spks is None else spks.float().contiguous()
        cond = None if cond is None else cond.float().contiguous()
        mask = None if mask is None else mask.float().contiguous()
        out = x if out_alias_x else torch.empty_like(x)
        self._check(self.lib.mossdec_estimator(self.h, B2, T, _ptr(x), _ptr(mask), _ptr(mu), _ptr(t), _ptr(spks),
                                               _ptr(cond), _ptr(out), int(chunk), self._stream()), "estimator")
        return out

    def flow_solve(self, lens, mu, spks, cond, noise=None, n_steps=10, cfg_rate=0.7, temperature=1.0, chunk=0,
                   time_major=False):
        """mu / cond: ragged fp32 CUDA buffers ([80, T_b] blocks back to back); returns mel in the same layout."""
        B = len(lens)
        mu, cond, spks = mu.float().contiguous(), cond.float().contiguous(), spks.float().contiguous()
        out = torch.empty_like(mu)
        nlen = 0
        if noise is not None:
            noise = noise.float().contiguous()
            nlen = noise.shape[-1]
        self._check(self.lib.mossdec_flow_solve(self.h, B, _i32_array(lens), _ptr(mu), _ptr(spks), _ptr(cond),
                                                _ptr(noise), nlen, int(n_steps), float(cfg_rate), float(temperature),
                                                int(chunk), _ptr(out), int(bool(time_major)), self._stream()), "flow_solve")
        return out

    def hift(self, lens, mel, cache_source=None, cache_len=None, nsf_noise=None, seed=0, want_source=True,
             time_major=False):
        B = len(lens)
        mel = mel.float().contiguous()
        total = int(sum(lens)) * 480
        wav = torch.empty(total, dtype=torch.float32, device=mel.device)
        src = torch.empty(total, dtype=torch.float32, device=mel.device) if want_source else None
        cstride = 0
        clen = None
        if cache_source is not None:
            cache_source = cache_source.float().contiguous()
            cstride = cache_source.shape[-1]
            clen = _i32_array(cache_len)
        if nsf_noise is not None:
            nsf_noise = nsf_noise.float().contiguous()
        self._check(self.lib.mossdec_hift(self.h, B, _i32_array(lens), _ptr(mel), _ptr(cache_source), cstride, clen,
                                          _ptr(nsf_noise), _ptr(wav), _ptr(src), int(seed), int(bool(time_major)),
                                          self._stream()), "hift")
        return wav, src

    def hift_skip(self, lens, skip, mel, nsf_noise=None, seed=0, want_source=False):
        """Vocoder over the generated part of flow results: block b of `mel` (time-major) holds skip[b] + lens[b] frames."""
        total = int(sum(lens)) * 480
        wav = torch.empty(total, dtype=torch.float32, device=mel.device)
        src = torch.empty(total, dtype=torch.float32, device=mel.device) if want_source else None
        if nsf_noise is not None:
            nsf_noise = nsf_noise.float().contiguous()
        self._check(self.lib.mossdec_hift_skip(self.h, len(lens), _i32_array(lens), _i32_array(skip), _ptr(mel), _ptr(nsf_noise),
                                               _ptr(wav), _ptr(src), int(seed), self._stream()), "hift_skip")
        return wav, src

    def prefix_reserve(self, n_slots, prefix_frames, n_steps=10):
        """Allocate the frozen-prefix cache of streaming solves (mossdec_prefix_reserve)."""
        self._check(self.lib.mossdec_prefix_reserve(self.h, int(n_slots), int(prefix_frames), int(n_steps)), "prefix_reserve")

    def token2mel(self, n_tok, tokens, prompt_frames, prompt_feat, embedding, ctx_len, chunk_tokens, est_chunk_frames,
                  n_steps=10, cfg_rate=0.7, prefix=None):
        """CausalMaskedDiffWithXvec.inference for a ragged batch (mossdec_token2mel): tokens int64 CUDA (ragged),
        prompt_feat fp32 CUDA ragged [F_b, 80], embedding fp32 CUDA [B, 192] -> mel fp32 ragged time-major [sum T_b, 80]."""
        B = len(n_tok)
        frames = sum(4 * (int(n) - ctx_len) for n in n_tok)
        out = torch.empty(frames, 80, dtype=torch.float32, device=tokens.device)
        if prefix is not None and prefix["mode"] != 0:
            # prefix = {"mode": 1 record | 2 reuse, "slots": [B], "tags": [B]}; in reuse mode the frozen frames of `out` stay unwritten
            tags = (ctypes.c_uint64 * B)(*[int(t) & 0xFFFFFFFFFFFFFFFF for t in prefix["tags"]])
            self._check(self.lib.mossdec_token2mel_prefix(
                self.h, B, _i32_array(n_tok), _ptr(tokens), _i32_array(prompt_frames), _ptr(prompt_feat), _ptr(embedding),
                int(ctx_len), int(chunk_tokens), int(est_chunk_frames), int(n_steps), float(cfg_rate), int(prefix["mode"]),
                _i32_array(prefix["slots"]), tags, _ptr(out), self._stream()), "token2mel_prefix")
            return out
        self._check(self.lib.mossdec_token2mel(self.h, B, _i32_array(n_tok), _ptr(tokens), _i32_array(prompt_frames),
                                               _ptr(prompt_feat), _ptr(embedding), int(ctx_len), int(chunk_tokens),
                                               int(est_chunk_frames), int(n_steps), float(cfg_rate), _ptr(out), self._stream()),
                    "token2mel")
        return out

    def fade_in_out(self, speech, prev_tail, window2n):
        n = prev_tail.numel()
        self._check(self.lib.mossdec_fade_in_out(self.h, _ptr(speech), _ptr(prev_tail), _ptr(window2n), n,
                                                 self._stream()), "fade_in_out")

    def mel_spectrogram(self, wav):
        """matcha mel_spectrogram (audio.py:45-82; n_fft 1920, hop 480, 80 mels, 24 kHz): wav fp32 CUDA [L] -> [80, L // 480]."""
        wav = wav.float().contiguous()
        L = wav.numel()
        out = torch.empty(80, max(L // 480, 0), device=wav.device, dtype=torch.float32)
        T = self.lib.mossdec_mel_spectrogram(self.h, _ptr(wav), L, _ptr(out), self._stream())
        if T < 0:
            raise RuntimeError(f"mel_spectrogram failed: {self.lib.mossdec_last_error(self.h).decode()}")
        assert T == out.shape[1], (T, out.shape)
        return out

    MEL_STATIC_MAX_FRAMES = 512

    def mel_spectrogram_static(self, wav, out):
        """mossdec_mel_spectrogram_static: the same result on the library's static device scratch -- the native call allocates
        nothing, so `out` ([80, L // 480] fp32 CUDA) comes from the caller; L // 480 <= MEL_STATIC_MAX_FRAMES."""
        L = wav.numel()
        if wav.dtype != torch.float32 or not wav.is_contiguous() or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError("mel_spectrogram_static takes contiguous fp32 tensors (it does not copy or allocate)")
        if out.numel() != 80 * (L // 480):
            raise ValueError(f"out must hold 80 x {L // 480} floats")
        T = self.lib.mossdec_mel_spectrogram_static(self.h, _ptr(wav), L, _ptr(out), self._stream())
        if T < 0:
            raise RuntimeError(f"mel_spectrogram_static failed: {self.lib.mossdec_last_error(self.h).decode()}")
        return out

    # ---- cosyvoice1 / Matcha non-causal building blocks; same call shapes as the reference modules
    def cv1_resnet_block(self, prefix, x, mask, time_emb=None):
        """ResnetBlock1D.forward(x, mask, time_emb) (matcha decoder.py:56-61): x [B, Cin, T], mask [B, 1, T] -> [B, 256, T]."""
        B, Cin, T = x.shape
        x, mask = x.float().contiguous(), mask.float().contiguous()
        te = None if time_emb is None else time_emb.float().contiguous()
        out = torch.empty(B, 256, T, device=x.device, dtype=torch.float32)
        self._check(self.lib.mossdec_cv1_resnet_block(self.h, prefix.encode(), B, T, Cin, _ptr(x), _ptr(mask), _ptr(te),
                                                      0 if te is None else te.shape[1], _ptr(out), self._stream()),
                    "cv1_resnet_block")
        return out

    def cv1_block1d(self, prefix, x, mask):
        """Block1D.forward(x, mask) (matcha decoder.py:41-43): x [B, Cin, T], mask [B, 1, T] -> [B, 256, T]."""
        B, Cin, T = x.shape
        x, mask = x.float().contiguous(), mask.float().contiguous()
        out = torch.empty(B, 256, T, device=x.device, dtype=torch.float32)
        self._check(self.lib.mossdec_cv1_block1d(self.h, prefix.encode(), B, T, Cin, _ptr(x), _ptr(mask), _ptr(out),
                                                 self._stream()), "cv1_block1d")
        return out

    def cv1_downsample(self, prefix, x):
        """Downsample1D.forward (matcha decoder.py:69-70): [B, C, T] -> [B, C, (T - 1) // 2 + 1]."""
        B, C, T = x.shape
        x = x.float().contiguous()
        out = torch.empty(B, C, (T - 1) // 2 + 1, device=x.device, dtype=torch.float32)
        self._check(self.lib.mossdec_cv1_downsample(self.h, prefix.encode(), B, T, C, _ptr(x), _ptr(out), self._stream()),
                    "cv1_downsample")
        return out

    def cv1_upsample(self, prefix, x):
        """Upsample1D.forward with ConvTranspose1d(C, C, 4, 2, 1) (matcha decoder.py:147-150): [B, C, T] -> [B, C, 2 T]."""
        B, C, T = x.shape
        x = x.float().contiguous()
        out = torch.empty(B, C, 2 * T, device=x.device, dtype=torch.float32)
        self._check(self.lib.mossdec_cv1_upsample(self.h, prefix.encode(), B, T, C, _ptr(x), _ptr(out), self._stream()),
                    "cv1_upsample")
        return out

    def cv1_estimator(self, prefix, n_blocks, n_mid, x, mask, mu, t, spks=None, cond=None):
        """ConditionalDecoder.forward of the non-causal cosyvoice / cosyvoice1 decoder (flow/decoder.py:206-291), channels [256, 256]."""
        B, _, T = x.shape
        f = lambda a: None if a is None else a.float().contiguous()
        x, mask, mu, t, spks, cond = f(x), f(mask), f(mu), f(t), f(spks), f(cond)
        out = torch.empty_like(x)
        self._check(self.lib.mossdec_cv1_estimator(self.h, prefix.encode(), int(n_blocks), int(n_mid), B, T, _ptr(x), _ptr(mask),
                                                   _ptr(mu), _ptr(t), _ptr(spks), _ptr(cond), _ptr(out), self._stream()),
                    "cv1_estimator")
        return out

    def debug_read(self, name, max_floats):
        buf = torch.empty(int(max_floats), dtype=torch.float32)
        n = self.lib.mossdec_debug_read(self.h, name.encode(), ctypes.c_void_p(buf.data_ptr()), int(max_floats))
        if n < 0:
            raise RuntimeError(f"debug_read({name}) failed: {self.lib.mossdec_last_error(self.h).decode()}")
        return buf[:n]

    def profile(self, on=True):
        self.lib.mossdec_profile(self.h, int(bool(on)))

    def profile_read(self):
        out = (ctypes.c_double * 9)()
        self._check(self.lib.mossdec_profile_read(self.h, out), "profile_read")
        return {"gemm_ms": out[0], "gemm_launches": int(out[1]), "gemm_hw_flops": out[2], "attn_ms": out[3],
                "attn_launches": int(out[4]), "ffn_ms": out[6], "ffn_launches": int(out[7]), "ffn_hw_flops": out[8]}

    def reserve(self, flow_rows=0, flow_seqs=1, hift_frames=0):
        """Pre-size the workspaces (mossdec_reserve) so the first decode of that size does not allocate."""
        self._check(self.lib.mossdec_reserve(self.h, int(flow_rows), int(flow_seqs), int(hift_frames)), "reserve")

    def set_graph_mode(self, mode):
        self.lib.mossdec_set_graph_mode(self.h, int(mode))

    def graph_replays(self):
        return int(self.lib.mossdec_graph_replays(self.h))

    @property
    def device_index(self):
        if self._device_index is None:
            self._device_index = int(self.lib.mossdec_device(self.h))
        return self._device_index

    BANDWIDTH_KERNELS = ("euler_cfg_kernel", "flow_prepare_kernel", "hift_source_kernel", "hift_stft_kernel",
                         "hift_fuse_snake_kernel", "hift_istft_kernel")

    def profile_bandwidth(self):
        """{kernel: {ms, launches, bytes}} of the bandwidth-class kernels since the last profile_read()."""
        out = {}
        for name in self.BANDWIDTH_KERNELS:
            v = (ctypes.c_double * 3)()
            self._check(self.lib.mossdec_profile_named(self.h, name.encode(), v), "profile_named")
            out[name] = {"ms": v[0], "launches": int(v[1]), "bytes": v[2]}
        return out

    def set_hift_group_frames(self, frames):
        self.lib.mossdec_set_hift_group_frames(self.h, int(frames))

    def launch_count(self):
        return int(self.lib.mossdec_launch_count(self.h))

    def reset_launch_count(self):
        self.lib.mossdec_reset_launch_count(self.h)

"""Drop-in mirror of the reference's `GLM_modules/flow_inference.py` (AudioDecoder) on top of libmossdec.

Same constructor and method signatures, argument meaning, return shapes and per-uuid cache behaviour as the reference
(`AudioDecoder.__init__` flow_inference.py:49-92, `token2wav` :94-123, `token2wav_streaming` :130-164,
`stream_inference` :166-243), so `GLM4Encoder.decode / decode_streaming` (whisper_encoder_decoder.py:240-264,302-309)
run unchanged.  New capability on top of it: `token2wav_batch` / `decode_batch` run a ragged batch of independent
utterances through ONE flow solve and ONE vocoder pass (the reference asserts batch 1, flow.py:248), and
`streaming.StreamScheduler` steps many concurrent streams together.

There is no CPU or PyTorch fallback: constructing the decoder without a CUDA device or without the built
libmossdec.so raises.
"""
import os
import re
import uuid as _uuid
from collections import defaultdict

import numpy as np
import torch

from . import lib, mel_frontend, synth, weights

TOKEN_MEL_RATIO = 4
PRE_LOOKAHEAD = 3


_DEFAULTS = {"static_chunk_size": 25, "inference_cfg_rate": 0.7, "sample_rate": 24000, "n_timesteps": 10,
             "estimator_chunk_frames": None}


def _read_config(config_path):
    """The reference instantiates its models from a hyperpyyaml object graph (flow_inference.py:53-55).  Only scalar
    hyper-parameters matter here; they are read from the yaml text -- plain scalars and hyperpyyaml `!ref <name>`
    aliases of plain scalars.  A key that is present but cannot be resolved raises (no silent default); absent keys take
    the pinned defaults of SURVEY.md 8(a0).  `estimator_chunk_frames` (mel frames of the estimator's block-causal
    attention, decoder.py:439-440) defaults to static_chunk_size * token_mel_ratio, as upstream CosyVoice2 configures it."""
    cfg = dict(_DEFAULTS)
    if config_path is None:
        return cfg
    if not os.path.exists(config_path):
        raise FileNotFoundError(f"config_path {config_path!r} does not exist")   # the reference's open() raises
    text = open(config_path).read()
    scalars = {m.group(1): m.group(2) for m in re.finditer(r"^\s*([A-Za-z_][\w]*)\s*:\s*([-+0-9.eE]+)\s*(?:#.*)?$", text, re.M)}

    def resolve(key, cast):
        m = re.search(rf"^\s*{key}\s*:\s*(.+?)\s*(?:#.*)?$", text, re.M)
        if not m:
            return None
        val = m.group(1).strip()
        ref = re.fullmatch(r"!ref\s*<\s*([\w]+)\s*>", val)
        if ref:
            if ref.group(1) not in scalars:
                raise ValueError(f"{config_path}: `{key}: {val}` refers to `{ref.group(1)}`, which is not a plain scalar")
            val = scalars[ref.group(1)]
        try:
            return cast(float(val))
        except ValueError:
            raise ValueError(f"{config_path}: cannot parse `{key}: {val}` as a number") from None

    for key, cast in (("static_chunk_size", int), ("inference_cfg_rate", float), ("sample_rate", int),
                      ("n_timesteps", int), ("estimator_chunk_frames", int)):
        v = resolve(key, cast)
        if v is not None:
            cfg[key] = v
    return cfg


def _load_checkpoint(path, what):
    """torch.load of a reference-format checkpoint; a path that is given but missing raises like the reference's
    torch.load (flow_inference.py:59-61) instead of silently decoding with random weights."""
    if not os.path.exists(path):
        raise FileNotFoundError(f"{what} checkpoint {path!r} does not exist (pass None for the synthetic random-init weights)")
    return torch.load(path, map_location="cpu")


class _MissingCampplus:
    """Stands in for `campplus_session` when it cannot be built; any use raises with the reason (the reference fails in
    its constructor instead, flow_inference.py:86-89)."""

    def __init__(self, why):
        self.why = why

    def __getattr__(self, name):
        raise RuntimeError(f"campplus_session is unavailable: {self.why}")


def _make_campplus_session(campplus_model):
    if campplus_model is None:
        return _MissingCampplus("no campplus_model path was given")
    try:
        import onnxruntime
    except ImportError:
        return _MissingCampplus("onnxruntime is not installed")
    if not os.path.exists(campplus_model):
        raise FileNotFoundError(f"campplus_model {campplus_model!r} does not exist")
    option = onnxruntime.SessionOptions()
    option.graph_optimization_level = onnxruntime.GraphOptimizationLevel.ORT_ENABLE_ALL
    option.intra_op_num_threads = 1
    return onnxruntime.InferenceSession(campplus_model, sess_options=option, providers=["CPUExecutionProvider"])


class _FlowFacade:
    """What callers read from `decoder.flow` (flow_inference.py:77,141,191) + an `inference` with the reference's
    signature (flow.py:237-283)."""
    token_mel_ratio = TOKEN_MEL_RATIO
    pre_lookahead_len = PRE_LOOKAHEAD
    input_frame_rate = 12.5

    def __init__(self, dec):
        self._dec = dec

    @torch.inference_mode()
    def inference(self, token, token_len, prompt_token, prompt_token_len, prompt_feat, prompt_feat_len, embedding,
                  streaming, finalize):
        assert token.shape[0] == 1
        r = self._dec.flow_batch([token], [prompt_token], [prompt_feat], embedding, streaming, finalize)
        full = r["mel"].t().unsqueeze(0).contiguous()             # [1, 80, T]
        mel_len1 = int(r["prompt_frames"][0])
        feat_return = full[:, :, mel_len1:]
        assert feat_return.shape[2] == full.shape[2] - mel_len1
        return feat_return.float(), full.float()


class _HiftFacade:
    """`decoder.hift.inference(speech_feat=, cache_source=)` (generator.py:570-582)."""

    def __init__(self, dec):
        self._dec = dec

    @torch.inference_mode()
    def inference(self, speech_feat, cache_source=torch.zeros(1, 1, 0), nsf_noise=None):
        assert speech_feat.shape[0] == 1
        dev = self._dec.device
        T = speech_feat.shape[2]
        mel = speech_feat[0].to(dev).float().contiguous()
        kw = {}
        if cache_source.shape[2] != 0:
            kw = dict(cache_source=cache_source.to(dev).reshape(1, -1).float(), cache_len=[cache_source.shape[2]])
        wav, src = self._dec.engine.hift([T], mel.reshape(-1), nsf_noise=self._dec._noise_for([T], nsf_noise),
                                         seed=self._dec._next_seed(), **kw)
        return wav.view(1, -1), src.view(1, 1, -1)


class AudioDecoder(torch.nn.Module):
    def __init__(self, config_path=None, flow_ckpt_path=None, hift_ckpt_path=None, campplus_model=None,
                 mel_cache_len=8, device="cuda", flow_state_dict=None, hift_state_dict=None, seed=0):
        super().__init__()
        if not torch.cuda.is_available():
            raise RuntimeError("AudioDecoder (mossdec) needs a CUDA device: there is no CPU path")
        self.device = torch.device(device if device != "cuda" else f"cuda:{torch.cuda.current_device()}")
        torch.cuda.set_device(self.device)
        self.scratch_configs = _read_config(config_path)
        # checkpoints in the reference's state_dict format (flow_inference.py:59-61).  Random-init weights of the reference
        # architecture only when NO path is given (tests / benchmarks: checkpoints are not available offline).
        if flow_state_dict is None:
            flow_state_dict = (_load_checkpoint(flow_ckpt_path, "flow") if flow_ckpt_path is not None
                               else synth.synth_flow_state_dict(seed))
        if hift_state_dict is None:
            hift_state_dict = (_load_checkpoint(hift_ckpt_path, "hift") if hift_ckpt_path is not None
                               else synth.synth_hift_state_dict(seed))
        self.engine = lib.Engine()
        tensors = weights.pack_estimator(flow_state_dict)
        tensors["flow.rand_noise"] = synth.synth_noise_bank()[0].contiguous()   # flow_matching.py:202-203
        tensors.update(weights.pack_hift(hift_state_dict))
        tensors.update(weights.pack_encoder(flow_state_dict))
        tensors.update(mel_frontend.pack_mel_tensors())
        self.engine.load_tensors(tensors)
        self.engine.finalize()
        self.flow = _FlowFacade(self)
        self.hift = _HiftFacade(self)
        self.sample_rate = self.scratch_configs["sample_rate"]
        # prompt mel extractor = matcha mel_spectrogram (flow_inference.py:64, whisper_encoder_decoder.py:192), on the device
        self.feat_extractor = mel_frontend.MelExtractor(self.engine, self.device)
        # campplus x-vector session, built exactly like the reference (flow_inference.py:86-89); the model itself is
        # out of scope (CPU ONNX, its embedding is an input of the path)
        self.campplus_session = _make_campplus_session(campplus_model)
        self.static_chunk_size = int(self.scratch_configs["static_chunk_size"])   # tokens (encoder chunk masks)
        ecf = self.scratch_configs["estimator_chunk_frames"]
        self.estimator_chunk_frames = int(ecf) if ecf is not None else self.static_chunk_size * TOKEN_MEL_RATIO
        self.n_timesteps = int(self.scratch_configs["n_timesteps"])
        self.inference_cfg_rate = float(self.scratch_configs["inference_cfg_rate"])

        # session state, same keys as the reference (flow_inference.py:70-74)
        self.mel_overlap_dict = defaultdict(lambda: None)
        self.hift_cache_dict = defaultdict(lambda: None)
        self.llm_end_dict = defaultdict(lambda: None)
        self.tts_speech_token_dict = defaultdict(lambda: None)
        self.flow_cache_dict = defaultdict(lambda: None)

        self.token_overlap_len = 3.5
        self.mel_overlap_len = int(self.token_overlap_len / self.flow.input_frame_rate * 24000 / 480 / 2)  # 7
        self.mel_window = np.hamming(2 * self.mel_overlap_len)
        self.mel_cache_len = mel_cache_len
        self.source_cache_len = int(self.mel_cache_len * 480)
        self.speech_window = np.hamming(2 * self.source_cache_len)
        self._speech_window_dev = torch.from_numpy(self.speech_window.astype(np.float32)).to(self.device)
        self._seed = 0
        self.last_h2d_bytes = 0

    def _next_seed(self):
        self._seed += 1
        return self._seed

    # Parity tests inject the reference's NSF noise draws (generator.py:334): a callable taking the list of vocoder
    # segment lengths (frames) and returning the ragged [sum 480 T_b, 9] noise on the device.  None -> device Philox.
    nsf_noise_hook = None

    def _noise_for(self, lens, nsf_noise=None):
        if nsf_noise is None and self.nsf_noise_hook is not None:
            nsf_noise = self.nsf_noise_hook(lens)
        return nsf_noise

    # ------------------------------------------------------------------------------------------ batched core
    @torch.inference_mode()
    def flow_batch(self, tokens, prompt_tokens, prompt_feats, embeddings, streaming, finalize, prefix=None):
        """CausalMaskedDiffWithXvec.inference (flow.py:237-283) for a list of utterances.

        tokens[b]: int [1, N_b]; prompt_tokens[b]: int [1, P_b]; prompt_feats[b]: float [1, F_b, 80];
        embeddings: float [B, 192].  Returns ragged time-major mel [sum T_b, 80] (prompt frames included), the
        per-utterance frame counts T_b and prompt frame counts F_b.  Host work: one concatenation per input kind into
        pinned memory; everything else is mossdec_token2mel.  prefix: frozen-prefix cache arguments of a streaming chunk
        call ({"mode", "slots", "tags"}, lib.Engine.token2mel); in reuse mode the frozen frames of "mel" are not written."""
        dev = self.device
        B = len(tokens)
        rows = [t.reshape(-1) for b in range(B) for t in (prompt_tokens[b], tokens[b])]
        n_tok = [int(rows[2 * b].numel() + rows[2 * b + 1].numel()) for b in range(B)]
        tok = torch.cat([r.long() for r in rows]).pin_memory()
        F_ = [int(pf.shape[1]) for pf in prompt_feats]
        feats = [pf.reshape(-1, 80) for pf in prompt_feats if pf.shape[1]]
        feat_cat = (torch.cat(feats, 0).float() if feats else torch.zeros(1, 80)).pin_memory()
        embeddings = embeddings.float().contiguous()
        self.last_h2d_bytes = tok.nbytes + feat_cat.nbytes + embeddings.nbytes   # what this call copies host -> device
        return self.flow_device(tok.to(dev, non_blocking=True), n_tok, feat_cat.to(dev, non_blocking=True), F_,
                                embeddings.to(dev, non_blocking=True), streaming, finalize, prefix)

    @torch.inference_mode()
    def flow_device(self, tok, n_tok, prompt_feat_cat, prompt_frames, embeddings, streaming, finalize, prefix=None):
        """Device-resident core of flow_batch: tok int64 [sum n_tok] (prompt + new tokens of every utterance back to back),
        n_tok host list, prompt_feat_cat fp32 [sum F_b, 80] (prompt mels back to back), prompt_frames host list F_b,
        embeddings fp32 [B, 192] -- all tensors already on the device."""
        ctx = 0 if finalize else PRE_LOOKAHEAD
        if min(n_tok) - ctx <= 0:
            raise ValueError("every utterance needs at least one token besides the look-ahead context")
        T = [(n - ctx) * TOKEN_MEL_RATIO for n in n_tok]
        F_ = list(prompt_frames)
        if any(f > t for f, t in zip(F_, T)):
            raise ValueError("prompt_feat longer than the decoded sequence")
        mel = self.engine.token2mel(n_tok, tok, F_, prompt_feat_cat, embeddings, ctx,
                                    self.static_chunk_size if streaming else 0,
                                    self.estimator_chunk_frames if streaming else 0,
                                    n_steps=self.n_timesteps, cfg_rate=self.inference_cfg_rate, prefix=prefix)
        return {"mel": mel, "frames": T, "prompt_frames": F_}

    @torch.inference_mode()
    def vocode_generated(self, r, nsf_noise=None, seed=None):
        """Run the vocoder on the generated part of a flow result (the prompt frames are passed over, flow.py:281)."""
        T, F_ = r["frames"], r["prompt_frames"]
        gen = [t - f for t, f in zip(T, F_)]
        wav, _ = self.engine.hift_skip(gen, F_, r["mel"], nsf_noise=self._noise_for(gen, nsf_noise),
                                       seed=self._next_seed() if seed is None else seed)
        return {"wav": wav, "wav_lens": [g * 480 for g in gen], "flow": r, "mel_lens": gen}

    def generated_mel(self, r):
        """Ragged time-major mel [sum 4 N_b, 80] of the generated frames of a flow result."""
        if not any(r["prompt_frames"]):
            return r["mel"]
        parts, off = [], 0
        for t, f in zip(r["frames"], r["prompt_frames"]):
            parts.append(r["mel"][off + f:off + t])
            off += t
        return torch.cat(parts, 0)

    @torch.inference_mode()
    def decode_batch(self, tokens, prompt_tokens=None, prompt_feats=None, embeddings=None, nsf_noise=None, seed=None,
                     want_mel=False):
        """token2wav (flow_inference.py:94-123) for B independent utterances in one pass.

        Returns {"wav": ragged fp32 [sum 1920 N_b], "wav_lens", "mel_lens", "flow": the flow result}; with want_mel also
        "mel": ragged time-major [sum 4 N_b, 80] (the generated frames gathered out of the flow result)."""
        B = len(tokens)
        prompt_tokens = prompt_tokens or [torch.zeros(1, 0, dtype=torch.int32)] * B
        prompt_feats = prompt_feats or [torch.zeros(1, 0, 80)] * B
        if embeddings is None:
            embeddings = torch.zeros(B, 192)
        r = self.flow_batch(tokens, prompt_tokens, prompt_feats, embeddings, streaming=False, finalize=True)
        out = self.vocode_generated(r, nsf_noise, seed)
        if want_mel:
            out["mel"] = self.generated_mel(r)
        return out

    def token2wav_batch(self, tokens, prompt_tokens=None, prompt_feats=None, embeddings=None, **kw):
        """List-of-utterances convenience wrapper: returns [(tts_speech [1, 1920 N_b], tts_mel [1, 80, 4 N_b])]."""
        r = self.decode_batch(tokens, prompt_tokens, prompt_feats, embeddings, want_mel=True, **kw)
        out, wo, mo = [], 0, 0
        for wl, ml in zip(r["wav_lens"], r["mel_lens"]):
            out.append((r["wav"][wo:wo + wl].view(1, -1), r["mel"][mo:mo + ml].t().unsqueeze(0)))
            wo += wl
            mo += ml
        return out

    # ------------------------------------------------------------------------------------------ reference API
    def token2wav(self, token, uuid, prompt_token=torch.zeros(1, 0, dtype=torch.int32),
                  prompt_feat=torch.zeros(1, 0, 80), embedding=torch.zeros(1, 192)):
        """flow_inference.py:94-123 (mel_overlap / hift caches of `uuid` are honoured and then dropped)."""
        tts_mel, _ = self.flow.inference(token=token, token_len=None, prompt_token=prompt_token, prompt_token_len=None,
                                         prompt_feat=prompt_feat, prompt_feat_len=None, embedding=embedding,
                                         streaming=False, finalize=True)
        if self.mel_overlap_dict[uuid] is not None:
            from .streaming import fade_in_out_mel
            tts_mel = fade_in_out_mel(tts_mel, self.mel_overlap_dict[uuid], self.mel_window)
        if self.hift_cache_dict[uuid] is not None:
            hift_cache_mel, hift_cache_source = self.hift_cache_dict[uuid]['mel'], self.hift_cache_dict[uuid]['source']
            tts_mel = torch.concat([hift_cache_mel, tts_mel], dim=2)
        else:
            hift_cache_source = torch.zeros(1, 1, 0)
        tts_speech, _ = self.hift.inference(speech_feat=tts_mel, cache_source=hift_cache_source)
        del self.hift_cache_dict[uuid]
        del self.mel_overlap_dict[uuid]
        return tts_speech, tts_mel

    def offline_inference(self, token):
        """flow_inference.py:125-128 (the reference passes a stale `finalize=` kwarg there and raises; this works)."""
        tts_speech, _ = self.token2wav(token, uuid=str(_uuid.uuid1()))
        return tts_speech.cpu()

    def token2wav_streaming(self, token, prompt_token, prompt_feat, token_offset, uuid, embedding=torch.zeros(1, 192),
                            finalize=False, stream=False, speed=1.0):
        from .streaming import token2wav_streaming
        return token2wav_streaming(self, token, prompt_token, prompt_feat, token_offset, uuid, embedding, finalize,
                                   stream, speed)

    def stream_inference(self, token, prompt_token=torch.zeros(1, 0, dtype=torch.int32),
                         prompt_feat=torch.zeros(1, 0, 80), embedding=torch.zeros(0, 192), block_size=8,
                         max_token_len=None):
        from .streaming import stream_inference
        return stream_inference(self, token, prompt_token, prompt_feat, embedding, block_size, max_token_len)

"""TEST INFRASTRUCTURE ONLY -- import shims that let the *unmodified* reference run on CPU.

The reference (alpacaking/MOSS-Speech-decoder-cosy, mounted read-only at /root/reference) imports a few
third-party packages that are not installed offline.  This module registers minimal stand-ins for them in
``sys.modules`` BEFORE anything from the reference is imported, so that

  * GLM_modules/flow_inference.py (AudioDecoder)
  * GLM_modules/cosyvoice/flow/{flow,flow_matching,decoder}.py
  * GLM_modules/cosyvoice/hifigan/{generator,f0_predictor}.py
  * GLM_modules/cosyvoice/transformer/upsample_encoder.py
  * GLM_modules/third_party/Matcha-TTS/matcha/models/components/{decoder,transformer,flow_matching}.py

run exactly as written.  The only arithmetic restated here lives in ``diffusers`` (pinned ==0.35.1 in the
reference's requirements.txt:32 and ==0.27.2 in GLM_modules/requirements.txt:3), whose call sites are
matcha/models/components/transformer.py:109-126,196-204,266-271 and decoder.py:92:

  Attention(query_dim, heads, dim_head, bias=False): to_q/to_k/to_v = Linear(no bias), to_out[0] = Linear(+bias),
      heads split, additive mask [B,T,T] broadcast over heads, F.scaled_dot_product_attention (scale 1/sqrt(d)).
  GELU(dim_in, dim_out): Linear then exact erf GELU.   LoRACompatibleLinear == nn.Linear.
  get_activation("silu") == nn.SiLU.

State-dict key names match diffusers' so real checkpoints would load.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference leg may import this package.
"""
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

REF_ROOT = os.environ.get("MOSS_REF_DIR", "/root/reference/GLM_modules")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "cosyvoice", "flow"))


class _Attention(nn.Module):
    """diffusers.models.attention_processor.Attention (self-attention subset used by matcha)."""

    def __init__(self, query_dim, cross_attention_dim=None, heads=8, dim_head=64, dropout=0.0, bias=False,
                 upcast_attention=False, **kwargs):
        super().__init__()
        inner = heads * dim_head
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(query_dim, inner, bias=bias)
        self.to_v = nn.Linear(query_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim, bias=True), nn.Dropout(dropout)])

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **kwargs):
        b, t, _ = hidden_states.shape
        h = self.heads
        q = self.to_q(hidden_states).view(b, t, h, -1).transpose(1, 2)
        k = self.to_k(hidden_states).view(b, t, h, -1).transpose(1, 2)
        v = self.to_v(hidden_states).view(b, t, h, -1).transpose(1, 2)
        if attention_mask is not None:
            # diffusers.prepare_attention_mask: [B, Tq, Tk] -> repeat_interleave over heads -> [B, H, Tq, Tk]
            attention_mask = attention_mask.view(b, 1, attention_mask.shape[-2], attention_mask.shape[-1])
            attention_mask = attention_mask.expand(b, h, -1, -1)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=attention_mask, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(b, t, -1).to(q.dtype)
        o = self.to_out[0](o)
        o = self.to_out[1](o)
        return o


class _GELU(nn.Module):
    def __init__(self, dim_in, dim_out, approximate="none", bias=True):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out, bias=bias)
        self.approximate = approximate

    def forward(self, x):
        return F.gelu(self.proj(x), approximate=self.approximate)


class _Unsupported(nn.Module):
    def __init__(self, *a, **k):
        raise NotImplementedError("not on the decode path")


def _get_activation(name):
    return {"silu": nn.SiLU(), "swish": nn.SiLU(), "mish": nn.Mish(), "gelu": nn.GELU(), "relu": nn.ReLU()}[name]


def _mod(name, **attrs):
    import importlib.machinery
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


_installed = False


def install(with_audio_decoder: bool = True):
    """Register the stand-in modules and put the reference on sys.path (idempotent)."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    if "diffusers" not in sys.modules:
        _mod("diffusers")
        _mod("diffusers.models")
        _mod("diffusers.models.attention", GEGLU=_Unsupported, GELU=_GELU, AdaLayerNorm=_Unsupported,
             AdaLayerNormZero=_Unsupported, ApproximateGELU=_Unsupported)
        _mod("diffusers.models.attention_processor", Attention=_Attention)
        _mod("diffusers.models.lora", LoRACompatibleLinear=nn.Linear)
        _mod("diffusers.models.activations", get_activation=_get_activation)
        _mod("diffusers.utils")
        _mod("diffusers.utils.torch_utils", maybe_allow_in_graph=lambda cls: cls)
    if "conformer" not in sys.modules:
        _mod("conformer", ConformerBlock=_Unsupported)
    if "omegaconf" not in sys.modules:
        class DictConfig(dict):
            def __getattr__(self, k):
                try:
                    return self[k]
                except KeyError as e:
                    raise AttributeError(k) from e
        _mod("omegaconf", DictConfig=DictConfig)
    matcha_root = os.path.join(REF_ROOT, "third_party", "Matcha-TTS")
    for p in (REF_ROOT, matcha_root):
        if p not in sys.path:
            sys.path.insert(0, p)
    # matcha.utils pulls lightning/hydra/rich: replace the package init, keep __path__ on the real dir
    if "matcha.utils" not in sys.modules:
        import importlib
        import logging
        importlib.import_module("matcha")
        mu = _mod("matcha.utils")
        mu.__path__ = [os.path.join(matcha_root, "matcha", "utils")]
        _mod("matcha.utils.pylogger", get_pylogger=lambda name=__name__: logging.getLogger(name))
    # class_utils.py:35,38 imports the text->token LLMs and CLI front ends only to list them in a registry; they are
    # out of scope and drag in transformers/Qwen2, so give the registry inert placeholders.
    if "cosyvoice.llm.llm" not in sys.modules:
        import importlib
        importlib.import_module("cosyvoice")
        for pkg in ("cosyvoice.llm", "cosyvoice.cli"):
            pm = _mod(pkg)
            pm.__path__ = [os.path.join(REF_ROOT, *pkg.split("."))]
        _mod("cosyvoice.llm.llm", TransformerLM=_Unsupported, Qwen2LM=_Unsupported)
        _mod("cosyvoice.cli.model", CosyVoiceModel=_Unsupported, CosyVoice2Model=_Unsupported)
    # matcha/utils/audio.py:4 imports librosa.filters.mel (librosa pinned 0.10.2 / 0.11.0, absent offline): the stand-in is
    # the restated Slaney filterbank of oracle/port.py, itself cross-checked against torchaudio's independent
    # implementation (tests/test_oracle_cpu.py::test_slaney_filterbank_matches_torchaudio)
    if "librosa" not in sys.modules:
        from . import port as _port
        _mod("librosa")
        _mod("librosa.filters", mel=lambda sr, n_fft, n_mels, fmin, fmax: _port.slaney_mel_filterbank(sr, n_fft, n_mels, fmin, fmax))
    if with_audio_decoder:
        if "hyperpyyaml" not in sys.modules:
            _mod("hyperpyyaml", load_hyperpyyaml=lambda f: _HYPER["cfg"])
        if "onnxruntime" not in sys.modules:
            class _SessOpt:
                graph_optimization_level = None
                intra_op_num_threads = 1

            class _GOL:
                ORT_ENABLE_ALL = 99

            class _Sess:
                def __init__(self, *a, **k):
                    pass
            _mod("onnxruntime", SessionOptions=_SessOpt, GraphOptimizationLevel=_GOL, InferenceSession=_Sess)
    _installed = True


# what the fake load_hyperpyyaml returns; filled by build_ref.make_audio_decoder
_HYPER = {"cfg": None}

"""Random-init checkpoints in the reference's state_dict format (no checkpoint ships with the reference).

`synth_flow_state_dict` / `synth_hift_state_dict` return {key: torch.Tensor(fp32, CPU)} with exactly the keys and
shapes of `flow.pt` / `hift.pt` (ckpt_schema.json, dumped from the reference classes by oracle/make_manifest.py).
Values come from numpy PCG64 streams keyed by (seed, crc32(key)), so they are identical on every machine and do not
depend on key order.  Scales are chosen so that every layer carries O(1) signals:

  * flow: Linear/Conv weights ~ N(0, gain / fan_in); LayerNorm gamma ~ 1 + 0.1 N; biases ~ 0.05 N.
  * hift: weight-norm direction v ~ N(0, 0.01) (the reference's own init, generator.py:83-84) and the gain g set per
    layer (w = g * v / |v|) so Snake / exp / sin stay in their non-linear range without clipping at +-0.99;
    Snake alpha ~ 1 + 0.3 U; the f0 classifier is scaled so |f0| straddles the 10 Hz voiced threshold.
"""
import json
import math
import os
import zlib

import numpy as np
import torch

_SCHEMA = None


def ckpt_schema():
    global _SCHEMA
    if _SCHEMA is None:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ckpt_schema.json")) as f:
            _SCHEMA = json.load(f)
    return _SCHEMA


def _rng(seed, key):
    return np.random.Generator(np.random.PCG64([seed, zlib.crc32(key.encode())]))


def _normal(seed, key, shape, std=1.0, mean=0.0):
    a = _rng(seed, key).standard_normal(size=tuple(shape), dtype=np.float32)
    return torch.from_numpy(a * np.float32(std) + np.float32(mean))


def _uniform(seed, key, shape):
    return torch.from_numpy(_rng(seed, key).random(size=tuple(shape), dtype=np.float32))


def synth_flow_state_dict(seed: int = 0):
    sd = {}
    for key, shape in ckpt_schema()["flow"].items():
        if key == "input_embedding.weight":
            t = _normal(seed, key, shape, 1.0)
        elif key.endswith("pos_bias_u") or key.endswith("pos_bias_v"):
            t = _normal(seed, key, shape, 0.1)
        elif len(shape) >= 2:
            fan_in = int(np.prod(shape[1:]))
            gain = 1.0
            if ".ff.net.0.proj." in key or ".feed_forward.w_1." in key or "block.0.weight" in key:
                gain = 2.0  # layers followed by GELU / Mish / ReLU keep unit variance
            t = _normal(seed, key, shape, math.sqrt(gain / fan_in))
        elif key.endswith("weight"):  # 1-D weights are LayerNorm gammas
            t = _normal(seed, key, shape, 0.1, 1.0)
        else:
            t = _normal(seed, key, shape, 0.05)
        sd[key] = t
    return sd


def synth_hift_state_dict(seed: int = 0):
    sd = {}
    for key, shape in ckpt_schema()["hift"].items():
        if key.endswith("parametrizations.weight.original1"):
            t = _normal(seed, key, shape, 0.01)
        elif key.endswith("parametrizations.weight.original0"):
            if key.startswith("ups."):
                stride = (8, 5, 3)[int(key.split(".")[1])]
                g = math.sqrt(stride / 2.0)
            elif key.startswith("conv_pre"):
                g = 1.0
            elif key.startswith("conv_post"):
                g = 0.3
            elif key.startswith("f0_predictor"):
                g = 1.2
            else:
                g = 0.7
            t = torch.full(tuple(shape), g, dtype=torch.float32)
        elif key.endswith(".alpha"):
            t = 1.0 + 0.3 * _uniform(seed, key, shape)
        elif key.startswith("source_downs") and key.endswith("weight"):
            t = _normal(seed, key, shape, 1.0)
            t = t * (4.0 / t.flatten(1).norm(dim=1).view(-1, 1, 1))
        elif key == "f0_predictor.classifier.weight":
            t = _normal(seed, key, shape, 3.0)
        elif key == "f0_predictor.classifier.bias":
            t = torch.full(tuple(shape), 12.0, dtype=torch.float32)
        elif key == "m_source.l_linear.weight":
            t = _normal(seed, key, shape, 0.6)
        else:  # biases
            t = _normal(seed, key, shape, 0.05)
        sd[key] = t
    return sd


def synth_noise_bank(n_feats: int = 80, length: int = 15000):
    """The reference's fixed CFM noise bank: set_all_random_seed(0); torch.randn([1, 80, 15000])
    (flow_matching.py:202-203).  torch's CPU generator is deterministic for a given torch build."""
    g = torch.Generator(device="cpu")
    g.manual_seed(0)
    return torch.randn([1, n_feats, length], generator=g)


def synth_cv1_blocks_state_dict(seed: int = 0, cin: int = 320, time_dim: int = 1024):
    """Random-init weights of the cosyvoice1 / Matcha building blocks: one ResnetBlock1D(cin, 256,
    time_dim), one ResnetBlock1D(256, 256, time_dim), a Block1D(256, 256) (the decoder's final_block), Downsample1D(256), Upsample1D(256) (matcha decoder.py:32-158)."""
    shapes = {}
    for name, ci in (("res_in", cin), ("res_mid", 256)):
        shapes.update({
            f"{name}.block1.block.0.weight": (256, ci, 3), f"{name}.block1.block.0.bias": (256,),
            f"{name}.block1.block.1.weight": (256,), f"{name}.block1.block.1.bias": (256,),
            f"{name}.block2.block.0.weight": (256, 256, 3), f"{name}.block2.block.0.bias": (256,),
            f"{name}.block2.block.1.weight": (256,), f"{name}.block2.block.1.bias": (256,),
            f"{name}.mlp.1.weight": (256, time_dim), f"{name}.mlp.1.bias": (256,),
            f"{name}.res_conv.weight": (256, ci, 1), f"{name}.res_conv.bias": (256,),
        })
    shapes.update({"final.block.0.weight": (256, 256, 3), "final.block.0.bias": (256,),
                   "final.block.1.weight": (256,), "final.block.1.bias": (256,)})
    shapes.update({"down.conv.weight": (256, 256, 3), "down.conv.bias": (256,),
                   "up.conv.weight": (256, 256, 4), "up.conv.bias": (256,)})
    sd = {}
    for key, shape in shapes.items():
        k = "cv1." + key
        if len(shape) >= 2:
            fan_in = int(np.prod(shape[1:]))
            gain = 2.0 if "block.0.weight" in key else 1.0
            sd[key] = _normal(seed, k, shape, math.sqrt(gain / fan_in))
        elif key.endswith("block.1.weight"):   # GroupNorm gamma
            sd[key] = _normal(seed, k, shape, 0.1, 1.0)
        else:
            sd[key] = _normal(seed, k, shape, 0.05)
    return sd

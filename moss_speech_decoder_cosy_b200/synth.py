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


def cv1_decoder_shapes(n_blocks: int = 4, n_mid: int = 12, in_channels: int = 320, out_channels: int = 80, C: int = 256,
                       heads: int = 8, head_dim: int = 64):
    """state_dict keys / shapes of the non-causal ConditionalDecoder (cosyvoice/flow/decoder.py:88-204, cosyvoice1/flow/decoder.py)
    with channels = [C, C] (oracle/make_golden_r2.py loads them with strict=True into the reference class)."""
    td, inner = 4 * C, heads * head_dim
    sh = {"time_mlp.linear_1.weight": (td, in_channels), "time_mlp.linear_1.bias": (td,),
          "time_mlp.linear_2.weight": (td, td), "time_mlp.linear_2.bias": (td,)}

    def resnet(p, cin):
        sh.update({f"{p}.block1.block.0.weight": (C, cin, 3), f"{p}.block1.block.0.bias": (C,),
                   f"{p}.block1.block.1.weight": (C,), f"{p}.block1.block.1.bias": (C,),
                   f"{p}.block2.block.0.weight": (C, C, 3), f"{p}.block2.block.0.bias": (C,),
                   f"{p}.block2.block.1.weight": (C,), f"{p}.block2.block.1.bias": (C,),
                   f"{p}.mlp.1.weight": (C, td), f"{p}.mlp.1.bias": (C,),
                   f"{p}.res_conv.weight": (C, cin, 1), f"{p}.res_conv.bias": (C,)})

    def tf(p):
        for j in range(n_blocks):
            q = f"{p}.{j}"
            sh.update({f"{q}.norm1.weight": (C,), f"{q}.norm1.bias": (C,), f"{q}.norm3.weight": (C,), f"{q}.norm3.bias": (C,),
                       f"{q}.attn1.to_q.weight": (inner, C), f"{q}.attn1.to_k.weight": (inner, C), f"{q}.attn1.to_v.weight": (inner, C),
                       f"{q}.attn1.to_out.0.weight": (C, inner), f"{q}.attn1.to_out.0.bias": (C,),
                       f"{q}.ff.net.0.proj.weight": (4 * C, C), f"{q}.ff.net.0.proj.bias": (4 * C,),
                       f"{q}.ff.net.2.weight": (C, 4 * C), f"{q}.ff.net.2.bias": (C,)})
    resnet("down_blocks.0.0", in_channels); tf("down_blocks.0.1")
    sh.update({"down_blocks.0.2.conv.weight": (C, C, 3), "down_blocks.0.2.conv.bias": (C,)})          # Downsample1D
    resnet("down_blocks.1.0", C); tf("down_blocks.1.1")
    sh.update({"down_blocks.1.2.weight": (C, C, 3), "down_blocks.1.2.bias": (C,)})                    # Conv1d k3 p1
    for i in range(n_mid):
        resnet(f"mid_blocks.{i}.0", C); tf(f"mid_blocks.{i}.1")
    resnet("up_blocks.0.0", 2 * C); tf("up_blocks.0.1")
    sh.update({"up_blocks.0.2.conv.weight": (C, C, 4), "up_blocks.0.2.conv.bias": (C,)})              # Upsample1D ConvTranspose
    resnet("up_blocks.1.0", 2 * C); tf("up_blocks.1.1")
    sh.update({"up_blocks.1.2.weight": (C, C, 3), "up_blocks.1.2.bias": (C,)})
    sh.update({"final_block.block.0.weight": (C, C, 3), "final_block.block.0.bias": (C,),
               "final_block.block.1.weight": (C,), "final_block.block.1.bias": (C,),
               "final_proj.weight": (out_channels, C, 1), "final_proj.bias": (out_channels,)})
    return sh


def synth_cv1_decoder_state_dict(seed: int = 0, n_blocks: int = 4, n_mid: int = 12):
    """Random-init weights of the assembled non-causal decoder, same scale rules as the flow checkpoint."""
    sd = {}
    for key, shape in cv1_decoder_shapes(n_blocks, n_mid).items():
        k = "cv1dec." + key
        if len(shape) >= 2:
            fan_in = int(np.prod(shape[1:]))
            gain = 2.0 if (".ff.net.0.proj." in key or "block.0.weight" in key) else 1.0
            sd[key] = _normal(seed, k, shape, math.sqrt(gain / fan_in))
        elif key.endswith("weight"):      # GroupNorm / LayerNorm gammas
            sd[key] = _normal(seed, k, shape, 0.1, 1.0)
        else:
            sd[key] = _normal(seed, k, shape, 0.05)
    return sd

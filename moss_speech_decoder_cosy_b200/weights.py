"""Weight packer: reference state_dicts (flow.pt / hift.pt key names) -> engine-ready tensors for libmossdec.

Layouts produced (all CPU tensors; bf16 for tensor-core operands, fp32 for biases / norms / small matrices):

  conv k taps, Cin -> Cout     : bf16 [Cout, k * Cin_pad], K index = tap * Cin_pad + c, Cin_pad = ceil64(Cin)
                                 (one 64-wide k-block never straddles two taps; extra columns are zero)
  Linear                        : bf16 [out, in]
  q|k|v                         : bf16 [3 * 512, 256], q rows pre-multiplied by 1/sqrt(head_dim) (exact: 1/8), columns
                                 scaled by norm1's gamma; bias = W . beta (the LayerNorm affine is folded; same for
                                 norm3 into ff1)
  ConvTranspose1d(stride s)     : bf16 [s * Cout, 3 * Cin]: row r * Cout + co, column (d + 1) * Cin + ci holds
                                 w[ci, co, r + p - s * d] for input offset d in {-1, 0, +1} (zero where the tap does
                                 not exist), so out[s * q + r] = sum_d x[q + d] . W_r,d    (polyphase form)
  weight_norm                   : folded, w = g * v / ||v|| with the norm over all dims but 0 (generator.py:26-29)

Reference layer -> engine name map is spelled out next to each block below (cites are to GLM_modules/).
"""
import math

import torch

BK = 64


def _ceil64(n):
    return (n + BK - 1) // BK * BK


def conv_weight_taps(w: torch.Tensor, cin_pad: int = None) -> torch.Tensor:
    """Conv1d weight [Cout, Cin, k] -> [Cout, k * Cin_pad] (tap-major K)."""
    cout, cin, k = w.shape
    cp = cin_pad or _ceil64(cin)
    out = torch.zeros(cout, k, cp, dtype=torch.float32)
    out[:, :, :cin] = w.permute(0, 2, 1)
    return out.reshape(cout, k * cp)


def fold_weight_norm(sd, prefix):
    """w = g * v / ||v||  (new-style parametrizations or old weight_g / weight_v)."""
    if prefix + ".parametrizations.weight.original0" in sd:
        g, v = sd[prefix + ".parametrizations.weight.original0"], sd[prefix + ".parametrizations.weight.original1"]
    elif prefix + ".weight_g" in sd:
        g, v = sd[prefix + ".weight_g"], sd[prefix + ".weight_v"]
    else:
        return sd[prefix + ".weight"].float()
    g, v = g.float(), v.float()
    norm = v.flatten(1).norm(dim=1).view(-1, *([1] * (v.ndim - 1)))
    return g * v / norm


def convtranspose_polyphase(w: torch.Tensor, stride: int, padding: int) -> torch.Tensor:
    """ConvTranspose1d weight [Cin, Cout, k] -> [stride * Cout, 3 * Cin] (see module docstring)."""
    cin, cout, k = w.shape
    out = torch.zeros(stride, cout, 3, cin, dtype=torch.float32)
    for r in range(stride):
        for di, d in enumerate((-1, 0, 1)):
            kk = r + padding - stride * d
            if 0 <= kk < k:
                out[r, :, di, :] = w[:, :, kk].t()
    return out.reshape(stride * cout, 3 * cin)


def _bf16(t):
    return t.to(torch.bfloat16).contiguous()


def _f32(t):
    return t.float().contiguous()


def pack_transformer_block(g, t, o, head_dim=64):
    """BasicTransformerBlock (matcha transformer.py:243-316 + diffusers Attention / GELU) keys under `t` (read through the
    getter g) -> engine tensors under `o`.  The affine part of norm1 / norm3 is folded into the Linear that consumes the
    normalised activation (W . (g * n + b) = (W diag g) . n + W . b), so the kernels' LayerNorm epilogues only normalise;
    1 / sqrt(head_dim) is folded into the q rows."""
    out = {}
    n1g, n1b = g(t + "norm1.weight"), g(t + "norm1.bias")
    n3g, n3b = g(t + "norm3.weight"), g(t + "norm3.bias")
    q = g(t + "attn1.to_q.weight") * (1.0 / math.sqrt(head_dim))
    wqkv = torch.cat([q, g(t + "attn1.to_k.weight"), g(t + "attn1.to_v.weight")], 0)
    out[o + "qkv.w"] = _bf16(wqkv * n1g[None, :])
    out[o + "qkv.b"] = _f32(wqkv @ n1b)
    out[o + "o.w"] = _bf16(g(t + "attn1.to_out.0.weight"))
    out[o + "o.b"] = _f32(g(t + "attn1.to_out.0.bias"))
    w1 = g(t + "ff.net.0.proj.weight")
    out[o + "ff1.w"] = _bf16(w1 * n3g[None, :])
    out[o + "ff1.b"] = _f32(g(t + "ff.net.0.proj.bias") + w1 @ n3b)
    out[o + "ff2.w"] = _bf16(g(t + "ff.net.2.weight"))
    out[o + "ff2.b"] = _f32(g(t + "ff.net.2.bias"))
    return out


def pack_estimator(sd, prefix="decoder.estimator.", n_mid=12, n_blocks=4, head_dim=64):
    """CausalConditionalDecoder (cosyvoice/flow/decoder.py:294-403) with channels=[256]."""
    out = {}
    g = lambda k: sd[prefix + k].float()
    out["est.time.w1"] = _f32(g("time_mlp.linear_1.weight"))   # matcha decoder.py:85
    out["est.time.b1"] = _f32(g("time_mlp.linear_1.bias"))
    out["est.time.w2"] = _f32(g("time_mlp.linear_2.weight"))   # matcha decoder.py:98
    out["est.time.b2"] = _f32(g("time_mlp.linear_2.bias"))
    groups = ["down_blocks.0"] + [f"mid_blocks.{i}" for i in range(n_mid)] + ["up_blocks.0"]
    tb_w, tb_b = [], []
    for gi, gp in enumerate(groups):
        r = f"{gp}.0."
        # CausalResnetBlock1D (decoder.py:81-85): block1/block2 = CausalConv1d k3 + LayerNorm + Mish
        out[f"est.res{gi}.c1.w"] = _bf16(conv_weight_taps(g(r + "block1.block.0.weight")))
        out[f"est.res{gi}.c1.b"] = _f32(g(r + "block1.block.0.bias"))
        out[f"est.res{gi}.ln1.g"] = _f32(g(r + "block1.block.2.weight"))
        out[f"est.res{gi}.ln1.b"] = _f32(g(r + "block1.block.2.bias"))
        out[f"est.res{gi}.c2.w"] = _bf16(conv_weight_taps(g(r + "block2.block.0.weight")))
        out[f"est.res{gi}.c2.b"] = _f32(g(r + "block2.block.0.bias"))
        out[f"est.res{gi}.ln2.g"] = _f32(g(r + "block2.block.2.weight"))
        out[f"est.res{gi}.ln2.b"] = _f32(g(r + "block2.block.2.bias"))
        out[f"est.res{gi}.rc.w"] = _bf16(conv_weight_taps(g(r + "res_conv.weight")))   # 1x1
        out[f"est.res{gi}.rc.b"] = _f32(g(r + "res_conv.bias"))
        tb_w.append(g(r + "mlp.1.weight"))                                              # matcha decoder.py:49
        tb_b.append(g(r + "mlp.1.bias"))
        for j in range(n_blocks):
            t = f"{gp}.1.{j}."
            o = f"est.tf{gi}.{j}."
            out.update(pack_transformer_block(g, t, o, head_dim))
    out["est.tbias.w"] = _f32(torch.cat(tb_w, 0))
    out["est.tbias.b"] = _f32(torch.cat(tb_b, 0))
    out["est.down.conv.w"] = _bf16(conv_weight_taps(g("down_blocks.0.2.weight")))     # decoder.py:348-350
    out["est.down.conv.b"] = _f32(g("down_blocks.0.2.bias"))
    out["est.up.conv.w"] = _bf16(conv_weight_taps(g("up_blocks.0.2.weight")))         # decoder.py:395-399
    out["est.up.conv.b"] = _f32(g("up_blocks.0.2.bias"))
    out["est.final.w"] = _bf16(conv_weight_taps(g("final_block.block.0.weight")))     # decoder.py:401
    out["est.final.b"] = _f32(g("final_block.block.0.bias"))
    out["est.final.ln.g"] = _f32(g("final_block.block.2.weight"))
    out["est.final.ln.b"] = _f32(g("final_block.block.2.bias"))
    out["est.proj.w"] = _bf16(conv_weight_taps(g("final_proj.weight")))               # decoder.py:402
    out["est.proj.b"] = _f32(g("final_proj.bias"))
    return out


HIFT_UPS = ((8, 16), (5, 11), (3, 7))          # (stride, kernel) of HiFTGenerator.ups, generator.py:437-451
HIFT_RES_K = (3, 7, 11)                        # resblock_kernel_sizes
HIFT_SRC_RES_K = (7, 7, 11)                    # source_resblock_kernel_sizes
HIFT_DIL = (1, 3, 5)
STFT_LD = 24                                   # 18 STFT channels padded to 24 (48-byte rows, TMA stride rule)


def _pad_k(w2d):
    n, k = w2d.shape
    kp = _ceil64(k)
    if kp == k:
        return w2d
    out = torch.zeros(n, kp, dtype=w2d.dtype)
    out[:, :k] = w2d
    return out


def source_down_weight(w: torch.Tensor) -> torch.Tensor:
    """Conv1d(18 -> C, k, stride) weight [C, 18, k] -> [C, ceil64(k * 24)], K index = tap * 24 + channel: the strided
    conv reads a window of k consecutive 24-wide STFT rows, i.e. one contiguous K-slab per output row."""
    c, cin, k = w.shape
    out = torch.zeros(c, k, STFT_LD, dtype=torch.float32)
    out[:, :, :cin] = w.permute(0, 2, 1)
    return _pad_k(out.reshape(c, k * STFT_LD))


def _pack_resblock(out, sd, src, dst, n_layers=3):
    """ResBlock (generator.py:46-101): convs1[l] (dilated), convs2[l], Snake alphas."""
    for l in range(n_layers):
        out[f"{dst}.c1.{l}.w"] = _bf16(conv_weight_taps(fold_weight_norm(sd, f"{src}.convs1.{l}")))
        out[f"{dst}.c1.{l}.b"] = _f32(sd[f"{src}.convs1.{l}.bias"])
        out[f"{dst}.c2.{l}.w"] = _bf16(conv_weight_taps(fold_weight_norm(sd, f"{src}.convs2.{l}")))
        out[f"{dst}.c2.{l}.b"] = _f32(sd[f"{src}.convs2.{l}.bias"])
        out[f"{dst}.a1.{l}"] = _f32(sd[f"{src}.activations1.{l}.alpha"])
        out[f"{dst}.a2.{l}"] = _f32(sd[f"{src}.activations2.{l}.alpha"])


def pack_hift(sd):
    """HiFTGenerator + ConvRNNF0Predictor (hifigan/generator.py:392-502, f0_predictor.py:22-53)."""
    out = {}
    # The f0 predictor runs in fp32 on the CUDA cores (csrc/f0_fp32.cu): the NSF phase integrates f0 over the whole
    # utterance (generator.py:280), and the tensor cores' truncating fp32 accumulation biases f0 by ~2e-5 relative.
    # Weights transposed to [3 * Cin, Cout], K index = tap * Cin + channel, tap j reads input row r + j - 1.
    for i, idx in enumerate((0, 2, 4, 6, 8)):                                     # f0_predictor.condnet
        w = fold_weight_norm(sd, f"f0_predictor.condnet.{idx}").float()           # [Cout, Cin, 3]
        out[f"hift.f0.c{i}.wt"] = _f32(w.permute(2, 1, 0).reshape(-1, w.shape[0]))
        out[f"hift.f0.c{i}.b"] = _f32(sd[f"f0_predictor.condnet.{idx}.bias"])
    out["hift.f0.cls.w"] = _f32(sd["f0_predictor.classifier.weight"].float().reshape(-1))     # [512]
    out["hift.f0.cls.b"] = _f32(sd["f0_predictor.classifier.bias"])
    out["hift.src.w"] = _f32(sd["m_source.l_linear.weight"].reshape(-1))          # generator.py:380
    out["hift.src.b"] = _f32(sd["m_source.l_linear.bias"])
    out["hift.pre.w"] = _bf16(conv_weight_taps(fold_weight_norm(sd, "conv_pre")))  # [512, 7 * 128]
    out["hift.pre.b"] = _f32(sd["conv_pre.bias"])
    for i, (s, k) in enumerate(HIFT_UPS):
        w = fold_weight_norm(sd, f"ups.{i}")                                       # [Cin, Cout, k], norm over dim 0 = Cin
        out[f"hift.up{i}.w"] = _bf16(convtranspose_polyphase(w, s, (k - s) // 2))
        out[f"hift.up{i}.b"] = _f32(sd[f"ups.{i}.bias"].float().repeat(s))
        out[f"hift.sd{i}.w"] = _bf16(source_down_weight(sd[f"source_downs.{i}.weight"].float()))
        out[f"hift.sd{i}.b"] = _f32(sd[f"source_downs.{i}.bias"])
        _pack_resblock(out, sd, f"source_resblocks.{i}", f"hift.srb{i}")
        for r in range(3):
            _pack_resblock(out, sd, f"resblocks.{3 * i + r}", f"hift.rb{i}.{r}")
    out["hift.post.w"] = _bf16(conv_weight_taps(fold_weight_norm(sd, "conv_post")))  # [18, 7 * 64]
    out["hift.post.b"] = _f32(sd["conv_post.bias"])
    return out


# ---- cosyvoice1 / Matcha non-causal building blocks (csrc/cv1_ops.cu) --------------------------------
def pack_cv1_resnet(sd, src, dst):
    """matcha ResnetBlock1D (decoder.py:46-61) state_dict keys under `src` -> engine tensors under `dst`.

      block{1,2}.block.0.{weight,bias}  Conv1d k3 pad 1     -> .c{1,2}.w bf16 [256, 3 * ceil64(Cin)] (tap-major K), .c{1,2}.b
      block{1,2}.block.1.{weight,bias}  GroupNorm(8, 256)   -> .gn{1,2}.g / .gn{1,2}.b
      mlp.1.{weight,bias}               Linear(time, 256)   -> .mlp.w fp32 [256, time], .mlp.b
      res_conv.{weight,bias}            Conv1d 1x1          -> .res.w bf16 [256, ceil64(Cin)], .res.b
    """
    out = {}
    for i in (1, 2):
        out[f"{dst}.c{i}.w"] = _bf16(conv_weight_taps(sd[f"{src}.block{i}.block.0.weight"].float()))
        out[f"{dst}.c{i}.b"] = _f32(sd[f"{src}.block{i}.block.0.bias"])
        out[f"{dst}.gn{i}.g"] = _f32(sd[f"{src}.block{i}.block.1.weight"])
        out[f"{dst}.gn{i}.b"] = _f32(sd[f"{src}.block{i}.block.1.bias"])
    out[f"{dst}.mlp.w"] = _f32(sd[f"{src}.mlp.1.weight"])
    out[f"{dst}.mlp.b"] = _f32(sd[f"{src}.mlp.1.bias"])
    out[f"{dst}.res.w"] = _bf16(conv_weight_taps(sd[f"{src}.res_conv.weight"].float()))
    out[f"{dst}.res.b"] = _f32(sd[f"{src}.res_conv.bias"])
    return out


def pack_cv1_block1d(sd, src, dst):
    """matcha Block1D (decoder.py:32-43) keys under `src` (block.0 = Conv1d k3 pad 1, block.1 = GroupNorm(8, 256))."""
    return {f"{dst}.c.w": _bf16(conv_weight_taps(sd[f"{src}.block.0.weight"].float())), f"{dst}.c.b": _f32(sd[f"{src}.block.0.bias"]),
            f"{dst}.gn.g": _f32(sd[f"{src}.block.1.weight"]), f"{dst}.gn.b": _f32(sd[f"{src}.block.1.bias"])}


def pack_cv1_downsample(sd, src, dst):
    """matcha Downsample1D (decoder.py:64-70) Conv1d(C, C, 3, 2, 1): tap-major [C, 3C]; the stride lives in the operand view."""
    return {f"{dst}.w": _bf16(conv_weight_taps(sd[f"{src}.conv.weight"].float())), f"{dst}.b": _f32(sd[f"{src}.conv.bias"])}


def pack_cv1_upsample(sd, src, dst):
    """matcha Upsample1D (decoder.py:120-158) ConvTranspose1d(C, C, 4, 2, 1): polyphase [2C, 3C], bias repeated per phase."""
    w = sd[f"{src}.conv.weight"].float()
    return {f"{dst}.w": _bf16(convtranspose_polyphase(w, 2, 1)), f"{dst}.b": _f32(sd[f"{src}.conv.bias"].float().repeat(2))}


def pack_encoder(sd, heads=8):
    """Token -> mu front end (csrc/engine_encoder.cu): input_embedding, spk_embed_affine_layer, encoder_proj (flow.py:180-183)
    and UpsampleConformerEncoder (upsample_encoder.py:105-242; key names: SURVEY.md appendix A)."""
    out = {}
    g = lambda k: sd[k].float()
    out["enc.emb"] = _bf16(g("input_embedding.weight"))
    out["spk.w"], out["spk.b"] = _f32(g("spk_embed_affine_layer.weight")), _f32(g("spk_embed_affine_layer.bias"))
    out["enc.proj.w"], out["enc.proj.b"] = _bf16(g("encoder_proj.weight")), _f32(g("encoder_proj.bias"))
    e = "encoder."
    for i, n in enumerate(("embed", "up_embed")):                                   # LinearNoSubsampling, subsampling.py:70-93
        out[f"enc.embed{i}.w"], out[f"enc.embed{i}.b"] = _bf16(g(e + f"{n}.out.0.weight")), _f32(g(e + f"{n}.out.0.bias"))
        out[f"enc.embed{i}.ln.g"], out[f"enc.embed{i}.ln.b"] = _f32(g(e + f"{n}.out.1.weight")), _f32(g(e + f"{n}.out.1.bias"))
    # PreLookaheadLayer (upsample_encoder.py:67-98): conv1 tap j reads token t + j, conv2 tap j reads token t + j - 2
    out["enc.pl.c1.w"] = _bf16(conv_weight_taps(g(e + "pre_lookahead_layer.conv1.weight")))
    out["enc.pl.c1.b"] = _f32(g(e + "pre_lookahead_layer.conv1.bias"))
    out["enc.pl.c2.w"] = _bf16(conv_weight_taps(g(e + "pre_lookahead_layer.conv2.weight")))
    out["enc.pl.c2.b"] = _f32(g(e + "pre_lookahead_layer.conv2.bias"))
    names = []
    for grp in ("encoders", "up_encoders"):
        i = 0
        while e + f"{grp}.{i}.norm_mha.weight" in sd:
            names.append(f"{grp}.{i}")
            i += 1
    if len(names) != 10:
        raise ValueError(f"expected 6 + 4 conformer layers, found {len(names)}")
    for l, n in enumerate(names):
        p, a, o = e + n + ".", e + n + ".self_attn.", f"enc.L{l}."
        dk = g(a + "linear_q.weight").shape[0] // heads
        sc = 1.0 / math.sqrt(dk)
        wq, bq = g(a + "linear_q.weight"), g(a + "linear_q.bias")
        u, v = g(a + "pos_bias_u").reshape(-1), g(a + "pos_bias_v").reshape(-1)    # [heads, dk] head-major == channel order
        # [(q + pos_bias_u) / sqrt(dk) | k | v | (q + pos_bias_v) / sqrt(dk)]   (attention.py:305-324)
        out[o + "qkv.w"] = _bf16(torch.cat([wq * sc, g(a + "linear_k.weight"), g(a + "linear_v.weight"), wq * sc], 0))
        out[o + "qkv.b"] = _f32(torch.cat([(bq + u) * sc, g(a + "linear_k.bias"), g(a + "linear_v.bias"), (bq + v) * sc], 0))
        out[o + "pos.wt"] = _f32(g(a + "linear_pos.weight").t())                    # [in, out], fp32 (csrc/f0_fp32.cu GEMM)
        out[o + "o.w"], out[o + "o.b"] = _bf16(g(a + "linear_out.weight")), _f32(g(a + "linear_out.bias"))
        out[o + "n1.g"], out[o + "n1.b"] = _f32(g(p + "norm_mha.weight")), _f32(g(p + "norm_mha.bias"))
        out[o + "n2.g"], out[o + "n2.b"] = _f32(g(p + "norm_ff.weight")), _f32(g(p + "norm_ff.bias"))
        out[o + "ff1.w"], out[o + "ff1.b"] = _bf16(g(p + "feed_forward.w_1.weight")), _f32(g(p + "feed_forward.w_1.bias"))
        out[o + "ff2.w"], out[o + "ff2.b"] = _bf16(g(p + "feed_forward.w_2.weight")), _f32(g(p + "feed_forward.w_2.bias"))
    # Upsample1D (upsample_encoder.py:37-64): nearest x s, left pad 2 s, conv k = 2 s + 1  ->  polyphase form over three
    # token rows: out[s q + r] = sum_d (sum_{j : (r + j) // s == d} W_j) x[q - 2 + d];  weight [(r, cout), (d, cin)]
    w = g(e + "up_layer.conv.weight")                                              # [Cout, Cin, 2 s + 1]
    s_ = (w.shape[2] - 1) // 2
    if s_ != 4:
        raise ValueError("the encoder front end is built for upsample stride 4 (token_mel_ratio 4)")
    wp = torch.zeros(s_, w.shape[0], 3, w.shape[1])
    for r in range(s_):
        for j in range(w.shape[2]):
            wp[r, :, (r + j) // s_, :] += w[:, :, j]
    out["enc.up.w"] = _bf16(wp.reshape(s_ * w.shape[0], 3 * w.shape[1]))
    out["enc.up.b"] = _f32(g(e + "up_layer.conv.bias").repeat(s_))
    out["enc.after.g"], out["enc.after.b"] = _f32(g(e + "after_norm.weight")), _f32(g(e + "after_norm.bias"))
    return out


def pack_cv1_decoder(sd, dst="cv1d", n_blocks=4, n_mid=12, head_dim=64):
    """The assembled non-causal ConditionalDecoder (cosyvoice/flow/decoder.py:88-204, channels = [256, 256]) -> engine tensors
    under `dst` for mossdec_cv1_estimator (csrc/cv1_ops.cu)."""
    out = {}
    g = lambda k: sd[k].float()
    for a, b in (("w1", "linear_1.weight"), ("b1", "linear_1.bias"), ("w2", "linear_2.weight"), ("b2", "linear_2.bias")):
        out[f"{dst}.time.{a}"] = _f32(g("time_mlp." + b))
    levels = [("d0", "down_blocks.0"), ("d1", "down_blocks.1")] + [(f"m{i}", f"mid_blocks.{i}") for i in range(n_mid)] + \
             [("u0", "up_blocks.0"), ("u1", "up_blocks.1")]
    for name, src in levels:
        out.update(pack_cv1_resnet(sd, f"{src}.0", f"{dst}.{name}.res"))
        for j in range(n_blocks):
            out.update(pack_transformer_block(g, f"{src}.1.{j}.", f"{dst}.{name}.tf.{j}.", head_dim))
    out.update(pack_cv1_downsample(sd, "down_blocks.0.2", f"{dst}.d0.down"))
    out.update(pack_cv1_upsample(sd, "up_blocks.0.2", f"{dst}.u0.up"))
    for name, src in (("d1", "down_blocks.1.2"), ("u1", "up_blocks.1.2")):                       # plain Conv1d(k3, p1) slots
        out[f"{dst}.{name}.conv.w"] = _bf16(conv_weight_taps(g(src + ".weight")))
        out[f"{dst}.{name}.conv.b"] = _f32(g(src + ".bias"))
    out.update(pack_cv1_block1d(sd, "final_block", f"{dst}.final"))
    out[f"{dst}.proj.w"] = _bf16(conv_weight_taps(g("final_proj.weight")))
    out[f"{dst}.proj.b"] = _f32(g("final_proj.bias"))
    return out

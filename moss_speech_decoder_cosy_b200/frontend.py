"""Token -> mu front end of `CausalMaskedDiffWithXvec.inference` (cosyvoice/flow/flow.py:237-272), batched.

Computes, for a ragged batch of utterances, everything the flow solver consumes:
  spks = Linear(192 -> 80)(L2-normalise(embedding))                          flow.py:250-251
  mu   = encoder_proj(UpsampleConformerEncoder(Embedding(clamp(token, 0))))  flow.py:254-265, upsample_encoder.py:244-307
  cond = prompt_feat in the first 4 P frames, zero elsewhere                 flow.py:268-270

This is SURVEY.md section 8(f) row N1 (about 1.5 % of the path's FLOPs, run once per call, not per Euler step): it is
written against torch's CUDA library ops (bf16 GEMMs via cuBLAS) and hands device buffers to libmossdec; the Euler
solve and the vocoder, i.e. everything `north_star` names, are the hand-written kernels.  There is no CPU path: the
module refuses to run without CUDA tensors.

The reference runs batch 1 only (flow.py:248).  A padded batch reproduces it per utterance because every op is
either pointwise, causal, or masked by key padding; the one look-ahead op (PreLookaheadLayer.conv1,
upsample_encoder.py:88-91) sees zeros after each utterance's last token, exactly like the reference's F.pad.
"""
import math

import torch
import torch.nn.functional as F


class EncoderWeights:
    """bf16/fp32 device copies of the flow.pt tensors used here (key names: SURVEY.md appendix A)."""

    def __init__(self, sd, device, dtype=torch.bfloat16):
        self.dtype = dtype
        f = lambda k: sd[k].to(device=device, dtype=torch.float32)
        h = lambda k: sd[k].to(device=device, dtype=dtype)
        self.emb = h("input_embedding.weight")
        self.spk_w, self.spk_b = f("spk_embed_affine_layer.weight"), f("spk_embed_affine_layer.bias")
        self.proj_w, self.proj_b = h("encoder_proj.weight"), f("encoder_proj.bias")
        e = "encoder."
        self.embed = [(h(e + f"{n}.out.0.weight"), h(e + f"{n}.out.0.bias"), f(e + f"{n}.out.1.weight"),
                       f(e + f"{n}.out.1.bias")) for n in ("embed", "up_embed")]
        self.pl = (h(e + "pre_lookahead_layer.conv1.weight"), h(e + "pre_lookahead_layer.conv1.bias"),
                   h(e + "pre_lookahead_layer.conv2.weight"), h(e + "pre_lookahead_layer.conv2.bias"))
        self.up = (h(e + "up_layer.conv.weight"), h(e + "up_layer.conv.bias"))
        self.after = (f(e + "after_norm.weight"), f(e + "after_norm.bias"))
        self.layers = []
        for grp, n in (("encoders", 6), ("up_encoders", 4)):
            ls = []
            i = 0
            while e + f"{grp}.{i}.norm_mha.weight" in sd:
                p = e + f"{grp}.{i}."
                a = p + "self_attn."
                qkv_w = torch.cat([sd[a + "linear_q.weight"], sd[a + "linear_k.weight"], sd[a + "linear_v.weight"]], 0)
                qkv_b = torch.cat([sd[a + "linear_q.bias"], sd[a + "linear_k.bias"], sd[a + "linear_v.bias"]], 0)
                ls.append(dict(
                    n1=(f(p + "norm_mha.weight"), f(p + "norm_mha.bias")),
                    n2=(f(p + "norm_ff.weight"), f(p + "norm_ff.bias")),
                    qkv_w=qkv_w.to(device=device, dtype=dtype), qkv_b=qkv_b.to(device=device, dtype=dtype),
                    out_w=h(a + "linear_out.weight"), out_b=h(a + "linear_out.bias"),
                    pos_w=h(a + "linear_pos.weight"), u=h(a + "pos_bias_u"), v=h(a + "pos_bias_v"),
                    w1=h(p + "feed_forward.w_1.weight"), b1=h(p + "feed_forward.w_1.bias"),
                    w2=h(p + "feed_forward.w_2.weight"), b2=h(p + "feed_forward.w_2.bias")))
                i += 1
            assert i == n or i > 0, grp
            self.layers.append(ls)
        self.heads = self.layers[0][0]["u"].shape[0]
        self.d_model = self.emb.shape[1]


def _rel_pos_table(T, d_model, device, dtype):
    """EspnetRelPositionalEncoding (embedding.py:224-252): rows m = 0 .. 2T-2 hold relative position T-1-m."""
    rel = torch.arange(T - 1, -T, -1, dtype=torch.float32, device=device).unsqueeze(1)
    div = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32, device=device) * -(math.log(10000.0) / d_model))
    pe = torch.zeros(2 * T - 1, d_model, device=device)
    pe[:, 0::2] = torch.sin(rel * div)
    pe[:, 1::2] = torch.cos(rel * div)
    return pe.to(dtype)


def _attn_mask(lens, T, chunk, device):
    """[B, 1, T, T] bool: key j visible to query i iff j < len_b and (chunk == 0 or j < (i // chunk + 1) * chunk)
    (mask.py:161-236 with static chunk, full left context)."""
    ar = torch.arange(T, device=device)
    m = (ar.unsqueeze(0) < lens.unsqueeze(1)).view(-1, 1, 1, T)
    if chunk > 0:
        lim = (ar // chunk + 1) * chunk
        m = m & (ar.view(1, 1, 1, T) < lim.view(1, 1, T, 1))
    return m


def _conformer_layer(x, L, pe, mask, heads):
    """ConformerEncoderLayer without macaron / conv module (encoder_layer.py:196-236) around
    RelPositionMultiHeadedAttention (attention.py:276-330)."""
    B, T, D = x.shape
    dk = D // heads
    y = F.layer_norm(x.float(), (D,), L["n1"][0], L["n1"][1], 1e-12).to(x.dtype)
    qkv = F.linear(y, L["qkv_w"], L["qkv_b"]).view(B, T, 3, heads, dk)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1].transpose(1, 2), qkv[:, :, 2].transpose(1, 2)
    p = F.linear(pe, L["pos_w"]).view(2 * T - 1, heads, dk).transpose(0, 1)       # [H, 2T-1, dk]
    q_u = (q + L["u"]).transpose(1, 2)                                              # [B, H, T, dk]
    q_v = (q + L["v"]).transpose(1, 2)
    bd = torch.matmul(q_v, p.transpose(-2, -1))                                     # [B, H, T, 2T-1]
    # rel_shift (attention.py:223-244): bd[i, j] <- raw[i, j + T - 1 - i]
    bd = F.pad(bd, (1, 0)).view(B, heads, 2 * T, T)[:, :, 1:].reshape(B, heads, T, 2 * T - 1)[..., :T]
    # softmax((q_u k^T + bd) / sqrt(dk)) v as ONE fused attention call: the positional term (and the -inf mask)
    # rides in as the additive bias, so the [B, H, T, T] content scores / probabilities are never materialised
    # (they were ~10 full passes over 0.4-0.9 GB tensors per mel-rate layer at batch 64).
    Tp = (T + 7) // 8 * 8                                                       # bias row stride: multiple of 8
    bias = torch.empty(B, heads, T, Tp, dtype=x.dtype, device=x.device)[..., :T]
    torch.mul(bd, 1.0 / math.sqrt(dk), out=bias)
    bias.masked_fill_(~mask, float("-inf"))
    o = F.scaled_dot_product_attention(q_u, k, v, attn_mask=bias, scale=1.0 / math.sqrt(dk))
    o = o.transpose(1, 2).reshape(B, T, D)
    x = x + F.linear(o, L["out_w"], L["out_b"])
    y = F.layer_norm(x.float(), (D,), L["n2"][0], L["n2"][1], 1e-12).to(x.dtype)
    return x + F.linear(F.silu(F.linear(y, L["w1"], L["b1"])), L["w2"], L["b2"])


@torch.inference_mode()
def encode(W: EncoderWeights, tokens, lens, ctx_len, chunk_tokens=0, max_batch_elems=1 << 28):
    """tokens: int64 [B, Lmax + ctx_len] (prompt + new tokens, right padded; the last `ctx_len` VALID tokens of each
    row are look-ahead context when ctx_len > 0, flow.py:262-263); lens: int64 [B] tokens INCLUDING context.
    Returns mu [B, 4 * (Lmax), 80] fp32 (rows beyond 4 * (lens - ctx_len) are padding) and the main lengths."""
    if not tokens.is_cuda:
        raise RuntimeError("mossdec front end runs on CUDA tensors only (no CPU path)")
    B, Ltot = tokens.shape
    main = lens - ctx_len
    T = int(main.max())
    # sub-batch so the [B, H, T, 2T-1] score tensors of the mel-rate layers stay bounded
    per = max(1, int(max_batch_elems // max(1, W.heads * (4 * T) * (8 * T))))
    if per < B:
        outs = [encode(W, tokens[i:i + per], lens[i:i + per], ctx_len, chunk_tokens, max_batch_elems)[0]
                for i in range(0, B, per)]
        Tm = max(o.shape[1] for o in outs)
        return torch.cat([F.pad(o, (0, 0, 0, Tm - o.shape[1])) for o in outs], 0), main
    dev, dt, D = tokens.device, W.dtype, W.d_model
    ar = torch.arange(Ltot, device=dev)
    valid = (ar.unsqueeze(0) < lens.unsqueeze(1)).unsqueeze(-1)                      # incl. context tokens
    x = W.emb[tokens.clamp(min=0)] * valid.to(dt)                                    # flow.py:256
    ew, eb, lg, lb = W.embed[0]
    x = F.layer_norm(F.linear(x, ew, eb).float(), (D,), lg, lb, 1e-5).to(dt) * math.sqrt(D)   # subsampling.py:85-93, embedding.py:262
    x = x * valid.to(dt)                      # zeros after each utterance's last (context) token == the reference's F.pad
    x = F.pad(x, (0, 0, 0, T + 3 - min(Ltot, T + 3)))[:, :T + 3]
    inputs = x[:, :T]
    c1w, c1b, c2w, c2b = W.pl
    y = F.leaky_relu(F.conv1d(x.transpose(1, 2), c1w, c1b))                          # upsample_encoder.py:88-91
    y = F.conv1d(F.pad(y, (2, 0)), c2w, c2b).transpose(1, 2)                         # :93-94
    x = y + inputs                                                                   # :98
    mask = _attn_mask(main, T, chunk_tokens, dev)
    pe = _rel_pos_table(T, D, dev, dt)
    for L in W.layers[0]:
        x = _conformer_layer(x, L, pe, mask, W.heads)
    # Upsample1D (upsample_encoder.py:60-64): nearest x4, left pad 8, conv k9
    uw, ub = W.up
    x = x.transpose(1, 2).repeat_interleave(4, dim=2)
    x = F.conv1d(F.pad(x, (8, 0)), uw, ub).transpose(1, 2)
    T4 = 4 * T
    ew, eb, lg, lb = W.embed[1]
    x = F.layer_norm(F.linear(x, ew, eb).float(), (D,), lg, lb, 1e-5).to(dt) * math.sqrt(D)
    mask = _attn_mask(main * 4, T4, chunk_tokens * 4, dev)
    pe = _rel_pos_table(T4, D, dev, dt)
    for L in W.layers[1]:
        x = _conformer_layer(x, L, pe, mask, W.heads)
    x = F.layer_norm(x.float(), (D,), W.after[0], W.after[1], 1e-5).to(dt)
    mu = F.linear(x, W.proj_w).float() + W.proj_b                                     # flow.py:265
    return mu, main


@torch.inference_mode()
def speaker_projection(W: EncoderWeights, embedding):
    """flow.py:250-251."""
    e = F.normalize(embedding.float(), dim=1)
    return F.linear(e, W.spk_w, W.spk_b)

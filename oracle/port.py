"""TEST INFRASTRUCTURE ONLY -- CPU restatement (plain PyTorch fp32 ops) of the reference's token->waveform path.

This is the travelling oracle: `/root/reference` does not exist on the GPU box, so tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / `--impl reference` leg use THIS restatement there.  It is pinned against outputs of the
unmodified reference (tests/golden/*.npz, produced in the build container by oracle/make_golden*.py from the
reference's own classes): see tests/test_oracle_cpu.py.  Nothing in the product imports it.

Every function cites the reference code it restates (paths relative to GLM_modules/).  Weights are read from
reference-format state_dicts (flow.pt / hift.pt key names).  Nothing here imports the product package: the encoder front
end (`encoder_forward`) is its own restatement of cosyvoice/transformer/upsample_encoder.py:244-307, pinned to taps of the
reference's UpsampleConformerEncoder (tests/golden/encoder_taps.npz).  All functions follow the device of their inputs, so
the same restatement also serves as the "eager PyTorch on the GPU" line of bench.py.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------------ estimator
def _sinusoidal(t, dim=320, scale=1000.0):
    """SinusoidalPosEmb, matcha/models/components/decoder.py:20-29."""
    half = dim // 2
    emb = math.log(10000) / (half - 1)
    emb = torch.exp(torch.arange(half, dtype=torch.float32, device=t.device) * -emb)
    emb = scale * t.unsqueeze(1) * emb.unsqueeze(0)
    return torch.cat([emb.sin(), emb.cos()], dim=-1)


def _causal_block(x, mask, sd, p):
    """CausalBlock1D, cosyvoice/flow/decoder.py:65-78: CausalConv1d(k3, left pad 2) -> LayerNorm(C) -> Mish, masked."""
    y = F.conv1d(F.pad(x * mask, (2, 0)), sd[p + "block.0.weight"], sd[p + "block.0.bias"])
    y = F.layer_norm(y.transpose(1, 2), (y.shape[1],), sd[p + "block.2.weight"], sd[p + "block.2.bias"]).transpose(1, 2)
    return F.mish(y) * mask


def _resnet(x, mask, temb, sd, p):
    """CausalResnetBlock1D, decoder.py:81-85 + matcha decoder.py:56-61."""
    h = _causal_block(x, mask, sd, p + "block1.")
    h = h + F.linear(F.mish(temb), sd[p + "mlp.1.weight"], sd[p + "mlp.1.bias"]).unsqueeze(-1)
    h = _causal_block(h, mask, sd, p + "block2.")
    return h + F.conv1d(x * mask, sd[p + "res_conv.weight"], sd[p + "res_conv.bias"])


def _transformer(x, bias, sd, p, heads=8):
    """BasicTransformerBlock, matcha transformer.py:262-314 with diffusers Attention / GELU (see ref_shims.py)."""
    B, T, C = x.shape
    y = F.layer_norm(x, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"])
    q = F.linear(y, sd[p + "attn1.to_q.weight"]).view(B, T, heads, -1).transpose(1, 2)
    k = F.linear(y, sd[p + "attn1.to_k.weight"]).view(B, T, heads, -1).transpose(1, 2)
    v = F.linear(y, sd[p + "attn1.to_v.weight"]).view(B, T, heads, -1).transpose(1, 2)
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=bias.unsqueeze(1))
    o = o.transpose(1, 2).reshape(B, T, -1)
    x = x + F.linear(o, sd[p + "attn1.to_out.0.weight"], sd[p + "attn1.to_out.0.bias"])
    y = F.layer_norm(x, (C,), sd[p + "norm3.weight"], sd[p + "norm3.bias"])
    y = F.gelu(F.linear(y, sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.0.proj.bias"]))
    return x + F.linear(y, sd[p + "ff.net.2.weight"], sd[p + "ff.net.2.bias"])


def _attn_bias(mask, chunk):
    """add_optional_chunk_mask + mask_to_bias, cosyvoice/utils/mask.py:161-236, common.py:160-168."""
    B, _, T = mask.shape
    m = mask.bool().expand(B, T, T).clone()
    if chunk > 0:
        ar = torch.arange(T, device=mask.device)
        m &= (ar.unsqueeze(0) < ((ar // chunk + 1) * chunk).unsqueeze(1)).unsqueeze(0)
    return (1.0 - m.float()) * -1.0e10


def estimator(sd, x, mask, mu, t, spks, cond, chunk=0, prefix="decoder.estimator.", n_mid=12, n_blocks=4):
    """CausalConditionalDecoder.forward, cosyvoice/flow/decoder.py:405-494 (channels=[256])."""
    sd = {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}
    temb = _sinusoidal(t)
    temb = F.linear(F.silu(F.linear(temb, sd["time_mlp.linear_1.weight"], sd["time_mlp.linear_1.bias"])),
                    sd["time_mlp.linear_2.weight"], sd["time_mlp.linear_2.bias"])
    x = torch.cat([x, mu, spks.unsqueeze(-1).expand(-1, -1, x.shape[-1]), cond], dim=1)   # decoder.py:425-431
    bias = _attn_bias(mask, chunk)

    def group(x, p):
        x = _resnet(x, mask, temb, sd, p + "0.")
        x = x.transpose(1, 2)
        for j in range(n_blocks):
            x = _transformer(x, bias, sd, p + f"1.{j}.")
        return x.transpose(1, 2)

    x = group(x, "down_blocks.0.")
    skip = x
    x = F.conv1d(F.pad(x * mask, (2, 0)), sd["down_blocks.0.2.weight"], sd["down_blocks.0.2.bias"])   # decoder.py:348-350
    for i in range(n_mid):
        x = group(x, f"mid_blocks.{i}.")
    x = group(torch.cat([x, skip], dim=1), "up_blocks.0.")                                            # decoder.py:477
    x = F.conv1d(F.pad(x * mask, (2, 0)), sd["up_blocks.0.2.weight"], sd["up_blocks.0.2.bias"])
    x = _causal_block(x, mask, sd, "final_block.")
    return F.conv1d(x * mask, sd["final_proj.weight"], sd["final_proj.bias"]) * mask


def solve_euler(sd, mu, spks, cond, noise, n_steps=10, cfg_rate=0.7, chunk=0):
    """CausalConditionalCFM.forward + ConditionalCFM.solve_euler, cosyvoice/flow/flow_matching.py:206-230, 76-127.
    mu, cond: [1, 80, T]; spks [1, 80]; noise [80, >=T] (the fixed bank)."""
    T = mu.shape[2]
    dev = mu.device
    x = noise[:, :T].unsqueeze(0).clone().to(dev)
    mask = torch.ones(1, 1, T, device=dev)
    t_span = torch.linspace(0, 1, n_steps + 1, device=dev)
    t_span = 1 - torch.cos(t_span * 0.5 * torch.pi)
    t, dt = t_span[0], t_span[1] - t_span[0]
    for step in range(1, n_steps + 1):
        x_in = torch.cat([x, x], 0)
        mu_in = torch.cat([mu, torch.zeros_like(mu)], 0)
        spk_in = torch.cat([spks, torch.zeros_like(spks)], 0)
        cond_in = torch.cat([cond, torch.zeros_like(cond)], 0)
        d = estimator(sd, x_in, mask.expand(2, -1, -1), mu_in, t.expand(2), spk_in, cond_in, chunk)
        d = (1.0 + cfg_rate) * d[:1] - cfg_rate * d[1:]
        x = x + dt * d
        t = t + dt
        if step < n_steps:
            dt = t_span[step + 1] - t
    return x.float()


# ------------------------------------------------------------------------------------------------ HiFT
def _wn(sd, p):
    """weight_norm fold, w = g * v / ||v|| over all dims but 0 (cosyvoice/hifigan/generator.py:26-29)."""
    if p + ".parametrizations.weight.original0" in sd:
        g, v = sd[p + ".parametrizations.weight.original0"], sd[p + ".parametrizations.weight.original1"]
        return g * v / v.flatten(1).norm(dim=1).view(-1, *([1] * (v.ndim - 1)))
    return sd[p + ".weight"]


def _snake(x, alpha):
    """Snake, cosyvoice/transformer/activation.py:73-84."""
    a = alpha.view(1, -1, 1)
    return x + (1.0 / (a + 1e-9)) * torch.sin(x * a) ** 2


def _resblock(sd, p, x, k, dils=(1, 3, 5)):
    """ResBlock.forward, generator.py:94-101."""
    for i, d in enumerate(dils):
        xt = _snake(x, sd[p + f".activations1.{i}.alpha"])
        xt = F.conv1d(xt, _wn(sd, p + f".convs1.{i}"), sd[p + f".convs1.{i}.bias"], padding=(k * d - d) // 2, dilation=d)
        xt = _snake(xt, sd[p + f".activations2.{i}.alpha"])
        xt = F.conv1d(xt, _wn(sd, p + f".convs2.{i}"), sd[p + f".convs2.{i}.bias"], padding=(k - 1) // 2)
        x = xt + x
    return x


def f0_predictor(sd, mel):
    """ConvRNNF0Predictor.forward, cosyvoice/hifigan/f0_predictor.py:55-60.  mel [B, 80, T] -> f0 [B, T]."""
    x = mel
    for idx in (0, 2, 4, 6, 8):
        p = f"f0_predictor.condnet.{idx}"
        x = F.elu(F.conv1d(x, _wn(sd, p), sd[p + ".bias"], padding=1))
    return torch.abs(F.linear(x.transpose(1, 2), sd["f0_predictor.classifier.weight"],
                              sd["f0_predictor.classifier.bias"]).squeeze(-1))


def nsf_source(sd, f0, noise, sr=24000.0, up=480, sine_amp=0.1, noise_std=0.003, thr=10.0):
    """f0_upsamp + SourceModuleHnNSF2 / SineGen2, generator.py:437,575, 261-283, 314-339, 383-384.
    f0 [T]; noise [480 T, 9] = the randn_like draw of generator.py:334 (injected).  Returns s [480 T]."""
    T = f0.shape[0]
    h = torch.arange(1, 10, dtype=torch.float32, device=f0.device)
    rad = ((f0.unsqueeze(1) * h) / sr) % 1                      # per frame: the x480 / (1/480) resampling pair is the identity
    phase = torch.cumsum(rad.double(), dim=0).float() * 2 * np.pi       # torch CPU cumsum accumulates in double
    ph = F.interpolate((phase * up).t().unsqueeze(0), scale_factor=float(up), mode="linear")[0].t()   # [480 T, 9]
    sines = torch.sin(ph) * sine_amp
    uv = (f0 > thr).float().repeat_interleave(up).unsqueeze(1)
    namp = uv * noise_std + (1 - uv) * sine_amp / 3
    sw = sines * uv + namp * noise
    return torch.tanh(F.linear(sw, sd["m_source.l_linear.weight"], sd["m_source.l_linear.bias"]))[:, 0]


def hift_decode(sd, mel, s, lrelu=0.1, limit=0.99):
    """HiFTGenerator.decode, generator.py:520-552.  mel [1, 80, T], s [1, 480 T] -> wav [1, 480 T]."""
    win = torch.from_numpy(__import__("scipy.signal", fromlist=["get_window"]).get_window("hann", 16, fftbins=True)
                           .astype(np.float32)).to(mel.device)
    spec = torch.view_as_real(torch.stft(s, 16, 4, 16, window=win, return_complex=True))
    s_stft = torch.cat([spec[..., 0], spec[..., 1]], dim=1)
    x = F.conv1d(mel, _wn(sd, "conv_pre"), sd["conv_pre.bias"], padding=3)
    ups, res_k, src_k = ((8, 16), (5, 11), (3, 7)), (3, 7, 11), (7, 7, 11)
    sdown = ((30, 15, 7), (6, 3, 1), (1, 1, 0))
    for i, (u, k) in enumerate(ups):
        x = F.leaky_relu(x, lrelu)
        x = F.conv_transpose1d(x, _wn(sd, f"ups.{i}"), sd[f"ups.{i}.bias"], stride=u, padding=(k - u) // 2)
        if i == 2:
            x = F.pad(x, (1, 0), mode="reflect")
        kk, st, pd = sdown[i]
        si = F.conv1d(s_stft, sd[f"source_downs.{i}.weight"], sd[f"source_downs.{i}.bias"], stride=st, padding=pd)
        si = _resblock(sd, f"source_resblocks.{i}", si, src_k[i])
        x = x + si
        xs = None
        for j in range(3):
            r = _resblock(sd, f"resblocks.{3 * i + j}", x, res_k[j])
            xs = r if xs is None else xs + r
        x = xs / 3
    x = F.leaky_relu(x)
    x = F.conv1d(x, _wn(sd, "conv_post"), sd["conv_post.bias"], padding=3)
    mag = torch.clip(torch.exp(x[:, :9]), max=1e2)
    ph = torch.sin(x[:, 9:])
    wav = torch.istft(torch.complex(mag * torch.cos(ph), mag * torch.sin(ph)), 16, 4, 16, window=win)
    return torch.clamp(wav, -limit, limit)


def hift_inference(sd, mel, noise, cache_source=None):
    """HiFTGenerator.inference, generator.py:570-582, with the NSF noise injected.  mel [80, T] -> (wav, source)."""
    f0 = f0_predictor(sd, mel.unsqueeze(0))[0]
    s = nsf_source(sd, f0, noise)
    if cache_source is not None and cache_source.numel():
        s = s.clone()
        s[:cache_source.numel()] = cache_source.reshape(-1)
    wav = hift_decode(sd, mel.unsqueeze(0), s.unsqueeze(0))[0]
    return wav, s


# ------------------------------------------------------------------------------------------------ token2wav
def noise_bank():
    """set_all_random_seed(0); torch.randn([1, 80, 15000]) (flow_matching.py:202-203)."""
    g = torch.Generator().manual_seed(0)
    return torch.randn([1, 80, 15000], generator=g)[0]


# ------------------------------------------------------------------------------------------------ encoder front end
def _rel_pos_table(T, d_model, device):
    """EspnetRelPositionalEncoding.extend_pe / position_encoding, cosyvoice/transformer/embedding.py:224-302: rows
    m = 0 .. 2T-2 hold the sinusoid of relative position T-1-m."""
    rel = torch.arange(T - 1, -T, -1, dtype=torch.float32, device=device).unsqueeze(1)
    div = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32, device=device) * -(math.log(10000.0) / d_model))
    pe = torch.zeros(2 * T - 1, d_model, device=device)
    pe[:, 0::2] = torch.sin(rel * div)
    pe[:, 1::2] = torch.cos(rel * div)
    return pe


def _enc_mask(T, chunk, device):
    """add_optional_chunk_mask with a static chunk and full left context (cosyvoice/utils/mask.py:161-236); batch 1, no
    padding.  [T, T] bool, key j visible to query i iff chunk == 0 or j < (i // chunk + 1) * chunk."""
    if chunk <= 0:
        return torch.ones(T, T, dtype=torch.bool, device=device)
    ar = torch.arange(T, device=device)
    return ar.unsqueeze(0) < ((ar // chunk + 1) * chunk).unsqueeze(1)


def _conformer_layer(x, sd, p, pe, mask, heads=8):
    """ConformerEncoderLayer without macaron / conv module, pre-norm (encoder_layer.py:196-236) around
    RelPositionMultiHeadedAttention (attention.py:215-330; LayerNorm eps 1e-12, encoder_layer.py:71-73)."""
    T, D = x.shape
    dk = D // heads
    a = p + "self_attn."
    y = F.layer_norm(x, (D,), sd[p + "norm_mha.weight"], sd[p + "norm_mha.bias"], 1e-12)
    q = F.linear(y, sd[a + "linear_q.weight"], sd[a + "linear_q.bias"]).view(T, heads, dk)
    k = F.linear(y, sd[a + "linear_k.weight"], sd[a + "linear_k.bias"]).view(T, heads, dk).transpose(0, 1)
    v = F.linear(y, sd[a + "linear_v.weight"], sd[a + "linear_v.bias"]).view(T, heads, dk).transpose(0, 1)
    pp = F.linear(pe, sd[a + "linear_pos.weight"]).view(2 * T - 1, heads, dk).transpose(0, 1)     # [H, 2T-1, dk]
    q_u = (q + sd[a + "pos_bias_u"]).transpose(0, 1)                                               # [H, T, dk]
    q_v = (q + sd[a + "pos_bias_v"]).transpose(0, 1)
    ac = torch.matmul(q_u, k.transpose(-2, -1))                                                    # [H, T, T]
    bd = torch.matmul(q_v, pp.transpose(-2, -1))                                                   # [H, T, 2T-1]
    # rel_shift (attention.py:223-244): bd[i, j] <- raw[i, j + T - 1 - i]
    bd = F.pad(bd, (1, 0)).view(heads, 2 * T, T)[:, 1:].reshape(heads, T, 2 * T - 1)[..., :T]
    scores = (ac + bd) / math.sqrt(dk)
    scores = scores.masked_fill(~mask, float("-inf"))
    attn = torch.softmax(scores, dim=-1).masked_fill(~mask, 0.0)                                   # attention.py:120-135
    o = torch.matmul(attn, v).transpose(0, 1).reshape(T, D)
    x = x + F.linear(o, sd[a + "linear_out.weight"], sd[a + "linear_out.bias"])
    y = F.layer_norm(x, (D,), sd[p + "norm_ff.weight"], sd[p + "norm_ff.bias"], 1e-12)
    f = p + "feed_forward."
    return x + F.linear(F.silu(F.linear(y, sd[f + "w_1.weight"], sd[f + "w_1.bias"])), sd[f + "w_2.weight"], sd[f + "w_2.bias"])


def encoder_forward(sd, tokens, ctx_len=0, chunk_tokens=0, taps=None):
    """`input_embedding` + UpsampleConformerEncoder.forward + `encoder_proj` (flow.py:254-265,
    upsample_encoder.py:244-307) for ONE utterance.  tokens: int64 [L] = prompt + new tokens; the last `ctx_len` of them
    are the look-ahead context (flow.py:262-263, 3 when finalize is False).  Returns mu [4 * (L - ctx_len), 80].
    `taps` (dict) receives intermediate activations for pinning."""
    dev = tokens.device
    D = sd["input_embedding.weight"].shape[1]
    e = "encoder."
    x = sd["input_embedding.weight"][tokens.clamp(min=0)]                                          # flow.py:256
    # LinearNoSubsampling + EspnetRelPositionalEncoding (subsampling.py:85-93, embedding.py:262); the context tokens go
    # through the same `embed` (upsample_encoder.py:283-286; the offset only moves pos_emb, which is not used for them)
    x = F.layer_norm(F.linear(x, sd[e + "embed.out.0.weight"], sd[e + "embed.out.0.bias"]), (D,),
                     sd[e + "embed.out.1.weight"], sd[e + "embed.out.1.bias"], 1e-5) * math.sqrt(D)
    T = tokens.shape[0] - ctx_len
    inputs = x[:T]
    # PreLookaheadLayer (upsample_encoder.py:81-98): conv1 k4 over [inputs | context | zero pad], leaky_relu, causal conv2 k3
    y = F.pad(x.t().unsqueeze(0), (0, 3 - ctx_len))
    y = F.leaky_relu(F.conv1d(y, sd[e + "pre_lookahead_layer.conv1.weight"], sd[e + "pre_lookahead_layer.conv1.bias"]))
    y = F.conv1d(F.pad(y, (2, 0)), sd[e + "pre_lookahead_layer.conv2.weight"], sd[e + "pre_lookahead_layer.conv2.bias"])
    x = y[0].t() + inputs
    if taps is not None:
        taps["pre_lookahead"] = x.clone()
    mask = _enc_mask(T, chunk_tokens, dev)
    pe = _rel_pos_table(T, D, dev)
    i = 0
    while e + f"encoders.{i}.norm_mha.weight" in sd:
        x = _conformer_layer(x, sd, e + f"encoders.{i}.", pe, mask)
        i += 1
    if taps is not None:
        taps["encoders"] = x.clone()
    # Upsample1D (upsample_encoder.py:37-64): nearest x stride, left pad 2 * stride, conv k = 2 * stride + 1
    stride = (sd[e + "up_layer.conv.weight"].shape[2] - 1) // 2
    y = x.t().unsqueeze(0).repeat_interleave(stride, dim=2)
    y = F.conv1d(F.pad(y, (2 * stride, 0)), sd[e + "up_layer.conv.weight"], sd[e + "up_layer.conv.bias"])
    x = y[0].t()
    if taps is not None:
        taps["up_layer"] = x.clone()
    T2 = stride * T
    x = F.layer_norm(F.linear(x, sd[e + "up_embed.out.0.weight"], sd[e + "up_embed.out.0.bias"]), (D,),
                     sd[e + "up_embed.out.1.weight"], sd[e + "up_embed.out.1.bias"], 1e-5) * math.sqrt(D)
    mask = _enc_mask(T2, chunk_tokens * stride, dev)
    pe = _rel_pos_table(T2, D, dev)
    i = 0
    while e + f"up_encoders.{i}.norm_mha.weight" in sd:
        x = _conformer_layer(x, sd, e + f"up_encoders.{i}.", pe, mask)
        i += 1
    x = F.layer_norm(x, (D,), sd[e + "after_norm.weight"], sd[e + "after_norm.bias"], 1e-5)
    if taps is not None:
        taps["final"] = x.clone()
    return F.linear(x, sd["encoder_proj.weight"], sd["encoder_proj.bias"])                       # flow.py:265


def speaker_projection(sd, embedding):
    """flow.py:250-251: L2-normalise, then spk_embed_affine_layer."""
    return F.linear(F.normalize(embedding.float(), dim=1), sd["spk_embed_affine_layer.weight"],
                    sd["spk_embed_affine_layer.bias"])


class PortDecoder:
    """AudioDecoder.token2wav (flow_inference.py:94-123) restated in fp32 (batch 1, like the reference), on the device of
    the weights (CPU for the oracle / CPU baseline, CUDA for bench.py's eager-PyTorch-on-the-GPU line)."""

    def __init__(self, flow_sd, hift_sd, static_chunk_tokens=25, device="cpu"):
        self.device = torch.device(device)
        self.fsd = {k: v.float().to(self.device) for k, v in flow_sd.items()}
        self.hsd = {k: v.float().to(self.device) for k, v in hift_sd.items()}
        self.bank = noise_bank().to(self.device)
        self.chunk = static_chunk_tokens

    @torch.inference_mode()
    def flow(self, token, prompt_token, prompt_feat, embedding, streaming=False, finalize=True):
        """CausalMaskedDiffWithXvec.inference, flow.py:237-283 -> (mel [80, 4N], full [80, T])."""
        dev = self.device
        tok = torch.cat([prompt_token.reshape(-1).long(), token.reshape(-1).long()]).to(dev)
        ctx = 0 if finalize else 3
        mu = encoder_forward(self.fsd, tok, ctx, self.chunk if streaming else 0).t().unsqueeze(0)
        T, F1 = mu.shape[2], prompt_feat.shape[1]
        cond = torch.zeros(1, 80, T, device=dev)
        cond[:, :, :F1] = prompt_feat.to(dev).transpose(1, 2)
        spks = speaker_projection(self.fsd, embedding.to(dev))
        full = solve_euler(self.fsd, mu, spks, cond, self.bank, chunk=self.chunk * 4 if streaming else 0)[0]
        return full[:, F1:], full

    @torch.inference_mode()
    def token2wav(self, token, prompt_token=None, prompt_feat=None, embedding=None, noise=None):
        prompt_token = torch.zeros(1, 0, dtype=torch.int32) if prompt_token is None else prompt_token
        prompt_feat = torch.zeros(1, 0, 80) if prompt_feat is None else prompt_feat
        embedding = torch.zeros(1, 192) if embedding is None else embedding
        mel, _ = self.flow(token, prompt_token, prompt_feat, embedding)
        if noise is None:
            noise = torch.randn(480 * mel.shape[1], 9)
        wav, _ = hift_inference(self.hsd, mel, noise.to(self.device))
        return wav.unsqueeze(0), mel.unsqueeze(0)


# ------------------------------------------------------------------------------------------------ prompt mel front end
def slaney_mel_filterbank(sr, n_fft, n_mels, fmin, fmax):
    """librosa.filters.mel(sr=, n_fft=, n_mels=, fmin=, fmax=) with its defaults htk=False, norm='slaney' (librosa 0.10 /
    0.11, pinned by the reference's requirements; call site matcha/utils/audio.py:53).  Published algorithm restated:
    Slaney mel scale (linear 200/3 Hz per mel below 1 kHz, log with step log(6.4)/27 above), n_mels + 2 equally spaced
    mel points, triangular weights from the two neighbouring ramps, each filter scaled by 2 / (f[i+2] - f[i]).
    Returns float32 [n_mels, 1 + n_fft // 2]."""
    f_sp, min_log_hz = 200.0 / 3, 1000.0
    min_log_mel, logstep = min_log_hz / f_sp, math.log(6.4) / 27.0

    def hz_to_mel(f):
        return min_log_mel + math.log(f / min_log_hz) / logstep if f >= min_log_hz else f / f_sp

    def mel_to_hz(m):
        return min_log_hz * math.exp(logstep * (m - min_log_mel)) if m >= min_log_mel else f_sp * m
    lo, hi = hz_to_mel(fmin), hz_to_mel(fmax)
    mel_f = np.array([mel_to_hz(lo + (hi - lo) * i / (n_mels + 1)) for i in range(n_mels + 2)], dtype=np.float64)
    freqs = np.linspace(0.0, sr / 2.0, 1 + n_fft // 2)
    w = np.zeros((n_mels, freqs.shape[0]), dtype=np.float64)
    for i in range(n_mels):
        lower = (freqs - mel_f[i]) / (mel_f[i + 1] - mel_f[i])
        upper = (mel_f[i + 2] - freqs) / (mel_f[i + 2] - mel_f[i + 1])
        w[i] = np.maximum(0.0, np.minimum(lower, upper)) * (2.0 / (mel_f[i + 2] - mel_f[i]))
    return w.astype(np.float32)


def mel_spectrogram(y, n_fft=1920, num_mels=80, sampling_rate=24000, hop_size=480, win_size=1920, fmin=0, fmax=8000):
    """matcha mel_spectrogram, third_party/Matcha-TTS/matcha/utils/audio.py:45-82 (center=False): y [1, L] -> [1, 80, T]."""
    fb = torch.from_numpy(slaney_mel_filterbank(sampling_rate, n_fft, num_mels, fmin, fmax)).to(y.device)
    pad = int((n_fft - hop_size) / 2)
    y = F.pad(y.unsqueeze(1), (pad, pad), mode="reflect").squeeze(1)
    spec = torch.view_as_real(torch.stft(y, n_fft, hop_length=hop_size, win_length=win_size,
                                         window=torch.hann_window(win_size).to(y.device), center=False, pad_mode="reflect",
                                         normalized=False, onesided=True, return_complex=True))
    spec = torch.sqrt(spec.pow(2).sum(-1) + 1e-9)
    return torch.log(torch.clamp(torch.matmul(fb, spec), min=1e-5))


# ------------------------------------------------------------------ cosyvoice1 / Matcha non-causal blocks
def cv1_block1d(x, mask, sd, p, groups=8):
    """Block1D, third_party/Matcha-TTS/matcha/models/components/decoder.py:32-43: Conv1d(k3, pad 1) -> GroupNorm(8) -> Mish."""
    y = F.conv1d(x * mask, sd[p + "block.0.weight"], sd[p + "block.0.bias"], padding=1)
    y = F.group_norm(y, groups, sd[p + "block.1.weight"], sd[p + "block.1.bias"], eps=1e-5)
    return F.mish(y) * mask


def cv1_resnet_block(sd, p, x, mask, temb):
    """ResnetBlock1D, matcha decoder.py:46-61."""
    h = cv1_block1d(x, mask, sd, p + "block1.")
    h = h + F.linear(F.mish(temb), sd[p + "mlp.1.weight"], sd[p + "mlp.1.bias"]).unsqueeze(-1)
    h = cv1_block1d(h, mask, sd, p + "block2.")
    return h + F.conv1d(x * mask, sd[p + "res_conv.weight"], sd[p + "res_conv.bias"])


def cv1_downsample(sd, p, x):
    """Downsample1D, matcha decoder.py:64-70: Conv1d(C, C, 3, stride 2, padding 1)."""
    return F.conv1d(x, sd[p + "conv.weight"], sd[p + "conv.bias"], stride=2, padding=1)


def cv1_upsample(sd, p, x):
    """Upsample1D (use_conv_transpose), matcha decoder.py:120-158: ConvTranspose1d(C, C, 4, stride 2, padding 1)."""
    return F.conv_transpose1d(x, sd[p + "conv.weight"], sd[p + "conv.bias"], stride=2, padding=1)


def cv1_decoder(sd, x, mask, mu, t, spks, cond, n_blocks=4, n_mid=12):
    """ConditionalDecoder.forward of the NON-causal decoder, cosyvoice/flow/decoder.py:206-291 (cosyvoice1/flow/decoder.py:151-222
    is the same graph), channels = [256, 256]: two resolutions, skip concatenation, GroupNorm resnets, key-padding attention."""
    temb = _sinusoidal(t)
    temb = F.linear(F.silu(F.linear(temb, sd["time_mlp.linear_1.weight"], sd["time_mlp.linear_1.bias"])),
                    sd["time_mlp.linear_2.weight"], sd["time_mlp.linear_2.bias"])
    x = torch.cat([x, mu, spks.unsqueeze(-1).expand(-1, -1, x.shape[-1]), cond], dim=1)

    def level(x, m, p):
        x = cv1_resnet_block(sd, p + ".0.", x, m, temb)
        bias = _attn_bias(m, 0)                                       # add_optional_chunk_mask(..., 0, 0, -1) + mask_to_bias
        x = x.transpose(1, 2)
        for j in range(n_blocks):
            x = _transformer(x, bias, sd, p + f".1.{j}.")
        return x.transpose(1, 2)

    m0 = mask
    x = level(x, m0, "down_blocks.0")
    skip0 = x
    x = cv1_downsample(sd, "down_blocks.0.2.", x * m0)
    m1 = m0[:, :, ::2]
    x = level(x, m1, "down_blocks.1")
    skip1 = x
    x = F.conv1d(x * m1, sd["down_blocks.1.2.weight"], sd["down_blocks.1.2.bias"], padding=1)
    for i in range(n_mid):
        x = level(x, m1, f"mid_blocks.{i}")
    x = level(torch.cat([x[:, :, :skip1.shape[-1]], skip1], dim=1), m1, "up_blocks.0")
    x = cv1_upsample(sd, "up_blocks.0.2.", x * m1)
    x = level(torch.cat([x[:, :, :skip0.shape[-1]], skip0], dim=1), m0, "up_blocks.1")
    x = F.conv1d(x * m0, sd["up_blocks.1.2.weight"], sd["up_blocks.1.2.bias"], padding=1)
    x = cv1_block1d(x, m0, sd, "final_block.")
    return F.conv1d(x * m0, sd["final_proj.weight"], sd["final_proj.bias"]) * m0

"""TEST INFRASTRUCTURE ONLY -- CPU restatement (plain PyTorch fp32 ops) of the reference's token->waveform path.

This is the travelling oracle: `/root/reference` does not exist on the GPU box, so tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / `--impl reference` leg use THIS restatement there.  It is pinned against outputs of the
unmodified reference (tests/golden/*.npz, produced in the build container by oracle/make_golden*.py from the
reference's own classes): see tests/test_oracle_cpu.py.  Nothing in the product imports it.

Every function cites the reference code it restates (paths relative to GLM_modules/).  Weights are read from
reference-format state_dicts (flow.pt / hift.pt key names).  The encoder front end reuses
moss_speech_decoder_cosy_b200.frontend.encode in fp32 (itself a restatement validated to 3e-6 against
cosyvoice/transformer/upsample_encoder.py:244-307).
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
    emb = torch.exp(torch.arange(half, dtype=torch.float32) * -emb)
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
        ar = torch.arange(T)
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
    x = noise[:, :T].unsqueeze(0).clone()
    mask = torch.ones(1, 1, T)
    t_span = torch.linspace(0, 1, n_steps + 1)
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
    h = torch.arange(1, 10, dtype=torch.float32)
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
                           .astype(np.float32))
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


class PortDecoder:
    """AudioDecoder.token2wav (flow_inference.py:94-123) restated on CPU fp32 (batch 1, like the reference)."""

    def __init__(self, flow_sd, hift_sd, static_chunk_tokens=25):
        from moss_speech_decoder_cosy_b200 import frontend
        self.fsd = {k: v.float() for k, v in flow_sd.items()}
        self.hsd = {k: v.float() for k, v in hift_sd.items()}
        self.enc = frontend.EncoderWeights(self.fsd, "cpu", torch.float32)
        self._frontend = frontend
        self.bank = noise_bank()
        self.chunk = static_chunk_tokens

    @torch.inference_mode()
    def flow(self, token, prompt_token, prompt_feat, embedding, streaming=False, finalize=True):
        """CausalMaskedDiffWithXvec.inference, flow.py:237-283 -> (mel [80, 4N], full [80, T])."""
        tok = torch.cat([prompt_token.reshape(-1).long(), token.reshape(-1).long()]).unsqueeze(0)
        ctx = 0 if finalize else 3
        mu, _ = self._frontend.encode(self.enc, tok, torch.tensor([tok.shape[1]]), ctx,
                                      self.chunk if streaming else 0, _allow_cpu=True)
        mu = mu.transpose(1, 2)
        T, F1 = mu.shape[2], prompt_feat.shape[1]
        cond = torch.zeros(1, 80, T)
        cond[:, :, :F1] = prompt_feat.transpose(1, 2)
        spks = self._frontend.speaker_projection(self.enc, embedding)
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
        wav, _ = hift_inference(self.hsd, mel, noise)
        return wav.unsqueeze(0), mel.unsqueeze(0)


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

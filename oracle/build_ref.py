"""TEST INFRASTRUCTURE ONLY -- builds the reference's own modules (random-init, pinned config) on CPU.

Config = SURVEY.md section 8(a0): CausalMaskedDiffWithXvec (flow.py:151) with UpsampleConformerEncoder
(upsample_encoder.py:105), CausalConditionalCFM (flow_matching.py:199) over CausalConditionalDecoder
(decoder.py:294, channels=[256], 12 mid blocks, 4 transformer blocks each) and HiFTGenerator
(generator.py:392, 24 kHz, ups [8,5,3]) with ConvRNNF0Predictor (f0_predictor.py:22).

The reference ships no checkpoint or config, so weights are random (torch.manual_seed) with two documented
calibrations that keep the fixtures meaningful:
  * f0_predictor.classifier is rescaled so that predicted f0 crosses the 10 Hz voiced threshold on a good share
    of frames (random-init gives f0 < 0.2 Hz => the harmonic branch would never be exercised);
  * HiFT conv weights are init N(0, 0.01) by the reference itself (generator.py:83-84,484-485).
"""
import math
import os
import tempfile

import torch

from . import ref_shims

DEFAULT_CHUNK_TOKENS = 25  # upstream CosyVoice2 static_chunk_size (tokens); MOSS scripts stream with 5


def build_flow(static_chunk_tokens: int = DEFAULT_CHUNK_TOKENS, seed: int = 0, vocab_size: int = 16384):
    ref_shims.install()
    from omegaconf import DictConfig
    from cosyvoice.transformer.upsample_encoder import UpsampleConformerEncoder
    from cosyvoice.flow.decoder import CausalConditionalDecoder
    from cosyvoice.flow.flow_matching import CausalConditionalCFM
    from cosyvoice.flow.flow import CausalMaskedDiffWithXvec

    cfm_params = DictConfig({'sigma_min': 1e-06, 'solver': 'euler', 't_scheduler': 'cosine',
                             'training_cfg_rate': 0.2, 'inference_cfg_rate': 0.7, 'reg_loss_type': 'l1'})
    # CausalConditionalCFM.__init__ reseeds every RNG to 0 (flow_matching.py:202) and draws the noise bank; build it
    # first with a placeholder estimator so that later weight init is controlled by `seed`.
    cfm = CausalConditionalCFM(in_channels=240, cfm_params=cfm_params, n_spks=1, spk_emb_dim=80, estimator=None)
    torch.manual_seed(seed)
    encoder = UpsampleConformerEncoder(
        input_size=512, output_size=512, attention_heads=8, linear_units=2048, num_blocks=6, dropout_rate=0.1,
        positional_dropout_rate=0.1, attention_dropout_rate=0.1, normalize_before=True, input_layer='linear',
        pos_enc_layer_type='rel_pos_espnet', selfattention_layer_type='rel_selfattn', use_cnn_module=False,
        macaron_style=False, static_chunk_size=static_chunk_tokens, upsample_stride=4)
    estimator = CausalConditionalDecoder(
        in_channels=320, out_channels=80, channels=[256], dropout=0.0, attention_head_dim=64, n_blocks=4,
        num_mid_blocks=12, num_heads=8, act_fn='gelu', static_chunk_size=static_chunk_tokens * 4,
        num_decoding_left_chunks=-1)
    cfm.estimator = estimator
    flow = CausalMaskedDiffWithXvec(
        input_size=512, output_size=80, spk_embed_dim=192, output_type='mel', vocab_size=vocab_size,
        input_frame_rate=12.5, only_mask_loss=True, token_mel_ratio=4, pre_lookahead_len=3,
        encoder=encoder, decoder=cfm)
    # kaiming/zero-bias init of the estimator leaves every LayerNorm at (1, 0) and biases at 0; perturb them so the
    # fixtures exercise gamma/beta/bias paths (deterministic under `seed`).
    g = torch.Generator().manual_seed(seed + 17)
    with torch.no_grad():
        for name, p in estimator.named_parameters():
            if p.ndim == 1:
                if ('norm' in name or '.block.2.' in name) and name.endswith('weight'):
                    p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
                else:
                    p.copy_(0.05 * torch.randn(p.shape, generator=g))
    return flow.eval()


def build_hift(seed: int = 0, calibrate_f0: bool = True):
    ref_shims.install()
    from cosyvoice.hifigan.generator import HiFTGenerator
    from cosyvoice.hifigan.f0_predictor import ConvRNNF0Predictor

    torch.manual_seed(seed + 1)
    f0p = ConvRNNF0Predictor(num_class=1, in_channels=80, cond_channels=512)
    hift = HiFTGenerator(
        in_channels=80, base_channels=512, nb_harmonics=8, sampling_rate=24000, nsf_alpha=0.1, nsf_sigma=0.003,
        nsf_voiced_threshold=10, upsample_rates=[8, 5, 3], upsample_kernel_sizes=[16, 11, 7],
        istft_params={'n_fft': 16, 'hop_len': 4}, resblock_kernel_sizes=[3, 7, 11],
        resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]], source_resblock_kernel_sizes=[7, 7, 11],
        source_resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]], lrelu_slope=0.1, audio_limit=0.99,
        f0_predictor=f0p)
    g = torch.Generator().manual_seed(seed + 23)
    with torch.no_grad():
        # Snake alphas default to exactly 1; spread them so per-channel alpha indexing is tested.
        for name, p in hift.named_parameters():
            if name.endswith('.alpha'):
                p.copy_(1.0 + 0.3 * torch.rand(p.shape, generator=g))
        # N(0, 0.01) conv weights make every activation ~1e-3 and the waveform ~0.  Every weight-normed layer has
        # w = g * v / |v|, so setting the gain g directly fixes each layer's signal gain regardless of fan-in and keeps
        # O(1) signals through Snake / exp / sin (deterministic; no clipping at the +-0.99 limiter).
        params = dict(hift.named_parameters())
        for name, p in params.items():
            if not name.endswith('parametrizations.weight.original0') or name.startswith('f0_predictor'):
                continue
            if name.startswith('ups.'):
                stride = (8, 5, 3)[int(name.split('.')[1])]
                p.fill_(math.sqrt(stride / 2.0))
            elif name.startswith('conv_pre'):
                p.fill_(1.0)
            elif name.startswith('conv_post'):
                p.fill_(0.5)
            else:
                p.fill_(0.7)
        for name, p in params.items():
            if name.startswith('source_downs') and name.endswith('weight'):
                p.mul_(4.0 / p.flatten(1).norm(dim=1).view(-1, 1, 1))
            if name.endswith('.bias') and not name.startswith('f0_predictor'):
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
        if calibrate_f0:
            # shift/scale the classifier so |f0| spans roughly 0..400 Hz on N(−5, 2) log-mel-like inputs
            hift.f0_predictor.classifier.weight.mul_(1500.0)
            hift.f0_predictor.classifier.bias.fill_(12.0)
    return hift.eval()


def make_audio_decoder(flow, hift, device="cpu"):
    """Instantiate the reference's real AudioDecoder (flow_inference.py:49) around the given modules."""
    ref_shims.install(with_audio_decoder=True)
    import flow_inference  # noqa: reference module
    tmp = tempfile.mkdtemp(prefix="mossdec_ref_")
    torch.save(flow.state_dict(), os.path.join(tmp, "flow.pt"))
    torch.save(hift.state_dict(), os.path.join(tmp, "hift.pt"))
    cfg_path = os.path.join(tmp, "config.yaml")
    with open(cfg_path, "w") as f:
        f.write("# placeholder: hyperpyyaml is shimmed\n")
    ref_shims._HYPER["cfg"] = {"flow": flow, "hift": hift, "sample_rate": 24000, "feat_extractor": None}
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        dec = flow_inference.AudioDecoder(cfg_path, os.path.join(tmp, "flow.pt"), os.path.join(tmp, "hift.pt"),
                                          campplus_model=os.path.join(tmp, "campplus.onnx"), mel_cache_len=8,
                                          device=device)
    return dec

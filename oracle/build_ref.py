"""TEST INFRASTRUCTURE ONLY -- builds the reference's own modules (random-init, pinned config) on CPU.

Config = SURVEY.md section 8(a0): CausalMaskedDiffWithXvec (flow.py:151) with UpsampleConformerEncoder
(upsample_encoder.py:105), CausalConditionalCFM (flow_matching.py:199) over CausalConditionalDecoder
(decoder.py:294, channels=[256], 12 mid blocks, 4 transformer blocks each) and HiFTGenerator
(generator.py:392, 24 kHz, ups [8,5,3]) with ConvRNNF0Predictor (f0_predictor.py:22).

The reference ships no checkpoint or config, so the modules are loaded (strict=True) with the product's deterministic
random-init checkpoint (moss_speech_decoder_cosy_b200/synth.py), whose scales keep every layer in its non-linear
range (voiced and unvoiced frames, no clipping).
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
    # weights: the product's deterministic random-init checkpoint (strict load also pins the key/shape schema)
    from moss_speech_decoder_cosy_b200 import synth
    flow.load_state_dict(synth.synth_flow_state_dict(seed), strict=True)
    return flow.eval()


def build_hift(seed: int = 0):
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
    from moss_speech_decoder_cosy_b200 import synth
    hift.load_state_dict(synth.synth_hift_state_dict(seed), strict=True)
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

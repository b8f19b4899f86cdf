"""TEST INFRASTRUCTURE ONLY -- golden vectors for the cosyvoice1 / Matcha non-causal building blocks.

    python -m oracle.make_golden_cv1

Runs the UNMODIFIED reference modules ResnetBlock1D / Block1D, Downsample1D and Upsample1D
(GLM_modules/third_party/Matcha-TTS/matcha/models/components/decoder.py:32-158) on CPU in fp32 with the product's
deterministic random-init weights (synth.synth_cv1_blocks_state_dict, seed 0) and seeded inputs, and writes
tests/golden/cv1_blocks.npz (inputs + outputs only; the weights are regenerated from the seed on the GPU box).
"""
import torch

from . import ref_shims
from .make_golden import _save


def _sub(sd, prefix):
    return {k[len(prefix) + 1:]: v for k, v in sd.items() if k.startswith(prefix + ".")}


def main():
    ref_shims.install()
    from matcha.models.components.decoder import Block1D, Downsample1D, ResnetBlock1D, Upsample1D
    from moss_speech_decoder_cosy_b200 import synth

    sd = synth.synth_cv1_blocks_state_dict(0)
    res_in = ResnetBlock1D(320, 256, 1024).eval()
    res_mid = ResnetBlock1D(256, 256, 1024).eval()
    down, up = Downsample1D(256).eval(), Upsample1D(256, use_conv_transpose=True).eval()
    final = Block1D(256, 256).eval()
    final.load_state_dict(_sub(sd, "final"), strict=True)
    res_in.load_state_dict(_sub(sd, "res_in"), strict=True)
    res_mid.load_state_dict(_sub(sd, "res_mid"), strict=True)
    down.load_state_dict(_sub(sd, "down"), strict=True)
    up.load_state_dict(_sub(sd, "up"), strict=True)

    g = torch.Generator().manual_seed(4242)
    out = {}
    with torch.no_grad():
        # ragged dense batch: utterance 1 is padded (mask 0 after 61 frames), odd and even lengths
        for tag, B, T, valid in (("a", 2, 77, (77, 61)), ("b", 2, 130, (130, 97))):
            x = torch.randn(B, 320, T, generator=g)
            mask = torch.zeros(B, 1, T)
            for b, v in enumerate(valid):
                mask[b, :, :v] = 1.0
            temb = torch.randn(B, 1024, generator=g)
            h = res_in(x, mask, temb)
            h2 = res_mid(h, mask, temb)
            d = down(h2 * mask)
            u = up(d)
            fb = final(h2, mask)
            out.update({f"{tag}_x": x, f"{tag}_mask": mask, f"{tag}_temb": temb, f"{tag}_res_in": h, f"{tag}_res_mid": h2,
                        f"{tag}_down": d, f"{tag}_up": u, f"{tag}_final": fb})
    _save("cv1_blocks.npz", **out)


def decoder_golden(n_blocks=2, n_mid=2):
    """The assembled non-causal ConditionalDecoder (cosyvoice/flow/decoder.py:88-291, UNMODIFIED) with channels = [256, 256],
    8 heads x 64, GELU, `n_blocks` transformer blocks per level and `n_mid` mid blocks, random-init weights of
    synth.synth_cv1_decoder_state_dict; ragged masks, odd and even lengths.  -> tests/golden/cv1_decoder.npz"""
    import numpy as np
    ref_shims.install()
    from cosyvoice.flow.decoder import ConditionalDecoder
    from moss_speech_decoder_cosy_b200 import synth
    dec = ConditionalDecoder(in_channels=320, out_channels=80, channels=[256, 256], dropout=0.0, attention_head_dim=64,
                             n_blocks=n_blocks, num_mid_blocks=n_mid, num_heads=8, act_fn="gelu").eval()
    dec.load_state_dict(synth.synth_cv1_decoder_state_dict(0, n_blocks, n_mid), strict=True)
    from . import cases
    out = {"n_blocks": np.int64(n_blocks), "n_mid": np.int64(n_mid)}
    for tag, T, lens in (("a", 77, (77, 61)), ("b", 130, (130, 97))):
        c = cases.estimator_inputs(T, 8800 + T, lens=list(lens))
        with torch.inference_mode():
            y = dec(c["x"], c["mask"], c["mu"], c["t"], c["spks"], c["cond"])
        out.update({f"{tag}_T": np.int64(T), f"{tag}_lens": np.asarray(lens), f"{tag}_out": y,
                    f"{tag}_fp": cases.fingerprint(c["x"], c["cond"], c["spks"])})
    _save("cv1_decoder.npz", **out)


if __name__ == "__main__":
    import sys
    if "decoder" in sys.argv[1:]:
        decoder_golden()
    else:
        main()

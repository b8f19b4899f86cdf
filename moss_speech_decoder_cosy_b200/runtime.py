"""Engine construction helpers shared by the AudioDecoder mirror, the tests and bench.py."""
import torch

from . import lib, synth, weights


def build_engine(flow_sd=None, hift_sd=None, seed=0, with_hift=True):
    """Create a libmossdec engine on the current CUDA device and upload packed weights.

    flow_sd / hift_sd: reference-format state_dicts (flow.pt / hift.pt); None -> deterministic random init."""
    eng = lib.Engine()
    if flow_sd is None:
        flow_sd = synth.synth_flow_state_dict(seed)
    tensors = weights.pack_estimator(flow_sd)
    tensors["flow.rand_noise"] = synth.synth_noise_bank()[0].contiguous()
    if with_hift and hasattr(weights, "pack_hift"):
        if hift_sd is None:
            hift_sd = synth.synth_hift_state_dict(seed)
        tensors.update(weights.pack_hift(hift_sd))
    eng.load_tensors(tensors)
    eng.finalize()
    return eng

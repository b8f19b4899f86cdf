"""TEST INFRASTRUCTURE ONLY -- dumps the checkpoint schema (state_dict key -> shape) of the reference modules.

Run in the build container (needs /root/reference):  python -m oracle.make_manifest
Writes moss_speech_decoder_cosy_b200/ckpt_schema.json, the key/shape interface that `flow.pt` / `hift.pt` checkpoints of
the reference follow (flow_inference.py:59-61).  The product uses it to validate real checkpoints and to synthesise
random-init weights on machines where the reference is absent.
"""
import json
import os

from . import build_ref


def main():
    flow = build_ref.build_flow()
    hift = build_ref.build_hift()
    schema = {
        "flow": {k: list(v.shape) for k, v in flow.state_dict().items()},
        "hift": {k: list(v.shape) for k, v in hift.state_dict().items()},
    }
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "moss_speech_decoder_cosy_b200",
                       "ckpt_schema.json")
    with open(out, "w") as f:
        json.dump(schema, f, indent=0, sort_keys=True)
    print("wrote", out, len(schema["flow"]), len(schema["hift"]))


if __name__ == "__main__":
    main()

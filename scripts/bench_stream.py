"""Config-3 style measurement only: p50 of a chunk step for N concurrent streams (see bench.py:bench_config3)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder  # noqa: E402

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    dec = AudioDecoder(seed=0)
    r = bench.bench_config3(dec, n, n_tokens=70)
    print(json.dumps({"env": {k: v for k, v in os.environ.items() if k.startswith("MOSSDEC")}, **r}))

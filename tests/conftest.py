import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def engine():
    """Engine with the seed-0 random-init checkpoint (estimator + HiFT) loaded."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200 import runtime
    return runtime.build_engine(seed=0)

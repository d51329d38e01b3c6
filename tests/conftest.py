import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# The unmodified reference installed by scripts/install_ref.sh (git-ignored, travels to the GPU box): when it is there
# its scorer interfaces are the ones llm_guided_asr_b200.scorer_interface re-exports, so the reference's own
# BeamSearch / BatchBeamSearch can drive this package's scorers (INTEGRATION.md level 1, tests/test_gpu_reference_driving.py)
_REF = os.path.join(ROOT, "baseline", "_ref")
if os.path.isdir(os.path.join(_REF, "espnet", "nets")) and _REF not in sys.path:
    sys.path.insert(1, _REF)
    sys.dont_write_bytecode = True


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)

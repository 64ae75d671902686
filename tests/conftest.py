import os
import sys
from pathlib import Path

import pytest

os.environ.setdefault("LLM_OTEL_EXPORT", "0")  # no collector in the test environment

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "agentic-traffic-testing_b200"
for p in (str(ROOT), str(PKG)):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: minutes of CPU oracle time; skipped unless ATX_RUN_SLOW=1")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)

import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


try:        # property tests draw the same examples on every run: a red test is then a reproducible one
    from hypothesis import settings as _hyp_settings
    _hyp_settings.register_profile("fedb200", derandomize=True, deadline=None, database=None)
    _hyp_settings.load_profile("fedb200")
except ImportError:      # hypothesis is optional; tests/test_properties.py is skipped without it
    pass


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `pytest -m gpu`)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords or "multigpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def cuda_lib():
    """Loads libfedb200.so; on a GPU box a missing library is a hard failure, never a skip."""
    import torch
    from fedb200.ops import _lib
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return _lib.lib()

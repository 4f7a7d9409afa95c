"""pytest configuration: the `gpu` marker and repo-root imports.

`-m "not gpu"` : oracle vs golden vectors, planner/host logic (incl. world_size-2 gloo), C-ABI
                 symbol checks. No CUDA device is touched.
`-m gpu`       : parity tests proper — every call goes through the C ABI of libelb200.so.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def grid1():
    """A 1 x 1 El::Grid on cuda:0 (single-process GPU tests)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from elemental_b200 import el
    torch.cuda.set_device(0)
    g = el.Grid(height=1, rank=0, size=1, device=0)
    yield g
    g.sync()

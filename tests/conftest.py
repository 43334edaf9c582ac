import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_SRC = "/root/reference/src"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")
    config.addinivalue_line("markers", "reference: differential test against the reference checkout")


def pytest_collection_modifyitems(config, items):
    import torch
    has_gpu = torch.cuda.is_available()
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)
    # the C-ABI no-allocation test must be collected last
    items.sort(key=lambda it: 1 if "test_zz_" in it.nodeid else 0)


@pytest.fixture(scope="session")
def reference_modules():
    """Import the reference's own modules (read-only checkout) for differential tests."""
    if not os.path.isdir(REFERENCE_SRC):
        pytest.skip("reference checkout not present on this box")
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import importlib
    mods = {}
    for name in ("Model", "Trainer", "Evaluator", "DataLoader"):
        mods[name] = importlib.import_module(name)
    return mods


@pytest.fixture(scope="session")
def tiny_clients():
    """Six small synthetic clients, prepared (split + standardised) once per session."""
    import numpy as np
    from fedmse_decentralized_b200.data import SyntheticSpec, make_clients, prepare_from_raw
    spec = SyntheticSpec(n_normal=300, n_abnormal=260, n_test_normal=60, jitter_rows=10)
    raw = make_clients(6, spec, seed=7)
    clients, dev = prepare_from_raw(raw, True, seed=3)
    return clients, dev

import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_VAL = "/root/reference/data/val.tfrecords"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def val_tfrecords():
    """The reference's bundled validation file, or a byte-identical-format synthetic
    stand-in when the reference tree is not mounted (the GPU box)."""
    if os.path.exists(REFERENCE_VAL):
        return REFERENCE_VAL
    local = os.path.join(ROOT, "tests", "data", "val_sample.tfrecords")
    if os.path.exists(local):
        return local
    pytest.skip("no TFRecord sample available")

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
    # the GPU box has neither: a Criteo-shaped synthetic file of the same size written by the framework's own generator
    import tempfile
    from deepfm_b200.data.tfrecord import generate_synthetic
    path = os.path.join(tempfile.gettempdir(), "dfm_val_standin_10000.tfrecords")
    if not os.path.exists(path):
        generate_synthetic(path + ".tmp", 10000, 39, 117581, n_dense=13, seed=2553, pos_rate=0.2553, zipf_s=1.05)
        os.replace(path + ".tmp", path)
    return path

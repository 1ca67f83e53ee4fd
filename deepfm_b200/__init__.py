"""Import shim.

The framework's source tree lives in the directory named after the project,
``deepfm-tensorflow-distributed-training-on-amazon-sagemaker_b200/`` -- a name
Python cannot import because of the hyphens.  This package gives it the
importable name ``deepfm_b200`` by pointing ``__path__`` at that directory, so
``import deepfm_b200.models`` resolves to ``<that directory>/models``.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
PACKAGE_DIR = _os.path.join(_ROOT, "deepfm-tensorflow-distributed-training-on-amazon-sagemaker_b200")
__path__.insert(0, PACKAGE_DIR)
REPO_ROOT = _ROOT

__version__ = "0.1.0"

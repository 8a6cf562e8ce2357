import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_DIR = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long-running")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def reference_modules():
    """The reference's own model.py / utils.py as oracles. Skipped when the
    reference tree is not mounted (e.g. on the GPU box)."""
    ref_dir = REFERENCE_DIR
    if not os.path.isdir(ref_dir):
        ref_dir = os.path.join(ROOT, "baseline", "_ref")   # the pip --target install of the reference (travels to the GPU box)
    if not os.path.exists(os.path.join(ref_dir, "model.py")):
        pytest.skip("reference tree not available")
    import importlib.util

    mods = {}
    for name in ("model", "utils"):
        spec = importlib.util.spec_from_file_location(f"_ref_{name}", os.path.join(ref_dir, f"{name}.py"))
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        mods[name] = m
    return mods


TINY_ARGS = [
    "--tokenizer-name-or-path", "offline", "--vocab-size", "512", "--dim", "64", "--n-layers", "2",
    "--n-heads", "4", "--n-kv-heads", "2", "--multiple-of", "32", "--sequence-length", "32",
    "--batch-size", "2", "--model-dtype", "fp32", "--device", "cpu", "--synthetic-docs", "64",
    "--learning-rate", "1e-3", "--lr-warmup-steps", "5",
]

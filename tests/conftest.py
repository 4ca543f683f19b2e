import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def load_golden(name):
    from safetensors import safe_open
    f = safe_open(os.path.join(GOLDEN, name), "pt", "cpu")
    return {k: f.get_tensor(k) for k in f.keys()}, (f.metadata() or {})


def rel_err(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    denom = b.norm().item()
    return (a - b).norm().item() / (denom if denom > 0 else 1.0)


@pytest.fixture(scope="session")
def reference_lora():
    """The real reference module, when the build container has it mounted."""
    import importlib.util
    path = os.path.join(REFERENCE, "lora_diffusion", "lora.py")
    if not os.path.exists(path):
        pytest.skip("reference not mounted")
    spec = importlib.util.spec_from_file_location("ref_lora", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod

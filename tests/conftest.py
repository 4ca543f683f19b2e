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


def max_abs_err(a, b):
    """Largest element error, relative to the largest reference element: one wrong element (a TMA-clipped column,
    a stale tile corner) shows up here while it vanishes in a Frobenius ratio over millions of elements."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    denom = b.abs().max().item()
    return (a - b).abs().max().item() / (denom if denom > 0 else 1.0)


def worst_row_err(a, b):
    """max over rows of |a_i - b_i|_2 relative to the RMS row norm of the reference: a wrong last ragged row (or a
    wrong row block) reads ~1 here, whatever the size of the tensor."""
    a = a.detach().double().cpu().reshape(-1, a.shape[-1])
    b = b.detach().double().cpu().reshape(-1, b.shape[-1])
    rms = b.norm(dim=1).pow(2).mean().sqrt().item()
    return (a - b).norm(dim=1).max().item() / (rms if rms > 0 else 1.0)


def check_close(a, b, tol, what=""):
    """The BASELINE bar (`tol` on the Frobenius ratio) plus the two outlier metrics: worst row within 4x and
    largest element error within 4x of `tol` (a single element of a 16-bit result may sit a rounding step away
    from the fp32 reference: half an ulp of the largest element is 2e-3 for bf16)."""
    errs = {"rel": rel_err(a, b), "row": worst_row_err(a, b), "max": max_abs_err(a, b)}
    assert errs["rel"] < tol and errs["row"] < 4 * tol and errs["max"] < 4 * tol, (what, tol, errs)
    return errs


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

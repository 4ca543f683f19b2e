"""Recipe for `oracle/_ref/` — TEST INFRASTRUCTURE, not product code.

The reference's hot path is one pure-Python file, `lora_diffusion/lora.py`, whose imports (json, math, numpy, PIL,
torch, safetensors) are all present in this image; the package around it is not importable (`lora_diffusion/__init__.py`
pulls `diffusers`).  This script puts that ONE file, unmodified, under `oracle/_ref/` (git-ignored, so it never enters
the history; not gpurun-ignored, so it travels to the GPU box) so that `bench.py`'s CPU legs time the REFERENCE ITSELF
(`cpu_baseline.kind = "reference"`) and the tests can differential-check the oracle against it on the GPU box as well.

Run by `__graft_entry__.build()` whenever `/root/reference` is mounted; a no-op elsewhere (the GPU box uses the copy that
travelled).  `load()` returns the module or None.
"""
import importlib.util
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/lora_diffusion/lora.py"
DST_DIR = os.path.join(HERE, "_ref")
DST = os.path.join(DST_DIR, "reference_lora.py")


def build():
    if not os.path.exists(SRC):
        return os.path.exists(DST)
    os.makedirs(DST_DIR, exist_ok=True)
    if not os.path.exists(DST) or os.path.getmtime(DST) < os.path.getmtime(SRC) or \
            open(DST, "rb").read() != open(SRC, "rb").read():
        shutil.copyfile(SRC, DST)
    return True


def load():
    """The reference module (its own code, executed as is), or None when `oracle/_ref/` was never built."""
    if not os.path.exists(DST):
        return None
    spec = importlib.util.spec_from_file_location("reference_lora", DST)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print("oracle/_ref built" if build() else "reference not mounted: oracle/_ref not built")

"""The C-ABI library loads without a GPU and exports every symbol include/b200lora.h declares;
argument validation runs before any CUDA call."""
import ctypes
import os
import re

from diffusion_finetuning_b200 import _native as nat

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "b200lora.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200lora_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_header_symbols():
    import __graft_entry__ as g
    g.build()
    lib = nat.lib()
    names = declared_functions()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200lora.h but not exported"
    assert set(nat.SIGNATURES) == set(names), "ctypes table and header must list the same functions"
    assert lib.b200lora_version() == 100


def test_argument_validation_needs_no_gpu():
    lib = nat.lib()
    rc = lib.b200lora_lora_gemm(None, 0, None, 0, 0, None, 0, None, 0, 0, None, None, 0, 0, 0, None, 1.0, None, 0,
                                None, 4, 8, 8, 2, nat.BF16, nat.IMPL_AUTO, None)
    assert rc == -6 and b"null" in lib.b200lora_last_error()
    buf = ctypes.create_string_buffer(64)
    p = ctypes.cast(buf, ctypes.c_void_p)
    rc = lib.b200lora_lora_gemm(p, 8, p, 8, 0, None, 0, p, 8, 1, None, p, 2, 1, 7, None, 1.0, p, 8, None, 4, 8, 8, 2,
                                nat.BF16, nat.IMPL_AUTO, None)
    assert rc == -4                                   # bad dtype enum
    rc = lib.b200lora_merge(p, p, p, p, 1.0, 4, 6, 2, nat.BF16, nat.BF16, None)
    assert rc == -3                                   # K % 8 != 0 for 16-bit weights
    assert lib.b200lora_lora_grads_workspace(4096, 320, 320, 4) > 0
    assert lib.b200lora_pack_elems(320, 640) == 80 * 960

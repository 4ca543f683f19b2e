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


def test_round2_entry_points_validate_without_a_gpu():
    """The entry points added in round 2 reject bad arguments before any CUDA work: null pointers, dtypes the path does
    not take, sizes beyond the built-in scratch array."""
    lib = nat.lib()
    buf = ctypes.create_string_buffer(256)
    p = ctypes.cast(buf, ctypes.c_void_p)
    # GEGLU epilogue: 16-bit io only, d_ff % 32, rank <= 16
    assert lib.b200lora_linear_fwd_geglu(None, p, p, 0, 1.0, None, p, None, 8, 64, 64, 4, nat.BF16, None) == -6
    assert lib.b200lora_linear_fwd_geglu(p, p, p, 0, 1.0, None, p, None, 8, 64, 64, 4, nat.F32, None) == -4
    assert lib.b200lora_linear_fwd_geglu(p, p, p, 0, 1.0, None, p, None, 8, 64, 80, 4, nat.BF16, None) == -2
    assert lib.b200lora_linear_fwd_geglu(p, p, p, 0, 1.0, None, p, None, 8, 64, 64, 17, nat.BF16, None) == -2
    assert lib.b200lora_geglu_bwd(p, p, None, 8, 64, nat.BF16, None) == -6
    assert lib.b200lora_geglu_bwd(p, p, p, 8, 60, nat.BF16, None) == -2          # d_ff % 8 for 16-bit vectors
    # fp32 dX store: 16-bit io only
    assert lib.b200lora_linear_bwd_dx_f32out(p, p, p, 1.0, None, None, 8, 64, 64, 4, nat.BF16, None) == -6
    assert lib.b200lora_linear_bwd_dx_f32out(p, p, p, 1.0, p, None, 8, 64, 64, 4, nat.F32, None) == -4
    # pack + cast: the rider's pointers are checked with the pack's
    assert lib.b200lora_pack_lora_cast(p, p, nat.F32, None, 0, p, 64, 64, 4, nat.BF16, None, p, 128, None) == -6
    assert b"pack + cast" in lib.b200lora_last_error()
    # the no-allocation reductions: a shape whose workspace exceeds the built-in array is refused with both sizes named
    M, K, N, r = 8192, 8192, 8192, 16
    assert lib.b200lora_lora_grads_workspace(M, K, N, r) > nat.STATIC_SCRATCH_BYTES
    assert lib.b200lora_lora_grads_static(p, p, p, p, 1.0, p, p, M, K, N, r, nat.BF16, None) == -2
    assert str(nat.STATIC_SCRATCH_BYTES).encode() in lib.b200lora_last_error()


def test_merge_batch_plan_tiling_is_host_only():
    """b200lora_merge_batch_plan (host code): tile counts of the streamed kernel's layout — whole-row tiles of R rows
    (R = what fits 32 KB, a multiple of 4, at most 128), rows longer than 4 KB cut into equal column blocks of 8 rows —
    and of the register-staged layout ([32 rows x 512 bytes]) when a rank above 8 is present."""
    lib = nat.lib()

    class Item(ctypes.Structure):
        _fields_ = [("Win", ctypes.c_void_p), ("Wout", ctypes.c_void_p), ("up", ctypes.c_void_p), ("down", ctypes.c_void_p),
                    ("N", ctypes.c_int32), ("K", ctypes.c_int32), ("r", ctypes.c_int32), ("reserved", ctypes.c_int32)]

    def plan(shapes, dt):
        n = len(shapes)
        items = (Item * n)()
        for i, (N, K, r) in enumerate(shapes):
            items[i] = Item(0x1000, 0x1000, 0x1000, 0x1000, N, K, r, 0)      # aligned dummies: the plan never dereferences
        nbytes = lib.b200lora_merge_batch_workspace(n)
        table = ctypes.create_string_buffer(nbytes)
        tiles, max_r = ctypes.c_int64(0), ctypes.c_int32(0)
        rc = lib.b200lora_merge_batch_plan(ctypes.cast(items, ctypes.c_void_p), n, dt, ctypes.cast(table, ctypes.c_void_p),
                                           nbytes, ctypes.byref(tiles), ctypes.byref(max_r))
        assert rc == 0, lib.b200lora_last_error()
        return tiles.value, max_r.value

    ceil = lambda a, b: (a + b - 1) // b
    # fp16: 640-byte rows -> R = 48; 2560-byte rows -> R = 12; fp32 K = 1280: 5120-byte rows -> 2 column blocks x 8 rows
    assert plan([(320, 320, 4)], nat.F16) == (ceil(320, 48), 4)
    assert plan([(1280, 1280, 4)], nat.F16) == (ceil(1280, 12), 4)
    assert plan([(1280, 1280, 4)], nat.F32) == (ceil(1280, 8) * 2, 4)
    assert plan([(1000, 8, 1)], nat.F16) == (ceil(1000, 128), 1)                  # 16-byte rows: the 128-row cap
    assert plan([(3, 20000, 2)], nat.F16) == (1 * ceil(40000, 4096), 2)           # one 8-row panel, 10 column blocks
    assert plan([(320, 320, 4), (77, 768, 8)], nat.F32) == (ceil(320, 24) + ceil(77, 8), 8)
    # a rank above 8 selects the register-staged kernel's layout for the whole table
    assert plan([(320, 320, 16)], nat.F16) == (ceil(320, 32) * ceil(320 * 2, 512), 16)

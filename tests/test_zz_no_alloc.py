"""`b200lora_lora_grads_static` — the no-allocation variant of the deterministic gradient reductions (its partial sums
live in a device array built into the library; include/b200lora.h).  No counterpart in the reference: the check is
against the existing entry point `b200lora_lora_grads(deterministic=1, caller workspace)`, which runs the same kernels
on the same split plan, so the results must be BIT-identical (and within the reduction tolerance of the fp64 product,
as tests/test_gpu_linear.py::test_gradient_reductions_all_modes asks).  Collected last on purpose
(file name)."""
import pytest
import torch

from conftest import rel_err
from diffusion_finetuning_b200 import _native as nat

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL_RED = 1e-4      # fp32 accumulation over the tokens, same bound as test_gradient_reductions_all_modes


def _case(shape, dtype, seed):
    M, K, N, r = shape
    g = torch.Generator(DEV).manual_seed(seed)
    mk = lambda *s: torch.randn(*s, device=DEV, generator=g).to(dtype)
    return mk(M, N), mk(M, K), mk(M, r), mk(M, r)


def _existing(dy, x, T, dT, s, shape, dtype, init):
    M, K, N, r = shape
    lib = nat.lib()
    nbytes = int(lib.b200lora_lora_grads_workspace(M, K, N, r))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
    dA = torch.full((r, K), init, dtype=torch.float32, device=DEV)
    dB = torch.full((N, r), init, dtype=torch.float32, device=DEV)
    nat.check(lib.b200lora_lora_grads(nat.ptr(dy), nat.ptr(x), nat.ptr(T), nat.ptr(dT), s, nat.ptr(dA), nat.ptr(dB),
                                      M, K, N, r, nat.dtype_code(dtype), 1, nat.ptr(ws), nbytes,
                                      nat.stream_ptr(torch.device(DEV))), "b200lora_lora_grads")
    torch.cuda.synchronize()
    return dA, dB


def _static(dy, x, T, dT, s, shape, dtype, init, watch_memory=False):
    M, K, N, r = shape
    lib = nat.lib()
    dA = torch.full((r, K), init, dtype=torch.float32, device=DEV)      # the caller's outputs: allocated by the caller
    dB = torch.full((N, r), init, dtype=torch.float32, device=DEV)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(DEV)[0]
    nat.check(lib.b200lora_lora_grads_static(nat.ptr(dy), nat.ptr(x), nat.ptr(T), nat.ptr(dT), s, nat.ptr(dA),
                                             nat.ptr(dB), M, K, N, r, nat.dtype_code(dtype),
                                             nat.stream_ptr(torch.device(DEV))), "b200lora_lora_grads_static")
    torch.cuda.synchronize()
    if watch_memory:        # kernels already loaded by an earlier call: the device's free memory must not have moved
        assert torch.cuda.mem_get_info(DEV)[0] >= free0
    return dA, dB


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_variant_equals_workspace_variant_twice_in_a_row(dtype):
    # two calls in a row on DIFFERENT shapes, the larger first: the second call finds the first one's partial sums in
    # the shared array and must not read any of them
    shapes = [(8192, 320, 640, 4), (1000, 128, 72, 8)]
    s = 0.5
    for i, shape in enumerate(shapes):
        dy, x, T, dT = _case(shape, dtype, seed=20 + i)
        dA_s, dB_s = _static(dy, x, T, dT, s, shape, dtype, init=0.0, watch_memory=(i > 0))
        dA_e, dB_e = _existing(dy, x, T, dT, s, shape, dtype, init=0.0)
        assert torch.equal(dA_s, dA_e) and torch.equal(dB_s, dB_e), (shape, rel_err(dA_s, dA_e), rel_err(dB_s, dB_e))
        dB_ref = (s * dy.double().t() @ T.double()).float()
        dA_ref = (s * dT.double().t() @ x.double()).float()
        assert rel_err(dA_s, dA_ref) < TOL_RED and rel_err(dB_s, dB_ref) < TOL_RED
    # same shape twice, accumulating into non-zero outputs (the contract of the existing entry point: it ADDS)
    shape = shapes[0]
    dy, x, T, dT = _case(shape, dtype, seed=31)
    a0, b0 = _static(dy, x, T, dT, s, shape, dtype, init=1.0)
    a1, b1 = _static(dy, x, T, dT, s, shape, dtype, init=1.0, watch_memory=True)
    ae, be = _existing(dy, x, T, dT, s, shape, dtype, init=1.0)
    assert torch.equal(a0, a1) and torch.equal(b0, b1)
    assert torch.equal(a0, ae) and torch.equal(b0, be)


def test_static_scratch_variant_rejects_sizes_beyond_its_array():
    lib = nat.lib()
    M, K, N, r = 8192, 8192, 8192, 16
    assert int(lib.b200lora_lora_grads_workspace(M, K, N, r)) > nat.STATIC_SCRATCH_BYTES
    tiny = torch.zeros(64, dtype=torch.float32, device=DEV)       # never dereferenced: the size check comes first
    p = nat.ptr(tiny)
    rc = lib.b200lora_lora_grads_static(p, p, p, p, 1.0, p, p, M, K, N, r, nat.dtype_code(torch.bfloat16),
                                        nat.stream_ptr(torch.device(DEV)))
    assert rc != 0
    msg = lib.b200lora_last_error().decode()
    assert "scratch" in msg and str(nat.STATIC_SCRATCH_BYTES) in msg, msg
    torch.cuda.synchronize()
    assert float(tiny.abs().sum()) == 0.0

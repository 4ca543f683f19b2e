"""`torch.ops.b200lora.*` (diffusion_finetuning_b200/torch_ops.py): the fused LoRA-linear kernels registered with
torch.library — same numbers as the module path, opcheck-clean (schema, fake kernels, autograd registration), traceable
by torch.compile without graph breaks, autocast rule for bf16 AND fp16."""
import pytest
import torch

import diffusion_finetuning_b200 as L
from conftest import rel_err
from diffusion_finetuning_b200 import torch_ops
from oracle import lora_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(dtype, M=(3, 70), K=320, N=192, r=4, bias=True, seed=0):
    g = torch.Generator().manual_seed(seed)
    t = lambda *s, k=1.0: (torch.randn(*s, generator=g) * k).to(dtype).to(DEV)
    return dict(x=t(*M, K), w0=t(N, K, k=K ** -0.5), bias=t(N, k=0.1) if bias else None, A=t(r, K, k=1 / r),
                B=t(N, r, k=0.02), dy=t(*M, N))


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2), (torch.float16, 2e-2)])
def test_registered_op_matches_oracle_and_module_path(dtype, tol):
    c = _case(dtype)
    x = c["x"].clone().requires_grad_(True)
    A = c["A"].clone().requires_grad_(True)
    B = c["B"].clone().requires_grad_(True)
    y, T = torch.ops.b200lora.lora_linear(x, c["w0"], c["bias"], A, B, 0.7)
    y.backward(c["dy"])
    f = lambda t: None if t is None else t.detach().float().cpu()
    y_ref = O.lora_linear_forward(f(c["x"]), f(c["w0"]), f(c["bias"]), f(c["A"]), f(c["B"]), 0.7)
    dx, dA, dB = O.lora_linear_grads(f(c["x"]), f(c["w0"]), f(c["bias"]), f(c["A"]), f(c["B"]), 0.7, f(c["dy"]))
    assert y.shape == c["dy"].shape and T.shape == (210, 4)
    assert rel_err(y, y_ref) < tol and rel_err(x.grad, dx) < tol
    assert rel_err(A.grad, dA) < tol and rel_err(B.grad, dB) < tol
    # the module path (autograd.Function with caches) gives the same numbers
    m = L.LoraInjectedLinear(320, 192, bias=True, r=4).to(DEV).to(dtype)
    with torch.no_grad():
        m.linear.weight.copy_(c["w0"]); m.linear.bias.copy_(c["bias"])
        m.lora_down.weight.copy_(c["A"]); m.lora_up.weight.copy_(c["B"])
    m.scale = 0.7
    assert rel_err(m(c["x"]), y.detach().float()) < (1e-6 if dtype == torch.float32 else 1e-3)


def test_opcheck_and_compile_without_graph_breaks():
    c = _case(torch.bfloat16, M=(64,), K=128, N=64)
    args = (c["x"].requires_grad_(True), c["w0"], c["bias"], c["A"].requires_grad_(True), c["B"].requires_grad_(True), 1.0)
    torch.library.opcheck(torch.ops.b200lora.lora_linear, args,
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))

    def fn(x, w0, b, A, B):
        y, _ = torch.ops.b200lora.lora_linear(x, w0, b, A, B, 0.5)
        return torch.nn.functional.gelu(y.float()).sum()

    compiled = torch.compile(fn, backend="aot_eager", fullgraph=True)     # fullgraph: a graph break would raise
    x = c["x"].detach().clone().requires_grad_(True)
    A = c["A"].detach().clone().requires_grad_(True)
    want = fn(x, c["w0"], c["bias"], A, c["B"].detach())
    want.backward()
    gx, gA = x.grad.clone(), A.grad.clone()
    x.grad = A.grad = None
    got = compiled(x, c["w0"], c["bias"], A, c["B"].detach())
    got.backward()
    assert torch.allclose(got, want, rtol=1e-3) and rel_err(x.grad, gx) < 1e-3 and rel_err(A.grad, gA) < 1e-2


@pytest.mark.parametrize("amp", [torch.bfloat16, torch.float16])
def test_autocast_rule_with_fp32_masters(amp):
    """The reference's mixed-precision mode: fp32 parameters, autocast region (bf16 under accelerate, fp16 in
    cli_lora_pti.py:200).  Output in the autocast dtype, gradients of the fp32 masters in fp32."""
    c = _case(torch.float32, M=(2, 77), K=768, N=320)
    x = c["x"].clone().requires_grad_(True)
    A = c["A"].clone().requires_grad_(True)
    B = c["B"].clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=amp):
        y = torch_ops.lora_linear_op(x, c["w0"], c["bias"], A, B, 1.0)
    assert y.dtype == amp
    y.float().square().mean().backward()
    assert A.grad.dtype == torch.float32 and B.grad.dtype == torch.float32 and x.grad.dtype == torch.float32
    f = lambda t: t.detach().float().cpu()
    xr, Ar, Br = f(c["x"]).requires_grad_(True), f(c["A"]).requires_grad_(True), f(c["B"]).requires_grad_(True)
    yr = O.lora_linear_forward(xr, f(c["w0"]), f(c["bias"]), Ar, Br, 1.0)
    yr.square().mean().backward()
    assert rel_err(y, yr) < 2e-2 and rel_err(x.grad, xr.grad) < 2e-2
    assert rel_err(A.grad, Ar.grad) < 2e-2 and rel_err(B.grad, Br.grad) < 2e-2

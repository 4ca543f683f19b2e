"""`torch.ops.b200lora.*` — the fused LoRA-linear kernels registered with `torch.library` on top of the same C-ABI
("a thin torch custom-op shim": autograd formula, autocast rule, fake kernels for shape inference), so
the layer is visible to `torch.compile` / `torch.export` as two opaque ops instead of a Python `autograd.Function`.

    y, T = torch.ops.b200lora.lora_linear(x, w0, bias, lora_down, lora_up, scale)      # lora.py:49-50
    dx, dA, dB = torch.ops.b200lora.lora_linear_backward(dy, x, T, w0, lora_down, lora_up, scale, need_dx)

`LoraInjectedLinear.forward` keeps calling `ops.lora_linear` (the autograd.Function with the per-module caches of the
frozen weight's cast / transpose and the flat-arena gradient sink); setting `B200LORA_TORCH_OP=1` — or calling
`lora_linear_op` — routes a layer through the registered op instead.  The op is stateless: it re-packs the LoRA factors
and re-derives `W0ᵀ` per call (two small / one weight-sized extra launches), which is the price of having no Python
object in its signature.  CUDA only, like everything else here: the fake kernels exist for tracing, not as a fallback.
"""
from typing import Optional, Tuple

import torch

from . import _native as nat
from . import ops


def _tc(cdt, K, N, r, device):
    return ops._tc_ok(cdt, K, N, r, device, ops._impl_from_env())


@torch.library.custom_op("b200lora::lora_linear", mutates_args=(), device_types="cuda")
def lora_linear(x: torch.Tensor, w0: torch.Tensor, bias: Optional[torch.Tensor], lora_down: torch.Tensor,
                lora_up: torch.Tensor, scale: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """y = x·W0ᵀ + b + s·(x·Aᵀ)·Bᵀ in one kernel; also returns T = x·Aᵀ [M, r] (saved for the backward)."""
    lib = nat.lib()
    cdt = x.dtype
    N, K = w0.shape
    r = lora_down.shape[0]
    if x.shape[-1] != K or lora_down.shape != (r, K) or lora_up.shape != (N, r):
        raise RuntimeError(f"b200lora::lora_linear: shapes x {tuple(x.shape)}, W0 {(N, K)}, down {tuple(lora_down.shape)}, "
                           f"up {tuple(lora_up.shape)} do not fit")
    if w0.dtype != cdt:
        raise RuntimeError(f"b200lora::lora_linear: x is {cdt} but W0 is {w0.dtype} (under autocast the registered rule casts both)")
    if r > ops.MAX_RANK:
        raise nat.B200LoraError(f"LoRA rank {r} is above the {ops.MAX_RANK} the CUDA kernels support")
    lead = x.shape[:-1]
    x2 = x.reshape(-1, K).contiguous()
    M = x2.shape[0]
    w = w0.contiguous()
    A, B = lora_down.contiguous(), lora_up.contiguous()
    b = None if bias is None else bias.contiguous()
    y = torch.empty((M, N), dtype=cdt, device=x.device)
    T = torch.empty((M, r), dtype=cdt, device=x.device)
    if M == 0:
        return y.reshape(*lead, N), T
    use_tc = _tc(cdt, K, N, r, x.device)
    with torch.cuda.device(x.device):
        pack = None
        if use_tc:
            pack = torch.empty(80 * (K + N), dtype=cdt, device=x.device)
            nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), nat.dtype_code(A.dtype), nat.ptr(b),
                                             nat.dtype_code(b.dtype) if b is not None else 0, nat.ptr(pack), K, N, r,
                                             nat.dtype_code(cdt), nat.stream_ptr(x.device)), "b200lora_pack_lora")
        nat.check(lib.b200lora_linear_fwd(nat.ptr(x2), nat.ptr(w), nat.ptr(b), nat.dtype_code(b.dtype) if b is not None else 0,
                                          nat.ptr(A), nat.ptr(B), nat.dtype_code(A.dtype), nat.ptr(pack), float(scale),
                                          nat.ptr(y), nat.ptr(T), M, K, N, r, nat.dtype_code(cdt),
                                          nat.IMPL_TC if use_tc else nat.IMPL_SIMT, nat.stream_ptr(x.device)),
                  "b200lora_linear_fwd")
    return y.reshape(*lead, N), T


@lora_linear.register_fake
def _(x, w0, bias, lora_down, lora_up, scale):
    N, K = w0.shape
    r = lora_down.shape[0]
    M = 1
    for d in x.shape[:-1]:
        M = M * d
    return x.new_empty((*x.shape[:-1], N)), x.new_empty((M, r))


@torch.library.custom_op("b200lora::lora_linear_backward", mutates_args=(), device_types="cuda")
def lora_linear_backward(dy: torch.Tensor, x: torch.Tensor, T: torch.Tensor, w0: torch.Tensor, lora_down: torch.Tensor,
                         lora_up: torch.Tensor, scale: float, need_dx: bool) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """dX = dY·W0 + s·(dY·B)·A (one kernel), dA = s·(dY·B)ᵀ·X and dB = s·dYᵀ·T (one reduction launch); dA / dB in fp32."""
    lib = nat.lib()
    cdt = dy.dtype
    N, K = w0.shape
    r = lora_down.shape[0]
    dy2 = dy.reshape(-1, N).contiguous()
    x2 = x.reshape(-1, K).contiguous()
    M = dy2.shape[0]
    A, B = lora_down.contiguous(), lora_up.contiguous()
    dx = torch.empty((M, K), dtype=cdt, device=dy.device)
    dT = torch.empty((M, r), dtype=cdt, device=dy.device)
    dA = torch.zeros((r, K), dtype=torch.float32, device=dy.device)
    dB = torch.zeros((N, r), dtype=torch.float32, device=dy.device)
    if M == 0:
        return dx.zero_().reshape(x.shape), dA, dB
    use_tc = _tc(cdt, K, N, r, dy.device)
    st = nat.stream_ptr(dy.device)
    io, ab = nat.dtype_code(cdt), nat.dtype_code(A.dtype)
    with torch.cuda.device(dy.device):
        if use_tc:
            wt = w0.t().contiguous()                                   # [K, N]: the W_NK operand of the transposed problem
            pack = torch.empty(80 * (K + N), dtype=cdt, device=dy.device)
            nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), ab, None, 0, nat.ptr(pack), K, N, r, io, st),
                      "b200lora_pack_lora")
            rc = lib.b200lora_linear_bwd_dx(nat.ptr(dy2), nat.ptr(wt), nat.W_NK, nat.ptr(A), nat.ptr(B), ab, nat.ptr(pack),
                                            float(scale), nat.ptr(dx), nat.ptr(dT), M, K, N, r, io, nat.IMPL_TC, st)
        else:
            w = w0.contiguous()
            rc = lib.b200lora_linear_bwd_dx(nat.ptr(dy2), nat.ptr(w), nat.W_KN, nat.ptr(A), nat.ptr(B), ab, None,
                                            float(scale), nat.ptr(dx), nat.ptr(dT), M, K, N, r, io, nat.IMPL_SIMT, st)
        nat.check(rc, "b200lora_linear_bwd_dx")
        nat.check(lib.b200lora_lora_grads(nat.ptr(dy2), nat.ptr(x2), nat.ptr(T), nat.ptr(dT), float(scale), nat.ptr(dA),
                                          nat.ptr(dB), M, K, N, r, io, 0, None, 0, st), "b200lora_lora_grads")
    return dx.reshape(x.shape), dA, dB


@lora_linear_backward.register_fake
def _(dy, x, T, w0, lora_down, lora_up, scale, need_dx):
    N, K = w0.shape
    r = lora_down.shape[0]
    return x.new_empty(x.shape), x.new_empty((r, K), dtype=torch.float32), x.new_empty((N, r), dtype=torch.float32)


def _setup_context(ctx, inputs, output):
    x, w0, bias, A, B, scale = inputs
    _, T = output
    ctx.save_for_backward(x, T, w0, A, B)
    ctx.scale = float(scale)
    ctx.has_bias = bias is not None


def _backward(ctx, dy, _dT):
    x, T, w0, A, B = ctx.saved_tensors
    need_x, need_w, need_b, need_A, need_B = ctx.needs_input_grad[:5]
    dx, dA, dB = torch.ops.b200lora.lora_linear_backward(dy.contiguous(), x, T, w0, A, B, ctx.scale, bool(need_x))
    N, K = w0.shape
    dW = db = None
    if need_w:          # unfrozen base: never the case in the reference's flows; served by a library GEMM
        dW = (dy.reshape(-1, N).t() @ x.reshape(-1, K)).to(w0.dtype)
    if need_b and ctx.has_bias:
        db = dy.reshape(-1, N).sum(0)
    return (dx if need_x else None, dW, db, dA.to(A.dtype) if need_A else None, dB.to(B.dtype) if need_B else None, None)


lora_linear.register_autograd(_backward, setup_context=_setup_context)


def _autocast_rule(x, w0, bias, lora_down, lora_up, scale):
    """Under autocast every floating operand is cast to the autocast dtype (fp32 masters stay fp32 outside; autograd
    carries the gradients back through the casts), as `F.linear`'s own rule does for the reference's three linears."""
    dt = torch.get_autocast_dtype("cuda")
    c = lambda t: None if t is None else t.to(dt)
    with torch.autocast("cuda", enabled=False):
        return torch.ops.b200lora.lora_linear(c(x), c(w0), c(bias), c(lora_down), c(lora_up), scale)


# the autocast kernel: registered on the dispatch key itself, because torch.library.register_autocast only offers a FIXED
# cast dtype and the reference runs under both bf16 (accelerate) and fp16 (cli_lora_pti.py:200) autocast
_FRAGMENT = torch.library.Library("b200lora", "FRAGMENT")
_FRAGMENT.impl("lora_linear", _autocast_rule, "AutocastCUDA")


def lora_linear_op(x, w0, bias, lora_down, lora_up, scale):
    """Functional form through the registered op (same contract as `ops.lora_linear`, minus caches and sinks)."""
    nat.require_cuda(x, "b200lora::lora_linear")
    return torch.ops.b200lora.lora_linear(x, w0, bias, lora_down, lora_up, float(scale))[0]

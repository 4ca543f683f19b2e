"""LoRA-Conv2d on the implicit-GEMM path.

    y = conv(x; W0, b) + s * conv1x1(conv(x; A); B)          (upstream LoraInjectedConv2d)

Forward: ONE launch of the tcgen05 CTA-pair kernel over NHWC activations (TMA im2col loads; the
rank rows of A ride as 16 extra accumulator columns, B and the bias are the tail MMA step).
Backward: `zcat` (dT = dY·B and Z = [dY | s·dT] in one pass) -> dX = conv(Z; [W0 ; A] flipped and
transposed) on the same kernel -> dB (token reduction) -> dA (im2col gather reduction).
There is no cuDNN / eager fallback: unsupported geometry raises.
"""
import torch
import torch.nn as nn

from . import _native as nat
from . import ops


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


def _ceil64(c):
    return (c + 63) // 64 * 64


def _nhwc(x, dtype):
    """[N,C,H,W] (any memory format) -> contiguous [N,H,W,C] in `dtype`: zero-copy when `x` is already channels_last in
    that dtype, otherwise ONE strided-read / dense-write pass that changes layout and dtype together (`.to(dtype)` on
    the permuted view keeps the NCHW strides, so cast-then-`.contiguous()` was two passes over the activation)."""
    t = x.permute(0, 2, 3, 1)
    if t.dtype == dtype and t.is_contiguous():
        return t
    out = torch.empty(t.shape, dtype=dtype, device=x.device)
    out.copy_(t)
    return out


class _ConvState:
    """Per-module cache of the derived operands (same invalidation rule as ops.BaseWeightCache)."""

    def __init__(self):
        self.store = {}
        self.grad_on = True        # grad mode at apply time (set by lora_conv2d; needs_input_grad ignores it)

    @staticmethod
    def key(t):
        return (t.data_ptr(), t._version, tuple(t.shape), t.dtype, t.device)

    def base(self, w, cdt, mode):
        k = ("base", mode, cdt)
        hit = self.store.get(k)
        if hit is not None and hit[0] == self.key(w):
            return hit[1]
        Cout, Cin, R, S = w.shape
        rows, cols = (Cout, R * S * _ceil64(Cin)) if mode == 0 else (Cin, R * S * (_ceil64(Cout) + 64))
        out = torch.zeros(rows, cols, dtype=cdt, device=w.device)
        wc = w.detach().contiguous()
        with torch.cuda.device(w.device):
            rc = nat.lib().b200lora_conv2d_repack_weight(nat.ptr(wc), nat.dtype_code(wc.dtype), nat.ptr(out),
                                                         Cout, Cin, R, S, mode, nat.dtype_code(cdt),
                                                         nat.stream_ptr(w.device))
        nat.check(rc, "b200lora_conv2d_repack_weight")
        self.store[k] = (self.key(w), out)
        self.store.pop(("pack", cdt), None)      # the backward operand carries LoRA columns: re-pack
        return out

    def pack(self, A, B, bias, w, cdt, need_bwd, live=False):
        """(A_pack [16,R*S*Cp], Bn_pack [Cout,64], Wb or None); Wb's LoRA columns refreshed here."""
        k = ("pack", cdt)
        key = (self.key(A), self.key(B), None if bias is None else self.key(bias), self.key(w), need_bwd)
        hit = self.store.get(k)
        if live:                                # being trained: never cached (ops.BaseWeightCache.lora_pack)
            self.store.pop(k, None)
        elif hit is not None and hit[0] == key:
            return hit[1]
        Cout, Cin, R, S = w.shape
        r = A.shape[0]
        a_pack = torch.empty(16, R * S * _ceil64(Cin), dtype=cdt, device=w.device)
        bn_pack = torch.empty(Cout, 64, dtype=cdt, device=w.device)
        wb = self.base(w, cdt, 1) if need_bwd else None
        Ac, Bc = A.detach().contiguous(), B.detach().reshape(Cout, r).contiguous()
        bc = None if bias is None else bias.detach().contiguous()
        with torch.cuda.device(w.device):
            rc = nat.lib().b200lora_conv2d_pack_lora(
                nat.ptr(Ac), nat.ptr(Bc), nat.dtype_code(Ac.dtype), nat.ptr(bc),
                nat.dtype_code(bc.dtype) if bc is not None else 0, nat.ptr(a_pack), nat.ptr(bn_pack), nat.ptr(wb),
                Cout, Cin, R, S, r, nat.dtype_code(cdt), nat.stream_ptr(w.device))
        nat.check(rc, "b200lora_conv2d_pack_lora")
        out = (a_pack, bn_pack, wb)
        if not live:
            self.store[k] = (key, out)
        return out

    def zeros(self, n, cdt, device):
        k = ("zeros", cdt)
        hit = self.store.get(k)
        if hit is None or hit.numel() < n or hit.device != device:
            hit = self.store[k] = torch.zeros(n, dtype=cdt, device=device)
        return hit


def _gemm(x_nhwc, wk, has_bias, s1, s2, scale, T_out, Cout, geo, r, cdt, kind="conv_fwd", c_alg=None):
    """One implicit-GEMM launch.  `kind` / `c_alg` only label it for an installed launch profiler
    (ops.PROFILER): "<kind>:<taps>", M = output pixels, K = taps * (algorithmic input channels), N = Cout."""
    N, H, W, C = x_nhwc.shape
    R, S, sh, sw, ph, pw, dh, dw = geo
    Ho = (H + 2 * ph - dh * (R - 1) - 1) // sh + 1
    Wo = (W + 2 * pw - dw * (S - 1) - 1) // sw + 1
    y = torch.empty((N, Ho, Wo, Cout), dtype=cdt, device=x_nhwc.device)
    with torch.cuda.device(x_nhwc.device), ops._Prof(f"{kind}:{R * S}", N * Ho * Wo, R * S * (c_alg or C), Cout, r,
                                                     x_nhwc.element_size()):
        rc = nat.lib().b200lora_conv2d_gemm(nat.ptr(x_nhwc), nat.ptr(wk), 1 if has_bias else 0, nat.ptr(s1),
                                            nat.ptr(s2), float(scale), nat.ptr(y), nat.ptr(T_out), N, H, W, C,
                                            Cout, R, S, sh, sw, ph, pw, dh, dw, r, nat.dtype_code(cdt),
                                            nat.stream_ptr(x_nhwc.device))
    nat.check(rc, "b200lora_conv2d_gemm")
    return y


class _LoraConv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w0, bias, A, B, scale, geo, state):
        nat.require_cuda(x, "LoraInjectedConv2d.forward")
        cdt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        if cdt not in (torch.bfloat16, torch.float16):
            raise nat.B200LoraError(
                "LoraInjectedConv2d runs on the tcgen05 implicit-GEMM path, which is 16-bit io only: run the "
                "model in bf16/fp16 or under torch.autocast (there is no fp32 convolution fallback)")
        Cout, Cin, R, S = w0.shape
        r = A.shape[0]
        if Cin % 8 or Cout % 8:
            raise nat.B200LoraError(f"LoraInjectedConv2d: channel counts must be multiples of 8 (got {Cin}, {Cout})")
        if r > 16:
            raise nat.B200LoraError("LoraInjectedConv2d: the tcgen05 path supports r <= 16")
        need_grad = state.grad_on and any(ctx.needs_input_grad[i] for i in (0, 3, 4))
        xh = _nhwc(x, cdt)
        wk = state.base(w0, cdt, 0)
        a_pack, bn_pack, wb = state.pack(A, B, bias, w0, cdt, need_grad and geo[2] == 1 and geo[3] == 1,
                                         live=state.grad_on and (ctx.needs_input_grad[3] or ctx.needs_input_grad[4]))
        M_in = xh.shape[0]
        T = None
        if need_grad:
            H, W = xh.shape[1], xh.shape[2]
            Ho = (H + 2 * geo[4] - geo[6] * (R - 1) - 1) // geo[2] + 1
            Wo = (W + 2 * geo[5] - geo[7] * (S - 1) - 1) // geo[3] + 1
            T = torch.empty((M_in * Ho * Wo, r), dtype=cdt, device=x.device)
        y = _gemm(xh, wk, bias is not None, a_pack, bn_pack, scale, T, Cout, geo, r, cdt)
        if need_grad:
            ctx.save_for_backward(xh, T, A, B, w0)
            ctx.meta = (geo, float(scale), cdt, state, x.dtype, wb, r, bias is not None)
        return y.permute(0, 3, 1, 2)        # NCHW view of the NHWC result (channels_last)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dy):
        xh, T, A, B, w0 = ctx.saved_tensors
        geo, scale, cdt, state, x_dtype, wb, r, has_bias = ctx.meta
        R, S, sh, sw, ph, pw, dh, dw = geo
        if sh != 1 or sw != 1:
            raise nat.B200LoraError("LoraInjectedConv2d backward supports stride 1 only (SD-1.5 resnet convs); "
                                    "there is no fallback for strided LoRA convolutions")
        lib = nat.lib()
        dev = dy.device
        Cout, Cin = w0.shape[0], w0.shape[1]
        need_x, _, _, need_A, need_B = ctx.needs_input_grad[:5]
        dyh = _nhwc(dy, cdt)
        N, Ho, Wo, _ = dyh.shape
        M = N * Ho * Wo
        Cz = _ceil64(Cout) + 64
        Bc = B.detach().reshape(Cout, r).contiguous()
        stream = nat.stream_ptr(dev)
        Z = torch.empty((N, Ho, Wo, Cz), dtype=cdt, device=dev)
        dT = torch.empty((M, r), dtype=cdt, device=dev)
        with torch.cuda.device(dev):
            rc = lib.b200lora_conv2d_zcat(nat.ptr(dyh), nat.ptr(Bc), nat.dtype_code(Bc.dtype), scale, nat.ptr(Z),
                                          nat.ptr(dT), M, Cout, r, nat.dtype_code(cdt), stream)
            nat.check(rc, "b200lora_conv2d_zcat")
        dx = None
        if need_x:
            # dX = conv(Z; Wb) with the flipped geometry: pad' = (R-1)*dil - pad
            geo_b = (R, S, 1, 1, (R - 1) * dh - ph, (S - 1) * dw - pw, dh, dw)
            if geo_b[4] < 0 or geo_b[5] < 0:
                raise nat.B200LoraError("LoraInjectedConv2d backward needs padding <= (kernel-1)*dilation")
            zs1 = state.zeros(16 * R * S * Cz, cdt, dev)
            zs2 = state.zeros(Cin * 64, cdt, dev)
            dxh = _gemm(Z, wb, False, zs1, zs2, 0.0, None, Cin, geo_b, r, cdt, kind="conv_bwd_dx", c_alg=Cout)
            dx = dxh.permute(0, 3, 1, 2)
            if dx.dtype != x_dtype:
                dx = dx.to(x_dtype)
        dA = dB = None
        with torch.cuda.device(dev):
            if need_B:
                dB32 = torch.zeros((Cout, r), dtype=torch.float32, device=dev)
                rc = lib.b200lora_lora_grads(nat.ptr(dyh), None, nat.ptr(T), None, scale, None, nat.ptr(dB32), M,
                                             Cin, Cout, r, nat.dtype_code(cdt), 0, None, 0, stream)
                nat.check(rc, "b200lora_lora_grads(conv dB)")
                dB = dB32.to(B.dtype).reshape(B.shape)
            if need_A:
                dA32 = torch.zeros(A.shape, dtype=torch.float32, device=dev)
                Nn, H, W, _ = xh.shape
                rc = lib.b200lora_conv2d_grad_a(nat.ptr(xh), nat.ptr(dT), scale, nat.ptr(dA32), Nn, H, W, Cin, R, S,
                                            sh, sw, ph, pw, dh, dw, r, nat.dtype_code(cdt), stream)
                nat.check(rc, "b200lora_conv2d_grad_a")
                dA = dA32.to(A.dtype)
        return dx, None, None, dA, dB, None, None, None


def _outside_kernel(module, x, cdt):
    """Why this call cannot take the tcgen05 implicit-GEMM kernel, or None.  The kernel covers what the SD-1.5 resnets
    use (16-bit io, groups 1, zero padding, r <= 16, channel counts that are multiples of 8, stride 1 when training)."""
    conv = module.conv
    if cdt not in (torch.bfloat16, torch.float16):
        return "fp32 io (the implicit-GEMM kernel is 16-bit)"
    if conv.groups != 1:
        return "groups != 1"
    if isinstance(conv.padding, str) or conv.padding_mode != "zeros":
        return "non-zero padding mode"
    if module.training and module.dropout.p > 0.0:
        return "active dropout on the LoRA branch"
    if module.r > 16:
        return "rank above 16"
    if conv.in_channels % 8 or conv.out_channels % 8:
        return "channel counts not multiples of 8"
    if torch.is_grad_enabled() and _pair(conv.stride) != (1, 1) and (
            x.requires_grad or module.lora_down.weight.requires_grad or module.lora_up.weight.requires_grad):
        return "strided convolution in training"
    return None


def _library_path(module, x, cdt):
    """Upstream `LoraInjectedConv2d.forward` evaluated by cuDNN (three library convolutions, autograd backward) for the
    geometries listed in `_outside_kernel`: the layer stays a drop-in for any nn.Conv2d the injector can meet, on the
    GPU, in the call's own precision.  NOT a CPU / no-extension fallback: CUDA tensors only, and the extension must load."""
    nat.require_cuda(x, "LoraInjectedConv2d.forward")
    nat.lib()
    conv = module.conv
    A, B = module.lora_down.weight, module.lora_up.weight
    if not isinstance(module.selector, nn.Identity):
        sel = module.selector.weight.reshape(module.r, module.r).to(A.dtype)
        A = torch.einsum("ij,jckl->ickl", sel, A)
    base = conv(x)                                        # honours padding_mode / groups / stride / autocast
    t = conv._conv_forward(x, A, None)                    # lora_down: the base conv's geometry (incl. padding mode)
    up = torch.nn.functional.conv2d(t, B)                 # lora_up: 1x1
    return base + module.dropout(up) * module.scale


def lora_conv2d(module, x):
    """Forward of LoraInjectedConv2d (called by the module)."""
    conv = module.conv
    cdt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
    why = _outside_kernel(module, x, cdt)
    if why is not None:
        module._last_path = "cudnn: " + why
        return _library_path(module, x, cdt)
    module._last_path = "tcgen05 implicit GEMM"
    A = module.lora_down.weight
    if not isinstance(module.selector, nn.Identity):
        sel = module.selector.weight.reshape(module.r, module.r).to(A.dtype)     # diagonal r x r
        A = torch.einsum("ij,jckl->ickl", sel, A)
    kh, kw = _pair(conv.kernel_size)
    sh, sw = _pair(conv.stride)
    ph, pw = _pair(conv.padding)
    dh, dw = _pair(conv.dilation)
    state = getattr(module, "_conv_state", None)
    if state is None:
        state = module._conv_state = _ConvState()
    state.grad_on = torch.is_grad_enabled()
    return _LoraConv2dFn.apply(x, conv.weight, conv.bias, A, module.lora_up.weight, module.scale,
                               (kh, kw, sh, sw, ph, pw, dh, dw), state)

"""Fused sibling projections: `LoraInjectedLinear` layers that their caller
evaluates on the SAME input tensor — attention `to_q/to_k/to_v` on `hidden_states`, cross-attention
`to_k/to_v` on the text context, CLIP `q_proj/k_proj/v_proj` — run as ONE LoRA GEMM over their
concatenated outputs.

The modules stay what the reference makes them (`lora.py:32-50`): separate `nn.Module`s with their own
`linear / lora_down / lora_up / scale`, called one after the other by stock attention code.  The group
object hooks their `forward`: the first sibling called on a tensor launches the fused kernel for all
of them and parks the other outputs; the next siblings, called with the same tensor object, pick
their slice up.  Nothing about the caller changes.

  forward : Y_cat = X·W_catᵀ + [T | 1]·[B_bd ; bias]ᵀ,  T = X·A_catᵀ  (rank rows of all siblings ride
            together, r_total <= 16; each layer's `scale` is folded into its block of B_bd)
  backward: ONE dX kernel over the K-concatenation of the siblings' output gradients (three tensor
            maps, no concatenation copy; autograd's two dX accumulation adds disappear) and ONE
            launch for all 2·nseg gradient reductions.
Opt-in: `fuse_sibling_projections(model)`.  A group that turns out NOT to be called on one tensor
(outputs parked and never collected) switches itself off and the layers run unfused.
"""
import ctypes

import torch
import torch.nn as nn

from . import _native as nat
from . import ops

SIBLING_NAMES = (("to_q", "to_k", "to_v"), ("q_proj", "k_proj", "v_proj"))


def _key(t):
    return (t.data_ptr(), t._version, tuple(t.shape), t.dtype, t.device)


class _Pending:
    __slots__ = ("x", "version", "grad_mode", "outs")

    def __init__(self, x, outs):
        self.x, self.version, self.grad_mode, self.outs = x, x._version, torch.is_grad_enabled(), outs


class SiblingGroup:
    def __init__(self, layers):
        self.layers = list(layers)
        self.pending = None
        self.wasted = 0
        self.grad_on = True         # grad mode at apply time (member_forward)
        self.enabled = True
        self.fused_calls = 0
        self._store = {}

    # ---- eligibility (checked per call: weights, dtypes and scales may change between calls) ----
    def eligible(self, x):
        if not self.enabled or not x.is_cuda:
            return False
        L0 = self.layers[0]
        N, K = L0.linear.weight.shape
        if N % 64 or K % 8 or x.shape[-1] != K:
            return False
        r_total = 0
        for m in self.layers:
            if m.linear.weight.shape != (N, K) or m.linear.weight.requires_grad or m.linear.weight.device != x.device:
                return False
            if m.linear.bias is not None and m.linear.bias.requires_grad:
                return False
            if not isinstance(m.selector, nn.Identity) or (m.training and m.dropout.p > 0.0):
                return False
            if m.lora_down.weight.dtype != m.lora_up.weight.dtype:
                return False
            if m.lora_down.weight.dtype != self.layers[0].lora_down.weight.dtype:
                return False
            if m.linear.bias is not None and m.linear.bias.dtype != m.linear.weight.dtype:
                return False
            r_total += m.lora_down.weight.shape[0]
        if r_total > 16:
            return False
        cdt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        if cdt not in (torch.bfloat16, torch.float16):
            return False
        if not torch.is_autocast_enabled("cuda") and any(m.linear.weight.dtype != cdt for m in self.layers):
            return False
        return ops._is_sm100(x.device) and ops._impl_from_env() != nat.IMPL_SIMT

    # ---- derived operands, cached per version of their sources ----
    def w_cat(self, cdt, transposed):
        slot = ("wt" if transposed else "w", cdt)
        key = tuple(_key(m.linear.weight) for m in self.layers)
        hit = self._store.get(slot)
        if hit is not None and hit[0] == key:
            return hit[1]
        with torch.no_grad():
            w = torch.cat([m.linear.weight.detach().to(cdt) for m in self.layers], 0)     # [Ncat, K]
            out = w.t().contiguous() if transposed else w.contiguous()
        self._store[slot] = (key, out)
        return out

    def pack(self, cdt, live=False, cast=None):
        """(pack buffer, a bias tensor or None — only its presence matters to the GEMM call).  `cast` = (fp32 input,
        destination in `cdt`): the input's conversion rides in the pack launch (ops.BaseWeightCache.lora_pack)."""
        Ls = self.layers
        key = tuple((_key(m.lora_down.weight), _key(m.lora_up.weight), float(m.scale),
                     None if m.linear.bias is None else _key(m.linear.bias)) for m in Ls) + (cdt,)
        hit = self._store.get("pack")
        if live:                                # being trained: never cached (ops.BaseWeightCache.lora_pack)
            self._store.pop("pack", None)
        elif hit is not None and hit[0] == key:
            return hit[1]
        N, K = Ls[0].linear.weight.shape
        dev = Ls[0].linear.weight.device
        n = len(Ls)
        rs = [m.lora_down.weight.shape[0] for m in Ls]
        # the pack of the concatenated problem straight from the members' factors (A rows stacked, B block-
        # diagonal with each scale folded in, biases concatenated): ONE launch, no cat / zeros / slice launches
        As = [m.lora_down.weight.detach().contiguous() for m in Ls]
        Bs = [m.lora_up.weight.detach().contiguous() for m in Ls]
        bs = [None if m.linear.bias is None else m.linear.bias.detach().contiguous() for m in Ls]
        bias_any = next((t for t in bs if t is not None), None)
        buf = torch.empty(80 * (K + N * n), dtype=cdt, device=dev)
        vp = ctypes.c_void_p
        with torch.cuda.device(dev):
            head = ((vp * n)(*[t.data_ptr() for t in As]), (vp * n)(*[t.data_ptr() for t in Bs]),
                    (vp * n)(*[None if t is None else t.data_ptr() for t in bs]),
                    (ctypes.c_float * n)(*[float(m.scale) for m in Ls]), (ctypes.c_int * n)(*rs), n,
                    nat.dtype_code(As[0].dtype), nat.dtype_code(bias_any.dtype) if bias_any is not None else 0,
                    nat.ptr(buf), K, N, nat.dtype_code(cdt))
            if cast is not None:
                rc = nat.lib().b200lora_pack_lora_fused_cast(*head, nat.ptr(cast[0]), nat.ptr(cast[1]), cast[0].numel(),
                                                             nat.stream_ptr(dev))
            else:
                rc = nat.lib().b200lora_pack_lora_fused(*head, nat.stream_ptr(dev))
        nat.check(rc, "b200lora_pack_lora(fused)")
        out = (buf, bias_any)
        if not live:
            self._store["pack"] = (key, out)
        return out

    # ---- the hook ----
    def member_forward(self, layer, x):
        idx = self.layers.index(layer)
        p = self.pending
        if (p is not None and p.x is x and p.version == x._version and p.grad_mode == torch.is_grad_enabled()
                and p.outs[idx] is not None):
            y = p.outs[idx]
            p.outs[idx] = None
            if all(o is None for o in p.outs):
                self.pending = None
            return y
        if p is not None:
            # outputs computed for another tensor were never collected: the siblings are not
            # called on one input here — stop fusing after a few such misses
            self.wasted += 1
            self.pending = None
            if self.wasted >= 3:
                self.enabled = False
                return None
        if not self.eligible(x):
            return None
        flat = []
        for m in self.layers:
            flat += [m.lora_down.weight, m.lora_up.weight]
        self.grad_on = torch.is_grad_enabled()      # needs_input_grad inside the Function ignores the grad mode
        outs = list(_FusedSiblingFn.apply(x, self, *flat))
        self.fused_calls += 1
        y = outs[idx]
        outs[idx] = None
        self.pending = _Pending(x, outs)
        return y


def _rows(t, N, cdt):
    """[.., N] gradient -> 2-D view usable as a TMA source (unit inner stride, 16-byte pitch)."""
    t2 = t.reshape(-1, N)
    if t2.dtype != cdt:
        t2 = t2.to(cdt)
    if t2.stride(1) != 1 or t2.stride(0) % 8 or t2.data_ptr() % 16 or t2.stride(0) < N:
        t2 = t2.contiguous()
    return t2


class _FusedSiblingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group, *ab):
        lib = nat.lib()
        Ls = group.layers
        n = len(Ls)
        N, K = Ls[0].linear.weight.shape
        rs = [ab[2 * i].shape[0] for i in range(n)]
        r_total, Ncat = sum(rs), N * n
        cdt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        lead = x.shape[:-1]
        x2 = x.reshape(-1, K)
        M = x2.shape[0]
        need_grad = group.grad_on and any(ctx.needs_input_grad)
        live = group.grad_on and any(ctx.needs_input_grad[2:])
        # fp32 input while the factors are trained: its conversion rides in the pack launch (see ops._LoraLinearFn)
        ride = x2.dtype == torch.float32 and live and M > 0 and x2.is_contiguous() and ops._castride_enabled()
        x_f32 = x2 if ride else None
        if ride:
            x2 = torch.empty((M, K), dtype=cdt, device=x.device)
        else:
            if x2.dtype != cdt:
                x2 = x2.to(cdt)
            x2 = x2.contiguous()
        w = group.w_cat(cdt, False)
        pack, bias_flag = group.pack(cdt, live=live, cast=(x_f32, x2) if ride else None)
        y = torch.empty((M, Ncat), dtype=cdt, device=x.device)
        T = torch.empty((M, r_total), dtype=cdt, device=x.device) if need_grad else None
        if M > 0:
            with torch.cuda.device(x.device), ops._Prof("fwd", M, K, Ncat, r_total, x2.element_size()):
                rc = lib.b200lora_linear_fwd(
                    nat.ptr(x2), nat.ptr(w), nat.ptr(bias_flag),          # the bias VALUES ride in the pack
                    nat.dtype_code(bias_flag.dtype) if bias_flag is not None else 0, None, None,
                    nat.dtype_code(ab[0].dtype), nat.ptr(pack), 1.0, nat.ptr(y), nat.ptr(T), M, K, Ncat, r_total,
                    nat.dtype_code(cdt), nat.IMPL_TC, nat.stream_ptr(x.device))
                nat.check(rc, "b200lora_linear_fwd(fused)")
        if need_grad:
            ctx.save_for_backward(x2, T)
            ctx.meta = (group, lead, M, K, N, rs, cdt, x.dtype, pack, [float(m.scale) for m in Ls],
                        [m._grad_sink for m in Ls], [ab[2 * i].dtype for i in range(n)])
        y3 = y.reshape(*lead, Ncat)
        return tuple(y3[..., i * N:(i + 1) * N] for i in range(n))

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, *dys):
        x2, T = ctx.saved_tensors
        group, lead, M, K, N, rs, cdt, x_dtype, pack, scales, sinks, ab_dtypes = ctx.meta
        lib = nat.lib()
        n = len(rs)
        r_total = sum(rs)
        dev = x2.device
        d2 = [_rows(d, N, cdt) if d is not None else torch.zeros((M, N), dtype=cdt, device=dev) for d in dys]
        need_x = ctx.needs_input_grad[0]
        need_ab = any(ctx.needs_input_grad[2:])
        io = nat.dtype_code(cdt)
        stream = nat.stream_ptr(dev)
        dyp = (ctypes.c_void_p * n)(*[d.data_ptr() for d in d2])
        ldp = (ctypes.c_int64 * n)(*[d.stride(0) for d in d2])
        f32out = need_x and x_dtype == torch.float32 and ops._f32out_enabled()     # see ops._LoraLinearFn.backward
        dx = torch.empty((M, K), dtype=torch.float32 if f32out else cdt, device=dev)
        dT = torch.empty((M, r_total), dtype=cdt, device=dev)
        grads = [None] * (2 * n)
        if M > 0:
            with torch.cuda.device(dev):
                wt = group.w_cat(cdt, True)
                with ops._Prof("bwd_dx", M, K, N * n, r_total, x2.element_size()):
                    entry = lib.b200lora_linear_bwd_dx_fused_f32out if f32out else lib.b200lora_linear_bwd_dx_fused
                    rc = entry(dyp, ldp, n, nat.ptr(wt), nat.ptr(pack), nat.ptr(dx), nat.ptr(dT), M, K, N, r_total, io,
                               stream)
                    nat.check(rc, "b200lora_linear_bwd_dx_fused")
                if need_ab:
                    use_sink = all(s is not None for s in sinks) and all(ctx.needs_input_grad[2:])
                    if use_sink:
                        dA32 = [s.dA for s in sinks]
                        dB32 = [s.dB for s in sinks]
                    else:
                        dA32 = [torch.zeros((rs[i], K), dtype=torch.float32, device=dev) for i in range(n)]
                        dB32 = [torch.zeros((N, rs[i]), dtype=torch.float32, device=dev) for i in range(n)]
                    dap = (ctypes.c_void_p * n)(*[t.data_ptr() for t in dA32])
                    dbp = (ctypes.c_void_p * n)(*[t.data_ptr() for t in dB32])
                    sc = (ctypes.c_float * n)(*scales)
                    re = (ctypes.c_int * n)(*rs)
                    with ops._Prof("grads", M, K, N * n, r_total, x2.element_size()):
                        rc = lib.b200lora_lora_grads_fused(dyp, ldp, n, nat.ptr(x2), nat.ptr(T), nat.ptr(dT), sc,
                                                           dap, dbp, M, K, N, re, io, stream)
                        nat.check(rc, "b200lora_lora_grads_fused")
                    if not use_sink:
                        for i in range(n):
                            if ctx.needs_input_grad[2 + 2 * i]:
                                grads[2 * i] = dA32[i].to(ab_dtypes[i])
                            if ctx.needs_input_grad[3 + 2 * i]:
                                grads[2 * i + 1] = dB32[i].to(ab_dtypes[i])
        if need_x:
            dx = dx.reshape(*lead, K)
            if dx.dtype != x_dtype:
                dx = dx.to(x_dtype)
        else:
            dx = None
        return (dx, None, *grads)


def fuse_sibling_projections(model, names=SIBLING_NAMES):
    """Group the injected q/k/v-style layers of every attention module of `model`; returns the groups.
    Children of one parent whose attribute names come from one tuple of `names` and whose base weights
    have the same shape form a group (self-attention: q,k,v; cross-attention: k,v)."""
    from .lora import LoraInjectedLinear
    groups = []
    for parent in model.modules():
        for fam in names:
            kids = [getattr(parent, nm, None) for nm in fam]
            kids = [k for k in kids if isinstance(k, LoraInjectedLinear)]
            by_shape = {}
            for k in kids:
                by_shape.setdefault(tuple(k.linear.weight.shape), []).append(k)
            for same in by_shape.values():
                if 2 <= len(same) <= 3:
                    g = SiblingGroup(same)
                    for k in same:
                        k._sibling_group = g
                    groups.append(g)
    return groups


def unfuse_sibling_projections(model):
    for m in model.modules():
        if getattr(m, "_sibling_group", None) is not None:
            m._sibling_group = None


# ---------------------------------------------------------------------------------------------
# GEGLU: activation in the projection kernel's epilogue
# ---------------------------------------------------------------------------------------------
def fuse_geglu(model, class_names=("GEGLU",)):
    """Route every `GEGLU` module whose `proj` is a LoraInjectedLinear through `proj.forward_geglu`: the module's
    `hidden, gate = proj(x).chunk(2, dim=-1); hidden * gelu(gate)` becomes ONE kernel in the forward (the activation
    rides in the GEMM epilogue; the `[M, 2*d_ff]` pre-activation is only written when a backward will read it) and
    one elementwise kernel + the layer's usual backward in the backward.  Calls the fused path cannot take (fp32
    compute, active dropout, d_ff % 64 != 0) fall through to the module's own forward.  Returns the patched modules."""
    from .lora import LoraInjectedLinear
    patched = []
    for m in model.modules():
        if type(m).__name__ not in class_names or not isinstance(getattr(m, "proj", None), LoraInjectedLinear):
            continue
        if getattr(m, "_b200_geglu_orig", None) is not None:
            patched.append(m)
            continue
        orig = m.forward

        def forward(x, *args, _m=m, _orig=orig, **kwargs):
            out = _m.proj.forward_geglu(x) if not args and not kwargs else None
            if out is None:
                return _orig(x, *args, **kwargs)
            _m.fused_calls = getattr(_m, "fused_calls", 0) + 1
            return out

        m._b200_geglu_orig = orig
        m.forward = forward
        patched.append(m)
    return patched


def unfuse_geglu(model):
    for m in model.modules():
        orig = getattr(m, "_b200_geglu_orig", None)
        if orig is not None:
            m.forward = orig
            m._b200_geglu_orig = None


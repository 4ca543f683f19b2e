"""Functional layer over the C-ABI: the fused LoRA-linear autograd op, the weight merge,
the inspect metric and the SVD-distillation epilogue.

Reference semantics being replaced (lora_diffusion/lora.py in the reference):
  forward   lora.py:49-50   linear(x) + lora_up(lora_down(x)) * scale
  backward  autograd of the above (dX, dA, dB; W0 is frozen in every reference flow)
  merge     lora.py:423     weight + alpha * (up @ down).type(weight.dtype)
  inspect   lora.py:744-746 (ups @ downs).flatten().abs().mean()
  distill   cli_svd.py:69-84
"""
import os

import torch

from . import _native as nat


# --------------------------------------------------------------------------------------------
# frozen-base shadow copies
# --------------------------------------------------------------------------------------------
class BaseWeightCache:
    """Per-module cache of derived copies of the FROZEN base weight W0:
      ("cast", dtype)   W0 cast to the compute dtype (autocast with fp32 masters: the reference
                        re-casts the frozen weight on every call)
      ("t", dtype)      W0ᵀ [K,N] contiguous in the compute dtype (B operand of the dX GEMM)
    Entries are keyed on (data_ptr, _version, shape, dtype, device) of the source Parameter, so
    re-assigning `linear.weight` (weight_apply_lora, monkeypatch_*) or an in-place update through
    autograd-visible ops invalidates them.  In-place edits through `.data` do not bump
    `_version`; call `invalidate()` after such an edit.
    """

    def __init__(self):
        self._store = {}

    @staticmethod
    def _key(w):
        return (w.data_ptr(), w._version, tuple(w.shape), w.dtype, w.device)

    def get(self, w, kind, dtype):
        slot = (kind, dtype)
        key = self._key(w)
        hit = self._store.get(slot)
        if hit is not None and hit[0] == key:
            return hit[1]
        with torch.no_grad():
            base = w.detach()
            if kind == "cast":
                out = base.to(dtype).contiguous()
            elif kind == "t":
                out = base.to(dtype).t().contiguous()
            else:
                raise KeyError(kind)
        self._store[slot] = (key, out)
        return out

    def lora_pack(self, A, B, bias, dtype, live=False, cast=None):
        """The packed LoRA operands of the tcgen05 kernel (b200lora_pack_lora): one buffer of
        80·(K+N) elements in `dtype` holding the rank rows and tail blocks of forward and backward
        (and the frozen bias, which the forward adds through the tail MMA).

        `live` = the factors are being TRAINED (the caller's autograd node needs their gradient): the pack
        is rebuilt on every call and never cached: fused/capturable optimizers (`torch.optim.AdamW(fused=True)`) update
        parameters without bumping `_version`, so no key can tell that a step happened.  One tiny launch
        per layer per step; inside a captured step it is part of the graph.  Otherwise (inference) the
        pack is cached and re-built when a factor changes: `monkeypatch_*` swap the Parameter, in-place
        autograd-visible edits bump `_version`."""
        key = (self._key(A), self._key(B), None if bias is None else self._key(bias), dtype)
        hit = self._store.get("pack")
        if live:
            self._store.pop("pack", None)     # whatever was cached predates the coming optimizer steps
        elif hit is not None and hit[0] == key:
            return hit[1]
        r, K = A.shape
        N = B.shape[0]
        buf = torch.empty(80 * (K + N), dtype=dtype, device=A.device)
        Ac, Bc = A.detach().contiguous(), B.detach().contiguous()
        bc = None if bias is None else bias.detach().contiguous()
        with torch.cuda.device(A.device):
            if cast is not None:          # (x fp32 contiguous, x_out in `dtype`): the input's conversion rides in this launch
                rc = nat.lib().b200lora_pack_lora_cast(nat.ptr(Ac), nat.ptr(Bc), nat.dtype_code(Ac.dtype),
                                                       nat.ptr(bc), nat.dtype_code(bc.dtype) if bc is not None else 0,
                                                       nat.ptr(buf), K, N, r, nat.dtype_code(dtype), nat.ptr(cast[0]),
                                                       nat.ptr(cast[1]), cast[0].numel(), nat.stream_ptr(A.device))
            else:
                rc = nat.lib().b200lora_pack_lora(nat.ptr(Ac), nat.ptr(Bc), nat.dtype_code(Ac.dtype),
                                                  nat.ptr(bc), nat.dtype_code(bc.dtype) if bc is not None else 0,
                                                  nat.ptr(buf), K, N, r, nat.dtype_code(dtype),
                                                  nat.stream_ptr(A.device))
        nat.check(rc, "b200lora_pack_lora")
        if not live:
            self._store["pack"] = (key, buf)
        return buf

    def invalidate(self):
        self._store.clear()


def _impl_from_env():
    v = os.environ.get("B200LORA_IMPL", "auto").lower()
    return {"auto": nat.IMPL_AUTO, "simt": nat.IMPL_SIMT, "tc": nat.IMPL_TC}.get(v, nat.IMPL_AUTO)


_CC_CACHE = {}


def _is_sm100(device):
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _CC_CACHE:
        _CC_CACHE[idx] = torch.cuda.get_device_capability(idx)[0] == 10
    return _CC_CACHE[idx]


def _tc_ok(cdt, K, N, r, device, impl):
    """Mirror of tc_applicable() in csrc/api.cu (decides which layout of W0 the dX call gets)."""
    if impl == nat.IMPL_SIMT:
        return False
    return cdt in (torch.bfloat16, torch.float16) and 1 <= r <= 16 and K % 8 == 0 and N % 8 == 0 \
        and _is_sm100(device)


def _compute_dtype(x, w0):
    if torch.is_autocast_enabled("cuda"):
        return torch.get_autocast_dtype("cuda")
    if x.dtype != w0.dtype:
        raise RuntimeError(
            f"LoraInjectedLinear: input dtype {x.dtype} and weight dtype {w0.dtype} differ "
            "(outside autocast the reference's F.linear raises here too)")
    return x.dtype


# Largest LoRA rank the forward / backward kernels take (csrc/linear_simt.cu keeps the rank slices in shared memory).
# r <= 16 runs on the tcgen05 kernels, 17..64 on the CUDA-core kernels with the reductions in windows of 16 ranks.
MAX_RANK = 64

# Optional launch profiler (bench.py installs one): an object with begin(kind, M, K, N, r, esize)
# and end(), called on either side of every hot-path launch on the launching stream.
PROFILER = None


class _Prof:
    """`with _Prof(kind, ...)`: brackets one launch with the installed profiler, if any."""
    __slots__ = ("args", "p")

    def __init__(self, *args):
        self.args = args
        self.p = PROFILER

    def __enter__(self):
        if self.p is not None:
            self.p.begin(*self.args)

    def __exit__(self, *exc):
        if self.p is not None:
            self.p.end()
        return False


def _castride_enabled():
    """fp32 -> 16-bit input conversion inside the pack launch (B200LORA_CASTRIDE=0 keeps the ATen cast: A/B)."""
    return os.environ.get("B200LORA_CASTRIDE", "1") != "0"


def _f32out_enabled():
    """fp32 dX straight from the pair kernel's epilogue (B200LORA_F32OUT=0 keeps the 16-bit store + ATen cast: A/B)."""
    return os.environ.get("B200LORA_F32OUT", "1") != "0" and os.environ.get("B200LORA_PAIR", "1") != "0"


class GradSink:
    """Where the gradient reductions of one injected layer land: fp32 slices of a flat gradient
    arena (dp.LoraGradArena).  With a sink the dA/dB kernel ADDS straight into the arena (no
    zero-fill, cast or autograd accumulate launches) and the autograd Function returns no gradient
    for the factors.  "This slice is final for the step" — what fires a bucket's all-reduce — is
    NOT signalled from here: autograd runs the parameter's AccumulateGrad node (and with it the
    arena's post-accumulate hook) exactly once per backward, after EVERY use of the layer has run
    its backward, also when the gradient it receives is undefined.  A layer that runs twice in one
    step (a text encoder called twice, shared modules) is therefore exchanged after its last
    contribution, never between two of them."""
    __slots__ = ("dA", "dB")

    def __init__(self, dA, dB):
        self.dA, self.dB = dA, dB


class _LoraLinearFn(torch.autograd.Function):
    """y = x·W0ᵀ + b + s·(x·Aᵀ)·Bᵀ as ONE kernel; backward = fused dX kernel + token reductions."""

    @staticmethod
    def forward(ctx, x, w0, bias, A, B, scale, cache, impl, deterministic, sink=None, grad_on=True, geglu=False):
        nat.require_cuda(x, "LoraInjectedLinear.forward")
        for name, t in (("linear.weight", w0), ("lora_down.weight", A), ("lora_up.weight", B)):
            if t.device != x.device:
                raise RuntimeError(f"{name} is on {t.device} but the input is on {x.device}")
        lib = nat.lib()
        cdt = _compute_dtype(x, w0)
        N, K = w0.shape
        r = A.shape[0]
        if A.shape != (r, K) or B.shape != (N, r):
            raise RuntimeError(f"LoRA shapes do not match the base: down {tuple(A.shape)}, "
                               f"up {tuple(B.shape)}, base {(N, K)}")
        if x.shape[-1] != K:
            raise RuntimeError(f"input feature size {x.shape[-1]} != in_features {K}")
        if A.dtype != B.dtype:
            raise RuntimeError("lora_down and lora_up must share a dtype")
        if r > MAX_RANK:
            raise nat.B200LoraError(
                f"LoRA rank {r} is above the {MAX_RANK} the CUDA kernels support (1..16 on the tcgen05 path, 17..{MAX_RANK} "
                "on the CUDA-core path); the module can still be built, saved, loaded and merged at any rank")
        lead = x.shape[:-1]
        x2 = x.reshape(-1, K)
        M = x2.shape[0]
        # fp32 input of a layer whose factors are trained on the tcgen05 path: its conversion rides in the pack launch
        # (rebuilt on every call anyway) instead of an ATen cast in front of it
        ride = (x2.dtype == torch.float32 and cdt != torch.float32 and M > 0 and grad_on
                and (ctx.needs_input_grad[3] or ctx.needs_input_grad[4]) and x2.is_contiguous()
                and _tc_ok(cdt, K, N, r, x.device, impl) and _castride_enabled())
        x_f32 = x2 if ride else None
        if ride:
            x2 = torch.empty((M, K), dtype=cdt, device=x.device)
        else:
            if x2.dtype != cdt:
                x2 = x2.to(cdt)
            x2 = x2.contiguous()
        if cache is None:
            cache = BaseWeightCache()
        w = w0.detach() if (w0.dtype == cdt and w0.is_contiguous()) else cache.get(w0, "cast", cdt)
        Ac = A.detach().contiguous()
        Bc = B.detach().contiguous()
        bc = bias.detach().contiguous() if bias is not None else None
        use_tc = _tc_ok(cdt, K, N, r, x.device, impl)
        # needs_input_grad reflects requires_grad only; `grad_on` is the grad mode at apply time
        need_grad = grad_on and any(ctx.needs_input_grad[i] for i in (0, 1, 2, 3, 4))
        # geglu: w0 is a GEGLU projection [2*d_ff, K]; the kernel's epilogue also writes hidden * gelu(gate) [M, d_ff].
        # The pre-activation y is only materialised when a backward will read it.
        y = torch.empty((M, N), dtype=cdt, device=x.device) if (need_grad or not geglu) else None
        T = torch.empty((M, r), dtype=cdt, device=x.device) if need_grad else None
        pack = None
        if use_tc and M > 0:
            pack = cache.lora_pack(A, B, bias, cdt, live=grad_on and (ctx.needs_input_grad[3] or ctx.needs_input_grad[4]),
                                   cast=(x_f32, x2) if ride else None)
        out = None
        if geglu:
            if not use_tc:
                raise nat.B200LoraError("the GEGLU epilogue needs the tcgen05 path (lora_linear_geglu checks this)")
            out = torch.empty((M, N // 2), dtype=cdt, device=x.device)
            if M > 0:
                with torch.cuda.device(x.device), _Prof("fwd", M, K, N, r, x2.element_size()):
                    rc = lib.b200lora_linear_fwd_geglu(nat.ptr(x2), nat.ptr(w), nat.ptr(pack), 1 if bc is not None else 0,
                                                       float(scale), nat.ptr(y), nat.ptr(out), nat.ptr(T), M, K, N // 2, r,
                                                       nat.dtype_code(cdt), nat.stream_ptr(x.device))
                    nat.check(rc, "b200lora_linear_fwd_geglu")
        elif M > 0:
            with torch.cuda.device(x.device), _Prof("fwd", M, K, N, r, x2.element_size()):
                rc = lib.b200lora_linear_fwd(
                    nat.ptr(x2), nat.ptr(w), nat.ptr(bc), nat.dtype_code(bc.dtype) if bc is not None else 0,
                    nat.ptr(Ac), nat.ptr(Bc), nat.dtype_code(Ac.dtype), nat.ptr(pack), float(scale),
                    nat.ptr(y), nat.ptr(T), M, K, N, r, nat.dtype_code(cdt),
                    nat.IMPL_TC if use_tc else nat.IMPL_SIMT, nat.stream_ptr(x.device))
                nat.check(rc, "b200lora_linear_fwd")
        if need_grad:
            ctx.save_for_backward(x2, T, A, B, w0, w)
            ctx.meta = (lead, M, K, N, r, float(scale), cdt, use_tc, cache, bool(deterministic),
                        x.dtype, bias is not None, pack, sink)
        if geglu:
            out = out.reshape(*lead, N // 2)
            if y is None:
                return out
            ctx.mark_non_differentiable(out)
            return y.reshape(*lead, N), out
        return y.reshape(*lead, N)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dy, _dout=None):
        x2, T, A, B, w0, w = ctx.saved_tensors
        (lead, M, K, N, r, scale, cdt, use_tc, cache, deterministic, x_dtype, has_bias, pack,
         sink) = ctx.meta
        lib = nat.lib()
        dev = dy.device
        dy2 = dy.reshape(-1, N)
        if dy2.dtype != cdt:
            dy2 = dy2.to(cdt)
        dy2 = dy2.contiguous()
        need_x, need_w, need_b, need_A, need_B = ctx.needs_input_grad[:5]
        Ac = A.detach().contiguous()
        Bc = B.detach().contiguous()
        ab_code = nat.dtype_code(Ac.dtype)
        io_code = nat.dtype_code(cdt)
        dx = None
        dT = None
        stream = nat.stream_ptr(dev)
        with torch.cuda.device(dev):
            if (need_x or need_A) and M > 0:
                # dX = dY·W0 + s·(dY·B)·A and dT = dY·B in one kernel.  When only dA is needed
                # (frozen input, e.g. attn2.to_k on the text context) dX is still produced: dT
                # falls out of the same pass.
                # fp32 input under autocast (LayerNorm outputs are fp32): autograd wants an fp32 dX — the pair kernel
                # stores its fp32 accumulator directly, no bf16 -> fp32 cast launch behind it
                f32out = use_tc and need_x and x_dtype == torch.float32 and _f32out_enabled()
                dx = torch.empty((M, K), dtype=torch.float32 if f32out else cdt, device=dev)
                dT = torch.empty((M, r), dtype=cdt, device=dev)
                if use_tc:
                    wt = cache.get(w0, "t", cdt)
                prof = _Prof("bwd_dx", M, K, N, r, dy2.element_size())
                prof.__enter__()
                if f32out:
                    rc = lib.b200lora_linear_bwd_dx_f32out(nat.ptr(dy2), nat.ptr(wt), nat.ptr(pack), scale, nat.ptr(dx),
                                                           nat.ptr(dT), M, K, N, r, io_code, stream)
                elif use_tc:
                    rc = lib.b200lora_linear_bwd_dx(nat.ptr(dy2), nat.ptr(wt), nat.W_NK, nat.ptr(Ac),
                                                    nat.ptr(Bc), ab_code, nat.ptr(pack), scale, nat.ptr(dx),
                                                    nat.ptr(dT), M, K, N, r, io_code, nat.IMPL_TC, stream)
                else:
                    rc = lib.b200lora_linear_bwd_dx(nat.ptr(dy2), nat.ptr(w), nat.W_KN, nat.ptr(Ac),
                                                    nat.ptr(Bc), ab_code, None, scale, nat.ptr(dx),
                                                    nat.ptr(dT), M, K, N, r, io_code, nat.IMPL_SIMT, stream)
                nat.check(rc, "b200lora_linear_bwd_dx")
                prof.__exit__()
            dA = dB = None
            use_sink = sink is not None and need_A and need_B
            if (need_A or need_B):
                if use_sink:
                    dA32, dB32 = sink.dA, sink.dB      # fp32 arena slices; the kernel adds into them
                else:
                    dA32 = torch.zeros((r, K), dtype=torch.float32, device=dev) if need_A else None
                    dB32 = torch.zeros((N, r), dtype=torch.float32, device=dev) if need_B else None
                if M > 0:
                    ws = None
                    ws_bytes = 0
                    if deterministic:
                        ws_bytes = int(lib.b200lora_lora_grads_workspace(M, K, N, r))
                        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
                    with _Prof("grads", M, K, N, r, dy2.element_size()):
                        rc = lib.b200lora_lora_grads(nat.ptr(dy2), nat.ptr(x2), nat.ptr(T), nat.ptr(dT),
                                                     scale, nat.ptr(dA32), nat.ptr(dB32), M, K, N, r,
                                                     io_code, 1 if deterministic else 0, nat.ptr(ws),
                                                     ws_bytes, stream)
                        nat.check(rc, "b200lora_lora_grads")
                if not use_sink:                       # (with a sink the grads live in the arena, not in autograd)
                    dA = dA32.to(A.dtype) if need_A else None
                    dB = dB32.to(B.dtype) if need_B else None
        dW = db = None
        if need_w:
            # Unfrozen base: never the case in the reference's flows (unet.requires_grad_(False),
            # train_lora_dreambooth.py:595).  Served by a library GEMM, not a hot-path kernel.
            dW = (dy2.t() @ x2).to(w0.dtype)
        if need_b and has_bias:
            db = dy2.sum(0)
        if need_x and dx is None:            # empty batch: no launch, but autograd still gets a [0, K] gradient
            dx = torch.zeros((M, K), dtype=cdt, device=dev)
        if dx is not None and need_x:
            dx = dx.reshape(*lead, K)
            if dx.dtype != x_dtype:
                dx = dx.to(x_dtype)
        else:
            dx = None
        return dx, dW, db, dA, dB, None, None, None, None, None, None, None


def lora_linear(x, w0, bias, lora_down, lora_up, scale, cache=None, impl=None, deterministic=None,
                sink=None):
    """Functional form of LoraInjectedLinear.forward (reference lora.py:49-50).
    `sink` (GradSink, optional) routes dA/dB straight into a flat gradient arena."""
    if impl is None:
        impl = _impl_from_env()
    if deterministic is None:
        deterministic = os.environ.get("B200LORA_DETERMINISTIC", "0") == "1"
    return _LoraLinearFn.apply(x, w0, bias, lora_down, lora_up, scale, cache, impl, deterministic,
                               sink, torch.is_grad_enabled())


class _GegluAct(torch.autograd.Function):
    """The autograd node between the pre-activation `y` = [hidden | gate] and `out` = hidden * gelu(gate), which the
    GEMM epilogue has ALREADY computed: forward hands `out` through (no launch), backward is one elementwise kernel
    (b200lora_geglu_bwd) whose result feeds the layer's usual backward."""

    @staticmethod
    def forward(ctx, y, out):
        ctx.save_for_backward(y)
        return out.view_as(out)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dout):
        (y,) = ctx.saved_tensors
        lib = nat.lib()
        d_ff = y.shape[-1] // 2
        y2 = y.reshape(-1, 2 * d_ff)
        do2 = dout.reshape(-1, d_ff)
        if do2.dtype != y2.dtype:
            do2 = do2.to(y2.dtype)
        do2 = do2.contiguous()
        dy = torch.empty_like(y2)
        if y2.shape[0] > 0:
            with torch.cuda.device(y.device):
                nat.check(lib.b200lora_geglu_bwd(nat.ptr(do2), nat.ptr(y2), nat.ptr(dy), y2.shape[0], d_ff,
                                                 nat.dtype_code(y2.dtype), nat.stream_ptr(y.device)), "b200lora_geglu_bwd")
        return dy.reshape(y.shape), None


def geglu_fusable(x, w0, lora_down, impl=None):
    """Can `hidden, gate = proj(x).chunk(2, -1); hidden * gelu(gate)` run with the activation in the GEMM epilogue?
    (16-bit compute dtype on the CTA-pair tcgen05 kernel, d_ff % 64 == 0, rank <= 16.)"""
    if not x.is_cuda or os.environ.get("B200LORA_PAIR", "1") == "0":
        return False
    N, K = w0.shape
    r = lora_down.shape[0]
    cdt = _compute_dtype(x, w0)
    return N % 128 == 0 and r <= 16 and _tc_ok(cdt, K, N, r, x.device, _impl_from_env() if impl is None else impl)


def lora_linear_geglu(x, w0, bias, lora_down, lora_up, scale, cache=None, impl=None, deterministic=None, sink=None):
    """`h, g = lora_linear(x, ...).chunk(2, dim=-1); h * gelu(g)` (the GEGLU module around reference lora.py:49-50) with
    the activation applied in the projection kernel's epilogue (b200lora_linear_fwd_geglu).  Call `geglu_fusable` first."""
    if impl is None:
        impl = _impl_from_env()
    if deterministic is None:
        deterministic = os.environ.get("B200LORA_DETERMINISTIC", "0") == "1"
    res = _LoraLinearFn.apply(x, w0, bias, lora_down, lora_up, scale, cache, impl, deterministic, sink,
                              torch.is_grad_enabled(), True)
    if isinstance(res, tuple):
        y, out = res
        return _GegluAct.apply(y, out)
    return res


# --------------------------------------------------------------------------------------------
# merge / inspect / distill
# --------------------------------------------------------------------------------------------
def _cuda_device():
    if not torch.cuda.is_available():
        raise nat.B200LoraError(
            "no CUDA device: the b200lora kernels are the only compute path (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def merge_lora_weight(weight, up, down, alpha=1.0):
    """W + alpha * (up @ down).type(W.dtype)  (reference lora.py:423) on the merge kernel.

    Returns a new tensor with the dtype and device of `weight`.  Host-resident weights (the
    reference CLI merges on the CPU, cli_lora_add.py:74-78) are staged through the GPU.
    """
    lib = nat.lib()
    home = weight.device
    dev = home if weight.is_cuda else _cuda_device()
    with torch.no_grad():
        w = weight.detach().to(dev).contiguous()
        u = up.detach().to(dev).contiguous()
        d = down.detach().to(dev).contiguous()
        if d.dtype != u.dtype:
            d = d.to(u.dtype)
        N, K = w.shape
        r = u.shape[1]
        if u.shape != (N, r) or d.shape != (r, K):
            raise RuntimeError(f"merge: shapes up {tuple(u.shape)} / down {tuple(d.shape)} do not "
                               f"match weight {(N, K)}")
        out = torch.empty_like(w)
        with torch.cuda.device(dev):
            rc = lib.b200lora_merge(nat.ptr(w), nat.ptr(out), nat.ptr(u), nat.ptr(d), float(alpha),
                                    N, K, r, nat.dtype_code(w.dtype), nat.dtype_code(u.dtype),
                                    nat.stream_ptr(dev))
        nat.check(rc, "b200lora_merge")
        return out if home == dev else out.to(home)


class MergePlan:
    """The problem table of a whole-model merge, resident on the device (b200lora_merge_batch_plan / _run):
    every (Win, Wout, up, down) of one (device, weight dtype, LoRA dtype) group.  `run(alpha)` is ONE launch;
    the plan can be re-run (a sweep over alpha: Win = pristine base, Wout = live weight)."""

    def __init__(self, wins, wouts, ups, downs):
        import ctypes
        lib = nat.lib()
        n = len(wins)
        if n == 0:
            raise RuntimeError("merge: empty plan")
        self.dev = wins[0].device
        self.wdt, self.ldt = wins[0].dtype, ups[0].dtype
        self.keep = (list(wins), list(wouts), list(ups), list(downs))      # the table holds raw pointers
        items = (nat.MergeItem * n)()
        for i, (wi, wo, u, d) in enumerate(zip(wins, wouts, ups, downs)):
            N, K = wi.shape
            r = u.shape[1]
            if u.shape != (N, r) or d.shape != (r, K) or wo.shape != wi.shape:
                raise RuntimeError(f"merge: shapes up {tuple(u.shape)} / down {tuple(d.shape)} do not "
                                   f"match weight {(N, K)}")
            for t in (wi, wo, u, d):
                if t.device != self.dev or not t.is_contiguous():
                    raise RuntimeError("merge: every tensor of a plan must be contiguous and on the plan's device")
            if wi.dtype != self.wdt or wo.dtype != self.wdt or u.dtype != self.ldt or d.dtype != self.ldt:
                raise RuntimeError("merge: a plan holds ONE weight dtype and ONE LoRA dtype")
            items[i] = nat.MergeItem(wi.data_ptr(), wo.data_ptr(), u.data_ptr(), d.data_ptr(), N, K, r, 0)
        nbytes = int(lib.b200lora_merge_batch_workspace(n))
        host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        tiles, max_r = ctypes.c_int64(0), ctypes.c_int(0)
        rc = lib.b200lora_merge_batch_plan(items, n, nat.dtype_code(self.wdt), ctypes.c_void_p(host.data_ptr()), nbytes,
                                           ctypes.byref(tiles), ctypes.byref(max_r))
        if rc != 0:
            nat.check(rc, "b200lora_merge_batch_plan")
        self.n, self.tiles, self.max_r = n, tiles.value, max_r.value
        self.table = host.to(self.dev, non_blocking=True)
        self._host = host              # pinned source of the async copy: keep it until the plan dies
        self.bytes = sum(2 * wi.numel() * wi.element_size() + (u.numel() + d.numel()) * u.element_size()
                         for wi, u, d in zip(wins, ups, downs))             # algorithmic HBM bytes of one run

    def run(self, alpha):
        with torch.cuda.device(self.dev):
            rc = nat.lib().b200lora_merge_batch_run(nat.ptr(self.table), self.n, self.tiles, self.max_r, float(alpha),
                                                    nat.dtype_code(self.wdt), nat.dtype_code(self.ldt),
                                                    nat.stream_ptr(self.dev))
        nat.check(rc, "b200lora_merge_batch_run")


def merge_lora_weights(weights, ups, downs, alpha=1.0, inplace=False):
    """`W_i + alpha * (up_i @ down_i).type(W_i.dtype)` for a whole list of matrices — the loop of
    `weight_apply_lora` (reference lora.py:410-424) — in ONE launch per (device, weight dtype, LoRA dtype) group
    (`MergePlan`: a device-side problem table and a persistent grid over all tiles of all matrices).
    Returns new tensors (dtype/device of each weight), or writes through `weights` when `inplace`."""
    nat.lib()
    n = len(weights)
    if not (len(ups) == n and len(downs) == n):
        raise RuntimeError("merge: weights / ups / downs differ in length")
    outs = [None] * n
    groups = {}
    with torch.no_grad():
        staged = []
        for i, (weight, up, down) in enumerate(zip(weights, ups, downs)):
            home = weight.device
            dev = home if weight.is_cuda else _cuda_device()
            w = weight.detach().to(dev)
            if not w.is_contiguous():
                w = w.contiguous()
            u = up.detach().to(dev).contiguous()
            d = down.detach().to(dev).contiguous()
            if d.dtype != u.dtype:
                d = d.to(u.dtype)
            if w.dim() != 2 or u.dim() != 2 or d.dim() != 2 or u.shape != (w.shape[0], u.shape[1]) \
                    or d.shape != (u.shape[1], w.shape[1]):
                raise RuntimeError(f"merge: shapes up {tuple(u.shape)} / down {tuple(d.shape)} do not "
                                   f"match weight {tuple(w.shape)}")
            same = inplace and w.data_ptr() == weight.data_ptr()
            out = w if same else torch.empty_like(w)
            staged.append((w, u, d, out, home, dev, same))
            groups.setdefault((dev, w.dtype, u.dtype), []).append(i)
        for idx in groups.values():
            MergePlan([staged[i][0] for i in idx], [staged[i][3] for i in idx], [staged[i][1] for i in idx],
                      [staged[i][2] for i in idx]).run(alpha)
        for i, (w, u, d, out, home, dev, same) in enumerate(staged):
            if inplace and not same:
                weights[i].data.copy_(out)
                outs[i] = weights[i]
            else:
                outs[i] = out if home == dev else out.to(home)
    return outs


def lora_abs_mean(up, down):
    """mean(|up @ down|) (reference lora.py:744-746) without materialising the product."""
    lib = nat.lib()
    dev = up.device if up.is_cuda else _cuda_device()
    with torch.no_grad():
        u = up.detach().to(dev).contiguous()
        d = down.detach().to(dev).to(u.dtype).contiguous()
        N, r = u.shape
        K = d.shape[1]
        acc = torch.zeros(1, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.b200lora_absmean_accum(nat.ptr(u), nat.ptr(d), nat.ptr(acc), N, K, r,
                                            nat.dtype_code(u.dtype), nat.stream_ptr(dev))
        nat.check(rc, "b200lora_absmean_accum")
        return acc.item() / float(N * K)


def svd_distill_matrix(delta, rank=4, clamp_quantile=0.99, driver=None):
    """Rank-r factorisation of one weight difference (reference cli_svd.py:69-84).

    The SVD stays on cuSOLVER through `torch.linalg.svd`, economy size (`full_matrices=False`: only the top-r
    triplets are kept, the reference's full `U [N, N]` — 419 MB at N = 10240 — is wasted work).  `driver=None` is
    torch's default for CUDA tensors, i.e. exactly the call the reference makes on `device="cuda:0"`
    (cli_svd.py:69: Jacobi `gesvdj`, QR `gesvd` if it does not converge); `driver="gesvd"` forces the QR driver, whose
    fp32 factors sit closer to LAPACK's (1e-6 vs 5e-5 in the rank-4 product) at twice the time.  U·diag(S), the
    symmetric clamp at the `clamp_quantile` quantile of [U, Vh] and the slicing run in the epilogue kernel.
    Returns (U [N,r], Vh [r,K]) fp32 — the `[up, down]` pair the reference saves.
    `clamp_quantile=None` skips the clamp (not a reference mode; used to test the factorisation,
    because the clamp threshold depends on the SVD's arbitrary sign convention).
    """
    lib = nat.lib()
    dev = delta.device if delta.is_cuda else _cuda_device()
    with torch.no_grad():
        mat = delta.detach().to(dev).float()
        U, S, Vh = torch.linalg.svd(mat, full_matrices=False, driver=driver)
        U = U[:, :rank].contiguous()
        S = S[:rank].contiguous()
        Vh = Vh[:rank, :].contiguous()
        N, K = mat.shape
        stream = nat.stream_ptr(dev)
        with torch.cuda.device(dev):
            rc = lib.b200lora_svd_epilogue(nat.ptr(U), nat.ptr(S), nat.ptr(Vh), -1.0, 1, N, K, rank, stream)
            nat.check(rc, "b200lora_svd_epilogue(scale)")
            if clamp_quantile is not None:
                dist = torch.cat([U.flatten(), Vh.flatten()])
                hi = torch.quantile(dist, clamp_quantile).item()
                rc = lib.b200lora_svd_epilogue(nat.ptr(U), nat.ptr(S), nat.ptr(Vh), float(hi), 0, N, K, rank, stream)
                nat.check(rc, "b200lora_svd_epilogue(clamp)")
        return U, Vh

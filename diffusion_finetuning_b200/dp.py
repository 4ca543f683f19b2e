"""Data-parallel exchange step of LoRA training: one flat fp32 gradient arena, split into a few
buckets that are all-reduced (averaged) over NCCL on a side stream WHILE the rest of the backward
runs.

What it replaces in the reference: `accelerator.prepare(unet, ...)` + `accelerator.backward(loss)`
(training_scripts/train_lora_dreambooth.py:751-757, 877), i.e. torch DDP over the trainable
parameters only.  The base model is frozen (train_lora_dreambooth.py:595), so the payload is the
LoRA A/B gradients: 3.2 MB (attention, r=4) to 20 MB (default targets, r=16) in fp32.  Stock DDP
puts all of it in ONE 25 MB bucket that fires after the last gradient, so nothing overlaps.  Here:

  * every injected layer owns two slices of ONE flat fp32 arena, laid out in reverse module order
    (the order the backward produces them); the dA/dB reduction kernel ADDS straight into those
    slices (ops.GradSink), so no per-layer zero-fill / cast / accumulate kernels run;
  * the arena is cut into buckets of >= `bucket_bytes`; a bucket's all-reduce is enqueued on a side
    stream the moment its last slice is final, gated by an event on the compute stream;
  * `finish()` joins the side stream, fires whatever did not fire (unused layers), and leaves
    `p.grad` of every parameter pointing at the averaged values, ready for any torch optimizer.

The path shards by batch only: there is no other collective.  With world_size 1
(or no process group) the arena still removes the per-layer gradient plumbing; no collective runs.
Works on CPU tensors with the `gloo` backend (tests), where the "side stream" is the backend's
own worker thread.
"""
from typing import Iterable, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from . import ops
from .lora import LoraInjectedConv2d, LoraInjectedLinear


class _Bucket:
    __slots__ = ("lo", "hi", "expected", "pending", "work", "fired")

    def __init__(self, lo):
        self.lo, self.hi = lo, lo
        self.expected = 0      # parameters whose slices live in this bucket
        self.pending = 0
        self.work = None
        self.fired = False


class LoraGradArena:
    """Flat gradient arena + bucketed, overlapped all-reduce for the trainable LoRA parameters.

    Usage per step:
        arena.zero()                    # instead of optimizer.zero_grad()
        loss.backward()                 # buckets fire as their layers finish
        arena.finish()                  # join; p.grad = averaged gradients
        arena.clip_grad_norm_(1.0)      # optional, one pass over the arena
        optimizer.step()
    """

    def __init__(self, model: nn.Module, extra_params: Iterable[nn.Parameter] = (),
                 bucket_bytes: int = 1 << 20, process_group=None, average: bool = True):
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.average = average
        layers = [m for m in model.modules() if isinstance(m, (LoraInjectedLinear, LoraInjectedConv2d))]
        if not layers and not list(extra_params):
            raise ValueError("No lora injected.")
        # (param, owning Linear layer or None), in the order the backward finishes them
        entries = []
        for layer in reversed(layers):
            entries.append((layer.lora_up.weight, layer))
            entries.append((layer.lora_down.weight, layer))
        for p in extra_params:
            entries.append((p, None))
        entries = [(p, l) for p, l in entries if p.requires_grad]
        self.params: List[nn.Parameter] = [p for p, _ in entries]
        if not self.params:
            raise ValueError("no trainable LoRA parameter")
        dev = self.params[0].device
        self.device = dev
        total = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self._lowp = {}          # dtype -> flat low-precision copy for non-fp32 parameters
        self.buckets: List[_Bucket] = []
        self._slot = {}          # id(param) -> (offset, numel, bucket index)
        cap = max(1, bucket_bytes // 4)
        b = _Bucket(0)
        off = 0
        for p, _ in entries:
            if b.hi - b.lo >= cap:
                self.buckets.append(b)
                b = _Bucket(off)
            self._slot[id(p)] = (off, p.numel(), len(self.buckets))
            off += p.numel()
            b.hi = off
            b.expected += 1
        self.buckets.append(b)
        self._hooks = []
        self._sync = True
        self._comm = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self._comm_used = False
        self._compute_stream = None
        for p, layer in entries:
            o, n, bi = self._slot[id(p)]
            view = self.flat[o:o + n].view_as(p)
            if p.dtype == torch.float32:
                p.grad = view
            else:
                lp = self._lowp.get(p.dtype)
                if lp is None:
                    lp = self._lowp[p.dtype] = torch.zeros(total, dtype=p.dtype, device=dev)
                p.grad = lp[o:o + n].view_as(p)
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(bi)))
        for layer in layers:
            if isinstance(layer, LoraInjectedLinear) and id(layer.lora_up.weight) in self._slot \
                    and id(layer.lora_down.weight) in self._slot:
                oB, nB, bB = self._slot[id(layer.lora_up.weight)]
                oA, nA, bA = self._slot[id(layer.lora_down.weight)]
                layer._grad_sink = ops.GradSink(
                    self.flat[oA:oA + nA].view_as(layer.lora_down.weight),
                    self.flat[oB:oB + nB].view_as(layer.lora_up.weight))
        self._layers = layers
        # bench.py: a callable returning a fresh recordable CUDA event; finish() then stamps one event before the
        # join of the exchange stream and one after it (exposed = non-overlapped all-reduce time per step)
        self.exposure_probe = None
        self.exposure_events = []
        self.zero()

    # ------------------------------------------------------------------ bookkeeping
    def _make_hook(self, bi):
        def hook(param):
            # Runs once per backward per parameter, after every use of the layer has been differentiated —
            # also when the Function returned no gradient because its kernel added straight into the arena
            # (ops.GradSink); it is the ONLY "slice final" signal.  Otherwise autograd accumulated into p.grad
            # (a view of the arena for fp32 parameters; for low-precision parameters the value is folded into
            # the fp32 arena here)
            # A gradient that came THROUGH autograd (no sink: the LoRA convolutions, low-precision parameters) was
            # added into p.grad by the AccumulateGrad node on the node's own stream — the stream this hook runs under,
            # which for a node that outlived a captured step is not the step's stream.  The bucket's event is recorded
            # on the step's stream, so that stream first waits for the accumulation (found at N = 2 / 8 on config 3:
            # the `lora_up` factors of five resnet convolutions diverged between replicas under the graph, never eagerly).
            cs = self._compute_stream
            if cs is not None:
                cur = torch.cuda.current_stream(self.device)
                if cur != cs:
                    cs.wait_stream(cur)
            if param.dtype != torch.float32:
                o, n, _ = self._slot[id(param)]
                with self._on_compute_stream():
                    self.flat[o:o + n].add_(param.grad.reshape(-1).float())
                    param.grad.zero_()      # folded; finish() writes the reduced value back
            self._ready(bi)
        return hook

    def _on_compute_stream(self):
        """The stream the step's kernels run on (taken in `zero()` on the caller's thread).  The hook itself runs on
        the autograd thread under the stream its AccumulateGrad node was CREATED on — and that node outlives a step
        whenever something keeps the step's loss alive (dp.GraphedStep keeps the captured one): in a later eager step
        the hook would then see a stale side stream, the bucket's event would not cover the gradient kernels and the
        all-reduce would race with them (found on 2 GPUs: replicas drifted apart bit by bit)."""
        import contextlib
        if self._compute_stream is None:
            return contextlib.nullcontext()
        return torch.cuda.stream(self._compute_stream)

    def _ready(self, bi):
        b = self.buckets[bi]
        b.pending -= 1
        if b.pending == 0 and self._sync:
            self._fire(b)

    def _fire(self, b):
        if b.fired:
            return
        b.fired = True
        if self.world == 1:
            return
        view = self.flat[b.lo:b.hi]
        nccl = self._comm is not None
        op = dist.ReduceOp.AVG if (nccl and self.average) else dist.ReduceOp.SUM
        if nccl:
            self._comm_used = True
            ev = torch.cuda.Event()
            ev.record(self._compute_stream or torch.cuda.current_stream(self.device))    # see _on_compute_stream
            with torch.cuda.stream(self._comm):
                self._comm.wait_event(ev)
                b.work = dist.all_reduce(view, op=op, group=self.group, async_op=True)
        else:
            b.work = dist.all_reduce(view, op=op, group=self.group, async_op=True)

    # ------------------------------------------------------------------ public
    def zero(self):
        """Zero the arena and re-arm the buckets (replaces optimizer.zero_grad()).  Call it on the stream the step
        runs on: that stream is what the bucket events are recorded on."""
        self._compute_stream = torch.cuda.current_stream(self.device) if self._comm is not None else None
        self.flat.zero_()
        for lp in self._lowp.values():
            lp.zero_()
        for b in self.buckets:
            b.pending, b.work, b.fired = b.expected, None, False

    class _NoSync:
        def __init__(self, arena):
            self.arena = arena

        def __enter__(self):
            self.arena._sync = False

        def __exit__(self, *exc):
            self.arena._sync = True
            for b in self.arena.buckets:      # re-arm: the next backward is the one that fires
                b.pending = b.expected
            return False

    def no_sync(self):
        """Gradient accumulation: backward passes inside this context only accumulate; the
        all-reduce happens at `finish()` (accelerate's `accumulate`, train_lora_dreambooth.py:817)."""
        return LoraGradArena._NoSync(self)

    def finish(self):
        """Join the exchange: every bucket reduced, gradients averaged, `p.grad` final."""
        for b in self.buckets:
            if not b.fired:
                self._fire(b)
        probe = self.exposure_probe if (self._comm is not None and self._comm_used) else None
        if probe is not None:
            e0 = probe()
            e0.record(torch.cuda.current_stream(self.device))
        for b in self.buckets:
            if b.work is not None:
                b.work.wait()
                b.work = None
        if self._comm is not None and self._comm_used:
            # (only when something ran there this step: under CUDA-graph capture a stream that never
            # joined the capture must not be waited on)
            torch.cuda.current_stream(self.device).wait_stream(self._comm)
            self._comm_used = False
        if probe is not None:
            e1 = probe()
            e1.record(torch.cuda.current_stream(self.device))
            self.exposure_events.append((e0, e1))
        if self.world > 1 and self.average and self._comm is None:
            self.flat.mul_(1.0 / self.world)        # gloo has no AVG
        for dt, lp in self._lowp.items():
            lp.copy_(self.flat)

    def clip_grad_norm_(self, max_norm: float):
        """Global L2 clip over the arena (clip_grad_norm_ over all LoRA parameters,
        train_lora_dreambooth.py:878-884 / cli_lora_pti.py:448-450) in one pass."""
        total = torch.linalg.vector_norm(self.flat)
        coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        self.flat.mul_(coef)
        for lp in self._lowp.values():
            lp.mul_(coef.to(lp.dtype))
        return total

    def broadcast_params(self, src: int = 0):
        """Replicas must start identical: lora_down is random-initialised per process
        (lora.py:46), so rank `src`'s values are broadcast once (what DDP does at wrap time)."""
        if self.world == 1:
            return
        with torch.no_grad():
            flat = torch.cat([p.detach().reshape(-1).float() for p in self.params])
            dist.broadcast(flat, src=src, group=self.group)
            off = 0
            for p in self.params:
                n = p.numel()
                p.copy_(flat[off:off + n].view_as(p).to(p.dtype))
                off += n

    @property
    def payload_bytes(self):
        return self.flat.numel() * 4

    def detach(self):
        """Undo the wiring (hooks and sinks); gradients stay where they are."""
        for h in self._hooks:
            h.remove()
        self._hooks = []
        for layer in self._layers:
            if isinstance(layer, LoraInjectedLinear):
                layer._grad_sink = None


class GraphedStep:
    """A whole training step — arena.zero(), forward, backward (with the bucketed exchange on its side
    stream), clip, optimizer — captured ONCE as a CUDA graph and replayed.

    Why: the LoRA step of an SD-1.5 UNet is ~3000 kernel launches; the host needs ~50 ms to enqueue them
    against ~55 ms of GPU work, so any host jitter (8 ranks sharing the cores of one node) shows up as
    step time.  Replaying a graph takes the host out of the loop; the kernels, their streams and their
    programmatic-launch edges are the ones the eager step issued.

    `step_fn(batch_dict) -> loss tensor` must be free of host synchronisation (no .item(), optimizer
    built with capturable=True) and shape-static.  Inputs are copied into static buffers before every
    replay; the returned loss tensor is overwritten by each replay.  The LoRA packs are rebuilt by every
    training forward (ops.BaseWeightCache.lora_pack: never cached while the factors are trained), so the
    pack launches are part of the captured step and every replay packs the live parameters.
    """

    def __init__(self, step_fn, example_batch: dict, warmup: int = 3, before_capture=None):
        dev = next(iter(example_batch.values())).device
        self.static = {k: v.clone() for k, v in example_batch.items()}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                step_fn(self.static)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        from . import _native as nat
        self.graph = torch.cuda.CUDAGraph()
        if before_capture is not None:      # e.g. a launch profiler dropping what the warm-up steps recorded
            before_capture()
        n0 = nat.LAUNCHES
        with torch.cuda.graph(self.graph):
            self.loss = step_fn(self.static)
        self.kernel_launches = nat.LAUNCHES - n0     # C-ABI launches captured = launched per replay

    def __call__(self, batch: dict):
        for k, v in batch.items():
            self.static[k].copy_(v, non_blocking=True)
        self.graph.replay()
        return self.loss


class ArenaAdamW:
    """Global-norm clip + AdamW over the flat arenas in one C-ABI call.

    Replaces `clip_grad_norm_(params, max_norm)` + `torch.optim.AdamW(params).step()`
    (train_lora_dreambooth.py:676, 878-888; cli_lora_pti.py:448-451) for the LoRA parameters of a
    `LoraGradArena`.  The parameters are re-homed into ONE flat fp32 buffer laid out exactly like the
    gradient arena (`p.data` becomes a view of it; the Parameter objects, and so every reference the
    modules, the optimizer-free training loop and `save_lora_weight` hold, stay the same), so the
    update is two kernels over contiguous memory instead of ~250 small tensors.  The step count lives
    on the device: the whole step stays CUDA-graph capturable.

    fp32 parameters only (fp32 masters under autocast, the reference's mixed-precision mode); anything
    else raises — there is no per-tensor fallback.
    """

    def __init__(self, arena: "LoraGradArena", lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, max_grad_norm: Optional[float] = None):
        from . import _native as nat
        if arena.device.type != "cuda":
            raise nat.B200LoraError("ArenaAdamW runs on the b200lora kernels: CUDA arena required (no CPU fallback)")
        for p in arena.params:
            if p.dtype != torch.float32:
                raise nat.B200LoraError(f"ArenaAdamW needs fp32 LoRA parameters (got {p.dtype}); keep fp32 masters "
                                        "and run the model under autocast")
        self.arena = arena
        self.betas, self.eps, self.weight_decay = (float(betas[0]), float(betas[1])), float(eps), float(weight_decay)
        # torch layout (one group) so LR schedulers that walk `param_groups` drive this optimizer too
        self.param_groups = [{"params": arena.params, "lr": float(lr), "initial_lr": float(lr)}]
        self.max_grad_norm = None if max_grad_norm is None else float(max_grad_norm)
        n = arena.flat.numel()
        self.flat_p = torch.empty(n, dtype=torch.float32, device=arena.device)
        with torch.no_grad():
            for p in arena.params:                     # same offsets as the gradient arena
                o, cnt, _ = arena._slot[id(p)]
                self.flat_p[o:o + cnt].copy_(p.detach().reshape(-1))
                p.data = self.flat_p[o:o + cnt].view_as(p)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=arena.device)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=arena.device)
        self.scratch = torch.zeros(2, dtype=torch.float32, device=arena.device)   # [sum g^2, step count]

    @property
    def lr(self) -> float:
        return float(self.param_groups[0]["lr"])

    @lr.setter
    def lr(self, value):
        self.param_groups[0]["lr"] = float(value)

    def _check_homes(self):
        # monkeypatch_* / weight loading re-assign `.weight` with fresh Parameters: the arena (and this
        # optimizer) must be rebuilt after that, exactly as a torch optimizer would have to be
        for p in self.arena.params:
            o, cnt, _ = self.arena._slot[id(p)]
            if p.data_ptr() != self.flat_p.data_ptr() + 4 * o:
                raise RuntimeError("a LoRA parameter no longer lives in the ArenaAdamW parameter arena "
                                   "(its .data was replaced); rebuild LoraGradArena and ArenaAdamW")

    @torch.no_grad()
    def step(self):
        """Call after `arena.finish()`.  Clips (if max_grad_norm was given) and updates in place."""
        from . import _native as nat
        self._check_homes()
        dev = self.arena.device
        with torch.cuda.device(dev):
            rc = nat.lib().b200lora_arena_adamw(
                nat.ptr(self.flat_p), nat.ptr(self.arena.flat), nat.ptr(self.exp_avg), nat.ptr(self.exp_avg_sq),
                self.flat_p.numel(), self.lr, self.betas[0], self.betas[1], self.eps, self.weight_decay,
                self.max_grad_norm if self.max_grad_norm is not None else 0.0, nat.ptr(self.scratch),
                nat.stream_ptr(dev))
        nat.check(rc, "b200lora_arena_adamw")
        # the kernel wrote through raw pointers: tell autograd's version counters, which key the
        # per-layer LoRA pack caches (ops.BaseWeightCache.lora_pack, fused.SiblingGroup.pack)
        for p in self.arena.params:
            torch.autograd.graph.increment_version(p)

    def grad_norm(self):
        """L2 norm of the gradients seen by the last step (before clipping), as a device tensor."""
        return self.scratch[0].sqrt()

    def zero_grad(self, set_to_none: bool = False):
        self.arena.zero()

    def state_dict(self):
        return {"exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "step": self.scratch[1:2],
                "lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay,
                "max_grad_norm": self.max_grad_norm}

    def load_state_dict(self, sd):
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.scratch[1:2].copy_(sd["step"])
        for k in ("lr", "betas", "eps", "weight_decay", "max_grad_norm"):
            if k in sd:
                setattr(self, k, sd[k])

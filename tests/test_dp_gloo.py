"""Host logic of the data-parallel exchange step (diffusion_finetuning_b200/dp.py) on CPU with the
gloo backend, world_size 2: bucketing, ready-counting, averaging, initial broadcast, clip.
The LoRA layers' own forward is CUDA-only, so the loss here is a torch restatement over the
injected layers' parameters (oracle math) — what is under test is the exchange, not the kernel."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class CrossAttention(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.to_q = nn.Linear(d, d, bias=False)
        self.to_k = nn.Linear(d, d, bias=False)
        self.to_v = nn.Linear(d, d, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(d, d)])


def _model(seed):
    import diffusion_finetuning_b200 as L
    torch.manual_seed(0)
    m = nn.Sequential(CrossAttention(32), CrossAttention(32), CrossAttention(32))
    m.requires_grad_(False)
    L.inject_trainable_lora(m, {"CrossAttention"}, r=4)
    g = torch.Generator().manual_seed(seed)         # per-rank init on purpose (lora.py:46)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, L.LoraInjectedLinear):
                mod.lora_down.weight.copy_(torch.randn(4, 32, generator=g) / 4)
                mod.lora_up.weight.copy_(torch.randn(32, 4, generator=g) * 0.02)
    return m


def _loss(m, x):
    """sum over layers of the oracle forward (lora.py:49-50), chained"""
    import diffusion_finetuning_b200 as L
    from oracle import lora_oracle as O
    h = x
    for mod in m.modules():
        if isinstance(mod, L.LoraInjectedLinear):
            h = torch.tanh(O.lora_linear_forward(h, mod.linear.weight, mod.linear.bias, mod.lora_down.weight,
                                                 mod.lora_up.weight, mod.scale))
    return (h ** 2).mean()


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from diffusion_finetuning_b200 import dp
        m = _model(seed=10 + rank)
        arena = dp.LoraGradArena(m, bucket_bytes=2048)
        assert len(arena.buckets) > 2
        arena.broadcast_params(0)
        p0 = torch.cat([p.detach().reshape(-1) for p in arena.params])
        gathered = [torch.empty_like(p0) for _ in range(world)]
        dist.all_gather(gathered, p0)
        assert all(torch.equal(gathered[0], g) for g in gathered), "replicas differ after broadcast"
        g = torch.Generator().manual_seed(99)
        x_all = torch.randn(world * 6, 32, generator=g)
        # two micro-steps with accumulation, then the exchange
        arena.zero()
        xs = x_all[rank * 6:(rank + 1) * 6]
        with arena.no_sync():
            (_loss(m, xs[:3]) * 0.5).backward()
        (_loss(m, xs[3:]) * 0.5).backward()
        arena.finish()
        got = torch.cat([p.grad.reshape(-1) for p in arena.params]).clone()
        norm = arena.clip_grad_norm_(1e-3)
        clipped = torch.cat([p.grad.reshape(-1) for p in arena.params]).clone()
        if rank == 0:
            torch.save({"got": got, "norm": norm, "clipped": clipped, "x": x_all,
                        "state": {k: v.clone() for k, v in m.state_dict().items()}}, out)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.timeout(180)
def test_dp2_gradients_equal_single_process_on_concatenated_batch(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    blob = torch.load(out)
    from diffusion_finetuning_b200 import dp
    m = _model(seed=10)                      # rank 0's values == the broadcast values
    m.load_state_dict(blob["state"])
    arena = dp.LoraGradArena(m, bucket_bytes=2048)
    x = blob["x"]
    # mean over ranks of (mean of two half-batch losses) == mean over the four quarter losses
    loss = sum(_loss(m, x[i * 3:(i + 1) * 3]) for i in range(4)) / 4
    arena.zero()
    loss.backward()
    arena.finish()
    want = torch.cat([p.grad.reshape(-1) for p in arena.params])
    assert want.abs().max() > 0
    rel = (blob["got"] - want).norm() / want.norm()
    assert rel < 1e-5, rel
    assert abs(blob["norm"].item() - want.norm().item()) / want.norm().item() < 1e-5
    assert abs(blob["clipped"].norm().item() - 1e-3) < 1e-6


def test_arena_layout_and_low_precision_params():
    from diffusion_finetuning_b200 import dp
    m = _model(seed=1).to(torch.bfloat16)
    arena = dp.LoraGradArena(m, bucket_bytes=1 << 20)
    assert len(arena.buckets) == 1 and arena.payload_bytes == 4 * sum(p.numel() for p in arena.params)
    x = torch.randn(4, 32).to(torch.bfloat16)
    arena.zero()
    _loss(m, x).float().backward()
    arena.finish()
    for p in arena.params:
        assert p.grad.dtype == torch.bfloat16 and p.grad.shape == p.shape
    assert sum(p.grad.float().abs().sum() for p in arena.params) > 0
    # every parameter got a disjoint slice, in reverse module order
    offs = sorted(arena._slot.values())
    assert offs[0][0] == 0 and all(a[0] + a[1] == b[0] for a, b in zip(offs, offs[1:]))


# ---------------------------------------------------------------------------------------------------
# GradSink protocol (ops.GradSink): the kernel adds into the arena and the Function returns NO gradient for
# the factors; the arena's post-accumulate hook — once per backward per parameter, after every use of the
# layer — is the only "slice final" signal.  A layer that runs TWICE in one step must be exchanged after its
# LAST backward, and a sink layer must count once, not twice (round 1 signalled from the sink AND the hook:
# buckets fired after half of their layers).
# ---------------------------------------------------------------------------------------------------
class _SinkLinear(torch.autograd.Function):
    """What ops._LoraLinearFn does with a sink, with the oracle's math in place of the CUDA kernels."""
    pending = {}        # id(module) -> backward calls still to come (test instrumentation)

    @staticmethod
    def forward(ctx, x, mod, A, B):
        from oracle import lora_oracle as O
        ctx.mod = mod
        ctx.save_for_backward(x)
        _SinkLinear.pending[id(mod)] = _SinkLinear.pending.get(id(mod), 0) + 1
        return O.lora_linear_forward(x, mod.linear.weight, mod.linear.bias, A.detach(), B.detach(), mod.scale)

    @staticmethod
    def backward(ctx, dy):
        from oracle import lora_oracle as O
        (x,) = ctx.saved_tensors
        mod = ctx.mod
        dx, dA, dB = O.lora_linear_grads(x, mod.linear.weight.detach(), None, mod.lora_down.weight.detach(),
                                         mod.lora_up.weight.detach(), mod.scale, dy)
        mod._grad_sink.dA.add_(dA)
        mod._grad_sink.dB.add_(dB)
        _SinkLinear.pending[id(mod)] -= 1
        return dx, None, None, None


def _layers(m):
    import diffusion_finetuning_b200 as L
    return [mod for mod in m.modules() if isinstance(mod, L.LoraInjectedLinear)]


def _loss_twice(m, x, through_sink):
    """every layer is applied TWICE (weights shared between the two uses)"""
    from oracle import lora_oracle as O
    h = x
    for _ in range(2):
        for mod in _layers(m):
            if through_sink:
                h = torch.tanh(_SinkLinear.apply(h, mod, mod.lora_down.weight, mod.lora_up.weight))
            else:
                h = torch.tanh(O.lora_linear_forward(h, mod.linear.weight, mod.linear.bias, mod.lora_down.weight,
                                                     mod.lora_up.weight, mod.scale))
    return (h ** 2).mean()


def _worker_twice(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from diffusion_finetuning_b200 import dp
        m = _model(seed=10 + rank)
        arena = dp.LoraGradArena(m, bucket_bytes=2048)
        assert len(arena.buckets) > 2
        arena.broadcast_params(0)
        early = []
        orig_fire = arena._fire

        def spy(b):
            bi = arena.buckets.index(b)
            for layer in _layers(m):
                if arena._slot[id(layer.lora_up.weight)][2] == bi or arena._slot[id(layer.lora_down.weight)][2] == bi:
                    early.append(_SinkLinear.pending.get(id(layer), 0) > 0)
            return orig_fire(b)
        arena._fire = spy
        g = torch.Generator().manual_seed(7)
        x_all = torch.randn(world * 4, 32, generator=g)
        arena.zero()
        x = x_all[rank * 4:(rank + 1) * 4].clone().requires_grad_(True)
        _loss_twice(m, x, through_sink=True).backward()
        fired_in_backward = sum(b.fired for b in arena.buckets)
        arena.finish()
        assert early and not any(early), "a bucket was exchanged while one of its layers still had a backward to run"
        assert fired_in_backward == len(arena.buckets), "every bucket should fire from inside the backward"
        got = arena.flat.clone()
        if rank == 0:
            torch.save({"got": got, "x": x_all, "state": {k: v.clone() for k, v in m.state_dict().items()}}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_layer_used_twice_per_step_is_exchanged_after_its_last_backward(tmp_path):
    out = str(tmp_path / "twice.pt")
    mp.spawn(_worker_twice, args=(2, _free_port(), out), nprocs=2, join=True)
    blob = torch.load(out)
    m = _model(seed=10)
    m.load_state_dict(blob["state"])
    params = [p for mod in reversed(_layers(m)) for p in (mod.lora_up.weight, mod.lora_down.weight)]
    x = blob["x"]
    loss = (_loss_twice(m, x[:4], False) + _loss_twice(m, x[4:], False)) / 2
    grads = torch.autograd.grad(loss, params)
    want = torch.cat([g.reshape(-1) for g in grads])
    rel = (blob["got"] - want).norm() / want.norm()
    assert rel < 1e-5, rel

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

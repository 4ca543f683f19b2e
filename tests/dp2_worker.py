"""torchrun worker of tests/test_gpu_dp2.py: DP-N gradients after the bucketed NCCL all-reduce == single-process
gradients on the concatenated batch (reference semantics: accelerate DDP averaging,
train_lora_dreambooth.py:751-757, 877), through the REAL kernel path: ops.GradSink (the reduction kernels add into
the flat arena), buckets fired from inside the backward on the exchange stream, and the same step replayed as a
captured CUDA graph (dp.GraphedStep).  Also: after 3 optimizer steps every rank holds bit-identical parameters.

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 tests/dp2_worker.py OUT.json
"""
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def build(device, fuse, extended=False):
    import diffusion_finetuning_b200 as L
    from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
    torch.manual_seed(0)
    unet = UNet2DConditionModel.tiny().to(device)
    unet.requires_grad_(False)
    torch.manual_seed(1)
    if extended:        # resnet convolutions too: their factor gradients come back THROUGH autograd (no GradSink)
        L.inject_trainable_lora_extended(unet, r=4)
    else:
        L.inject_trainable_lora(unet, r=4)
    g = torch.Generator().manual_seed(2)
    with torch.no_grad():
        for m in unet.modules():
            if isinstance(m, (L.LoraInjectedLinear, L.LoraInjectedConv2d)):
                m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(device))
    if fuse:
        L.fuse_sibling_projections(unet)
        L.fuse_geglu(unet)
    return unet


def batch(n, seed, device):
    g = torch.Generator().manual_seed(seed)
    return {"latents": torch.randn(n, 4, 32, 32, generator=g).to(device), "t": torch.randint(0, 1000, (n,), generator=g).to(device),
            "ctx": torch.randn(n, 77, 32, generator=g).to(device), "target": torch.randn(n, 4, 32, 32, generator=g).to(device)}


def loss_of(unet, b, autocast):
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        pred = unet(b["latents"], b["t"], b["ctx"])
    return F.mse_loss(pred.float(), b["target"].float(), reduction="mean")


def rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


def main():
    out_path = sys.argv[1]
    os.environ.setdefault("B200LORA_DETERMINISTIC", "1")     # reductions without atomics: run-to-run identical
    torch.backends.cudnn.allow_tf32 = False                  # the stock convs between the LoRA layers in true fp32:
    torch.backends.cuda.matmul.allow_tf32 = False            # a batch of 4 and two batches of 2 must agree to 1e-5
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", device_id=device)
    from diffusion_finetuning_b200 import dp
    res = {}
    per = 2
    full = batch(per * world, 77, device)
    mine = {k: v[rank * per:(rank + 1) * per].contiguous() for k, v in full.items()}
    for name, autocast, fuse, tol, ext in (("fp32", False, False, 1e-5, False), ("bf16_fused", True, True, 2e-2, False),
                                           ("bf16_extended", True, True, 3e-2, True)):
        # single process, concatenated batch (same seeds on every rank -> same model)
        ref = build(device, fuse, ext)
        ref_arena = dp.LoraGradArena(ref, bucket_bytes=4096, process_group=None)
        ref_arena.world = 1
        ref_arena.zero()
        loss_of(ref, full, autocast).backward()
        torch.cuda.synchronize(device)
        want = ref_arena.flat.clone()
        ref_arena.detach()
        # DP over the shards: eager, buckets fire inside the backward
        unet = build(device, fuse, ext)
        arena = dp.LoraGradArena(unet, bucket_bytes=4096)
        assert len(arena.buckets) > 2
        arena.broadcast_params(0)
        arena.zero()
        loss_of(unet, mine, autocast).backward()
        fired = sum(b.fired for b in arena.buckets)
        arena.finish()
        torch.cuda.synchronize(device)
        got = arena.flat.clone()
        res[name] = {"eager_rel": rel(got, want), "tol": tol, "buckets": len(arena.buckets),
                     "fired_in_backward": fired, "grad_norm": want.norm().item()}
        # the same step as a captured graph (exchange stream + NCCL inside the graph)

        def step(b):
            arena.zero()
            loss = loss_of(unet, b, autocast)
            loss.backward()
            arena.finish()
            return loss
        graphed = dp.GraphedStep(step, mine)
        graphed(mine)
        torch.cuda.synchronize(device)
        res[name]["graph_rel"] = rel(arena.flat, want)
        res[name]["graph_vs_eager_rel"] = rel(arena.flat, got)
        # three eager + three replayed optimizer steps on rank-specific batches: replicas stay bit-identical (the eager
        # steps run with the captured step's autograd nodes still alive — the stale-stream case of dp.LoraGradArena)
        opt = torch.optim.AdamW(arena.params, lr=1e-3, fused=True)
        for i in range(6):
            b = batch(per, 100 + 10 * i + rank, device)
            if i < 3:
                step(b)
            else:
                graphed(b)
            arena.clip_grad_norm_(1.0)
            opt.step()
        flat = torch.cat([p.detach().reshape(-1) for p in arena.params]).view(torch.int32)
        hi, lo = flat.clone(), flat.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        res[name]["replicas_bit_identical"] = bool((hi == lo).all().item())
        del graphed
        arena.detach()
    import gc
    gc.collect()
    torch.cuda.synchronize(device)
    dist.barrier()
    if rank == 0:
        json.dump(res, open(out_path, "w"), indent=1)
        print(json.dumps(res))
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

"""2-rank diagnosis of replica divergence under the captured step: after every step compare the exchanged gradient
arena and the parameters across ranks, per bucket.   torchrun --nproc-per-node 2 tools/dp_debug.py [--no-graph]"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.argv = [sys.argv[0]] + sys.argv[1:]
import bench  # noqa: E402


def differing(t):
    v = t.detach().reshape(-1).float().view(torch.int32)
    hi, lo = v.clone(), v.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    return hi != lo


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--clip-off", action="store_true")
    a = ap.parse_args()
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    tr = bench.Trainer(dev, 4, capturable=not a.no_graph)
    pool = bench.make_batches(4, 8, 64, 1234 + rank, dev)
    if not a.no_graph:
        tr.capture(pool[0])
    d0 = differing(torch.cat([p.detach().reshape(-1) for p in tr.arena.params]))
    if rank == 0:
        print(f"after capture/warm-up: params differing {int(d0.sum())}", flush=True)
    for i in range(a.steps):
        tr.step(pool[i % 4])
        torch.cuda.synchronize(dev)
        dg = differing(tr.arena.flat)
        dp = differing(torch.cat([p.detach().reshape(-1) for p in tr.arena.params]))
        if rank == 0:
            sizes = [b.hi - b.lo for b in tr.arena.buckets]
            per_g = [int(dg[b.lo:b.hi].sum()) for b in tr.arena.buckets]
            print(f"step {i}: grads differing per bucket {per_g} of {sizes}; params differing {int(dp.sum())}", flush=True)
            if dg.any():
                idx = dg.nonzero().reshape(-1)
                print(f"   first/last differing grad index {int(idx[0])} .. {int(idx[-1])}; "
                      f"runs: {int((idx[1:] - idx[:-1] > 1).sum()) + 1}", flush=True)
    tr.release()
    del tr
    import gc
    gc.collect()
    torch.cuda.synchronize(dev)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Kernel-level breakdown of the training step (torch.profiler, CUDA activities).
    python tools/prof_step.py [--config 3]
"""
import argparse, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from torch.profiler import profile, ProfilerActivity
ap = argparse.ArgumentParser(); ap.add_argument("--config", type=int, default=2); ap.add_argument("--rows", type=int, default=30)
a = ap.parse_args()
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
import __graft_entry__ as e; e.build()
tr = bench.Trainer(dev, 16 if a.config == 3 else 4, extended=a.config == 3)
pool = bench.make_batches(2, 8, 64, 1234, dev)
for i in range(3): tr.step(pool[i % 2])
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(3): tr.step(pool[i % 2])
    torch.cuda.synchronize()
ev = prof.key_averages()
tot = sum(x.self_device_time_total for x in ev)
print(f"total kernel time per step: {tot / 3 / 1e3:.2f} ms")
for x in sorted(ev, key=lambda x: -x.self_device_time_total)[:a.rows]:
    print(f"{x.self_device_time_total / tot * 100:6.2f}%  {x.self_device_time_total / 3 / 1e3:8.3f} ms/step  n={x.count // 3:5d}  avg {x.self_device_time_total / max(1, x.count):8.1f} us  {x.key[:110]}")

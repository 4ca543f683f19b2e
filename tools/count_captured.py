"""Which C-ABI calls does one captured training step contain?  (bench.py's Trainer, config 2)"""
import collections, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from diffusion_finetuning_b200 import _native as nat
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
import __graft_entry__ as e; e.build()
tr = bench.Trainer(dev, 4, capturable=True)
pool = bench.make_batches(2, 8, 64, 1234, dev)
counts = collections.Counter(); orig = nat.check
def check(rc, what):
    counts[what] += 1
    return orig(rc, what)
nat.check = check
import diffusion_finetuning_b200.ops as ops, diffusion_finetuning_b200.fused as fused
for _ in range(3): tr.eager_step(pool[0])
torch.cuda.synchronize(); counts.clear()
tr.eager_step(pool[1]); torch.cuda.synchronize()
print("eager step:", dict(counts)); counts.clear()
tr.capture(pool[0])
print("warmup(3)+capture:", dict(counts), "captured launches:", tr.graphed.kernel_launches)

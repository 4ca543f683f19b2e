"""How far ahead of the GPU does the host run?  CPU time to ENQUEUE one training step vs its GPU time."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
import __graft_entry__ as e; e.build()
tr = bench.Trainer(dev, 4)
pool = bench.make_batches(2, 8, 64, 1234, dev)
for i in range(4): tr.step(pool[i % 2])
torch.cuda.synchronize()
cpu, gpu = [], []
for i in range(6):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    tr.step(pool[i % 2])
    e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize()
    cpu.append((t1 - t0) * 1e3); gpu.append(e0.elapsed_time(e1))
print("cpu enqueue ms", [round(c, 1) for c in cpu]); print("gpu ms", [round(g, 1) for g in gpu]); print("cores", os.cpu_count())

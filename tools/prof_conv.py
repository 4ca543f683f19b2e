"""Kernel-level breakdown of one LoRA-Conv2d forward+backward (torch.profiler, CUDA activities)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffusion_finetuning_b200 as L
from torch.profiler import profile, ProfilerActivity
ci, co, H, ks, r = [int(v) for v in (sys.argv[1:6] if len(sys.argv) > 5 else (320, 320, 64, 3, 16))]
dev, dt = torch.device("cuda:0"), torch.bfloat16
m = L.LoraInjectedConv2d(ci, co, ks, padding=ks // 2, r=r).to(dev).to(dt).to(memory_format=torch.channels_last)
with torch.no_grad():
    m.lora_up.weight.normal_(0, 0.02)
m.conv.requires_grad_(False)
x = torch.randn(8, ci, H, H, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last).requires_grad_(True)
dy = torch.randn(8, co, H, H, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
for _ in range(3):
    x.grad = None; m(x).backward(dy)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(5):
        x.grad = None; m(x).backward(dy)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=70))

"""K4: LoRA-Conv2d on BASELINE config 3's resnet shapes: our implicit-GEMM path vs the stock
PyTorch/cuDNN evaluation of the same module (conv + down conv + 1x1 up + mul + add), bf16 channels_last,
batch 8, r = 16.  CUDA events, L2-flushing pool of inputs, fwd and fwd+bwd.
    python tools/bench_conv.py --json gpurun_out/conv.json
"""
import argparse, json, os, sys
import torch
import torch.nn as nn
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffusion_finetuning_b200 as L

# (C_in, C_out, H, ksize, count): conv1 / conv2 / conv_shortcut of the 22 ResnetBlock2D
SHAPES = [(320, 320, 64, 3, 6), (640, 640, 32, 3, 5), (1280, 1280, 16, 3, 5), (1280, 1280, 8, 3, 12),
          (320, 640, 32, 3, 1), (640, 1280, 16, 3, 1), (2560, 1280, 8, 3, 3), (2560, 1280, 16, 3, 2),
          (1920, 1280, 16, 3, 1), (1920, 640, 32, 3, 1), (1280, 640, 32, 3, 1), (960, 640, 32, 3, 1),
          (960, 320, 64, 3, 1), (640, 320, 64, 3, 2), (640, 320, 64, 1, 2), (2560, 1280, 8, 1, 3)]


class Stock(nn.Module):
    """the upstream module evaluated by stock PyTorch ops"""
    def __init__(self, m):
        super().__init__()
        self.conv, self.lora_down, self.lora_up, self.scale = m.conv, m.lora_down, m.lora_up, m.scale
    def forward(self, x):
        return self.conv(x) + self.lora_up(self.lora_down(x)) * self.scale


def timeit(fn, iters):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


ap = argparse.ArgumentParser()
ap.add_argument("--r", type=int, default=16); ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--iters", type=int, default=20); ap.add_argument("--json", default=None)
ap.add_argument("--hw", type=int, default=0, help="only the shapes at this spatial size (0 = all)")
ap.add_argument("--no-torch", action="store_true", help="skip the stock arm (tile-width A/Bs via B200LORA_BN)")
a = ap.parse_args()
if a.hw:
    SHAPES = [s for s in SHAPES if s[2] == a.hw]
dev = torch.device("cuda:0"); dt = torch.bfloat16
rows = []
for (ci, co, H, ks, cnt) in SHAPES:
    torch.manual_seed(0)
    m = L.LoraInjectedConv2d(ci, co, ks, padding=ks // 2, r=a.r).to(dev).to(dt).to(memory_format=torch.channels_last)
    with torch.no_grad():
        m.lora_up.weight.normal_(0, 0.02)
    m.conv.requires_grad_(False)
    st = Stock(m)
    pool = [torch.randn(a.batch, ci, H, H, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last).requires_grad_(True)
            for _ in range(4)]
    dy = torch.randn(a.batch, co, H, H, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
    i = [0]
    def fwd(mod):
        with torch.no_grad():
            return mod(pool[i[0] % 4])
    def fb(mod):
        i[0] += 1
        x = pool[i[0] % 4]; x.grad = None
        mod(x).backward(dy)
    flops = 2.0 * a.batch * H * H * co * ci * ks * ks
    r_ = {"Cin": ci, "Cout": co, "H": H, "k": ks, "count": cnt,
          "fwd_us": timeit(lambda: fwd(m), a.iters),
          "torch_fwd_us": float("nan") if a.no_torch else timeit(lambda: fwd(st), a.iters),
          "fwdbwd_us": timeit(lambda: fb(m), a.iters),
          "torch_fwdbwd_us": float("nan") if a.no_torch else timeit(lambda: fb(st), a.iters)}
    r_["fwd_tflops"] = flops / r_["fwd_us"] / 1e6
    r_["fwdbwd_tflops"] = 2 * flops / r_["fwdbwd_us"] / 1e6       # fwd + dX (W0 frozen: no dW)
    r_["speedup_fwd"] = r_["torch_fwd_us"] / r_["fwd_us"]; r_["speedup_fwdbwd"] = r_["torch_fwdbwd_us"] / r_["fwdbwd_us"]
    rows.append(r_); print(json.dumps({k: (round(v, 2) if isinstance(v, float) else v) for k, v in r_.items()}), flush=True)
    del m, st, pool, dy
tot = lambda k: sum(r[k] * r["count"] for r in rows)
summ = {"resnet_convs_fwd_ms": tot("fwd_us") / 1e3, "torch_fwd_ms": tot("torch_fwd_us") / 1e3,
        "resnet_convs_fwdbwd_ms": tot("fwdbwd_us") / 1e3, "torch_fwdbwd_ms": tot("torch_fwdbwd_us") / 1e3}
print(json.dumps(summ))
if a.json:
    json.dump({"r": a.r, "batch": a.batch, "rows": rows, "summary": summ}, open(a.json, "w"), indent=1)

"""GEGLU projection: activation in the GEMM epilogue (fuse_geglu) vs our layer + the module's chunk / gelu / mul,
forward and forward+backward, SD-1.5 feed-forward shapes at batch 8 x 64x64 latents, bf16, r = 16 (config 3).
    python tools/bench_geglu.py > profiles/r02_geglu_bench.json"""
import json
import sys

import torch

sys.path.insert(0, ".")
import diffusion_finetuning_b200 as L  # noqa: E402
from diffusion_finetuning_b200.unet_sd15 import GEGLU  # noqa: E402

DEV = torch.device("cuda:0")


def timed(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


rows = []
for M, K, d_ff in ((32768, 320, 1280), (8192, 640, 2560), (2048, 1280, 5120)):
    torch.manual_seed(0)
    m = torch.nn.Sequential(GEGLU(K, d_ff)).to(DEV).to(torch.bfloat16)
    m.requires_grad_(False)
    L.inject_trainable_lora(m, {"GEGLU"}, r=16)
    with torch.no_grad():
        m[0].proj.lora_up.weight.normal_(0, 0.02)
    x = torch.randn(8, M // 8, K, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    dout = torch.randn(8, M // 8, d_ff, device=DEV, dtype=torch.bfloat16)

    def fwd():
        with torch.no_grad():
            return m(x)

    def fwdbwd():
        x.grad = None
        m(x).backward(dout)

    row = {"M": M, "K": K, "d_ff": d_ff, "r": 16}
    row["unfused_fwd_us"], row["unfused_fwdbwd_us"] = timed(fwd), timed(fwdbwd)
    L.fuse_geglu(m)
    row["fused_fwd_us"], row["fused_fwdbwd_us"] = timed(fwd), timed(fwdbwd)
    row["speedup_fwd"] = round(row["unfused_fwd_us"] / row["fused_fwd_us"], 3)
    row["speedup_fwdbwd"] = round(row["unfused_fwdbwd_us"] / row["fused_fwdbwd_us"], 3)
    flops = 2.0 * M * K * 2 * d_ff
    row["fused_fwd_tflops"] = round(flops / row["fused_fwd_us"] / 1e6, 1)
    rows.append({k: (round(v, 1) if isinstance(v, float) else v) for k, v in row.items()})
print(json.dumps({"what": "GEGLU.proj (LoRA r=16, bf16) + chunk/gelu/mul: epilogue-fused vs unfused, microseconds per call, "
                          "back-to-back launches from Python (host-enqueue bound on the small shapes)", "rows": rows}, indent=1))

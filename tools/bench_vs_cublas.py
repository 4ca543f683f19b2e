"""Standalone launch duration of the fused LoRA GEMM against the library GEMM of the same shape (torch.matmul -> cuBLAS
nvjet kernels), event-bracketed SINGLE launches separated by a small elementwise kernel (no back-to-back overlap, as in the
training step where stock kernels sit between our launches).  Answers: how much of the ~17 us of an N = K launch is the
price of being a 148-CTA TMA/tcgen05 kernel on a 6.7-GFLOP problem at all, and how much is ours.
    python tools/bench_vs_cublas.py > profiles/r02_gemm_vs_cublas.json"""
import json
import sys

import torch

sys.path.insert(0, ".")
import diffusion_finetuning_b200 as L  # noqa: E402

DEV = torch.device("cuda:0")


def single(fn, spacer, n=30, warm=3):
    for _ in range(warm):
        fn(); spacer()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        spacer()
        a.record(); fn(); b.record()
        small.add_(1.0)
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


rows = []
junk = torch.zeros(1 << 27, device=DEV)          # 512 MB: ~170 us per pass, so the host has enqueued the bracketed launch
small = torch.zeros(1 << 18, device=DEV)         # long before the device reaches it (device-side bracket, no host latency)


def spacer():
    junk.add_(1.0)
    small.add_(1.0)                              # the kernel right before ours is a SMALL one, like the casts in the step
for M, K, N in ((32768, 320, 320), (8192, 640, 640), (2048, 1280, 1280), (512, 1280, 1280), (616, 768, 1280),
                (32768, 320, 960), (8192, 640, 1920), (32768, 320, 2560), (2048, 1280, 10240)):
    torch.manual_seed(0)
    m = L.LoraInjectedLinear(K, N, bias=False, r=4).to(DEV).to(torch.bfloat16)
    m.requires_grad_(False)
    with torch.no_grad():
        m.lora_up.weight.normal_(0, 0.02)
    x = torch.randn(M, K, device=DEV, dtype=torch.bfloat16)
    w = m.linear.weight
    y = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    with torch.no_grad():
        ours = single(lambda: m(x), spacer)                       # inference: the pack is cached, ONE launch
        lib = single(lambda: torch.matmul(x, w.t(), out=y), spacer)
        ref3 = single(lambda: torch.nn.functional.linear(x, w) + torch.nn.functional.linear(
            torch.nn.functional.linear(x, m.lora_down.weight), m.lora_up.weight) * m.scale, spacer)
    fl = 2.0 * M * K * N
    rows.append({"M": M, "K": K, "N": N, "ours_fused_us": round(ours, 1), "cublas_base_gemm_only_us": round(lib, 1),
                 "stock_composition_5_launches_us": round(ref3, 1), "ideal_us_at_sustained_peak": round(fl / 1393.1e6, 1),
                 "ours_tflops": round(fl / ours / 1e6), "cublas_tflops": round(fl / lib / 1e6)})
print(json.dumps({"what": "median event-bracketed duration of ONE launch (a small elementwise kernel before and after), bf16, r = 4, "
                          "inference (cached pack): the fused LoRA layer vs torch.matmul of the base GEMM alone vs the reference composition",
                  "rows": rows}, indent=1))

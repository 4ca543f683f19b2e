"""Run one C-ABI kernel a few times (for ncu / sanitizer captures).
    python tools/run_one.py --what fwd --M 32768 --K 320 --N 2560 --r 4 --reps 3
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_finetuning_b200 import _native as nat  # noqa: E402


CENSUS = [(32768, 320, 320), (32768, 320, 2560), (8192, 640, 640), (8192, 640, 5120), (2048, 1280, 1280),
          (2048, 1280, 10240), (512, 1280, 1280), (512, 1280, 10240), (616, 768, 320), (616, 768, 640),
          (616, 768, 1280)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--census", action="store_true", help="every SURVEY section-8 shape, fwd + bdx + grads + merge")
    ap.add_argument("--what", default="fwd", choices=["fwd", "bdx", "grads", "merge", "all"])
    ap.add_argument("--M", type=int, default=32768)
    ap.add_argument("--K", type=int, default=320)
    ap.add_argument("--N", type=int, default=320)
    ap.add_argument("--r", type=int, default=4)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    if a.census:
        for (M, K, N) in CENSUS:
            run(M, K, N, a.r, "all", a.reps)
    else:
        run(a.M, a.K, a.N, a.r, a.what, a.reps)
    print("ok")


def run(M, K, N, r, what, reps):
    dev = torch.device("cuda:0")
    lib = nat.lib()
    dt, io, st = torch.bfloat16, nat.BF16, nat.stream_ptr(dev)
    x = torch.randn(M, K, device=dev, dtype=dt)
    w = (torch.rand(N, K, device=dev) * 2 - 1).mul_(K ** -0.5).to(dt)
    wt = w.t().contiguous()
    b = torch.zeros(N, device=dev, dtype=dt)
    A = (torch.randn(r, K, device=dev) / r).to(dt)
    B = (torch.randn(N, r, device=dev) * 0.02).to(dt)
    dy = torch.randn(M, N, device=dev, dtype=dt)
    y = torch.empty(M, N, device=dev, dtype=dt)
    T = torch.empty(M, r, device=dev, dtype=dt)
    dx = torch.empty(M, K, device=dev, dtype=dt)
    dT = torch.empty(M, r, device=dev, dtype=dt)
    dA = torch.zeros(r, K, device=dev)
    dB = torch.zeros(N, r, device=dev)
    ap_ = bp_ = torch.empty(80 * (K + N), dtype=dt, device=dev)
    nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), io, nat.ptr(b), io, nat.ptr(ap_), K, N, r, io, st), "pack")
    w32 = torch.randn(N, K, device=dev)
    for _ in range(reps):
        if what in ("fwd", "all"):
            nat.check(lib.b200lora_linear_fwd(nat.ptr(x), nat.ptr(w), nat.ptr(b), io, nat.ptr(A), nat.ptr(B), io,
                                              nat.ptr(ap_), 1.0, nat.ptr(y), nat.ptr(T), M, K, N, r, io, nat.IMPL_TC, st), "fwd")
        if what in ("bdx", "all"):
            nat.check(lib.b200lora_linear_bwd_dx(nat.ptr(dy), nat.ptr(wt), nat.W_NK, nat.ptr(A), nat.ptr(B), io,
                                                 nat.ptr(bp_), 1.0, nat.ptr(dx), nat.ptr(dT), M, K, N, r, io, nat.IMPL_TC, st), "bdx")
        if what in ("grads", "all"):
            nat.check(lib.b200lora_lora_grads(nat.ptr(dy), nat.ptr(x), nat.ptr(T), nat.ptr(dT), 1.0, nat.ptr(dA),
                                              nat.ptr(dB), M, K, N, r, io, 0, None, 0, st), "grads")
        if what in ("merge", "all"):
            nat.check(lib.b200lora_merge(nat.ptr(w32), nat.ptr(w32), nat.ptr(B.float()), nat.ptr(A.float()), 0.5,
                                         N, K, r, nat.F32, nat.F32, st), "merge")
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()

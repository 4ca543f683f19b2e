"""One fused q/k/v block (forward, K-segmented dX, multi-problem reductions) per SD-1.5 resolution — for ncu.
    python tools/run_fused.py
"""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_finetuning_b200 import _native as nat
dev = torch.device("cuda:0"); lib = nat.lib(); st = nat.stream_ptr(dev); dt = torch.bfloat16; io = nat.BF16
for (M, K, N, n) in ((32768, 320, 320, 3), (8192, 640, 640, 3), (2048, 1280, 1280, 3), (616, 768, 640, 2)):
    r = 4; rt = r * n; Nc = N * n
    x = torch.randn(M, K, device=dev, dtype=dt)
    Wc = (torch.rand(Nc, K, device=dev) - 0.5).mul_(2 * K ** -0.5).to(dt); Wct = Wc.t().contiguous()
    Ac = (torch.randn(rt, K, device=dev) / r).to(dt); Bbd = torch.zeros(Nc, rt, device=dev, dtype=dt)
    for i in range(n): Bbd[i * N:(i + 1) * N, i * r:(i + 1) * r] = (torch.randn(N, r, device=dev) * 0.02).to(dt)
    dys = [torch.randn(M, N, device=dev, dtype=dt) for _ in range(n)]
    pack = torch.empty(80 * (K + Nc), dtype=dt, device=dev)
    nat.check(lib.b200lora_pack_lora(nat.ptr(Ac), nat.ptr(Bbd), io, None, 0, nat.ptr(pack), K, Nc, rt, io, st), "pack")
    y = torch.empty(M, Nc, device=dev, dtype=dt); T = torch.empty(M, rt, device=dev, dtype=dt)
    dx = torch.empty(M, K, device=dev, dtype=dt); dT = torch.empty(M, rt, device=dev, dtype=dt)
    dA = [torch.zeros(r, K, device=dev) for _ in range(n)]; dB = [torch.zeros(N, r, device=dev) for _ in range(n)]
    dyp = (ctypes.c_void_p * n)(*[d.data_ptr() for d in dys]); ldp = (ctypes.c_int64 * n)(*[N] * n)
    dap = (ctypes.c_void_p * n)(*[t.data_ptr() for t in dA]); dbp = (ctypes.c_void_p * n)(*[t.data_ptr() for t in dB])
    sc = (ctypes.c_float * n)(*[1.0] * n); re = (ctypes.c_int * n)(*[r] * n)
    nat.check(lib.b200lora_linear_fwd(nat.ptr(x), nat.ptr(Wc), None, 0, nat.ptr(Ac), nat.ptr(Bbd), io, nat.ptr(pack), 1.0,
                                      nat.ptr(y), nat.ptr(T), M, K, Nc, rt, io, nat.IMPL_TC, st), "fwd")
    nat.check(lib.b200lora_linear_bwd_dx_fused(dyp, ldp, n, nat.ptr(Wct), nat.ptr(pack), nat.ptr(dx), nat.ptr(dT), M, K, N, rt, io, st), "bdx")
    nat.check(lib.b200lora_lora_grads_fused(dyp, ldp, n, nat.ptr(x), nat.ptr(T), nat.ptr(dT), sc, dap, dbp, M, K, N, re, io, st), "grads")
    torch.cuda.synchronize()
print("ok")

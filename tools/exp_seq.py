import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_finetuning_b200 import _native as nat
dev = torch.device("cuda:0"); lib = nat.lib(); dt = torch.bfloat16; io = nat.BF16; st = nat.stream_ptr(dev)
def run(M, K, N, r=4):
    x = torch.randn(M, K, device=dev, dtype=dt); w = torch.randn(N, K, device=dev, dtype=dt)
    A = torch.randn(r, K, device=dev, dtype=dt); B = torch.randn(N, r, device=dev, dtype=dt)
    y = torch.empty(M, N, device=dev, dtype=dt); T = torch.empty(M, r, device=dev, dtype=dt)
    pack = torch.empty(80 * (K + N), dtype=dt, device=dev)
    nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), io, None, 0, nat.ptr(pack), K, N, r, io, st), "pack")
    rc = lib.b200lora_linear_fwd(nat.ptr(x), nat.ptr(w), None, 0, nat.ptr(A), nat.ptr(B), io, nat.ptr(pack), 1.0,
                                 nat.ptr(y), nat.ptr(T), M, K, N, r, io, nat.IMPL_TC, st)
    torch.cuda.synchronize()
    print(M, K, N, "rc", rc, lib.b200lora_last_error().decode())
for N in [int(a) for a in sys.argv[1:]]:
    run(512, 576, N)

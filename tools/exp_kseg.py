import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_finetuning_b200 import _native as nat
dev = torch.device("cuda:0"); lib = nat.lib(); st = nat.stream_ptr(dev); dt = torch.bfloat16; io = nat.BF16
def b2b(fn, reps=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(2000000); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps
for (M, K, N, n) in ((2048, 1280, 1280, 3), (8192, 640, 640, 3), (32768, 320, 320, 3)):
    r = 4; rt = r * n; Nc = N * n
    dys = [torch.randn(M, N, device=dev, dtype=dt) for _ in range(n)]
    dycat = torch.cat(dys, 1).contiguous()
    Wct = (torch.rand(K, Nc, device=dev) - 0.5).to(dt)
    Ac = torch.randn(rt, K, device=dev).to(dt); Bbd = torch.randn(Nc, rt, device=dev).to(dt)
    pack = torch.empty(80 * (K + Nc), dtype=dt, device=dev)
    nat.check(lib.b200lora_pack_lora(nat.ptr(Ac), nat.ptr(Bbd), io, None, 0, nat.ptr(pack), K, Nc, rt, io, st), "pack")
    dx = torch.empty(M, K, device=dev, dtype=dt); dT = torch.empty(M, rt, device=dev, dtype=dt)
    dyp = (ctypes.c_void_p * n)(*[d.data_ptr() for d in dys]); ldp = (ctypes.c_int64 * n)(*[N] * n)
    seg = lambda: nat.check(lib.b200lora_linear_bwd_dx_fused(dyp, ldp, n, nat.ptr(Wct), nat.ptr(pack), nat.ptr(dx), nat.ptr(dT), M, K, N, rt, io, st), "seg")
    plain = lambda: nat.check(lib.b200lora_linear_bwd_dx(nat.ptr(dycat), nat.ptr(Wct), nat.W_NK, nat.ptr(Ac), nat.ptr(Bbd), io, nat.ptr(pack), 1.0,
                                                         nat.ptr(dx), nat.ptr(dT), M, K, Nc, rt, io, nat.IMPL_TC, st), "plain")
    dx1 = dx.clone(); seg(); torch.cuda.synchronize(); a = dx.clone(); plain(); torch.cuda.synchronize()
    print(M, K, N, "seg", round(b2b(seg), 1), "plain", round(b2b(plain), 1), "maxdiff", float((a.float() - dx.float()).abs().max()))
    for pdl in ("0",):
        pass

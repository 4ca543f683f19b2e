"""Attention-block micro-benchmark: the q/k/v LoRA layers of one block as ONE fused
launch per pass, against the same layers launched one by one and against stock PyTorch on this GPU.
bf16, r = 4 per layer.  Timing: CUDA events; `b2b` = 20 launches back to back (launch latency
overlapped, operands L2-warm), `cold` = L2 flushed before every launch.
    python tools/bench_fused.py --json gpurun_out/fused.json
"""
import argparse, ctypes, json, os, sys
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_finetuning_b200 import _native as nat

BLOCKS = [(32768, 320, 320, 3, 5), (8192, 640, 640, 3, 5), (2048, 1280, 1280, 3, 5), (512, 1280, 1280, 3, 1),
          (616, 768, 320, 2, 5), (616, 768, 640, 2, 5), (616, 768, 1280, 2, 6)]   # (M, K, N, nseg, blocks in the UNet)
ap = argparse.ArgumentParser()
ap.add_argument("--r", type=int, default=4); ap.add_argument("--json", default=None)
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
dev = torch.device("cuda:0"); lib = nat.lib(); st = nat.stream_ptr(dev); dt = torch.bfloat16; io = nat.BF16
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
PEAKS = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) \
    if os.path.exists("MEASURED_PEAKS.json") else {}


def b2b(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(2000000); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


def cold(fn, reps):
    ts = []
    for _ in range(reps):
        flush.zero_(); torch.cuda._sleep(400000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]


rows = []
for (M, K, N, n, cnt) in BLOCKS:
    r = a.r; rt = r * n; Nc = N * n
    x = torch.randn(M, K, device=dev, dtype=dt)
    Ws = [(torch.rand(N, K, device=dev) * 2 - 1).mul_(K ** -0.5).to(dt) for _ in range(n)]
    As = [(torch.randn(r, K, device=dev) / r).to(dt) for _ in range(n)]
    Bs = [(torch.randn(N, r, device=dev) * 0.02).to(dt) for _ in range(n)]
    dys = [torch.randn(M, N, device=dev, dtype=dt) for _ in range(n)]
    Wc = torch.cat(Ws, 0).contiguous(); Wct = Wc.t().contiguous()
    Ac = torch.cat(As, 0).contiguous(); Bbd = torch.zeros(Nc, rt, device=dev, dtype=dt)
    for i in range(n): Bbd[i * N:(i + 1) * N, i * r:(i + 1) * r] = Bs[i]
    pack = torch.empty(80 * (K + Nc), dtype=dt, device=dev)
    nat.check(lib.b200lora_pack_lora(nat.ptr(Ac), nat.ptr(Bbd), io, None, 0, nat.ptr(pack), K, Nc, rt, io, st), "pack")
    y = torch.empty(M, Nc, device=dev, dtype=dt); T = torch.empty(M, rt, device=dev, dtype=dt)
    dx = torch.empty(M, K, device=dev, dtype=dt); dT = torch.empty(M, rt, device=dev, dtype=dt)
    dA = [torch.zeros(r, K, device=dev) for _ in range(n)]; dB = [torch.zeros(N, r, device=dev) for _ in range(n)]
    dyp = (ctypes.c_void_p * n)(*[d.data_ptr() for d in dys]); ldp = (ctypes.c_int64 * n)(*[N] * n)
    dap = (ctypes.c_void_p * n)(*[t.data_ptr() for t in dA]); dbp = (ctypes.c_void_p * n)(*[t.data_ptr() for t in dB])
    sc = (ctypes.c_float * n)(*[1.0] * n); re = (ctypes.c_int * n)(*[r] * n)
    f_fwd = lambda: nat.check(lib.b200lora_linear_fwd(nat.ptr(x), nat.ptr(Wc), None, 0, nat.ptr(Ac), nat.ptr(Bbd), io, nat.ptr(pack),
                                                      1.0, nat.ptr(y), nat.ptr(T), M, K, Nc, rt, io, nat.IMPL_TC, st), "fwd")
    f_bdx = lambda: nat.check(lib.b200lora_linear_bwd_dx_fused(dyp, ldp, n, nat.ptr(Wct), nat.ptr(pack), nat.ptr(dx), nat.ptr(dT),
                                                               M, K, N, rt, io, st), "bdx")
    f_gr = lambda: nat.check(lib.b200lora_lora_grads_fused(dyp, ldp, n, nat.ptr(x), nat.ptr(T), nat.ptr(dT), sc, dap, dbp,
                                                           M, K, N, re, io, st), "grads")
    # the same layers one launch each
    p1 = torch.empty(80 * (K + N), dtype=dt, device=dev); W0t = Ws[0].t().contiguous()
    nat.check(lib.b200lora_pack_lora(nat.ptr(As[0]), nat.ptr(Bs[0]), io, None, 0, nat.ptr(p1), K, N, r, io, st), "pack")
    y1 = torch.empty(M, N, device=dev, dtype=dt); T1 = torch.empty(M, r, device=dev, dtype=dt); dT1 = torch.empty(M, r, device=dev, dtype=dt)
    def u_fwd():
        for _ in range(n):
            nat.check(lib.b200lora_linear_fwd(nat.ptr(x), nat.ptr(Ws[0]), None, 0, nat.ptr(As[0]), nat.ptr(Bs[0]), io, nat.ptr(p1), 1.0,
                                              nat.ptr(y1), nat.ptr(T1), M, K, N, r, io, nat.IMPL_TC, st), "fwd1")
    def u_bdx():
        for _ in range(n):
            nat.check(lib.b200lora_linear_bwd_dx(nat.ptr(dys[0]), nat.ptr(W0t), nat.W_NK, nat.ptr(As[0]), nat.ptr(Bs[0]), io, nat.ptr(p1),
                                                 1.0, nat.ptr(dx), nat.ptr(dT1), M, K, N, r, io, nat.IMPL_TC, st), "bdx1")
    def u_gr():
        for _ in range(n):
            nat.check(lib.b200lora_lora_grads(nat.ptr(dys[0]), nat.ptr(x), nat.ptr(T1), nat.ptr(dT1), 1.0, nat.ptr(dA[0]), nat.ptr(dB[0]),
                                              M, K, N, r, io, 0, None, 0, st), "gr1")
    # stock PyTorch: 3 x (3 F.linear + mul + add), autograd backward
    xr = x.clone().requires_grad_(True)
    Ar = [t.clone().requires_grad_(True) for t in As]; Br = [t.clone().requires_grad_(True) for t in Bs]
    def t_fwd():
        return [F.linear(xr, Ws[i]) + F.linear(F.linear(xr, Ar[i]), Br[i]) * 1.0 for i in range(n)]
    def t_fb():
        outs = t_fwd()
        torch.autograd.backward(outs, dys)
    f_fwd(); f_bdx(); f_gr()
    fl_f = 2.0 * M * K * Nc + 2.0 * M * r * n * (K + N)
    fl_all = 2 * fl_f + 2.0 * M * r * n * (K + N)
    row = {"M": M, "K": K, "N": N, "nseg": n, "blocks": cnt}
    for tag, fn in (("fwd", f_fwd), ("bdx", f_bdx), ("grads", f_gr), ("unfused_fwd", u_fwd), ("unfused_bdx", u_bdx), ("unfused_grads", u_gr)):
        row[tag + "_b2b_us"] = b2b(fn, a.reps); row[tag + "_cold_us"] = cold(fn, 10)
    with torch.no_grad():
        row["torch_fwd_b2b_us"] = b2b(t_fwd, a.reps)
    row["torch_fwdbwd_b2b_us"] = b2b(t_fb, a.reps)
    row["fwd_tflops_b2b"] = fl_f / row["fwd_b2b_us"] / 1e6; row["fwd_tflops_cold"] = fl_f / row["fwd_cold_us"] / 1e6
    tot = row["fwd_b2b_us"] + row["bdx_b2b_us"] + row["grads_b2b_us"]
    row["fwdbwd_tflops_b2b"] = fl_all / tot / 1e6
    row["unfused_fwdbwd_tflops_b2b"] = fl_all / (row["unfused_fwd_b2b_us"] + row["unfused_bdx_b2b_us"] + row["unfused_grads_b2b_us"]) / 1e6
    row["torch_fwdbwd_tflops_b2b"] = fl_all / row["torch_fwdbwd_b2b_us"] / 1e6
    row["speedup_vs_torch_fwdbwd"] = row["torch_fwdbwd_b2b_us"] / tot
    rows.append(row); print(json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
if a.json:
    json.dump({"r": a.r, "peaks": PEAKS, "rows": rows}, open(a.json, "w"), indent=1)

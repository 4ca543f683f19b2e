"""Per-layer micro-benchmark on the SD-1.5 shape census: our fused kernels through
the C-ABI vs. the reference composition (3 F.linear + mul + add, autograd backward) in stock
PyTorch on the same GPU.  CUDA-event timing, L2 flushed between iterations.

    python tools/bench_layer.py [--iters 20] [--r 4] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from diffusion_finetuning_b200 import _native as nat  # noqa: E402

CENSUS = [  # (M, K, N, count)
    (32768, 320, 320, 30), (32768, 320, 2560, 5), (8192, 640, 640, 30), (8192, 640, 5120, 5),
    (2048, 1280, 1280, 30), (2048, 1280, 10240, 5), (512, 1280, 1280, 6), (512, 1280, 10240, 1),
    (616, 768, 320, 10), (616, 768, 640, 10), (616, 768, 1280, 12),
]
PEAK_TF, PEAK_GBS = 1629.1, 6539.9
if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
    _p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    PEAK_TF, PEAK_GBS = _p.get("bf16_tflops", PEAK_TF), _p.get("hbm_gbs", PEAK_GBS)


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        torch.cuda._sleep(400000)     # ~0.2 ms of GPU spin: the host enqueues fn() meanwhile
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--r", type=int, default=4)
    ap.add_argument("--json", default=None)
    ap.add_argument("--no-torch", action="store_true")
    ap.add_argument("--shapes", default=None, help="comma list of indices into CENSUS")
    ap.add_argument("--back2back", type=int, default=0,
                    help="also time N back-to-back launches (no L2 flush, launch latency overlapped)")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = nat.lib()
    dt = torch.bfloat16
    io = nat.BF16
    st = nat.stream_ptr(dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2
    r = args.r
    rows = []
    census = CENSUS if args.shapes is None else [CENSUS[int(i)] for i in args.shapes.split(",")]
    for (M, K, N, count) in census:
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

        a_pack = bt_pack = torch.empty(80 * (K + N), dtype=dt, device=dev)
        nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), io, nat.ptr(b), io, nat.ptr(a_pack), K, N, r, io,
                                         st), "pack")

        def fwd():
            nat.check(lib.b200lora_linear_fwd(nat.ptr(x), nat.ptr(w), nat.ptr(b), io, nat.ptr(A), nat.ptr(B), io,
                                              nat.ptr(a_pack), 1.0, nat.ptr(y), nat.ptr(T), M, K, N, r, io,
                                              nat.IMPL_TC, st), "fwd")

        def bdx():
            nat.check(lib.b200lora_linear_bwd_dx(nat.ptr(dy), nat.ptr(wt), nat.W_NK, nat.ptr(A), nat.ptr(B), io,
                                                 nat.ptr(bt_pack), 1.0, nat.ptr(dx), nat.ptr(dT), M, K, N, r, io,
                                                 nat.IMPL_TC, st), "bdx")

        def grads():
            nat.check(lib.b200lora_lora_grads(nat.ptr(dy), nat.ptr(x), nat.ptr(T), nat.ptr(dT), 1.0, nat.ptr(dA),
                                              nat.ptr(dB), M, K, N, r, io, 0, None, 0, st), "grads")

        t_f, t_b, t_g = timeit(fwd, args.iters, flush), timeit(bdx, args.iters, flush), timeit(grads, args.iters, flush)
        flops = 4.0 * M * K * N + 6.0 * M * r * (K + N)
        e = 2
        bytes_fb = e * (3 * M * K + 2 * M * N + 2 * N * K + 2 * M * r)
        ai = flops / bytes_fb
        ceil_tf = min(PEAK_TF, ai * PEAK_GBS / 1e3)
        row = {"M": M, "K": K, "N": N, "r": r, "count": count, "fwd_us": t_f * 1e6, "bwd_dx_us": t_b * 1e6,
               "grads_us": t_g * 1e6, "fwd_tflops": (2.0 * M * K * N + 2.0 * M * r * (K + N)) / t_f / 1e12,
               "fwdbwd_tflops": flops / (t_f + t_b + t_g) / 1e12, "ceiling_tflops": ceil_tf,
               "grads_gbs": e * (M * K + M * N + 2 * M * r) / t_g / 1e9}
        if args.back2back:
            def b2b(fn):
                fn(); torch.cuda.synchronize()
                a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda._sleep(2000000)
                a.record()
                for _ in range(args.back2back):
                    fn()
                b_.record()
                torch.cuda.synchronize()
                return a.elapsed_time(b_) * 1e3 / args.back2back
            row.update({"fwd_b2b_us": b2b(fwd), "bwd_dx_b2b_us": b2b(bdx), "grads_b2b_us": b2b(grads)})
        if not args.no_torch:
            xr = x.clone().requires_grad_(True)
            Ar, Br = A.clone().requires_grad_(True), B.clone().requires_grad_(True)

            def tfwd():
                return F.linear(xr, w, b) + F.linear(F.linear(xr, Ar), Br) * 1.0

            def tfb():
                out = tfwd()
                out.backward(dy)
                xr.grad = Ar.grad = Br.grad = None

            with torch.no_grad():
                t_tf = timeit(lambda: F.linear(x, w, b) + F.linear(F.linear(x, A), B) * 1.0, args.iters, flush)
            t_tfb = timeit(tfb, args.iters, flush)
            row.update({"torch_fwd_us": t_tf * 1e6, "torch_fwdbwd_us": t_tfb * 1e6,
                        "torch_fwdbwd_tflops": flops / t_tfb / 1e12,
                        "speedup_fwd": t_tf / t_f, "speedup_fwdbwd": t_tfb / (t_f + t_b + t_g)})
        row["frac_peak"] = row["fwdbwd_tflops"] / PEAK_TF
        row["frac_ceiling"] = row["fwdbwd_tflops"] / ceil_tf
        rows.append(row)
        print(json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
    if args.json:
        with open(args.json, "w") as fh:
            json.dump({"peak_tflops": PEAK_TF, "peak_gbs": PEAK_GBS, "rows": rows}, fh, indent=1)


if __name__ == "__main__":
    main()

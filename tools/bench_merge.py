"""K5: `weight_apply_lora` over the 144 SD-1.5-shaped matrices (lora.py:410-424, cli_lora_add.py:72-110).
GPU: b200lora_merge (one pass over W, in place) — GB/s of algorithmic traffic 2·e·N·K.
CPU: the oracle's restatement of the reference arithmetic on the host cores (bounded sample).
    python tools/bench_merge.py --json gpurun_out/merge.json
"""
import argparse, json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_finetuning_b200 import _native as nat

CENSUS = [(320, 320, 30), (2560, 320, 5), (640, 640, 30), (5120, 640, 5), (1280, 1280, 36), (10240, 1280, 6),
          (320, 768, 10), (640, 768, 10), (1280, 768, 12)]   # (N, K, count) = 144 matrices, 192.6 M elements

ap = argparse.ArgumentParser()
ap.add_argument("--r", type=int, default=4); ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--json", default=None); ap.add_argument("--cpu", action="store_true")
a = ap.parse_args()
dev = torch.device("cuda:0"); lib = nat.lib(); st = nat.stream_ptr(dev)
out = {"r": a.r, "rows": []}
for wdt in (torch.float32, torch.float16, torch.bfloat16):
    mats = []
    for (N, K, cnt) in CENSUS:
        for _ in range(cnt):
            mats.append((torch.randn(N, K, device=dev).mul_(0.03).to(wdt), (torch.randn(N, a.r, device=dev) * 0.05),
                         (torch.randn(a.r, K, device=dev) * 0.3)))
    elems = sum(w.numel() for w, _, _ in mats)
    code = nat.dtype_code(wdt)
    def run():
        for w, u, d in mats:
            N, K = w.shape
            nat.check(lib.b200lora_merge(nat.ptr(w), nat.ptr(w), nat.ptr(u), nat.ptr(d), 0.5, N, K, a.r, code, nat.F32, st), "merge")
    run(); torch.cuda.synchronize()
    ts = []
    for _ in range(a.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3)
    t = sorted(ts)[len(ts) // 2]
    byts = 2 * elems * mats[0][0].element_size()
    row = {"w_dtype": str(wdt).split(".")[1], "matrices": len(mats), "elements": elems, "ms": t * 1e3,
           "gbs": byts / t / 1e9, "frac_hbm_peak_6451": byts / t / 1e9 / 6451.2}
    # largest single matrix alone (launch overhead amortised)
    w, u, d = max(mats, key=lambda m: m[0].numel()); N, K = w.shape
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        nat.check(lib.b200lora_merge(nat.ptr(w), nat.ptr(w), nat.ptr(u), nat.ptr(d), 0.5, N, K, a.r, code, nat.F32, st), "merge")
    e1.record(); torch.cuda.synchronize()
    row["largest_gbs"] = 2 * w.numel() * w.element_size() * 20 / (e0.elapsed_time(e1) * 1e-3) / 1e9
    out["rows"].append(row); print(json.dumps(row), flush=True)
    del mats
if a.cpu:
    from oracle import lora_oracle as O
    torch.set_num_threads(os.cpu_count())
    mats = [(torch.randn(N, K) * 0.03, torch.randn(N, a.r) * 0.05, torch.randn(a.r, K) * 0.3) for (N, K, c) in CENSUS for _ in range(max(1, c // 5))]
    t0 = time.perf_counter()
    for w, u, d in mats:
        O.merge_weight(w, u, d, 0.5)
    t = time.perf_counter() - t0
    byts = sum(2 * w.numel() * 4 for w, _, _ in mats)
    out["cpu"] = {"cores": os.cpu_count(), "matrices": len(mats), "gbs": byts / t / 1e9, "kind": "oracle port of lora.py:423 on host cores"}
    print(json.dumps(out["cpu"]))
if a.json:
    json.dump(out, open(a.json, "w"), indent=1)

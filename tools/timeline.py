"""Per-CTA clock64 timeline of the tcgen05 kernel (DBG build): where a tile's time goes.
    python tools/timeline.py --M 32768 --K 320 --N 320 --cta 5
"""
import argparse, ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_finetuning_b200 import _native as nat

ap = argparse.ArgumentParser()
ap.add_argument("--M", type=int, default=32768); ap.add_argument("--K", type=int, default=320)
ap.add_argument("--N", type=int, default=320); ap.add_argument("--r", type=int, default=4)
ap.add_argument("--cta", type=int, default=5); ap.add_argument("--flush", action="store_true")
a = ap.parse_args()
dev = torch.device("cuda:0"); lib = nat.lib()
lib.b200lora_debug_timeline.argtypes = [ctypes.c_void_p, ctypes.c_int]; lib.b200lora_debug_timeline.restype = None
M, K, N, r = a.M, a.K, a.N, a.r
dt, io, st = torch.bfloat16, nat.BF16, nat.stream_ptr(dev)
x = torch.randn(M, K, device=dev, dtype=dt); w = (torch.rand(N, K, device=dev) - 0.5).to(dt)
b = torch.zeros(N, device=dev, dtype=dt)
A = (torch.randn(r, K, device=dev) / r).to(dt); B = (torch.randn(N, r, device=dev) * 0.02).to(dt)
y = torch.empty(M, N, device=dev, dtype=dt); T = torch.empty(M, r, device=dev, dtype=dt)
pack = torch.empty(80 * (K + N), dtype=dt, device=dev)
nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), io, nat.ptr(b), io, nat.ptr(pack), K, N, r, io, st), "pack")
def fwd():
    nat.check(lib.b200lora_linear_fwd(nat.ptr(x), nat.ptr(w), nat.ptr(b), io, nat.ptr(A), nat.ptr(B), io, nat.ptr(pack),
                                      1.0, nat.ptr(y), nat.ptr(T), M, K, N, r, io, nat.IMPL_TC, st), "fwd")
for _ in range(3): fwd()
torch.cuda.synchronize()
buf = torch.zeros(6 * 256, dtype=torch.int64, device=dev)
if a.flush:
    torch.empty(256 << 20, dtype=torch.uint8, device=dev).zero_()
lib.b200lora_debug_timeline(ctypes.c_void_p(buf.data_ptr()), a.cta)
fwd(); torch.cuda.synchronize()
lib.b200lora_debug_timeline(None, 0)
t = buf.cpu().view(6, 32, 8)
t0 = int(t[0, 0, 0])
names = {0: "setup", 1: "producer", 2: "mma", 3: "Twarp", 4: "epi_h0", 5: "epi_h1"}
keys = {0: ["start", "after_sync"], 1: ["tile_begin", "main_issued", "all_issued"],
        2: ["begin", "accempty_ok", "main_issued", "look_issued", "tsm_ok", "tails_issued"],
        3: ["wait_tfull", "tfull_ok", "tsm_arrive"], 4: ["wait_acc", "acc_ok", "done"], 5: ["wait_acc", "acc_ok", "done"]}
for role in range(6):
    for it in range(32):
        row = t[role, it]
        if int(row.max()) == 0: continue
        cells = []
        for k in range(8):
            v = int(row[k])
            if v and role == 2 and k == 6:
                cells.append(f"operand_wait={v / 1.9e3:5.2f}us")
            elif v:
                cells.append(f"{keys[role][k] if k < len(keys[role]) else 'k%d' % k}={(v - t0) / 1.9e3:7.2f}us")
        print(f"{names[role]:9s} it={it:2d} " + " ".join(cells))

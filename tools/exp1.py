import os, sys, time, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
DEV = "cuda:0"
# ---- SVD accuracy
g = torch.Generator().manual_seed(3)
up = torch.randn(320, 4, generator=g) * 0.05
down = torch.randn(4, 768, generator=g) * 0.5
delta = (up @ down + 1e-4 * torch.randn(320, 768, generator=g)).half()
mat = delta.float()
Uc, Sc, Vc = torch.linalg.svd(mat, full_matrices=False)
ref = (Uc[:, :4] * Sc[:4]) @ Vc[:4]
print("S cpu", Sc[:6])
for drv in (None, "gesvd", "gesvdj", "gesvda"):
    try:
        U, S, Vh = torch.linalg.svd(mat.to(DEV), full_matrices=False, driver=drv)
        got = ((U[:, :4] * S[:4]) @ Vh[:4]).cpu()
        got64 = ((U[:, :4].double() * S[:4].double()) @ Vh[:4].double()).cpu()
        print(drv, "rel prod err", ((got - ref).norm() / ref.norm()).item(), "f64 prod", ((got64 - ref.double()).norm() / ref.norm()).item(),
              "S err", ((S[:4].cpu() - Sc[:4]).abs().max()).item())
    except Exception as e:
        print(drv, "failed", e)
print("matmul tf32 flags", torch.backends.cuda.matmul.allow_tf32, torch.get_float32_matmul_precision())
a = torch.randn(320, 4, device=DEV); b = torch.randn(4, 768, device=DEV)
print("small matmul err", ((a @ b).cpu() - a.cpu() @ b.cpu()).abs().max().item())

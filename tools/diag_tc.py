"""Staged bring-up diagnostics for the tcgen05 kernel.  Each case runs in its own subprocess
under a timeout, so a trap or watchdog in one case cannot poison the next.

    python tools/diag_tc.py            # run all cases
    python tools/diag_tc.py --case 3   # one case, in-process
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# (name, M, K, N, r, scale, bias, dtype, pattern)
CASES = [
    ("base-only 1 tile, 1 kblock", 128, 64, 64, 4, 0.0, False, "bf16", "rand"),
    ("base-only identity W", 128, 64, 64, 4, 0.0, False, "bf16", "eye"),
    ("base-only 2 kblocks", 128, 128, 64, 4, 0.0, False, "bf16", "rand"),
    ("base+lora 1 tile", 128, 64, 64, 4, 1.0, False, "bf16", "rand"),
    ("lora-only (W0=0)", 128, 64, 64, 4, 1.0, False, "bf16", "zeroW"),
    ("bias, BN=160, 2 n tiles", 256, 320, 320, 4, 0.7, True, "bf16", "rand"),
    ("ragged M/N/K", 300, 200, 328, 8, 1.3, True, "bf16", "rand"),
    ("multi-tile persistent", 32768, 320, 320, 4, 1.0, True, "bf16", "rand"),
    ("fp16", 512, 640, 640, 16, 1.0, True, "fp16", "rand"),
]


def run_case(idx):
    import torch
    from diffusion_finetuning_b200 import _native as nat
    name, M, K, N, r, s, bias, dt, pattern = CASES[idx]
    dtype = torch.bfloat16 if dt == "bf16" else torch.float16
    g = torch.Generator().manual_seed(idx)
    x = torch.randn(M, K, generator=g).to(dtype)
    if pattern == "eye":
        w = torch.eye(N, K).to(dtype)
    elif pattern == "zeroW":
        w = torch.zeros(N, K).to(dtype)
    else:
        w = torch.empty(N, K).uniform_(-K ** -0.5, K ** -0.5, generator=g).to(dtype)
    b = torch.empty(N).uniform_(-0.1, 0.1, generator=g).to(dtype) if bias else None
    A = (torch.randn(r, K, generator=g) / r).to(dtype)
    B = (torch.randn(N, r, generator=g) * 0.5).to(dtype)
    xf, wf, Af, Bf = x.float(), w.float(), A.float(), B.float()
    T_ref = xf @ Af.t()
    y_ref = xf @ wf.t() + (b.float() if bias else 0) + s * (T_ref @ Bf.t())
    dev = torch.device("cuda:0")
    lib = nat.lib()
    xd, wd, Ad, Bd = x.to(dev), w.to(dev), A.to(dev), B.to(dev)
    bd = b.to(dev) if bias else None
    y = torch.full((M, N), float("nan"), dtype=dtype, device=dev)
    T = torch.full((M, r), float("nan"), dtype=dtype, device=dev)
    io = nat.dtype_code(dtype)
    a_pack = torch.empty(80 * (K + N), dtype=dtype, device=dev)
    nat.check(lib.b200lora_pack_lora(nat.ptr(Ad), nat.ptr(Bd), io, nat.ptr(bd), io, nat.ptr(a_pack), K, N, r, io,
                                     nat.stream_ptr(dev)), "pack")
    rc = lib.b200lora_linear_fwd(nat.ptr(xd), nat.ptr(wd), nat.ptr(bd), io, nat.ptr(Ad), nat.ptr(Bd), io,
                                 nat.ptr(a_pack), s, nat.ptr(y), nat.ptr(T), M, K, N, r, io, nat.IMPL_TC,
                                 nat.stream_ptr(dev))
    print(f"[{idx}] {name}: rc={rc} err='{lib.b200lora_last_error().decode()}'")
    try:
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        print(f"[{idx}] sync failed: {e}")
        try:
            print(f"[{idx}] watchdog code: 0x{lib.b200lora_watchdog_code():08x}")
        except Exception:  # noqa: BLE001
            pass
        return 1
    yc, Tc = y.float().cpu(), T.float().cpu()
    rel = lambda a, b_: ((a - b_).norm() / (b_.norm() + 1e-30)).item()
    ey, eT = rel(yc, y_ref), rel(Tc, T_ref)
    nan_y = int(torch.isnan(yc).sum())
    print(f"[{idx}] rel_err y={ey:.3e} T={eT:.3e} nan(y)={nan_y} nan(T)={int(torch.isnan(Tc).sum())}")
    ok = ey < 2e-2 and eT < 2e-2 and nan_y == 0
    if not ok:
        d = (yc - y_ref).abs()
        d = torch.nan_to_num(d, nan=1e9)
        bad = d > 0.05 * (y_ref.abs().max() + 1e-6)
        print(f"[{idx}] bad elements: {int(bad.sum())}/{bad.numel()}")
        rows = bad.any(1).nonzero().flatten()
        cols = bad.any(0).nonzero().flatten()
        print(f"[{idx}] bad rows (first 24): {rows[:24].tolist()} ... count {len(rows)}")
        print(f"[{idx}] bad cols (first 24): {cols[:24].tolist()} ... count {len(cols)}")
        print(f"[{idx}] y[0,:8]    = {yc[0, :8].tolist()}")
        print(f"[{idx}] yref[0,:8] = {y_ref[0, :8].tolist()}")
        print(f"[{idx}] y[1,:8]    = {yc[1, :8].tolist()}")
        print(f"[{idx}] yref[1,:8] = {y_ref[1, :8].tolist()}")
        print(f"[{idx}] T[0]={Tc[0].tolist()} Tref[0]={T_ref[0].tolist()}")
        # is y a permutation of the reference along k chunks?  correlation of y rows with ref rows
        if pattern == "eye":
            print(f"[{idx}] x[0,:16]={xf[0, :16].tolist()}")
            print(f"[{idx}] y[0,:16]={yc[0, :16].tolist()}")
    print(f"[{idx}] {'PASS' if ok else 'FAIL'}")
    return 0 if ok else 2


def main():
    if "--case" in sys.argv:
        sys.exit(run_case(int(sys.argv[sys.argv.index("--case") + 1])))
    fails = 0
    for i in range(len(CASES)):
        try:
            p = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", str(i)],
                               timeout=120, capture_output=True, text=True)
            sys.stdout.write(p.stdout)
            if p.returncode != 0:
                fails += 1
                sys.stdout.write(p.stderr[-1500:])
        except subprocess.TimeoutExpired:
            fails += 1
            print(f"[{i}] TIMEOUT")
        sys.stdout.flush()
    print(f"diag: {len(CASES) - fails}/{len(CASES)} cases passed")


if __name__ == "__main__":
    main()

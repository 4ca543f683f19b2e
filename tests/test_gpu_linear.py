"""Parity of the CUDA hot path with the oracle, called through the C-ABI (ctypes) and through the
autograd op.  Tolerances are BASELINE.json's: rel-err <= 1e-5 in fp32, <= 2e-2 in bf16/fp16
(Frobenius norm of the difference over the norm of the oracle result), dropout = 0."""
import pytest
import torch

import diffusion_finetuning_b200 as L
from conftest import check_close, load_golden, max_abs_err, rel_err, worst_row_err
from diffusion_finetuning_b200 import _native as nat
from diffusion_finetuning_b200 import ops
from oracle import lora_oracle as O
from oracle.make_golden import LINEAR_CASES

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = {torch.float32: 1e-5, torch.bfloat16: 2e-2, torch.float16: 2e-2}


def make_case(M, K, N, r, bias, dtype, seed=0, ab_dtype=None):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=g)
    w0 = torch.empty(N, K).uniform_(-K ** -0.5, K ** -0.5, generator=g)
    b = torch.empty(N).uniform_(-0.1, 0.1, generator=g) if bias else None
    A = torch.randn(r, K, generator=g) / r
    B = torch.randn(N, r, generator=g) * 0.02
    dy = torch.randn(M, N, generator=g)
    ab = ab_dtype or dtype
    q = lambda t, dt: None if t is None else t.to(dt)
    # the oracle sees exactly the values the kernel sees (inputs rounded to the io dtype)
    host = dict(x=q(x, dtype).float(), w0=q(w0, dtype).float(), b=None if b is None else q(b, dtype).float(),
                A=q(A, ab).to(dtype).float(), B=q(B, ab).to(dtype).float(), dy=q(dy, dtype).float())
    dev = dict(x=q(x, dtype).to(DEV), w0=q(w0, dtype).to(DEV), b=None if b is None else q(b, dtype).to(DEV),
               A=q(A, ab).to(DEV), B=q(B, ab).to(DEV), dy=q(dy, dtype).to(DEV))
    return host, dev


def run_cabi(dev, M, K, N, r, s, dtype, impl):
    """forward, backward-dX and the gradient reductions straight through the C functions"""
    lib = nat.lib()
    io = nat.dtype_code(dtype)
    ab = nat.dtype_code(dev["A"].dtype)
    st = nat.stream_ptr(torch.device(DEV))
    y = torch.empty(M, N, dtype=dtype, device=DEV)
    T = torch.empty(M, r, dtype=dtype, device=DEV)
    b = dev["b"]
    pack = None
    if impl == nat.IMPL_TC:
        assert lib.b200lora_pack_elems(K, N) == 80 * (K + N)
        pack = torch.empty(80 * (K + N), dtype=dtype, device=DEV)
        nat.check(lib.b200lora_pack_lora(nat.ptr(dev["A"]), nat.ptr(dev["B"]), ab, nat.ptr(b), io, nat.ptr(pack),
                                         K, N, r, io, st), "pack")
    nat.check(lib.b200lora_linear_fwd(nat.ptr(dev["x"]), nat.ptr(dev["w0"]), nat.ptr(b), io, nat.ptr(dev["A"]),
                                      nat.ptr(dev["B"]), ab, nat.ptr(pack), s, nat.ptr(y), nat.ptr(T), M, K, N, r,
                                      io, impl, st), "fwd")
    dx = torch.empty(M, K, dtype=dtype, device=DEV)
    dT = torch.empty(M, r, dtype=dtype, device=DEV)
    if impl == nat.IMPL_TC:
        w = dev["w0"].t().contiguous()
        lay = nat.W_NK
    else:
        w, lay = dev["w0"], nat.W_KN
    nat.check(lib.b200lora_linear_bwd_dx(nat.ptr(dev["dy"]), nat.ptr(w), lay, nat.ptr(dev["A"]), nat.ptr(dev["B"]),
                                         ab, nat.ptr(pack), s, nat.ptr(dx), nat.ptr(dT), M, K, N, r, io, impl, st),
              "bwd_dx")
    dA = torch.zeros(r, K, dtype=torch.float32, device=DEV)
    dB = torch.zeros(N, r, dtype=torch.float32, device=DEV)
    nat.check(lib.b200lora_lora_grads(nat.ptr(dev["dy"]), nat.ptr(dev["x"]), nat.ptr(T), nat.ptr(dT), s,
                                      nat.ptr(dA), nat.ptr(dB), M, K, N, r, io, 0, None, 0, st), "grads")
    torch.cuda.synchronize()
    return y, T, dx, dT, dA, dB


def check_against_oracle(host, got, s, tol, tag):
    y, T, dx, dT, dA, dB = got
    y_ref = O.lora_linear_forward(host["x"], host["w0"], host["b"], host["A"], host["B"], s)
    dx_ref, dA_ref, dB_ref = O.lora_linear_grads(host["x"], host["w0"], host["b"], host["A"], host["B"], s, host["dy"])
    pairs = {"y": (y, y_ref), "T": (T, host["x"] @ host["A"].t()), "dT": (dT, host["dy"] @ host["B"]),
             "dx": (dx, dx_ref)}
    if s != 0.0:
        pairs["dA"] = (dA, dA_ref)
        pairs["dB"] = (dB, dB_ref)
    else:
        assert dA.abs().max() == 0 and dB.abs().max() == 0
    errs = {k: rel_err(a, b) for k, (a, b) in pairs.items()}
    bad = {k: v for k, v in errs.items() if not v < tol}
    assert not bad, f"{tag}: rel-err over tolerance {tol}: {bad} (all: {errs})"
    # outliers the Frobenius ratio hides: a wrong ragged last row / clipped column / stale tile corner
    rows = {k: worst_row_err(a, b) for k, (a, b) in pairs.items()}
    peaks = {k: max_abs_err(a, b) for k, (a, b) in pairs.items()}
    bad = {k: (rows[k], peaks[k]) for k in pairs if not (rows[k] < 4 * tol and peaks[k] < 4 * tol)}
    assert not bad, f"{tag}: worst-row / max-element error over 4x{tol}: {bad}"


# (M, K, N, r, bias, scale): ragged M, N not a multiple of the tile, K not a multiple of 64
SHAPES = [
    (256, 64, 64, 4, False, 1.0),
    (128, 128, 160, 4, True, 0.7),
    (300, 320, 320, 4, True, 1.0),
    (77 * 2, 768, 320, 1, False, 1.0),
    (1000, 320, 2560, 16, True, 0.5),
    (130, 72, 40, 4, False, 0.0),
    (513, 200, 328, 8, True, 1.3),
    (4096, 640, 640, 4, False, 1.0),
]


@pytest.mark.parametrize("shape", SHAPES)
def test_simt_fp32_matches_oracle(shape):
    M, K, N, r, bias, s = shape
    host, dev = make_case(M, K, N, r, bias, torch.float32, seed=M + K)
    check_against_oracle(host, run_cabi(dev, M, K, N, r, s, torch.float32, nat.IMPL_SIMT), s, 1e-5, "simt/f32")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", SHAPES)
def test_tcgen05_matches_oracle(shape, dtype):
    M, K, N, r, bias, s = shape
    host, dev = make_case(M, K, N, r, bias, dtype, seed=M + N)
    check_against_oracle(host, run_cabi(dev, M, K, N, r, s, dtype, nat.IMPL_TC), s, 2e-2, f"tc/{dtype}")


@pytest.mark.parametrize("shape", SHAPES[:4])
def test_simt_bf16_matches_oracle(shape):
    M, K, N, r, bias, s = shape
    host, dev = make_case(M, K, N, r, bias, torch.bfloat16, seed=M)
    check_against_oracle(host, run_cabi(dev, M, K, N, r, s, torch.bfloat16, nat.IMPL_SIMT), s, 2e-2, "simt/bf16")


def test_tc_tighter_than_tolerance_and_fp32_masters():
    """fp32 accumulation end to end: the fused kernel should sit well inside the 2e-2 bar, and
    reads fp32 LoRA masters directly (autocast layout)."""
    M, K, N, r = 2048, 1280, 1280, 4
    host, dev = make_case(M, K, N, r, True, torch.bfloat16, seed=11, ab_dtype=torch.float32)
    y, T, dx, dT, dA, dB = run_cabi(dev, M, K, N, r, 1.0, torch.bfloat16, nat.IMPL_TC)
    y_ref = O.lora_linear_forward(host["x"], host["w0"], host["b"], host["A"], host["B"], 1.0)
    assert rel_err(y, y_ref) < 4e-3


def test_golden_reference_vectors_fp32():
    """The reference's own outputs (tests/golden/linear_cases) through the autograd op."""
    blob, _ = load_golden("linear_cases.safetensors")
    for i, (lead, K, N, r, bias, s) in enumerate(LINEAR_CASES):
        g = lambda k: blob.get(f"c{i}.{k}")
        m = L.LoraInjectedLinear(K, N, bias, r).to(DEV)
        with torch.no_grad():
            m.linear.weight.copy_(g("w0"))
            if bias:
                m.linear.bias.copy_(g("b"))
            m.lora_down.weight.copy_(g("A"))
            m.lora_up.weight.copy_(g("B"))
        m.linear.requires_grad_(False)
        m.scale = s
        x = g("x").to(DEV).requires_grad_(True)
        y = m(x)
        y.backward(g("dy").to(DEV))
        assert y.shape == g("y").shape
        assert rel_err(y, g("y")) < 1e-5 and rel_err(x.grad, g("dx")) < 1e-5
        if s != 0.0:
            assert rel_err(m.lora_down.weight.grad, g("dA")) < 1e-5
            assert rel_err(m.lora_up.weight.grad, g("dB")) < 1e-5


def test_module_modes_noncontiguous_nograd_autocast_deterministic():
    torch.manual_seed(3)
    m = L.LoraInjectedLinear(320, 640, bias=True, r=4).to(DEV)
    m.linear.requires_grad_(False)
    with torch.no_grad():
        m.lora_up.weight.normal_(0, 0.02)
    ref = O.OracleLoraLinear(320, 640, True, 4)
    ref.load_state_dict({k: v.cpu() for k, v in m.state_dict().items()})
    xb = torch.randn(4, 320, 50, device=DEV).transpose(1, 2)       # non-contiguous [4,50,320]
    assert not xb.is_contiguous()
    with torch.no_grad():
        y = m(xb)
    assert rel_err(y, ref(xb.cpu())) < 1e-5
    # autocast with fp32 masters (the reference's real mixed-precision mode)
    x = torch.randn(3, 77, 320, device=DEV, requires_grad=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = m(x)
    assert y.dtype == torch.bfloat16
    y.float().square().mean().backward()
    xr = x.detach().cpu().requires_grad_(True)
    yr = ref(xr)
    yr.square().mean().backward()
    assert rel_err(y, yr) < 2e-2 and rel_err(x.grad, xr.grad) < 2e-2
    assert m.lora_up.weight.grad.dtype == torch.float32
    assert rel_err(m.lora_up.weight.grad, ref.lora_up.weight.grad) < 2e-2
    assert rel_err(m.lora_down.weight.grad, ref.lora_down.weight.grad) < 2e-2
    # tune scale at run time, deterministic reductions are bit-reproducible
    L.tune_lora_scale(m, 0.25)
    outs = []
    for _ in range(2):
        m.zero_grad()
        x2 = torch.randn(2, 300, 320, device=DEV, generator=torch.Generator(DEV).manual_seed(5))
        ops.lora_linear(x2, m.linear.weight, m.linear.bias, m.lora_down.weight, m.lora_up.weight, m.scale,
                        deterministic=True).sum().backward()
        outs.append((m.lora_down.weight.grad.clone(), m.lora_up.weight.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


# SURVEY 8 census of the SD-1.5 UNet at batch 8 / 64x64 latents, at FULL size (M, K, N, r, bias): every
# distinct (M, K, N) the bench step launches as a single layer.  The oracle is fp32 CPU math on the inputs rounded
# to the io dtype (a CPU F.linear at (32768, 320, 2560) is 54 GFLOP: seconds).
CENSUS = [
    (32768, 320, 320, 4, False),       # attn1 q/k/v, attn2 q @ 64x64
    (32768, 320, 320, 4, True),        # to_out.0 carries a bias
    (32768, 320, 2560, 4, True),       # GEGLU proj @ 64x64
    (8192, 640, 640, 4, True),
    (8192, 640, 5120, 4, True),
    (2048, 1280, 1280, 4, True),
    (2048, 1280, 10240, 4, True),
    (512, 1280, 1280, 4, False),       # mid block @ 8x8
    (512, 1280, 10240, 16, True),
    (616, 768, 320, 4, False),         # attn2 to_k / to_v on the text context (8 x 77 tokens)
    (616, 768, 640, 4, False),
    (616, 768, 1280, 16, False),
    (8, 1280, 320, 16, True),          # time_emb_proj (config 3): M = batch
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", CENSUS)
def test_census_shapes_full_size_match_oracle(shape, dtype):
    """BASELINE full sizes through the C-ABI (forward, backward-dX, reductions) against the fp32 oracle: 2e-2 on
    the Frobenius ratio, plus the worst-row and max-element metrics."""
    M, K, N, r, bias = shape
    s = 0.7
    host, dev = make_case(M, K, N, r, bias, dtype, seed=M + N + r)
    check_against_oracle(host, run_cabi(dev, M, K, N, r, s, dtype, nat.IMPL_TC), s, 2e-2, f"census/{shape}/{dtype}")


def test_full_size_scale_linearity_bf16():
    """Size-independent property at the BASELINE shape: the branch is linear in `scale` (a runtime kernel argument)."""
    M, K, N, r = 32768, 320, 320, 4
    _, dev = make_case(M, K, N, r, True, torch.bfloat16, seed=1)
    y1, *_ = run_cabi(dev, M, K, N, r, 1.0, torch.bfloat16, nat.IMPL_TC)
    y0, *_ = run_cabi(dev, M, K, N, r, 0.0, torch.bfloat16, nat.IMPL_TC)
    y2, *_ = run_cabi(dev, M, K, N, r, 2.0, torch.bfloat16, nat.IMPL_TC)
    lhs = (y2.float() - y0.float())
    rhs = 2.0 * (y1.float() - y0.float())
    assert rel_err(lhs, rhs) < 5e-2      # branch is ~1e-2 of the output; bf16 rounding of y dominates


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(32768, 320, 640, 4), (5000, 200, 72, 12), (130, 64, 1280, 8)])
def test_gradient_reductions_all_modes(shape, dtype):
    """dB = s·dYᵀT, dA = s·dTᵀX through b200lora_lora_grads: atomic and deterministic modes, either
    output alone, accumulation into a non-zero buffer; the long case makes every CTA wrap its
    4-stage ring several times.  Reference: fp64 products of the same 16-bit inputs."""
    M, K, N, r = shape
    s = 0.5
    TOL_RED = 1e-4    # fp32 accumulation over up to 32768 tokens (BASELINE bar for 16-bit io: 2e-2)
    g = torch.Generator(DEV).manual_seed(11)
    dy = torch.randn(M, N, device=DEV, generator=g).to(dtype)
    x = torch.randn(M, K, device=DEV, generator=g).to(dtype)
    T = torch.randn(M, r, device=DEV, generator=g).to(dtype)
    dT = torch.randn(M, r, device=DEV, generator=g).to(dtype)
    dB_ref = (s * dy.double().t() @ T.double()).float()
    dA_ref = (s * dT.double().t() @ x.double()).float()
    lib, io, st = nat.lib(), nat.dtype_code(dtype), nat.stream_ptr(torch.device(DEV))
    ws_bytes = int(lib.b200lora_lora_grads_workspace(M, K, N, r))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)

    def run(want_a, want_b, det, init=0.0):
        dA = torch.full((r, K), init, dtype=torch.float32, device=DEV)
        dB = torch.full((N, r), init, dtype=torch.float32, device=DEV)
        nat.check(lib.b200lora_lora_grads(nat.ptr(dy), nat.ptr(x), nat.ptr(T), nat.ptr(dT), s,
                                          nat.ptr(dA) if want_a else None, nat.ptr(dB) if want_b else None,
                                          M, K, N, r, io, int(det), nat.ptr(ws) if det else None,
                                          ws_bytes if det else 0, st), "grads")
        torch.cuda.synchronize()
        return dA, dB

    for det in (False, True):
        dA, dB = run(True, True, det)
        assert rel_err(dA, dA_ref) < TOL_RED and rel_err(dB, dB_ref) < TOL_RED, (det, rel_err(dA, dA_ref), rel_err(dB, dB_ref))
        dA1, dB1 = run(True, False, det)
        assert rel_err(dA1, dA_ref) < TOL_RED and dB1.abs().max() == 0
        dA2, dB2 = run(False, True, det)
        assert rel_err(dB2, dB_ref) < TOL_RED and dA2.abs().max() == 0
    a0, b0 = run(True, True, True)
    a1, b1 = run(True, True, True)
    assert torch.equal(a0, a1) and torch.equal(b0, b1)          # deterministic mode is bit-reproducible
    dA, dB = run(True, True, False, init=1.0)                   # accumulates, does not overwrite
    assert rel_err(dA - 1.0, dA_ref) < 1e-4 and rel_err(dB - 1.0, dB_ref) < 1e-4


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(2048, 320, 320, 4), (777, 640, 1920, 16), (130, 1280, 1280, 8), (4100, 320, 72, 3)])
def test_backward_dx_fp32_output_equals_the_16bit_store(shape, dtype):
    """b200lora_linear_bwd_dx_f32out: the same kernel with the fp32 accumulator stored unrounded (what autograd needs when
    the layer's input was fp32 under autocast).  Rounded to the io dtype it must equal the 16-bit store BIT for bit, and it
    must agree with the fp32 oracle at least as well; dT is unchanged."""
    M, K, N, r = shape
    g = torch.Generator(DEV).manual_seed(21)
    dy = torch.randn(M, N, device=DEV, generator=g).to(dtype)
    w0 = (torch.randn(N, K, device=DEV, generator=g) * 0.05).to(dtype)
    A = (torch.randn(r, K, device=DEV, generator=g) / r).to(dtype)
    B = (torch.randn(N, r, device=DEV, generator=g) * 0.05).to(dtype)
    s = 0.7
    lib, io, st = nat.lib(), nat.dtype_code(dtype), nat.stream_ptr(torch.device(DEV))
    pack = torch.empty(80 * (K + N), dtype=dtype, device=DEV)
    nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), io, None, 0, nat.ptr(pack), K, N, r, io, st), "pack")
    wt = w0.t().contiguous()
    dx16 = torch.empty(M, K, dtype=dtype, device=DEV)
    dx32 = torch.full((M, K), float("nan"), dtype=torch.float32, device=DEV)
    dT16 = torch.empty(M, r, dtype=dtype, device=DEV)
    dT32 = torch.empty(M, r, dtype=dtype, device=DEV)
    nat.check(lib.b200lora_linear_bwd_dx(nat.ptr(dy), nat.ptr(wt), nat.W_NK, nat.ptr(A), nat.ptr(B), io, nat.ptr(pack), s,
                                         nat.ptr(dx16), nat.ptr(dT16), M, K, N, r, io, nat.IMPL_TC, st), "bwd_dx")
    nat.check(lib.b200lora_linear_bwd_dx_f32out(nat.ptr(dy), nat.ptr(wt), nat.ptr(pack), s, nat.ptr(dx32), nat.ptr(dT32),
                                                M, K, N, r, io, st), "bwd_dx_f32out")
    torch.cuda.synchronize()
    assert torch.isfinite(dx32).all()                      # every element written (ragged M / K tiles included)
    assert torch.equal(dx32.to(dtype), dx16)
    assert torch.equal(dT32, dT16)
    c = lambda t: t.detach().float().cpu()
    want = c(dy) @ c(w0) + ((c(dy) * s) @ c(B)).to(dtype).float() @ c(A)
    assert rel_err(dx32, want) < 2e-2 and rel_err(dx32, want) <= rel_err(dx16, want) + 1e-6


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_pack_with_input_cast_rider(dtype):
    """b200lora_pack_lora_cast / _fused_cast: the same pack as the plain entries, bit for bit, plus x_dst = x_src rounded to
    the io dtype (what `x.to(dtype)` gives), for sizes with and without a tail of fewer than 8 elements."""
    g = torch.Generator(DEV).manual_seed(4)
    lib, io, st = nat.lib(), nat.dtype_code(dtype), nat.stream_ptr(torch.device(DEV))
    K, N, r = 320, 640, 4
    A = torch.randn(r, K, device=DEV, generator=g) / r
    B = torch.randn(N, r, device=DEV, generator=g) * 0.05
    bias = torch.randn(N, device=DEV, generator=g)
    f32 = nat.dtype_code(torch.float32)
    for n_x in (8192 * 320, 777 * 72 + 5, 8):
        x = torch.randn(n_x, device=DEV, generator=g) * 3
        p0 = torch.empty(80 * (K + N), dtype=dtype, device=DEV)
        p1 = torch.full_like(p0, float("nan"))
        xd = torch.full((n_x,), float("nan"), dtype=dtype, device=DEV)
        nat.check(lib.b200lora_pack_lora(nat.ptr(A), nat.ptr(B), f32, nat.ptr(bias), f32, nat.ptr(p0), K, N, r, io, st), "pack")
        nat.check(lib.b200lora_pack_lora_cast(nat.ptr(A), nat.ptr(B), f32, nat.ptr(bias), f32, nat.ptr(p1), K, N, r, io,
                                              nat.ptr(x), nat.ptr(xd), n_x, st), "pack_cast")
        torch.cuda.synchronize()
        assert torch.equal(p0, p1) and torch.equal(xd, x.to(dtype))


def test_training_forward_with_fp32_input_uses_the_cast_rider_and_matches():
    """Under autocast with an fp32 input and trained factors the layer issues pack(+cast) + GEMM (2 launches, no ATen cast);
    outputs and gradients equal the B200LORA_CASTRIDE=0 path bit for bit (same conversion, same kernels)."""
    import os
    torch.manual_seed(0)
    m = L.LoraInjectedLinear(320, 640, bias=True, r=4).to(DEV)
    m.linear.requires_grad_(False)
    with torch.no_grad():
        m.lora_up.weight.normal_(0, 0.05)
    x = torch.randn(4, 500, 320, device=DEV, requires_grad=True)
    dy = torch.randn(4, 500, 640, device=DEV)
    res = []
    for flag in ("1", "0"):
        os.environ["B200LORA_CASTRIDE"] = flag
        try:
            x.grad = None
            m.lora_down.weight.grad = m.lora_up.weight.grad = None
            os.environ["B200LORA_DETERMINISTIC"] = "1"
            n0 = nat.LAUNCHES
            with torch.autocast("cuda", dtype=torch.bfloat16):
                y = m(x)
            fwd_launches = nat.LAUNCHES - n0
            y.backward(dy.to(y.dtype))
            res.append((y.detach().clone(), x.grad.clone(), m.lora_down.weight.grad.clone(), m.lora_up.weight.grad.clone(),
                        fwd_launches))
        finally:
            os.environ.pop("B200LORA_CASTRIDE", None)
            os.environ.pop("B200LORA_DETERMINISTIC", None)
    assert res[0][4] == 2 and res[1][4] == 2          # pack + GEMM either way; the ATen cast is not a C-ABI launch
    for a, b in zip(res[0][:4], res[1][:4]):
        assert torch.equal(a, b)
    assert res[0][1].dtype == torch.float32

"""Parity of LoraInjectedConv2d (implicit-GEMM tcgen05 path) with the oracle restatement
(oracle.lora_conv2d_forward + torch autograd on the CPU in fp32, inputs rounded to the io dtype).
The Conv2d surface is absent from the reference snapshot: parity is against the
restatement of upstream semantics — "unpinned by the reference".  Tolerance 2e-2 (bf16/fp16)."""
import pytest
import torch

import diffusion_finetuning_b200 as L
from conftest import max_abs_err, rel_err, worst_row_err
from oracle import lora_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

# (N, Cin, Cout, H, W, k, stride, pad, dil, r, bias, scale)
CASES = [
    (2, 64, 64, 16, 16, 3, 1, 1, 1, 4, True, 1.0),
    (1, 64, 128, 8, 8, 1, 1, 0, 1, 4, True, 0.7),          # conv_shortcut
    (2, 320, 320, 16, 16, 3, 1, 1, 1, 4, True, 1.0),       # SD-1.5 resnet conv
    (3, 96, 72, 10, 14, 3, 1, 1, 1, 16, False, 0.5),       # ragged pixels / channels, r = 16
    (1, 128, 64, 12, 12, 3, 1, 2, 2, 8, True, 1.0),        # dilation 2
    (2, 32, 40, 9, 9, 3, 1, 1, 1, 1, True, 1.3),           # Cin < 64 (zero-padded channel block)
    # BASELINE config 3: the wide resnet convs, r = 16 — many k-blocks per tap, Cin not a multiple of
    # the 5-stage ring, Cout split over several n tiles, the 1x1 shortcut with 2560 input channels
    (2, 1280, 1280, 8, 8, 3, 1, 1, 1, 16, True, 1.0),
    (2, 2560, 1280, 16, 16, 3, 1, 1, 1, 16, True, 1.0),
    (2, 1920, 640, 32, 32, 3, 1, 1, 1, 16, True, 0.7),
    (1, 960, 320, 64, 64, 3, 1, 1, 1, 16, True, 1.0),
    (2, 2560, 1280, 8, 8, 1, 1, 0, 1, 16, True, 1.0),      # conv_shortcut of the widest up block
]


def build(case, dtype):
    N, Cin, Cout, H, W, k, st, pad, dil, r, bias, s = case
    g = torch.Generator().manual_seed(N * 1000 + Cin + Cout)
    m = L.LoraInjectedConv2d(Cin, Cout, k, st, pad, dil, 1, bias, r=r, scale=s)
    with torch.no_grad():
        m.conv.weight.copy_(torch.empty_like(m.conv.weight).uniform_(-1, 1, generator=g) * (Cin * k * k) ** -0.5)
        if bias:
            m.conv.bias.copy_(torch.empty(Cout).uniform_(-0.1, 0.1, generator=g))
        m.lora_down.weight.copy_(torch.randn(m.lora_down.weight.shape, generator=g) / r * 0.3)
        m.lora_up.weight.copy_(torch.randn(m.lora_up.weight.shape, generator=g) * 0.05)
    m.conv.requires_grad_(False)
    x = torch.randn(N, Cin, H, W, generator=g)
    m = m.to(dtype)
    x = x.to(dtype)
    return m, x, g


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("case", CASES)
def test_conv_forward_backward_match_oracle(case, dtype):
    N, Cin, Cout, H, W, k, st, pad, dil, r, bias, s = case
    m, x, g = build(case, dtype)
    # oracle: fp32 math on the rounded values
    W0, b = m.conv.weight.detach().float(), None if m.conv.bias is None else m.conv.bias.detach().float()
    A = m.lora_down.weight.detach().float().clone().requires_grad_(True)
    B = m.lora_up.weight.detach().float().clone().requires_grad_(True)
    xr = x.float().clone().requires_grad_(True)
    y_ref = O.lora_conv2d_forward(xr, W0, b, A, B, s, st, pad, dil, 1)
    dy = torch.randn(y_ref.shape, generator=g).to(dtype)
    y_ref.backward(dy.float())

    md = m.to(DEV)
    xd = x.to(DEV).requires_grad_(True)
    y = md(xd)
    assert y.shape == y_ref.shape and y.dtype == dtype
    y.backward(dy.to(DEV))
    errs = {"y": rel_err(y, y_ref), "dx": rel_err(xd.grad, xr.grad),
            "dA": rel_err(md.lora_down.weight.grad, A.grad), "dB": rel_err(md.lora_up.weight.grad, B.grad)}
    bad = {k_: v for k_, v in errs.items() if not v < 2e-2}
    assert not bad, f"{case} {dtype}: {bad} (all {errs})"
    # outliers: one wrong output pixel / channel tile vanishes in the Frobenius ratio.  Rows = pixels over channels.
    nhwc = lambda t: t.detach().float().cpu().permute(0, 2, 3, 1)
    for name, a, b_ in (("y", nhwc(y), nhwc(y_ref)), ("dx", nhwc(xd.grad), nhwc(xr.grad))):
        assert worst_row_err(a, b_) < 8e-2 and max_abs_err(a, b_) < 8e-2, (case, dtype, name, worst_row_err(a, b_),
                                                                            max_abs_err(a, b_))


def test_conv_under_autocast_with_fp32_masters_and_channels_last():
    case = (2, 64, 64, 16, 16, 3, 1, 1, 1, 4, True, 1.0)
    m, x, g = build(case, torch.float32)
    md = m.to(DEV)
    xd = x.to(DEV).to(memory_format=torch.channels_last).requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = md(xd)
    assert y.dtype == torch.bfloat16
    y.float().square().mean().backward()
    assert md.lora_down.weight.grad.dtype == torch.float32 and md.lora_up.weight.grad.abs().sum() > 0
    y_ref = O.lora_conv2d_forward(x, m.conv.weight.cpu(), m.conv.bias.cpu(), m.lora_down.weight.cpu(),
                                  m.lora_up.weight.cpu(), 1.0, 1, 1, 1, 1)
    assert rel_err(y, y_ref) < 2e-2


def test_conv_fp32_without_autocast_fails_loudly():
    m = L.LoraInjectedConv2d(64, 64, 3, 1, 1, r=4).to(DEV)
    with pytest.raises(Exception, match="16-bit"):
        m(torch.randn(1, 64, 8, 8, device=DEV))


def test_extended_injection_runs_a_resnet_block():
    from diffusion_finetuning_b200.unet_sd15 import ResnetBlock2D
    torch.manual_seed(0)
    blk = torch.nn.Sequential()
    blk.add_module("res", ResnetBlock2D(64, 128, 64, groups=8))
    ref = torch.nn.Sequential()
    ref.add_module("res", ResnetBlock2D(64, 128, 64, groups=8))
    ref.load_state_dict(blk.state_dict())
    blk.requires_grad_(False)
    ref.requires_grad_(False)
    params, names = L.inject_trainable_lora_extended(blk, {"ResnetBlock2D"}, r=4)
    layers = O.inject(ref, {"ResnetBlock2D"}, r=4, convs=True)
    assert len(names) == len(layers) == 4       # conv1, time_emb_proj, conv2, conv_shortcut
    g = torch.Generator().manual_seed(5)
    ours = [mm for mm in blk.modules() if isinstance(mm, (L.LoraInjectedLinear, L.LoraInjectedConv2d))]
    for a, b_ in zip(ours, layers):
        with torch.no_grad():
            dn = torch.randn(b_.lora_down.weight.shape, generator=g) * 0.1
            up = torch.randn(b_.lora_up.weight.shape, generator=g) * 0.05
            a.lora_down.weight.copy_(dn); b_.lora_down.weight.copy_(dn)
            a.lora_up.weight.copy_(up); b_.lora_up.weight.copy_(up)
    x = torch.randn(2, 64, 12, 12, generator=g)
    temb = torch.randn(2, 64, generator=g)
    blk = blk.to(DEV)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = blk.res(x.to(DEV), temb.to(DEV))
    y_ref = ref.res(x, temb)
    assert rel_err(y, y_ref) < 3e-2
    y.float().mean().backward()
    y_ref.mean().backward()
    for a, b_ in zip(ours, layers):
        assert rel_err(a.lora_up.weight.grad, b_.lora_up.weight.grad) < 5e-2


def test_first_cuda_call_of_the_autograd_thread_is_ours():
    """Fresh process: the backward thread's very first CUDA work is our zcat launch (tensor-map
    encoding is a driver call and needs the context bound on that thread)."""
    import subprocess
    import sys
    code = (
        "import torch, diffusion_finetuning_b200 as L\n"
        "m = L.LoraInjectedConv2d(128, 128, 3, padding=1, r=4).cuda().to(torch.bfloat16)\n"
        "m.conv.requires_grad_(False)\n"
        "x = torch.randn(2, 128, 8, 8, device='cuda', dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)\n"
        "dy = torch.randn(2, 128, 8, 8, device='cuda', dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)\n"
        "m(x).backward(dy); torch.cuda.synchronize(); print('ok', float(x.grad.float().abs().sum()) > 0)\n")
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok True" in out.stdout, out.stderr[-2000:]

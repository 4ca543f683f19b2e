"""Regression: the forward must follow the LoRA factors through optimizers that do not bump `_version`
(`torch.optim.AdamW(fused=True)` updates parameters in place without touching autograd's version counter,
so a cache keyed on it never notices a training step).  Checked for the three packed paths: single Linear,
fused q/k/v siblings, Conv2d — in training mode AND in the no_grad call right after training."""
import pytest
import torch
import torch.nn as nn

import diffusion_finetuning_b200 as L
from conftest import rel_err
from diffusion_finetuning_b200 import _native as nat
from diffusion_finetuning_b200.unet_sd15 import CrossAttention
from oracle import lora_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _seed_up(mods):
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for m in mods:
            m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(DEV))


def _oracle_linear(m, x):
    c = lambda t: None if t is None else t.detach().float().cpu()
    return O.lora_linear_forward(c(x), c(m.linear.weight), c(m.linear.bias), c(m.lora_down.weight), c(m.lora_up.weight),
                                 m.scale)


def _train(params, fwd, steps=3):
    opt = torch.optim.AdamW(params, lr=2e-2, fused=True)
    v0 = [p._version for p in params]
    for _ in range(steps):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = fwd()
        y.float().square().mean().backward()
        opt.step()
        opt.zero_grad()
    return [p._version - v for p, v in zip(params, v0)]


def test_linear_follows_fused_adamw():
    torch.manual_seed(0)
    m = L.LoraInjectedLinear(320, 640, bias=True, r=4).to(DEV)
    m.linear.requires_grad_(False)
    _seed_up([m])
    x = torch.randn(2, 300, 320, device=DEV)
    B0 = m.lora_up.weight.detach().clone()
    launches = nat.LAUNCHES
    _train([m.lora_up.weight, m.lora_down.weight], lambda: m(x))
    assert rel_err(m.lora_up.weight, B0) > 0.5                   # the factors really moved
    assert nat.LAUNCHES - launches == 3 * 4                       # pack + fwd + dX + reductions, every step
    for grad in (True, False):
        with torch.set_grad_enabled(grad), torch.autocast("cuda", dtype=torch.bfloat16):
            y = m(x)
        assert rel_err(y, _oracle_linear(m, x)) < 2e-2, f"grad={grad}: forward used stale LoRA factors"
    # inference keeps its cache: two no_grad calls, one pack
    launches = nat.LAUNCHES
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        m(x)
        m(x)
    assert nat.LAUNCHES - launches == 2


def test_fused_siblings_follow_fused_adamw():
    torch.manual_seed(0)
    att = CrossAttention(320, None, 8, 40).to(DEV)
    att.requires_grad_(False)
    L.inject_trainable_lora(att, {"CrossAttention"}, r=4)
    layers = [m for m in att.modules() if isinstance(m, L.LoraInjectedLinear)]
    _seed_up(layers)
    groups = L.fuse_sibling_projections(att)
    assert len(groups) == 1
    x = torch.randn(2, 256, 320, device=DEV)
    params = [p for m in layers for p in (m.lora_up.weight, m.lora_down.weight)]
    _train(params, lambda: att(x))
    assert groups[0].enabled and groups[0].fused_calls >= 3
    for m in (att.to_q, att.to_k, att.to_v):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = m(x)                                              # first sibling launches the fused GEMM
        assert rel_err(y, _oracle_linear(m, x)) < 2e-2, "fused q/k/v forward used stale LoRA factors"


def test_conv_follows_fused_adamw():
    torch.manual_seed(0)
    conv = L.LoraInjectedConv2d(64, 128, 3, padding=1, r=4).to(DEV)
    conv.conv.requires_grad_(False)
    _seed_up([conv])
    x = torch.randn(2, 64, 16, 16, device=DEV)
    _train([conv.lora_up.weight, conv.lora_down.weight], lambda: conv(x))
    c = lambda t: t.detach().float().cpu()
    want = O.lora_conv2d_forward(c(x), c(conv.conv.weight), c(conv.conv.bias), c(conv.lora_down.weight),
                                 c(conv.lora_up.weight), 1.0, 1, 1, 1, 1)
    for grad in (True, False):
        with torch.set_grad_enabled(grad), torch.autocast("cuda", dtype=torch.bfloat16):
            y = conv(x)
        assert rel_err(y, want) < 3e-2, f"grad={grad}: conv forward used stale LoRA factors"

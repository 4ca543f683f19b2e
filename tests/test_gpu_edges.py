"""Edge cases of the module contract on the GPU: empty batches, `tune_lora_scale` between calls (scale is a
call-time argument, lora.py:597-600 — including for cached inference packs and fused siblings), re-assigned
`.lora_up.weight` Parameters (lora.py:531-536) under no_grad."""
import pytest
import torch

import diffusion_finetuning_b200 as L
from conftest import rel_err
from diffusion_finetuning_b200.unet_sd15 import CrossAttention
from oracle import lora_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _oracle(m, x):
    c = lambda t: None if t is None else t.detach().float().cpu()
    return O.lora_linear_forward(c(x), c(m.linear.weight), c(m.linear.bias), c(m.lora_down.weight), c(m.lora_up.weight),
                                 m.scale)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_empty_batch(dtype):
    m = L.LoraInjectedLinear(64, 128, bias=True, r=4).to(DEV).to(dtype)
    m.linear.requires_grad_(False)
    x = torch.zeros(0, 7, 64, device=DEV, dtype=dtype, requires_grad=True)
    y = m(x)
    assert y.shape == (0, 7, 128) and y.dtype == dtype
    y.sum().backward()
    assert x.grad.shape == x.shape
    assert m.lora_up.weight.grad is not None and float(m.lora_up.weight.grad.abs().sum()) == 0.0


def test_scale_and_swapped_weights_follow_at_call_time():
    torch.manual_seed(0)
    att = CrossAttention(320, None, 8, 40).to(DEV).to(torch.bfloat16)
    att.requires_grad_(False)
    L.inject_trainable_lora(att, {"CrossAttention"}, r=4)
    layers = [m for m in att.modules() if isinstance(m, L.LoraInjectedLinear)]
    with torch.no_grad():
        for m in layers:
            m.lora_up.weight.normal_(0, 0.05)
    x = torch.randn(2, 130, 320, device=DEV, dtype=torch.bfloat16)
    for fused in (False, True):
        if fused:
            L.fuse_sibling_projections(att)
        with torch.no_grad():
            for alpha in (1.0, 0.3, 0.0, 1.7):                    # the notebooks sweep alpha between calls
                L.tune_lora_scale(att, alpha)
                for m in (att.to_q, att.to_k, att.to_v, att.to_out[0]):
                    assert m.scale == alpha
                    assert rel_err(m(x), _oracle(m, x)) < 2e-2, (fused, alpha)
            # monkeypatch-style swap: a fresh Parameter object behind the same attribute
            L.tune_lora_scale(att, 1.0)
            new_up = torch.nn.Parameter(torch.randn(320, 4, device=DEV, dtype=torch.bfloat16) * 0.05)
            att.to_k.lora_up.weight = new_up
            for m in (att.to_q, att.to_k, att.to_v):
                assert rel_err(m(x), _oracle(m, x)) < 2e-2, fused


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("r", [32, 64, 20])
def test_ranks_above_16_train(r, dtype, tol):
    """The reference accepts any r <= min(in, out) (lora.py:36-39); `--lora_rank 32 / 64` are common.  Ranks above the
    16 the tcgen05 kernels carry run on the CUDA-core kernels, the token reductions in windows of 16 ranks (atomic and
    deterministic modes), `inspect_lora` on the any-rank abs-mean kernel."""
    from diffusion_finetuning_b200 import ops
    torch.manual_seed(r)
    m = L.LoraInjectedLinear(192, 136, bias=True, r=r).to(DEV).to(dtype)
    m.linear.requires_grad_(False)
    with torch.no_grad():
        m.lora_up.weight.normal_(0, 0.02)
    m.scale = 0.6
    x = torch.randn(3, 211, 192, device=DEV, dtype=dtype, requires_grad=True)
    dy = torch.randn(3, 211, 136, device=DEV, dtype=dtype)
    c = lambda t: t.detach().float().cpu()
    want_y = _oracle(m, x)
    dx, dA, dB = O.lora_linear_grads(c(x), c(m.linear.weight), None, c(m.lora_down.weight), c(m.lora_up.weight), 0.6, c(dy))
    for det in (False, True):
        m.zero_grad()
        x.grad = None
        y = ops.lora_linear(x, m.linear.weight, m.linear.bias, m.lora_down.weight, m.lora_up.weight, m.scale,
                            deterministic=det)
        y.backward(dy)
        assert rel_err(y, want_y) < tol and rel_err(x.grad, dx) < tol
        assert rel_err(m.lora_down.weight.grad, dA) < tol and rel_err(m.lora_up.weight.grad, dB) < tol, (
            det, rel_err(m.lora_down.weight.grad, dA), rel_err(m.lora_up.weight.grad, dB))
    moved = L.inspect_lora(torch.nn.Sequential(m))
    assert abs(moved["0"][0] - O.lora_abs_mean(c(m.lora_up.weight), c(m.lora_down.weight))) < (1e-6 if dtype == torch.float32 else 1e-4)


def test_rank_above_the_kernel_limit_fails_loudly_at_forward():
    m = L.LoraInjectedLinear(256, 256, r=96).to(DEV)          # constructing, saving, merging stay possible
    with pytest.raises(Exception, match="rank 96"):
        m(torch.randn(2, 256, device=DEV))

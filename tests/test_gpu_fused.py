"""Fused sibling projections: q/k/v layers called on one tensor run as one kernel; results
must equal the unfused per-layer path (which the other GPU tests pin against the oracle) and the oracle."""
import copy

import pytest
import torch

import diffusion_finetuning_b200 as L
from conftest import check_close, rel_err
from diffusion_finetuning_b200.unet_sd15 import CrossAttention
from oracle import lora_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _attn(query_dim, ctx_dim, heads, r, seed, bias_free=True):
    torch.manual_seed(seed)
    a = CrossAttention(query_dim, ctx_dim, heads, query_dim // heads)
    a.requires_grad_(False)
    return a


def _inject_pair(attn, r):
    ref = copy.deepcopy(attn)
    ref_layers = O.inject(ref, {"CrossAttention"}, r=r)
    ours = copy.deepcopy(attn)
    L.inject_trainable_lora(ours, {"CrossAttention"}, r=r)
    our_layers = [m for m in ours.modules() if isinstance(m, L.LoraInjectedLinear)]
    g = torch.Generator().manual_seed(5)
    with torch.no_grad():
        for a, b in zip(ref_layers, our_layers):
            a.lora_up.weight.normal_(0, 0.02, generator=g)
            b.lora_up.weight.copy_(a.lora_up.weight)
            b.lora_down.weight.copy_(a.lora_down.weight)
    return ref, ref_layers, ours.to(DEV), our_layers


@pytest.mark.parametrize("cross", [False, True])
@pytest.mark.parametrize("r", [4, 1])
def test_fused_attention_block_matches_oracle_and_unfused(cross, r):
    dim, ctx_dim, heads, B, Tq, Tk = 320, (768 if cross else None), 8, 2, 300, 77
    ref, ref_layers, ours, our_layers = _inject_pair(_attn(dim, ctx_dim, heads, r, 0), r)
    unf = copy.deepcopy(ours)
    unf_layers = [m for m in unf.modules() if isinstance(m, L.LoraInjectedLinear)]
    groups = L.fuse_sibling_projections(ours)
    assert len(groups) == 1 and len(groups[0].layers) == (2 if cross else 3)
    L.tune_lora_scale(ours, 0.7), L.tune_lora_scale(unf, 0.7)
    for m in ref_layers:
        m.scale = 0.7
    g = torch.Generator().manual_seed(9)
    x = torch.randn(B, Tq, dim, generator=g)
    ctx = torch.randn(B, Tk, ctx_dim, generator=g) if cross else None
    tgt = torch.randn(B, Tq, dim, generator=g)

    def run(mod, dev, autocast):
        xx = x.to(dev).detach().clone().requires_grad_(True)
        cc = None if ctx is None else ctx.to(dev).detach().clone().requires_grad_(True)
        with torch.autocast(dev.split(":")[0], dtype=torch.bfloat16, enabled=autocast):
            y = mod(xx, cc)
        torch.nn.functional.mse_loss(y.float(), tgt.to(dev)).backward()
        return y, xx.grad, (None if cc is None else cc.grad)

    # The oracle runs the whole block in FP32 on the CPU (the reference's fp32 path).  Between the LoRA layers sit
    # stock bf16 kernels (SDPA, softmax) on the GPU side, so this block-level comparison carries their rounding too;
    # the 2e-2 bar on the LoRA layers alone, against fp32 math on identical (rounded) inputs, is
    # test_fused_siblings_at_census_shapes_match_fp32_oracle below.
    y_ref, dx_ref, dc_ref = run(ref, "cpu", False)
    y_f, dx_f, dc_f = run(ours, DEV, True)
    y_u, dx_u, dc_u = run(unf, DEV, True)
    assert groups[0].fused_calls == 1 and groups[0].enabled and groups[0].pending is None
    tol = 2e-2
    assert rel_err(y_f, y_ref.float()) < tol and rel_err(dx_f, dx_ref) < tol
    if cross:
        assert rel_err(dc_f, dc_ref) < tol and rel_err(dc_f, dc_u) < 2e-2
    assert rel_err(y_f, y_u.float()) < 1e-2 and rel_err(dx_f, dx_u) < 2e-2
    for a, b, c in zip(ref_layers, our_layers, unf_layers):
        for nm in ("lora_up", "lora_down"):
            gf, gu, gr = getattr(b, nm).weight.grad, getattr(c, nm).weight.grad, getattr(a, nm).weight.grad
            assert gf is not None and gf.dtype == torch.float32
            assert rel_err(gf, gu) < 2e-2, (nm, rel_err(gf, gu))
            assert rel_err(gf, gr) < 2e-2, (nm, rel_err(gf, gr))


class _Siblings(torch.nn.Module):
    """to_q / to_k / to_v (or to_k / to_v) of one attention block, nothing else: the fused launch in isolation."""

    def __init__(self, K, N, names):
        super().__init__()
        for nm in names:
            setattr(self, nm, torch.nn.Linear(K, N, bias=False))


# (M, K, N, member names): the fused launches of one bench step at BASELINE size (SURVEY 8 census)
FUSED_CENSUS = [
    (32768, 320, 320, ("to_q", "to_k", "to_v")),
    (8192, 640, 640, ("to_q", "to_k", "to_v")),
    (2048, 1280, 1280, ("to_q", "to_k", "to_v")),
    (512, 1280, 1280, ("to_q", "to_k", "to_v")),
    (616, 768, 320, ("to_k", "to_v")),
    (616, 768, 640, ("to_k", "to_v")),
    (616, 768, 1280, ("to_k", "to_v")),
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("case", FUSED_CENSUS)
def test_fused_siblings_at_census_shapes_match_fp32_oracle(case, dtype):
    """ONE fused forward launch, ONE K-segmented dX launch and ONE multi-problem reduction launch at full size against
    fp32 CPU math (lora.py:49-50 and its autograd) on the same inputs rounded to the io dtype: 2e-2 (BASELINE), plus
    the worst-row / max-element metrics.  Different scale per member, separate output gradients per member."""
    M, K, N, names = case
    r = 4
    torch.manual_seed(M + N)
    blk = type("CrossAttention", (_Siblings,), {})(K, N, names)      # the reference matches targets by class NAME
    blk.requires_grad_(False)
    L.inject_trainable_lora(blk, {"CrossAttention"}, r=r)
    layers = [getattr(blk, nm) for nm in names]
    g = torch.Generator().manual_seed(7)
    rnd = lambda t: t.to(dtype).float()
    with torch.no_grad():
        for i, m in enumerate(layers):
            m.linear.weight.copy_(rnd(m.linear.weight))
            m.lora_up.weight.copy_(rnd(torch.randn(N, r, generator=g) * 0.02))
            m.lora_down.weight.copy_(rnd(m.lora_down.weight))
            m.scale = (1.0, 0.7, 0.35)[i]
    x = rnd(torch.randn(M, K, generator=g))
    dys = [rnd(torch.randn(M, N, generator=g)) for _ in names]
    want = []
    for m, dy in zip(layers, dys):
        W0, A, B = m.linear.weight.detach(), m.lora_down.weight.detach(), m.lora_up.weight.detach()
        y = O.lora_linear_forward(x, W0, None, A, B, m.scale)
        dx, dA, dB = O.lora_linear_grads(x, W0, None, A, B, m.scale, dy)
        want.append((y, dx, dA, dB))
    blk.to(DEV)
    groups = L.fuse_sibling_projections(blk)
    assert len(groups) == 1 and len(groups[0].layers) == len(names)
    xd = x.to(DEV).requires_grad_(True)
    with torch.autocast("cuda", dtype=dtype):
        ys = [m(xd) for m in layers]
    assert groups[0].fused_calls == 1
    loss = sum((y.float() * dy.to(DEV)).sum() for y, dy in zip(ys, dys))
    loss.backward()
    for i, (m, (y, dx, dA, dB)) in enumerate(zip(layers, want)):
        check_close(ys[i], y, 2e-2, f"y[{names[i]}]")
        check_close(m.lora_down.weight.grad, dA, 2e-2, f"dA[{names[i]}]")
        check_close(m.lora_up.weight.grad, dB, 2e-2, f"dB[{names[i]}]")
    check_close(xd.grad, sum(w[1] for w in want), 2e-2, "dx (sum over members)")


def test_group_switches_itself_off_when_inputs_differ_and_fp32_is_left_alone():
    _, _, ours, layers = _inject_pair(_attn(320, None, 8, 4, 1), 4)
    groups = L.fuse_sibling_projections(ours)
    q, k = ours.to_q, ours.to_k
    with torch.autocast("cuda", dtype=torch.bfloat16), torch.no_grad():
        for i in range(4):      # q and k on DIFFERENT tensors, v never called: parked outputs go to waste
            a = q(torch.randn(2, 64, 320, device=DEV))
            b = k(torch.randn(2, 64, 320, device=DEV))
            assert a.shape == b.shape == (2, 64, 320)
    assert not groups[0].enabled
    # fp32 without autocast: not eligible, the per-layer (SIMT) path answers
    _, _, ours32, _ = _inject_pair(_attn(320, None, 8, 4, 2), 4)
    g32 = L.fuse_sibling_projections(ours32)
    x = torch.randn(2, 64, 320, device=DEV)
    y = ours32(x)
    assert g32[0].fused_calls == 0 and y.dtype == torch.float32
    L.unfuse_sibling_projections(ours32)
    assert all(m._sibling_group is None for m in ours32.modules() if isinstance(m, L.LoraInjectedLinear))


def test_fused_grads_reach_the_gradient_arena():
    from diffusion_finetuning_b200 import dp
    _, _, ours, layers = _inject_pair(_attn(640, None, 8, 4, 3), 4)
    unf = copy.deepcopy(ours)
    L.fuse_sibling_projections(ours)
    outs = []
    for mod in (ours, unf):
        arena = dp.LoraGradArena(mod)
        arena.zero()
        x = torch.randn(4, 256, 640, device=DEV, generator=torch.Generator(DEV).manual_seed(1), requires_grad=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = mod(x)
        y.float().square().mean().backward()
        arena.finish()
        outs.append([p.grad.clone() for p in mod.parameters() if p.requires_grad])
    assert len(outs[0]) == len(outs[1]) == 8
    for a, b in zip(*outs):
        assert rel_err(a, b) < 2e-2

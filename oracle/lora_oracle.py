"""ORACLE — test infrastructure, NOT product code.

A CPU restatement, in plain torch ops, of the arithmetic of the reference's LoRA hot path
(levayz/diffusion_finetuning, `lora_diffusion/lora.py` unless another file is named).  Only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may
import this module, and only as the checker or the timed CPU baseline; the product package
`diffusion_finetuning_b200` never imports it.

Pinning: the reference ships NO tests or golden vectors for this path.  The oracle is
pinned by (a) `oracle/make_golden.py`, which imports the real reference `lora.py` in the build
container and commits its outputs on seeded inputs to `tests/golden/`, and (b)
`tests/test_oracle_golden.py` (oracle vs those vectors) and `tests/test_api_host.py::
test_against_reference_directly`, which runs against the reference itself whenever `/root/reference`
is present.  The Conv2d / selector / dropout surface does not exist in the
reference snapshot: for it the oracle restates upstream cloneofsimo/lora semantics as specified in
SURVEY §8(c) — **parity unpinned by the reference** for those rows.
"""
import json
from typing import List, Optional, Set

import torch
import torch.nn as nn
import torch.nn.functional as F


# ---------------------------------------------------------------------------------------------
# functional math
# ---------------------------------------------------------------------------------------------
def lora_linear_forward(x, W0, b, A, B, s):
    """lora.py:49-50 — evaluation order: linear -> down -> up -> *scale -> +.
    Intermediates are materialised in the tensors' dtype, as ATen does."""
    y0 = F.linear(x, W0, b)
    t = F.linear(x, A)
    u = F.linear(t, B)
    return y0 + u * s


def lora_linear_grads(x, W0, b, A, B, s, dy):
    """Closed-form backward of lora.py:50 (what autograd derives), in the inputs' dtype:
       dX = dY·W0 + s·(dY·B)·A ;  dB = s·dYᵀ·T ;  dA = s·(dY·B)ᵀ·X   with T = X·Aᵀ."""
    K = x.shape[-1]
    N = dy.shape[-1]
    x2, dy2 = x.reshape(-1, K), dy.reshape(-1, N)
    t = x2 @ A.t()
    dt = (dy2 * s) @ B
    dx = dy2 @ W0 + dt @ A
    dB = (dy2 * s).t() @ t
    dA = dt.t() @ x2
    return dx.reshape(x.shape), dA, dB


def merge_weight(W, up, down, alpha):
    """lora.py:423 — product in the LoRA dtype, cast to W's dtype, scaled, added."""
    return W + alpha * (up @ down).type(W.dtype)


def lora_abs_mean(up, down):
    """lora.py:744-746."""
    return (up @ down).flatten().abs().mean().item()


def svd_distill_matrix(mat, rank=4, clamp_quantile=0.99):
    """cli_svd.py:69-84 for one matrix: full SVD, top-r, U·diag(S), symmetric quantile clamp."""
    mat = mat.float()
    U, S, Vh = torch.linalg.svd(mat)
    U = U[:, :rank] @ torch.diag(S[:rank])
    Vh = Vh[:rank, :]
    if clamp_quantile is None:      # test-only: factorisation without the sign-dependent clamp
        return U, Vh
    hi = torch.quantile(torch.cat([U.flatten(), Vh.flatten()]), clamp_quantile)
    return U.clamp(-hi, hi), Vh.clamp(-hi, hi)


def lpl_mix(l1: List[torch.Tensor], l2: List[torch.Tensor], alpha: float):
    """cli_lora_add.py:49-58 — element-wise alpha*x1 + (1-alpha)*x2 over two .pt lists."""
    return [alpha * a + (1 - alpha) * b for a, b in zip(l1, l2)]


def lora_conv2d_forward(x, W0, b, A, B, s, stride=1, padding=0, dilation=1, groups=1):
    """SURVEY §8(c) (upstream LoraInjectedConv2d): conv + 1x1(conv_down) * scale."""
    y0 = F.conv2d(x, W0, b, stride, padding, dilation, groups)
    t = F.conv2d(x, A, None, stride, padding, dilation, groups)
    u = F.conv2d(t, B)
    return y0 + u * s


# ---------------------------------------------------------------------------------------------
# module-level restatement (for whole-model parity and the timed CPU baseline)
# ---------------------------------------------------------------------------------------------
class OracleLoraLinear(nn.Module):
    """Same composition as lora.py:32-50: three nn.Linear, rank check, init N(0, 1/r) / zeros."""

    def __init__(self, fin, fout, bias=False, r=4):
        super().__init__()
        if r > min(fin, fout):
            raise ValueError(f"LoRA rank {r} must be less or equal than {min(fin, fout)}")
        self.linear = nn.Linear(fin, fout, bias)
        self.lora_down = nn.Linear(fin, r, bias=False)
        self.lora_up = nn.Linear(r, fout, bias=False)
        self.scale = 1.0
        with torch.no_grad():
            self.lora_down.weight.normal_(0.0, 1.0 / r)
            self.lora_up.weight.zero_()

    def forward(self, x):
        return lora_linear_forward(x, self.linear.weight, self.linear.bias, self.lora_down.weight,
                                   self.lora_up.weight, self.scale)


class OracleLoraConv2d(nn.Module):
    def __init__(self, src: nn.Conv2d, r=4):
        super().__init__()
        self.conv = src
        self.lora_down = nn.Conv2d(src.in_channels, r, src.kernel_size, src.stride, src.padding,
                                   src.dilation, src.groups, bias=False)
        self.lora_up = nn.Conv2d(r, src.out_channels, 1, bias=False)
        self.scale = 1.0
        with torch.no_grad():
            self.lora_down.weight.normal_(0.0, 1.0 / r)
            self.lora_up.weight.zero_()

    def forward(self, x):
        c = self.conv
        return lora_conv2d_forward(x, c.weight, c.bias, self.lora_down.weight, self.lora_up.weight,
                                   self.scale, c.stride, c.padding, c.dilation, c.groups)


def walk_targets(model, ancestor_names: Set[str], kinds, skip_inside=(OracleLoraLinear, OracleLoraConv2d)):
    """lora.py:78-114 — ancestors by class NAME in model.modules() order, descendants by
    isinstance in named_modules() order, direct parent resolved by path, children of injected
    layers skipped.  Returns a list (the reference yields lazily while mutating; materialising
    first gives the same sequence because injected layers are skipped either way)."""
    out = []
    for anc in model.modules():
        if type(anc).__name__ not in ancestor_names:
            continue
        for full, mod in anc.named_modules():
            if not isinstance(mod, tuple(kinds)):
                continue
            *path, leaf = full.split(".")
            parent = anc
            for hop in path:
                parent = parent.get_submodule(hop)
            if isinstance(parent, tuple(skip_inside)):
                continue
            out.append((parent, leaf, mod))
    return out


def inject(model, targets: Set[str], r=4, convs=False):
    """lora.py:137-183 restated: returns the injected layers in discovery order."""
    kinds = [nn.Linear, nn.Conv2d] if convs else [nn.Linear]
    layers = []
    seen = set()
    for parent, leaf, mod in walk_targets(model, targets, kinds):
        if id(mod) in seen:      # the lazy reference walk never meets a module it already swapped
            continue
        seen.add(id(mod))
        if isinstance(mod, nn.Conv2d):
            new = OracleLoraConv2d(mod, r)
        else:
            new = OracleLoraLinear(mod.in_features, mod.out_features, mod.bias is not None, r)
            new.linear.weight = mod.weight
            if mod.bias is not None:
                new.linear.bias = mod.bias
        new.to(mod.weight.device).to(mod.weight.dtype)
        parent._modules[leaf] = new
        new.lora_up.weight.requires_grad = True
        new.lora_down.weight.requires_grad = True
        layers.append(new)
    return layers


def safeloras_layout(names_ranks_targets):
    """lora.py:228-258 — the key / metadata layout of a saved safetensors file, as data:
    input {name: (targets, [rank per pair])} -> (sorted tensor keys, metadata dict)."""
    keys, meta = [], {}
    for name, (targets, ranks) in names_ranks_targets.items():
        meta[name] = json.dumps(list(targets))
        for i, r in enumerate(ranks):
            meta[f"{name}:{i}:rank"] = str(r)
            keys += [f"{name}:{i}:up", f"{name}:{i}:down"]
    return sorted(keys), meta


# ---------------------------------------------------------------------------------------------
# textual-inversion step (cli_lora_pti.py) — restated; cli_lora_pti.py itself imports diffusers/fire/wandb
# at module top and cannot be imported here, so the optimizer IS torch.optim.AdamW as at :651-657
# ---------------------------------------------------------------------------------------------
def ti_make_optimizer(embedding: nn.Embedding, lr, weight_decay=0.0):
    """cli_lora_pti.py:651-657."""
    return torch.optim.AdamW(embedding.parameters(), lr=lr, betas=(0.9, 0.999), eps=1e-08, weight_decay=weight_decay)


@torch.no_grad()
def ti_post_step(embedding: nn.Embedding, orig_embeds_params, index_no_updates, lr, clip_ti_decay=True):
    """cli_lora_pti.py:315-346, after `optimizer.step(); optimizer.zero_grad()`: norm decay of the
    placeholder rows with lambda = min(1, 100·lr), then every other row restored from the clone taken
    before training (:280).  Returns (pre_norm, current_norm) as the reference prints them."""
    index_updates = ~index_no_updates
    w = embedding.weight
    pre_norm = w[index_updates, :].norm(dim=-1, keepdim=True)
    if clip_ti_decay:
        lambda_ = min(1.0, 100 * lr)
        w[index_updates] = F.normalize(w[index_updates, :], dim=-1) * (pre_norm + lambda_ * (0.4 - pre_norm))
    current_norm = w[index_updates, :].norm(dim=-1)
    w[index_no_updates] = orig_embeds_params[index_no_updates]
    return pre_norm.squeeze(-1), current_norm

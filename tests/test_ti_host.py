"""Host-side contract of the textual-inversion module (no GPU): validation errors, the CUDA-only rule,
and the oracle's restatement of cli_lora_pti.py:315-346 against its algebraic properties."""
import pytest
import torch
import torch.nn as nn

from diffusion_finetuning_b200 import _native as nat
from diffusion_finetuning_b200 import ti
from oracle import lora_oracle as O


def test_validation_and_cpu_refusal():
    emb = nn.Embedding(16, 8)
    with pytest.raises(ValueError):
        ti.PlaceholderTokenRows(emb, [])
    with pytest.raises(ValueError):
        ti.PlaceholderTokenRows(emb, [3, 3])
    with pytest.raises(ValueError):
        ti.PlaceholderTokenRows(emb, [16])
    with pytest.raises(ValueError):
        ti.PlaceholderTokenRows(nn.Embedding(16, 8, padding_idx=0), [1])
    pr = ti.PlaceholderTokenRows(emb, [14, 15]).install()
    assert pr.rows.shape == (2, 8) and pr.rows.dtype == torch.float32 and not emb.weight.requires_grad
    assert list(pr.state_dict()) == ["rows"] and len(list(pr.parameters())) == 1   # the table is not adopted
    with pytest.raises(nat.B200LoraError):          # no CPU fallback
        emb(torch.tensor([[1, 14]]))
    pr.remove()
    assert emb(torch.tensor([[1, 14]])).shape == (1, 2, 8)
    opt = ti.PlaceholderAdamW(pr, lr=1e-3)
    opt.step()                                      # no gradient yet: a no-op, like torch optimizers
    assert opt.state_dict()["step"] == 0


def test_c_abi_argument_checks_without_gpu():
    lib = nat.lib()
    assert lib.b200lora_ti_grad_rows(None, None, 0, 8, None, 1, None, nat.F32, None) == -6
    assert lib.b200lora_ti_step(None, 0, 8, None, 1, None, None, None, None, 1, 1e-3, 0.9, 0.999, 1e-8, 0.0, 0.1,
                                0.4, None, None, None, None) == -6


def test_oracle_post_step_properties():
    """lambda = min(1, 100·lr): lr >= 0.01 snaps the placeholder norms to 0.4 exactly; lr = 0 leaves
    them; every other row equals the clone taken before training."""
    torch.manual_seed(0)
    emb = nn.Embedding(32, 16)
    orig = emb.weight.data.clone()
    no_upd = torch.ones(32, dtype=torch.bool)
    no_upd[[30, 31]] = False
    with torch.no_grad():
        emb.weight.add_(0.1)
    drift = emb.weight.data.clone()
    pre, cur = O.ti_post_step(emb, orig, no_upd, lr=0.02)
    assert torch.allclose(cur, torch.full((2,), 0.4), atol=1e-6)
    assert torch.allclose(pre, drift[[30, 31]].norm(dim=-1))
    assert torch.equal(emb.weight.data[no_upd], orig[no_upd])
    with torch.no_grad():
        emb.weight.copy_(drift)
    pre, cur = O.ti_post_step(emb, orig, no_upd, lr=0.0)
    assert torch.allclose(emb.weight.data[[30, 31]], drift[[30, 31]], atol=1e-7)
    with torch.no_grad():
        emb.weight.copy_(drift)
    O.ti_post_step(emb, orig, no_upd, lr=0.02, clip_ti_decay=False)
    assert torch.equal(emb.weight.data[[30, 31]], drift[[30, 31]])

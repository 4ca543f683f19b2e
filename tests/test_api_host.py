"""Host-side logic of the drop-in API (no GPU): discovery order, injection contract, file formats,
monkeypatch family, error behaviour — against vectors produced by the reference
(tests/golden, oracle/make_golden.py) and against the reference itself when it is mounted."""
import copy
import itertools
import json
import os

import pytest
import torch
import torch.nn as nn

import diffusion_finetuning_b200 as L
from conftest import load_golden
from diffusion_finetuning_b200._native import B200LoraError
from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel


def tiny_unet():
    torch.manual_seed(0)
    unet = UNet2DConditionModel.tiny()
    unet.requires_grad_(False)
    return unet


def test_inject_matches_reference_order_names_and_init():
    blob, meta = load_golden("tiny_unet_step.safetensors")
    unet = tiny_unet()
    originals = {n: m.weight for n, m in unet.named_modules() if isinstance(m, nn.Linear)}
    torch.manual_seed(1)
    params, names = L.inject_trainable_lora(unet, r=4)
    assert names == json.loads(meta["names"])
    paths = [n for n, m in unet.named_modules() if isinstance(m, L.LoraInjectedLinear)]
    assert paths == json.loads(meta["paths"])
    assert len(params) == int(meta["n_param_groups"])
    flat = list(itertools.chain(*params))          # how every caller consumes the return value
    layers = [unet.get_submodule(p) for p in paths]
    for i, m in enumerate(layers):
        assert flat[2 * i] is m.lora_up.weight and flat[2 * i + 1] is m.lora_down.weight
        assert m.linear.weight is originals[paths[i]], "base Parameter must be shared, not copied"
        assert torch.equal(m.lora_down.weight, blob[f"down{i}"]), "same RNG consumption as reference"
        assert m.lora_up.weight.abs().max() == 0
        assert m.lora_up.weight.requires_grad and m.lora_down.weight.requires_grad
        assert not m.linear.weight.requires_grad
        assert m.scale == 1.0
    keys = set(unet.state_dict().keys())
    p0 = paths[0]
    assert {f"{p0}.linear.weight", f"{p0}.lora_down.weight", f"{p0}.lora_up.weight"} <= keys
    # a second inject finds nothing new (children of injected layers are skipped)
    assert L.inject_trainable_lora(unet, r=4)[1] == []


def test_rank_too_large_and_nothing_injected_errors():
    with pytest.raises(ValueError, match="LoRA rank 9 must be less or equal than 8"):
        L.LoraInjectedLinear(16, 8, r=9)
    with pytest.raises(ValueError, match="No lora injected."):
        L.extract_lora_ups_down(tiny_unet())
    with pytest.raises(AssertionError):
        L._text_lora_path("x.safetensors")


def test_cpu_forward_fails_loudly():
    m = L.LoraInjectedLinear(16, 16, r=2)
    with pytest.raises(B200LoraError, match="no CPU fallback"):
        m(torch.randn(2, 16))


def _load_golden_unet():
    blob, meta = load_golden("tiny_unet_step.safetensors")
    unet = tiny_unet()
    torch.manual_seed(1)
    L.inject_trainable_lora(unet, r=4)
    layers = [m for m in unet.modules() if isinstance(m, L.LoraInjectedLinear)]
    with torch.no_grad():
        for i, m in enumerate(layers):
            m.lora_up.weight.copy_(blob[f"up{i}"])
    return unet, layers


def test_pt_format_equals_reference_file(tmp_path, golden_dir):
    unet, _ = _load_golden_unet()
    out = tmp_path / "lora.pt"
    L.save_lora_weight(unet, str(out))
    mine = torch.load(out)
    ref = torch.load(os.path.join(golden_dir, "tiny_unet_lora.pt"))
    assert len(mine) == len(ref)
    for a, b in zip(mine, ref):
        assert a.dtype == torch.float16 and a.device.type == "cpu" and torch.equal(a, b)


def test_safetensors_format_equals_reference_file(tmp_path, golden_dir):
    from safetensors import safe_open
    unet, _ = _load_golden_unet()
    out = tmp_path / "lora.safetensors"
    L.save_safeloras({"unet": (unet, L.DEFAULT_TARGET_REPLACE)}, str(out))
    a = safe_open(str(out), "pt", "cpu")
    b = safe_open(os.path.join(golden_dir, "tiny_unet_lora.safetensors"), "pt", "cpu")
    assert sorted(a.keys()) == sorted(b.keys())
    ma, mb = a.metadata(), b.metadata()
    assert set(json.loads(ma.pop("unet"))) == set(json.loads(mb.pop("unet")))
    assert ma == mb
    for k in a.keys():
        assert torch.equal(a.get_tensor(k), b.get_tensor(k))


def test_parse_reference_shipped_safetensors_slice(golden_dir):
    with open(os.path.join(golden_dir, "disney_slice.json")) as fh:
        want = json.load(fh)
    import hashlib
    sha = lambda t: hashlib.sha256(t.detach().contiguous().numpy().tobytes()).hexdigest()[:16]
    path = os.path.join(golden_dir, "disney_slice.safetensors")
    loras, embeds = L.load_safeloras_both(path)
    assert set(loras) == set(want["parsed"])
    for name, (weights, ranks, targets) in loras.items():
        w = want["parsed"][name]
        assert ranks == w["ranks"] and sorted(targets) == w["targets"]
        assert [list(t.shape) for t in weights] == w["shapes"]
        assert [sha(t) for t in weights] == w["sha"]
        assert all(isinstance(t, nn.Parameter) for t in weights)
    assert {k: sha(v) for k, v in embeds.items()} == want["embeds"]
    assert want["full_file"]["unet_pairs"] == 144 and want["full_file"]["unet_params"] == 311616


def test_missing_metadata_raises(tmp_path):
    from safetensors.torch import save_file
    p = tmp_path / "bad.safetensors"
    save_file({"unet:0:up": torch.zeros(2, 1), "unet:0:down": torch.zeros(1, 2)}, str(p), {"x": "y"})
    with pytest.raises(ValueError, match="has no metadata"):
        L.load_safeloras(str(p))


class CrossAttention(nn.Module):
    def __init__(self):
        super().__init__()
        self.to_q = nn.Linear(16, 24, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(24, 16)])


def test_monkeypatch_family_and_scale():
    m = nn.Sequential(CrossAttention())
    w_q = m[0].to_q.weight
    ups = [torch.randn(24, 2), torch.randn(2, 16), torch.randn(16, 3), torch.randn(3, 24)]
    lst = [t.half() for t in ups]
    L.monkeypatch_or_replace_lora(m, lst, {"CrossAttention"}, r=[2, 3])
    assert lst == [], "the loras list is consumed (pop(0))"
    q, o = m[0].to_q, m[0].to_out[0]
    assert isinstance(q, L.LoraInjectedLinear) and q.linear.weight is w_q
    assert q.lora_up.weight.dtype == torch.float32 and q.lora_down.weight.shape == (2, 16)
    assert o.lora_down.weight.shape == (3, 24) and o.linear.bias is not None
    assert torch.equal(q.lora_up.weight, ups[0].half().float())
    # replace on already-injected layers
    L.monkeypatch_replace_lora(m, [torch.ones(24, 2), torch.ones(2, 16), torch.ones(16, 2), torch.ones(2, 24)],
                               {"CrossAttention"}, r=2)
    assert m[0].to_q.linear.weight is w_q and m[0].to_q.lora_up.weight.eq(1).all()
    # add: alpha*new + beta*old
    L.monkeypatch_add_lora(m, [torch.ones(24, 2), torch.ones(2, 16), torch.ones(16, 2), torch.ones(2, 24)],
                           {"CrossAttention"}, alpha=0.5, beta=2.0)
    assert torch.allclose(m[0].to_q.lora_up.weight, torch.full((24, 2), 2.5))
    L.tune_lora_scale(m, 0.3)
    assert m[0].to_q.scale == 0.3 and m[0].to_out[0].scale == 0.3
    assert len(L.extract_lora_ups_down(m, {"CrossAttention"})) == 2
    L.monkeypatch_remove_lora(m)
    assert type(m[0].to_q) is nn.Linear and m[0].to_q.weight is w_q
    assert type(m[0].to_out[0]) is nn.Linear and m[0].to_out[0].bias is not None


def test_extended_inject_wraps_convs():
    unet = tiny_unet()
    params, names = L.inject_trainable_lora_extended(unet, r=2)
    convs = [m for m in unet.modules() if isinstance(m, L.LoraInjectedConv2d)]
    lins = [m for m in unet.modules() if isinstance(m, L.LoraInjectedLinear)]
    assert len(convs) > 0 and len(lins) > 36 - 1      # attention linears + time_emb_proj
    c = convs[0]
    assert c.lora_down.kernel_size == c.conv.kernel_size and c.lora_up.kernel_size == (1, 1)
    assert c.lora_up.weight.abs().max() == 0 and len(params) == 2 * len(names)
    L.tune_lora_scale(unet, 0.5)
    assert c.scale == 0.5
    L.monkeypatch_remove_lora(unet)
    assert not any(isinstance(m, (L.LoraInjectedConv2d, L.LoraInjectedLinear)) for m in unet.modules())


def test_selector_and_realize():
    m = L.LoraInjectedLinear(8, 8, r=2)
    m.set_selector_from_diag(torch.tensor([1.0, 0.0]))
    assert torch.equal(m._down_weight()[1], torch.zeros(8))
    up, down = m.realize_as_lora()
    assert up.shape == (8, 2) and down.shape == (2, 8)


def test_against_reference_directly(reference_lora, tmp_path):
    """Differential run against the real reference (build container only)."""
    R = reference_lora
    a, b = tiny_unet(), tiny_unet()
    torch.manual_seed(5)
    pa, na = R.inject_trainable_lora(a, r=4)
    torch.manual_seed(5)
    pb, nb = L.inject_trainable_lora(b, r=4)
    assert na == nb and len(pa) == len(pb)
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa.keys()) == list(sb.keys())
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k
    R.save_lora_weight(a, str(tmp_path / "a.pt"))
    L.save_lora_weight(b, str(tmp_path / "b.pt"))
    for x, y in zip(torch.load(tmp_path / "a.pt"), torch.load(tmp_path / "b.pt")):
        assert torch.equal(x, y)
    # reference can load what we save, and vice versa
    R.monkeypatch_or_replace_lora(a, torch.load(tmp_path / "b.pt"), r=4)
    L.monkeypatch_or_replace_lora(b, torch.load(tmp_path / "a.pt"), r=4)
    R.tune_lora_scale(b, 0.25)       # class-name matching must find OUR class
    assert all(m.scale == 0.25 for m in b.modules() if isinstance(m, L.LoraInjectedLinear))


def test_fuse_geglu_patches_and_restores_only_injected_geglu_modules():
    """`fuse_geglu` routes a `GEGLU` module's forward through its LoRA-injected `proj`; modules whose `proj`
    is not injected are left alone, patching twice is idempotent, `unfuse_geglu` restores the module's own forward."""
    import diffusion_finetuning_b200 as L
    from diffusion_finetuning_b200.unet_sd15 import GEGLU

    class Block(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a = GEGLU(64, 128)
            self.b = GEGLU(64, 128)

    blk = Block()
    L.inject_trainable_lora(blk.a, {"GEGLU"}, r=4)          # only `a` gets a LoRA projection
    assert isinstance(blk.a.proj, L.LoraInjectedLinear) and not isinstance(blk.b.proj, L.LoraInjectedLinear)
    fwd_a, fwd_b = blk.a.forward, blk.b.forward
    patched = L.fuse_geglu(blk)
    assert patched == [blk.a] and blk.a.forward is not fwd_a and blk.b.forward == fwd_b
    assert L.fuse_geglu(blk) == [blk.a]                      # idempotent
    L.unfuse_geglu(blk)
    assert blk.a.forward == fwd_a
    # the un-injected module still runs on the CPU as plain PyTorch
    assert blk.b(torch.randn(3, 64)).shape == (3, 128)

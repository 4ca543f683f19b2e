"""Artefact tools on the GPU: the `lpl` mix is bit-identical to the reference's CPU expression
(cli_lora_add.py:49-58), the streamed writers produce the same files as lora.py:201-213 / 228-258."""
import os

import pytest
import torch

import diffusion_finetuning_b200 as L
from diffusion_finetuning_b200 import artefacts as A
from diffusion_finetuning_b200 import dp
from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
from oracle import lora_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _lists(dtype, n_layers=37):
    g = torch.Generator().manual_seed(5)
    l1, l2 = [], []
    for i in range(n_layers):
        N, K, r = 8 * (i % 5 + 1) + 1, 16 * (i % 3 + 1) + 3, 1 + i % 4     # odd sizes: unaligned tails
        for lst in (l1, l2):
            lst.append((torch.randn(N, r, generator=g) * 0.3).to(dtype))
            lst.append((torch.randn(r, K, generator=g) * 0.3).to(dtype))
    return l1, l2


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32, torch.bfloat16])
@pytest.mark.parametrize("alpha", [0.5, 0.3, 1.0, 0.0])
def test_lpl_mix_is_bit_identical_to_reference_expression(dtype, alpha, tmp_path):
    l1, l2 = _lists(dtype)
    p1, p2, out = tmp_path / "a.pt", tmp_path / "b.pt", tmp_path / "o.pt"
    torch.save(l1, p1)
    torch.save(l2, p2)
    A.lora_add(str(p1), str(p2), str(out), alpha=alpha, mode="lpl")
    got = torch.load(out)
    want = O.lpl_mix([t.clone() for t in l1], [t.clone() for t in l2], alpha)
    assert len(got) == len(want) == len(l1)
    for a, b in zip(got, want):
        assert a.dtype == b.dtype and a.device.type == "cpu" and torch.equal(a, b)


def test_lpl_with_text_lora_and_missing_sibling(tmp_path):
    l1, l2 = _lists(torch.float16, 3)
    for name, lst in (("a", l1), ("b", l2)):
        torch.save(lst, tmp_path / f"{name}.pt")
    torch.save(l1[:2], tmp_path / "a.text_encoder.pt")
    A.lora_add(str(tmp_path / "a.pt"), str(tmp_path / "b.pt"), str(tmp_path / "o.pt"), with_text_lora=True)
    assert os.path.exists(tmp_path / "o.pt") and not os.path.exists(tmp_path / "o.text_encoder.pt")
    torch.save(l2[:2], tmp_path / "b.text_encoder.pt")
    A.lora_add(str(tmp_path / "a.pt"), str(tmp_path / "b.pt"), str(tmp_path / "o.pt"), alpha=0.25, with_text_lora=True)
    got = torch.load(tmp_path / "o.text_encoder.pt")
    want = O.lpl_mix([t.clone() for t in l1[:2]], [t.clone() for t in l2[:2]], 0.25)
    assert all(torch.equal(a, b) for a, b in zip(got, want))


def test_upl_merges_on_the_gpu(tmp_path):
    """mode 'upl' with a stand-in pipeline object (diffusers is not installed): W += alpha·up@down for every
    target Linear (cli_lora_add.py:72-88, lora.py:410-424), fp32 LoRA so the bar is fp32 rounding."""
    torch.manual_seed(0)
    unet = UNet2DConditionModel.tiny().to(DEV)
    lins = [m for _, _, m in L._find_modules(unet, L.DEFAULT_TARGET_REPLACE, search_class=[torch.nn.Linear])]
    g = torch.Generator().manual_seed(1)
    loras = []
    for m in lins:
        loras += [torch.randn(m.out_features, 4, generator=g) * 0.05, torch.randn(4, m.in_features, generator=g) * 0.05]
    torch.save(loras, tmp_path / "l.pt")
    before = [m.weight.detach().cpu().clone() for m in lins]

    class Pipe:
        def __init__(self):
            self.unet, self.saved = unet, None

        def save_pretrained(self, path):
            self.saved = path

    pipe = A.lora_add("base", str(tmp_path / "l.pt"), str(tmp_path / "out"), alpha=0.7, mode="upl",
                      load_pipeline=lambda path: Pipe())
    assert pipe.saved == str(tmp_path / "out")
    lins2 = [m for _, _, m in L._find_modules(unet, L.DEFAULT_TARGET_REPLACE, search_class=[torch.nn.Linear])]
    for m, w0, (up, down) in zip(lins2, before, zip(loras[::2], loras[1::2])):
        want = O.merge_weight(w0, up, down, 0.7)
        assert torch.allclose(m.weight.detach().cpu(), want, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("arena_params", [False, True])
def test_streamed_writers_equal_the_reference_layout_writers(arena_params, tmp_path):
    torch.manual_seed(0)
    unet = UNet2DConditionModel.tiny().to(DEV)
    unet.requires_grad_(False)
    L.inject_trainable_lora(unet, r=4)
    with torch.no_grad():
        for m in unet.modules():
            if isinstance(m, L.LoraInjectedLinear):
                m.lora_up.weight.normal_(0, 0.02)
    if arena_params:                       # parameters re-homed into ONE flat buffer: no gather copy at all
        arena = dp.LoraGradArena(unet)
        opt = dp.ArenaAdamW(arena, lr=1e-3)
        w = [m.lora_up.weight for m in unet.modules() if isinstance(m, L.LoraInjectedLinear)]
        assert len({t.untyped_storage().data_ptr() for t in w}) == 1 and opt.flat_p.numel() > 0
    L.save_lora_weight(unet, str(tmp_path / "ref.pt"))
    A.save_lora_weight_streamed(unet, str(tmp_path / "ours.pt"))
    ref, ours = torch.load(tmp_path / "ref.pt"), torch.load(tmp_path / "ours.pt")
    assert len(ref) == len(ours) > 0
    for a, b in zip(ours, ref):
        assert a.dtype == torch.float16 and a.device.type == "cpu" and a.shape == b.shape and torch.equal(a, b)
        assert a.untyped_storage().nbytes() == a.numel() * 2       # own storage, like the reference's tensors
    embeds = {"<s1>": torch.randn(32, device=DEV)}
    L.save_safeloras_with_embeds({"unet": (unet, L.DEFAULT_TARGET_REPLACE)}, embeds, str(tmp_path / "ref.safetensors"))
    A.save_safeloras_streamed({"unet": (unet, L.DEFAULT_TARGET_REPLACE)}, embeds, str(tmp_path / "ours.safetensors"))
    from safetensors import safe_open
    fr, fo = safe_open(str(tmp_path / "ref.safetensors"), "pt"), safe_open(str(tmp_path / "ours.safetensors"), "pt")
    assert fr.metadata() == fo.metadata() and sorted(fr.keys()) == sorted(fo.keys())
    for k in fr.keys():
        assert torch.equal(fr.get_tensor(k), fo.get_tensor(k))


# ---------------------------------------------------------------------------------------------------
# against fixtures the REFERENCE computed (oracle/make_golden.py executes cli_lora_add.add and cli_svd.svd_distill)
# ---------------------------------------------------------------------------------------------------
def _numbered(blob, prefix):
    out = []
    while f"{prefix}.{len(out)}" in blob:
        out.append(blob[f"{prefix}.{len(out)}"])
    return out


def test_lpl_equals_reference_run_bit_for_bit(tmp_path):
    from conftest import load_golden
    blob, meta = load_golden("lpl_case.safetensors")
    for kind, suffix in (("unet", ".pt"), ("text", ".text_encoder.pt")):
        torch.save(_numbered(blob, f"{kind}.a"), tmp_path / f"a{suffix}")
        torch.save(_numbered(blob, f"{kind}.b"), tmp_path / f"b{suffix}")
    A.lora_add(str(tmp_path / "a.pt"), str(tmp_path / "b.pt"), str(tmp_path / "o.pt"), alpha=float(meta["alpha"]),
               mode="lpl", with_text_lora=True)
    for kind, suffix in (("unet", ".pt"), ("text", ".text_encoder.pt")):
        got, want = torch.load(tmp_path / f"o{suffix}"), _numbered(blob, f"{kind}.out")
        assert len(got) == len(want)
        for a, b in zip(got, want):
            assert a.dtype == b.dtype == torch.float16 and torch.equal(a, b)


class _Pipes:
    """Stand-in pipelines with the block class names cli_svd.py:46-54 matches on, filled from the fixture."""

    class CrossAttention(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.to_q = torch.nn.Linear(48, 48, bias=False)
            self.to_k = torch.nn.Linear(40, 48, bias=False)
            self.to_out = torch.nn.ModuleList([torch.nn.Linear(48, 48)])

    class GEGLU(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.proj = torch.nn.Linear(48, 192)

    class CLIPAttention(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.q_proj = torch.nn.Linear(32, 32)
            self.out_proj = torch.nn.Linear(32, 32)

    class Pipe:
        def __init__(self, blob, tag, device):
            self.unet = torch.nn.Sequential(_Pipes.CrossAttention(), _Pipes.GEGLU()).half()
            self.text_encoder = torch.nn.Sequential(_Pipes.CLIPAttention()).half()
            for which, model, targets in (("unet", self.unet, A.SVD_UNET_TARGETS), ("text", self.text_encoder, A.SVD_TEXT_TARGETS)):
                ws = A.extract_linear_weights(model, targets)
                src = _numbered(blob, f"{which}.{tag}")
                assert len(ws) == len(src)
                with torch.no_grad():
                    for w, s in zip(ws, src):
                        w.copy_(s)
            self.unet.to(device)
            self.text_encoder.to(device)


def _aligned(U, Vh, Ug, Vg):
    """cuSOLVER and LAPACK pick the sign of each singular pair independently: flip ours onto the fixture's."""
    sign = torch.sign((U * Ug).sum(0))
    sign[sign == 0] = 1
    return U * sign, Vh * sign[:, None]


def test_svd_distill_end_to_end_against_reference_run(tmp_path):
    """cli_svd.svd_distill (cli_svd.py:29-111): extraction order, fp16 difference, rank-4 factors, fp32 lists in
    `save_path` and `<save_path>.text_encoder.pt`, against what the reference itself saved for the same two pipelines.
    The clamp threshold is a quantile of the SIGNED entries, so it depends on the SVD's arbitrary sign choice per
    singular pair (cuSOLVER's differs from the CPU LAPACK's).  Hence two comparisons: (1) our UNCLAMPED factors,
    flipped onto the fixture's signs and clamped by the reference's rule (:79-84), must equal the fixture everywhere;
    (2) the files `svd_distill` wrote must agree with the fixture on every entry that neither clamp touched."""
    from conftest import load_golden, rel_err
    blob, meta = load_golden("svd_distill_case.safetensors")
    rank, q = int(meta["rank"]), float(meta["clamp_quantile"])
    pipes = {}
    out = tmp_path / "svd_distill.pt"
    got_unet, got_clip = A.svd_distill("tuned", "base", rank=rank, clamp_quantile=q, device=DEV, save_path=str(out),
                                       load_pipeline=lambda path, device: pipes.setdefault(path, _Pipes.Pipe(blob, path, device)))
    for which, path, got_live in (("unet", out, got_unet), ("text", tmp_path / "svd_distill.text_encoder.pt", got_clip)):
        got, want = torch.load(path), _numbered(blob, f"{which}.out")
        assert len(got) == len(want) == len(got_live)
        tuned, base = _numbered(blob, f"{which}.tuned"), _numbered(blob, f"{which}.base")
        raw = A.svd_distill_weights([t.to(DEV) for t in tuned], [b.to(DEV) for b in base], rank, None, DEV)
        for i in range(0, len(want), 2):
            Ug, Vg = want[i], want[i + 1]
            # (1) factorisation + the reference's clamp rule, sign-aligned: equal to the fixture
            U0, V0 = _aligned(raw[i].cpu(), raw[i + 1].cpu(), Ug, Vg)
            hi = torch.quantile(torch.cat([U0.flatten(), V0.flatten()]), q)
            assert torch.allclose(U0.clamp(-hi, hi), Ug, atol=3e-4 * Ug.abs().max().item())
            assert torch.allclose(V0.clamp(-hi, hi), Vg, atol=3e-4 * Vg.abs().max().item())
            # (2) the saved files: fp32, right shapes, equal wherever neither clamp bit
            U, Vh = got[i].float().cpu(), got[i + 1].float().cpu()
            assert got[i].dtype == torch.float32 and U.shape == Ug.shape and Vh.shape == Vg.shape
            U, Vh = _aligned(U, Vh, Ug, Vg)
            lim = min(max(U.abs().max(), Vh.abs().max()), max(Ug.abs().max(), Vg.abs().max())).item()
            for a, b in ((U, Ug), (Vh, Vg)):
                inside = (a.abs() < 0.999 * lim) & (b.abs() < 0.999 * lim)
                assert inside.float().mean() > 0.9
                assert (a[inside] - b[inside]).abs().max().item() < 3e-4 * max(1.0, b.abs().max().item())
            own_hi = torch.quantile(torch.cat([raw[i].flatten(), raw[i + 1].flatten()]).float().cpu(), q).item()
            both = torch.cat([got[i].flatten(), got[i + 1].flatten()]).float().cpu()
            assert both.max().item() <= own_hi * (1 + 1e-5) and both.min().item() >= -own_hi * (1 + 1e-5)   # clamp at OUR threshold
            delta = (tuned[i // 2] - base[i // 2]).float()
            assert rel_err(U @ Vh, delta) < 0.2                 # and it is a clamped rank-4 approximation of tuned - base

"""Generate tests/golden/* by EXECUTING THE REFERENCE (lora_diffusion/lora.py loaded by file path
from /root/reference) on seeded inputs.  Run in the build container only; the GPU box has no
/root/reference, so the committed fixtures are what travels.

    python oracle/make_golden.py

Test infrastructure — never imported by the product package.
"""
import hashlib
import importlib.util
import json
import os
import sys

import torch
import torch.nn as nn
from safetensors import safe_open
from safetensors.torch import save_file

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("B200LORA_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_lora", os.path.join(REF, "lora_diffusion", "lora.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


# (M-shape, K, N, r, bias, scale): small enough for the CPU, ragged on purpose
LINEAR_CASES = [
    ((3, 50), 64, 96, 4, False, 1.0),
    ((2, 77), 96, 64, 1, True, 0.7),
    ((130,), 160, 160, 16, True, 1.0),
    ((1, 7), 72, 40, 4, False, 0.0),
    ((5, 33), 128, 320, 8, True, 0.35),
]


def sha(t):
    return hashlib.sha256(t.detach().cpu().contiguous().numpy().tobytes()).hexdigest()[:16]


def golden_linear(ref):
    blob = {}
    for i, (lead, K, N, r, bias, s) in enumerate(LINEAR_CASES):
        g = torch.Generator().manual_seed(100 + i)
        m = ref.LoraInjectedLinear(K, N, bias, r)
        with torch.no_grad():
            m.linear.weight.copy_(torch.empty(N, K).uniform_(-K ** -0.5, K ** -0.5, generator=g))
            if bias:
                m.linear.bias.copy_(torch.empty(N).uniform_(-0.1, 0.1, generator=g))
            m.lora_down.weight.copy_(torch.randn(r, K, generator=g) / r)
            m.lora_up.weight.copy_(torch.randn(N, r, generator=g) * 0.02)   # non-zero up
        m.linear.requires_grad_(False)
        m.scale = s
        x = torch.randn(*lead, K, generator=g, requires_grad=True)
        dy = torch.randn(*lead, N, generator=g)
        y = m(x)
        y.backward(dy)
        pre = f"c{i}."
        blob[pre + "x"] = x.detach()
        blob[pre + "w0"] = m.linear.weight.detach()
        if bias:
            blob[pre + "b"] = m.linear.bias.detach()
        blob[pre + "A"] = m.lora_down.weight.detach()
        blob[pre + "B"] = m.lora_up.weight.detach()
        blob[pre + "dy"] = dy
        blob[pre + "y"] = y.detach()
        blob[pre + "dx"] = x.grad
        blob[pre + "dA"] = m.lora_down.weight.grad
        blob[pre + "dB"] = m.lora_up.weight.grad
    meta = {"cases": json.dumps(LINEAR_CASES)}
    save_file({k: v.contiguous() for k, v in blob.items()}, os.path.join(OUT, "linear_cases.safetensors"), meta)


def golden_tiny_unet(ref):
    """BASELINE config 1: inject r=4 into the tiny UNet, 32x32 latents, batch 2, fwd+bwd on CPU."""
    from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
    torch.manual_seed(0)
    unet = UNet2DConditionModel.tiny()
    unet.requires_grad_(False)
    torch.manual_seed(1)
    params, names = ref.inject_trainable_lora(unet, r=4)
    g = torch.Generator().manual_seed(2)
    layers = [m for m in unet.modules() if type(m).__name__ == "LoraInjectedLinear"]
    with torch.no_grad():
        for m in layers:
            m.lora_up.weight.copy_(torch.randn(m.lora_up.weight.shape, generator=g) * 0.02)
    lat = torch.randn(2, 4, 32, 32, generator=g)
    t = torch.randint(0, 1000, (2,), generator=g)
    ctx = torch.randn(2, 77, 32, generator=g)
    target = torch.randn(2, 4, 32, 32, generator=g)
    pred = unet(lat, t, ctx)
    loss = torch.nn.functional.mse_loss(pred, target)
    loss.backward()
    blob = {"latents": lat, "timesteps": t, "context": ctx, "target": target, "pred": pred.detach(),
            "loss": loss.detach().reshape(1)}
    paths = [n for n, m in unet.named_modules() if type(m).__name__ == "LoraInjectedLinear"]
    for i, m in enumerate(layers):
        blob[f"up{i}"] = m.lora_up.weight.detach()
        blob[f"down{i}"] = m.lora_down.weight.detach()
        blob[f"gup{i}"] = m.lora_up.weight.grad
        blob[f"gdown{i}"] = m.lora_down.weight.grad
    base_digest = hashlib.sha256()
    for k, v in unet.state_dict().items():
        if "lora_" not in k:
            base_digest.update(v.numpy().tobytes())
    meta = {"names": json.dumps(names), "paths": json.dumps(paths), "base_sha": base_digest.hexdigest(),
            "n_param_groups": str(len(params))}
    save_file({k: v.contiguous() for k, v in blob.items()}, os.path.join(OUT, "tiny_unet_step.safetensors"), meta)
    # .pt save format of the reference for the same model (fp16 list, up/down alternating)
    ref.save_lora_weight(unet, os.path.join(OUT, "tiny_unet_lora.pt"))
    ref.save_safeloras({"unet": (unet, ref.DEFAULT_TARGET_REPLACE)}, os.path.join(OUT, "tiny_unet_lora.safetensors"))


def golden_safeloras_fixture(ref):
    """A small slice of the reference's shipped example_loras/lora_disney.safetensors, re-saved
    through safetensors with the same key/metadata scheme, plus what the reference's
    parse_safeloras makes of it."""
    src = safe_open(os.path.join(REF, "example_loras", "lora_disney.safetensors"), "pt", "cpu")
    md = src.metadata()
    keep_t, keep_m = {}, {"unet": md["unet"], "text_encoder": md["text_encoder"]}
    for name, n in (("unet", 3), ("text_encoder", 2)):
        for i in range(n):
            for d in ("up", "down"):
                keep_t[f"{name}:{i}:{d}"] = src.get_tensor(f"{name}:{i}:{d}")
            keep_m[f"{name}:{i}:rank"] = md[f"{name}:{i}:rank"]
    for tok in ("<s1>", "<s2>"):
        keep_t[tok] = src.get_tensor(tok)
        keep_m[tok] = md[tok]
    path = os.path.join(OUT, "disney_slice.safetensors")
    save_file(keep_t, path, keep_m)
    f = safe_open(path, "pt", "cpu")
    parsed = ref.parse_safeloras(f)
    embeds = ref.parse_safeloras_embeds(f)
    digest = {
        "full_file": {"n_tensors": len(list(src.keys())), "n_metadata": len(md),
                      "unet_pairs": sum(1 for k in src.keys() if k.startswith("unet:") and k.endswith(":up")),
                      "text_pairs": sum(1 for k in src.keys() if k.startswith("text_encoder:") and k.endswith(":up")),
                      "unet_params": sum(src.get_tensor(k).numel() for k in src.keys() if k.startswith("unet:"))},
        "parsed": {name: {"ranks": ranks, "targets": sorted(targets),
                          "shapes": [list(w.shape) for w in weights],
                          "sha": [sha(w) for w in weights]}
                   for name, (weights, ranks, targets) in parsed.items()},
        "embeds": {k: sha(v) for k, v in embeds.items()},
    }
    with open(os.path.join(OUT, "disney_slice.json"), "w") as fh:
        json.dump(digest, fh, indent=1, sort_keys=True)


def golden_merge(ref):
    """weight_apply_lora (lora.py:410-424) on a stand-in CrossAttention, fp32 and fp16 LoRA."""
    class CrossAttention(nn.Module):
        def __init__(self):
            super().__init__()
            self.to_q = nn.Linear(64, 96, bias=False)
            self.to_out = nn.ModuleList([nn.Linear(96, 64)])

    blob = {}
    for tag, ldt in (("f32", torch.float32), ("f16", torch.float16)):
        torch.manual_seed(7)
        m = CrossAttention()
        g = torch.Generator().manual_seed(8)
        loras = []
        for lin in (m.to_q, m.to_out[0]):
            loras.append((torch.randn(lin.out_features, 4, generator=g) * 0.05).to(ldt))
            loras.append((torch.randn(4, lin.in_features, generator=g) * 0.5).to(ldt))
        for i, t in enumerate(loras):
            blob[f"{tag}.lora{i}"] = t.clone()
        blob[f"{tag}.w_q"] = m.to_q.weight.detach().clone()
        blob[f"{tag}.w_o"] = m.to_out[0].weight.detach().clone()
        ref.weight_apply_lora(m, list(loras), target_replace_module={"CrossAttention"}, alpha=0.8)
        blob[f"{tag}.merged_q"] = m.to_q.weight.detach().clone()
        blob[f"{tag}.merged_o"] = m.to_out[0].weight.detach().clone()
    save_file({k: v.contiguous() for k, v in blob.items()}, os.path.join(OUT, "merge_cases.safetensors"),
              {"alpha": "0.8"})


def golden_svd_fixture():
    """Layout digest of the reference's shipped cli_svd output (example_loras/analog_svd_distill.text_encoder.pt)."""
    lst = torch.load(os.path.join(REF, "example_loras", "analog_svd_distill.text_encoder.pt"), map_location="cpu")
    digest = {"n": len(lst), "dtypes": sorted({str(t.dtype) for t in lst}),
              "shape_up": list(lst[0].shape), "shape_down": list(lst[1].shape),
              "max_abs_up": max(float(t.abs().max()) for t in lst[0::2]),
              "max_abs_down": max(float(t.abs().max()) for t in lst[1::2])}
    with open(os.path.join(OUT, "svd_fixture.json"), "w") as fh:
        json.dump(digest, fh, indent=1, sort_keys=True)


# -------------------------------------------------------------------------------------------------
# Rows A12 / N3 / N4: the reference's CLI modules EXECUTED (oracle/ref_exec.py stands in for their
# diffusers / fire / wandb imports; the arithmetic that runs is the reference's own, unmodified)
# -------------------------------------------------------------------------------------------------
class _StandIn:
    """Small pipeline stand-ins whose block classes carry the names the reference matches on."""

    class CrossAttention(nn.Module):
        def __init__(self, d, ctx):
            super().__init__()
            self.to_q = nn.Linear(d, d, bias=False)
            self.to_k = nn.Linear(ctx, d, bias=False)
            self.to_out = nn.ModuleList([nn.Linear(d, d)])

    class GEGLU(nn.Module):
        def __init__(self, d):
            super().__init__()
            self.proj = nn.Linear(d, 4 * d)

    class CLIPAttention(nn.Module):
        def __init__(self, d):
            super().__init__()
            self.q_proj = nn.Linear(d, d)
            self.out_proj = nn.Linear(d, d)

    class Pipe:
        def __init__(self, seed, delta_seed=None):
            torch.manual_seed(seed)
            self.unet = nn.Sequential(_StandIn.CrossAttention(48, 40), _StandIn.GEGLU(48))
            self.text_encoder = nn.Sequential(_StandIn.CLIPAttention(32))
            if delta_seed is not None:           # "tuned" = base + a rank-3 update + small noise, per Linear
                g = torch.Generator().manual_seed(delta_seed)
                with torch.no_grad():
                    for m in list(self.unet.modules()) + list(self.text_encoder.modules()):
                        if isinstance(m, nn.Linear):
                            N, K = m.weight.shape
                            m.weight += (torch.randn(N, 3, generator=g) * 0.05) @ (torch.randn(3, K, generator=g) * 0.3) \
                                + 1e-3 * torch.randn(N, K, generator=g)
            self.unet.half()
            self.text_encoder.half()

        def to(self, *_a, **_k):
            return self


def golden_svd_distill(tmp):
    """cli_svd.svd_distill (cli_svd.py:29-111) run on two stand-in fp16 pipelines, on the CPU."""
    from oracle import ref_exec
    pipes = {"base": _StandIn.Pipe(21), "tuned": _StandIn.Pipe(21, delta_seed=22)}
    cli, _ = ref_exec.load_cli("cli_svd", pipeline_factory=lambda path, **kw: pipes[path])
    out = os.path.join(tmp, "svd_distill.pt")
    cli.svd_distill("tuned", "base", rank=4, clamp_quantile=0.99, device="cpu", save_path=out)
    blob = {}
    for which, model_of in (("unet", lambda p: p.unet), ("text", lambda p: p.text_encoder)):
        targets = ["CrossAttention", "Attention", "GEGLU"] if which == "unet" else ["CLIPAttention"]
        for tag in ("base", "tuned"):
            for i, w in enumerate(cli.extract_linear_weights(model_of(pipes[tag]), targets)):
                blob[f"{which}.{tag}.{i}"] = w.detach().clone()
        path = out if which == "unet" else cli._text_lora_path(out)
        for i, t in enumerate(torch.load(path)):
            blob[f"{which}.out.{i}"] = t.detach().clone()
    save_file({k: v.contiguous() for k, v in blob.items()}, os.path.join(OUT, "svd_distill_case.safetensors"),
              {"rank": "4", "clamp_quantile": "0.99"})


def golden_lpl(tmp):
    """cli_lora_add.add(mode="lpl", with_text_lora=True) (cli_lora_add.py:17-71) on two fp16 `.pt` lists."""
    from oracle import ref_exec
    cli, _ = ref_exec.load_cli("cli_lora_add")
    g = torch.Generator().manual_seed(31)
    blob = {}
    paths = {}
    for tag in ("a", "b"):
        for kind, shapes in (("unet", [(48, 4), (4, 48), (192, 4), (4, 48), (48, 1), (1, 40)]),
                             ("text", [(32, 4), (4, 32), (32, 4), (4, 32)])):
            lst = [(torch.randn(s, generator=g) * 0.1).half() for s in shapes]
            p = os.path.join(tmp, f"{tag}.pt" if kind == "unet" else f"{tag}.text_encoder.pt")
            torch.save(lst, p)
            paths[(tag, kind)] = p
            for i, t in enumerate(lst):
                blob[f"{kind}.{tag}.{i}"] = t.clone()
    out = os.path.join(tmp, "mix.pt")
    cli.add(paths[("a", "unet")], paths[("b", "unet")], out, alpha=0.3, mode="lpl", with_text_lora=True)
    for kind, p in (("unet", out), ("text", cli._text_lora_path(out))):
        for i, t in enumerate(torch.load(p)):
            blob[f"{kind}.out.{i}"] = t.detach().clone()
    save_file({k: v.contiguous() for k, v in blob.items()}, os.path.join(OUT, "lpl_case.safetensors"), {"alpha": "0.3"})


def golden_train_inversion(tmp):
    """cli_lora_pti.train_inversion (cli_lora_pti.py:250-405) run for 6 steps on a stand-in text encoder: its own
    optimizer recipe (:651-657), LambdaLR linear decay stepped first (:293), clip_ti_decay norm decay and the restore
    of every non-placeholder row (:315-346).  Only `loss_step` (the UNet denoising loss, :196-247) is replaced, by a
    quadratic loss over the embedded tokens, so the gradient rows are dense and known."""
    from oracle import ref_exec
    cli, _ = ref_exec.load_cli("cli_lora_pti")
    V, D, T, B, steps = 60, 24, 9, 3, 6
    rows = [57, 59]
    torch.manual_seed(41)

    class Text(nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = nn.Embedding(V, D)

        def get_input_embeddings(self):
            return self.emb

    class Unet(nn.Module):
        pass

    text = Text()
    with torch.no_grad():
        text.emb.weight.mul_(0.3)
    g = torch.Generator().manual_seed(42)
    batches = []
    for _ in range(steps):
        ids = torch.randint(0, V - 3, (B, T), generator=g)
        ids[:, 2], ids[:, 5] = rows[0], rows[1]
        ids[0, 7] = rows[1]
        batches.append({"input_ids": ids, "w": torch.randn(B, T, D, generator=g), "y": torch.randn(B, T, generator=g)})

    def loss_step(batch, unet, vae, text_encoder, scheduler, mixed_precision=False, **_kw):
        e = text_encoder.get_input_embeddings()(batch["input_ids"])
        return ((e * batch["w"]).sum(-1) - batch["y"]).pow(2).mean()

    cli.loss_step = loss_step
    table0 = text.emb.weight.detach().clone()
    index_no_updates = torch.ones(V, dtype=torch.bool)
    index_no_updates[rows] = False
    lr, wd = 5e-3, 1e-2
    opt = torch.optim.AdamW(text.get_input_embeddings().parameters(), lr=lr, betas=(0.9, 0.999), eps=1e-08,
                            weight_decay=wd)
    total = 8
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: max(0.0, (total - s) / total))    # diffusers "linear", no warm-up
    cli.train_inversion(Unet(), None, text, batches, steps, None, index_no_updates, opt, 10 ** 9, rows, ["<a>", "<b>"],
                        tmp, None, sched, tmp, accum_iter=1, log_wandb=False, mixed_precision=False, clip_ti_decay=True)
    blob = {"table0": table0, "table_final": text.emb.weight.detach().clone(), "rows": torch.tensor(rows)}
    for i, b in enumerate(batches):
        for k, v in b.items():
            blob[f"b{i}.{k}"] = v
    save_file({k: v.contiguous() for k, v in blob.items()}, os.path.join(OUT, "train_inversion_case.safetensors"),
              {"lr": str(lr), "weight_decay": str(wd), "steps": str(steps), "sched_total": str(total)})


def main():
    import tempfile
    os.makedirs(OUT, exist_ok=True)
    ref = load_reference()
    golden_linear(ref)
    golden_tiny_unet(ref)
    golden_safeloras_fixture(ref)
    golden_merge(ref)
    golden_svd_fixture()
    with tempfile.TemporaryDirectory() as tmp:
        golden_svd_distill(tmp)
        golden_lpl(tmp)
        golden_train_inversion(tmp)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()

"""Training driver on the GPU: LoraTrainer's step (arena exchange, global-norm clip, AdamW,
LR schedule order) against the same loop written with plain torch machinery on the same modules
(train_lora_dreambooth.py:876-889), eager and replayed as one CUDA graph."""
import itertools
import os
import subprocess
import sys

import pytest
import torch

import diffusion_finetuning_b200 as L
from conftest import rel_err
from diffusion_finetuning_b200 import train as T
from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _unet():
    torch.manual_seed(0)
    unet = UNet2DConditionModel.tiny().to(DEV)
    unet.requires_grad_(False)
    torch.manual_seed(1)
    params, _ = L.inject_trainable_lora(unet, r=4)
    g = torch.Generator().manual_seed(2)
    with torch.no_grad():
        for m in unet.modules():
            if isinstance(m, L.LoraInjectedLinear):
                m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(DEV))
    return unet, params


def _batches(n):
    g = torch.Generator().manual_seed(3)
    return [{"latents": torch.randn(2, 4, 32, 32, generator=g).to(DEV), "ctx": torch.randn(2, 77, 32, generator=g).to(DEV),
             "noise": torch.randn(2, 4, 32, 32, generator=g).to(DEV), "timesteps": torch.randint(0, 1000, (2,), generator=g).to(DEV)}
            for _ in range(n)]


def _cfg(tmp, **kw):
    base = dict(learning_rate=1e-3, lr_scheduler="linear", lr_warmup_steps=2, max_train_steps=6, max_grad_norm=1.0,
                save_steps=0, output_dir=str(tmp))
    base.update(kw)
    return T.TrainConfig(**base)


def _state(unet):
    return [p.detach().clone() for m in unet.modules() if isinstance(m, L.LoraInjectedLinear)
            for p in (m.lora_up.weight, m.lora_down.weight)]


@pytest.mark.parametrize("graph,arena_opt", [(False, False), (True, False), (False, True)])
def test_trainer_equals_plain_torch_loop(graph, arena_opt, tmp_path):
    """arena_opt: clip + AdamW over the flat arenas in one C-ABI call (dp.ArenaAdamW) instead of
    arena.clip_grad_norm_ + torch's fused AdamW; the LR schedule drives it through `param_groups`."""
    from diffusion_finetuning_b200 import dp
    data = _batches(6)
    unet, _ = _unet()
    tr = T.LoraTrainer(unet, _cfg(tmp_path, use_cuda_graph=graph, arena_optimizer=arena_opt))
    assert tr.graph_ok == graph and isinstance(tr.optimizer, dp.ArenaAdamW) == arena_opt
    losses = [tr.train_step(b).item() for b in data]
    # the reference's loop with stock pieces: autograd .grad, clip_grad_norm_, AdamW, LambdaLR
    ref_unet, params = _unet()
    opt = torch.optim.AdamW(itertools.chain(*params), lr=1e-3, betas=(0.9, 0.999), weight_decay=1e-2, eps=1e-8)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, T._schedule_factor("linear", 2, 6))
    ns = T.DDPMNoiseSchedule()
    ref_losses = []
    for b in data:
        noisy = ns.add_noise(b["latents"], b["noise"], b["timesteps"])
        pred = ref_unet(noisy, b["timesteps"], b["ctx"])
        loss = torch.nn.functional.mse_loss(pred.float(), b["noise"].float(), reduction="mean")
        loss.backward()
        torch.nn.utils.clip_grad_norm_([p for p in ref_unet.parameters() if p.requires_grad], 1.0)
        opt.step()
        sched.step()
        opt.zero_grad()
        ref_losses.append(loss.item())
    assert losses == pytest.approx(ref_losses, rel=2e-4)
    # AdamW's m/sqrt(v) turns last-bit gradient differences (atomic order, cuDNN algorithm choice) into
    # O(lr) parameter differences where a gradient is ~0: compare tensor-wise in norm, not element-wise
    for a, b in zip(_state(unet), _state(ref_unet)):
        assert rel_err(a, b) < 1e-2
    assert tr.lr_scheduler.get_last_lr() == pytest.approx(sched.get_last_lr())
    if graph:
        assert tr._graphed is not None and tr._graphed.kernel_launches > 0


def test_cli_synthetic_job_writes_reference_artefacts(tmp_path):
    out = tmp_path / "job"
    cmd = [sys.executable, "-m", "diffusion_finetuning_b200.train", "--synthetic", "--tiny", "--resolution", "256",
           "--train-batch-size", "2", "--max-train-steps", "6", "--save-steps", "4", "--lr-warmup-steps", "0",
           "--mixed-precision", "bf16", "--output-dir", str(out)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    names = sorted(os.listdir(out))
    assert names == ["lora_weight.pt", "lora_weight.safetensors", "lora_weight_e0_s4.pt"]
    assert "samples/s" in r.stdout
    loras = torch.load(out / "lora_weight.pt")
    assert len(loras) % 2 == 0 and all(t.dtype == torch.float16 for t in loras)
    assert any(t.abs().sum() > 0 for t in loras[::2])            # lora_up left zero => nothing trained


def test_config4_joint_unet_clip_and_placeholder_rows(tmp_path):
    """BASELINE config 4 in small: UNet LoRA + CLIP text-encoder LoRA + placeholder-token rows in ONE
    LoraTrainer (cli_lora_pti.perform_tuning with continue_inversion, :408-489, 699-751) against the same loop
    in plain torch.  The reference's loop trains every embedding row that occurs in a prompt (dense gradient, no
    restore in this phase); this package restricts training to the placeholder rows (ti.py), so the plain-torch
    side zeroes the other rows' gradient before the clip and restores them after the step."""
    from transformers import CLIPTextConfig, CLIPTextModel
    from diffusion_finetuning_b200 import ti
    V, rows = 200, [198, 199]
    ccfg = CLIPTextConfig(vocab_size=V, hidden_size=32, intermediate_size=64, num_hidden_layers=2,
                          num_attention_heads=4, max_position_embeddings=77)

    def build():
        unet, unet_params = _unet()
        torch.manual_seed(5)
        clip = CLIPTextModel(ccfg).to(DEV)
        clip.requires_grad_(False)
        torch.manual_seed(6)
        clip_params, _ = L.inject_trainable_lora(clip, {"CLIPAttention"}, r=4)
        g = torch.Generator().manual_seed(7)
        with torch.no_grad():
            for m in clip.modules():
                if isinstance(m, L.LoraInjectedLinear):
                    m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(DEV))
        return unet, unet_params, clip, clip_params

    g = torch.Generator().manual_seed(8)
    data = []
    for _ in range(4):
        ids = torch.randint(0, V - 2, (2, 77), generator=g)
        ids[:, 4], ids[:, 5] = rows[0], rows[1]
        data.append({"latents": torch.randn(2, 4, 32, 32, generator=g).to(DEV), "input_ids": ids.to(DEV),
                     "noise": torch.randn(2, 4, 32, 32, generator=g).to(DEV),
                     "timesteps": torch.randint(0, 1000, (2,), generator=g).to(DEV)})
    kw = dict(learning_rate=1e-3, learning_rate_text=5e-4, learning_rate_ti=2e-3, lr_scheduler="linear",
              lr_warmup_steps=0, max_train_steps=4, max_grad_norm=1.0, train_text_encoder=True, scheduler_first=True)
    unet, _, clip, _ = build()
    emb = clip.get_input_embeddings()
    table0 = emb.weight.detach().clone()
    pr = ti.PlaceholderTokenRows(emb, rows).install()
    tr = T.LoraTrainer(unet, _cfg(tmp_path, **kw), text_encoder=clip, placeholder_rows=pr)
    assert not tr.graph_ok and len(tr.lr_scheduler.get_last_lr()) == 3
    losses = [tr.train_step(b).item() for b in data]

    r_unet, r_up, r_clip, r_cp = build()
    r_emb = r_clip.get_input_embeddings()
    r_emb.weight.requires_grad_(True)
    orig = r_emb.weight.detach().clone()
    keep = torch.ones(V, dtype=torch.bool, device=DEV)
    keep[rows] = False
    opt = torch.optim.AdamW([{"params": itertools.chain(*r_up), "lr": 1e-3}, {"params": itertools.chain(*r_cp), "lr": 5e-4},
                             {"params": [r_emb.weight], "lr": 2e-3}], weight_decay=1e-2)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, T._schedule_factor("linear", 0, 4))
    ns = T.DDPMNoiseSchedule()
    ref_losses = []
    for b in data:
        sched.step()                                       # cli_lora_pti.py:434
        opt.zero_grad()
        noisy = ns.add_noise(b["latents"], b["noise"], b["timesteps"])
        pred = r_unet(noisy, b["timesteps"], r_clip(b["input_ids"])[0])
        loss = torch.nn.functional.mse_loss(pred.float(), b["noise"].float(), reduction="mean")
        loss.backward()
        r_emb.weight.grad[keep] = 0
        torch.nn.utils.clip_grad_norm_(itertools.chain(r_unet.parameters(), r_clip.parameters()), 1.0)   # :448-450
        opt.step()
        with torch.no_grad():
            r_emb.weight[keep] = orig[keep]
        ref_losses.append(loss.item())
    assert losses == pytest.approx(ref_losses, rel=5e-4)
    # (AdamW's m/sqrt(v) turns last-bit gradient differences — fp32 atomics in the reductions land in a different
    # order every run — into O(lr) moves where a gradient is ~0; measured 2e-3..4e-3 run to run)
    for a, b in zip(_state(unet) + _state(clip), _state(r_unet) + _state(r_clip)):
        assert rel_err(a, b) < 1e-2
    assert rel_err(emb.weight[rows], r_emb.weight[rows]) < 1e-4
    assert rel_err(emb.weight[rows] - table0[rows], r_emb.weight[rows] - orig[rows]) < 2e-2      # the UPDATE agrees
    assert torch.equal(emb.weight[keep], table0[keep])
    assert tr.lr_scheduler.get_last_lr() == pytest.approx(sched.get_last_lr())


def test_inversion_phase_through_the_trainer(tmp_path):
    """cli_lora_pti.train_inversion (:250-405) in small: frozen UNet WITHOUT LoRA, frozen CLIP, only the
    placeholder rows train (dense AdamW + norm decay + restore in the reference; sparse rows here), LR schedule
    stepped first (:293), lambda = min(1, 100·lr) following the schedule."""
    from transformers import CLIPTextConfig, CLIPTextModel
    from diffusion_finetuning_b200 import ti
    from oracle import lora_oracle as O
    V, rows = 200, [198, 199]
    ccfg = CLIPTextConfig(vocab_size=V, hidden_size=32, intermediate_size=64, num_hidden_layers=2,
                          num_attention_heads=4, max_position_embeddings=77)

    def build():
        torch.manual_seed(0)
        unet = UNet2DConditionModel.tiny().to(DEV)
        unet.requires_grad_(False)
        torch.manual_seed(5)
        clip = CLIPTextModel(ccfg).to(DEV)
        clip.requires_grad_(False)
        with torch.no_grad():
            clip.get_input_embeddings().weight[rows] *= 8.0       # norms away from 0.4: the decay matters
        return unet, clip

    g = torch.Generator().manual_seed(8)
    data = []
    for _ in range(5):
        ids = torch.randint(0, V - 2, (2, 77), generator=g)
        ids[:, 4], ids[:, 5] = rows[0], rows[1]
        data.append({"latents": torch.randn(2, 4, 32, 32, generator=g).to(DEV), "input_ids": ids.to(DEV),
                     "noise": torch.randn(2, 4, 32, 32, generator=g).to(DEV),
                     "timesteps": torch.randint(0, 1000, (2,), generator=g).to(DEV)})
    unet, clip = build()
    emb = clip.get_input_embeddings()
    table0 = emb.weight.detach().clone()
    pr = ti.PlaceholderTokenRows(emb, rows).install()
    cfg = _cfg(tmp_path, learning_rate=5e-3, lr_scheduler="linear", lr_warmup_steps=0, max_train_steps=5,
               max_grad_norm=0.0, adam_weight_decay=0.0, scheduler_first=True, clip_ti_decay=True)
    tr = T.LoraTrainer(unet, cfg, text_encoder=clip, placeholder_rows=pr)
    assert tr.optimizer is None and tr.ti_opt is not None
    losses = [tr.train_step(b).item() for b in data]

    r_unet, r_clip = build()
    r_emb = r_clip.get_input_embeddings()
    r_emb.weight.requires_grad_(True)
    orig = r_emb.weight.data.clone()                              # cli_lora_pti.py:280
    no_upd = torch.ones(V, dtype=torch.bool, device=DEV)
    no_upd[rows] = False
    opt = O.ti_make_optimizer(r_emb, lr=5e-3, weight_decay=0.0)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, T._schedule_factor("linear", 0, 5))
    ns = T.DDPMNoiseSchedule()
    ref_losses = []
    for b in data:
        sched.step()
        noisy = ns.add_noise(b["latents"], b["noise"], b["timesteps"])
        pred = r_unet(noisy, b["timesteps"], r_clip(b["input_ids"])[0])
        loss = torch.nn.functional.mse_loss(pred.float(), b["noise"].float(), reduction="mean")
        loss.backward()
        opt.step()
        opt.zero_grad()
        O.ti_post_step(r_emb, orig, no_upd, sched.get_last_lr()[0], clip_ti_decay=True)
        ref_losses.append(loss.item())
    assert losses == pytest.approx(ref_losses, rel=5e-4)
    assert rel_err(emb.weight[rows], r_emb.weight[rows]) < 1e-4
    assert torch.equal(emb.weight[no_upd], table0[no_upd])
    tr.save_final()                                               # nothing LoRA to write in this phase
    assert not any(n.startswith("lora_weight") for n in os.listdir(tmp_path))


def test_arena_optimizer_under_graph_needs_constant_lr_and_matches_eager(tmp_path):
    data = _batches(5)
    with pytest.raises(ValueError):
        T.LoraTrainer(_unet()[0], _cfg(tmp_path, use_cuda_graph=True, arena_optimizer=True, lr_scheduler="linear"))
    runs = []
    for graph in (False, True):
        unet, _ = _unet()
        tr = T.LoraTrainer(unet, _cfg(tmp_path, use_cuda_graph=graph, arena_optimizer=True, lr_scheduler="constant",
                                      lr_warmup_steps=0))
        losses = [tr.train_step(b).item() for b in data]
        runs.append((losses, _state(unet)))
    assert runs[0][0] == pytest.approx(runs[1][0], rel=2e-4)
    for a, b in zip(runs[0][1], runs[1][1]):
        assert rel_err(a, b) < 1e-2

"""Host logic of the training driver on CPU: LR schedules against torch's LambdaLR, the forward
diffusion restatement, the prior-preservation loss, and — gloo, world_size 2 — the loop itself (accumulation,
clip, AdamW groups, schedule, save cadence and names) against a single process on the concatenated batch.
The LoRA layers' forward is CUDA-only, so the loss is the oracle's restatement over the injected parameters."""
import math
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from diffusion_finetuning_b200 import train as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", ["constant", "constant_with_warmup", "linear", "cosine"])
def test_schedules_equal_torch_lambdalr(name):
    w = torch.nn.Parameter(torch.zeros(3))
    v = torch.nn.Parameter(torch.zeros(3))
    groups = lambda: [{"params": [w], "lr": 1e-4}, {"params": [v], "lr": 5e-5}]
    ours_opt, ref_opt = torch.optim.SGD(groups(), lr=1e-4), torch.optim.SGD(groups(), lr=1e-4)
    ours = T.get_scheduler(name, ours_opt, num_warmup_steps=4, num_training_steps=20)
    ref = torch.optim.lr_scheduler.LambdaLR(ref_opt, T._schedule_factor(name, 4, 20))
    for _ in range(24):
        assert ours.get_last_lr() == pytest.approx(ref.get_last_lr(), rel=1e-12)
        assert [g["lr"] for g in ours_opt.param_groups] == pytest.approx([g["lr"] for g in ref_opt.param_groups])
        ref_opt.step()
        ours.step()
        ref.step()
    # published diffusers.optimization definitions, spot values
    f = T._schedule_factor("linear", 4, 20)
    assert [f(0), f(2), f(4), f(12), f(20), f(25)] == [0.0, 0.5, 1.0, 0.5, 0.0, 0.0]
    c = T._schedule_factor("cosine", 0, 10)
    assert c(0) == 1.0 and abs(c(5) - 0.5) < 1e-12 and abs(c(10)) < 1e-12
    with pytest.raises(ValueError):
        T._schedule_factor("linear", 0, None)
    with pytest.raises(ValueError):
        T._schedule_factor("bogus", 0, 10)


def test_forward_diffusion_restatement():
    ns = T.DDPMNoiseSchedule()
    betas = torch.linspace(0.00085 ** 0.5, 0.012 ** 0.5, 1000) ** 2
    assert torch.allclose(ns.alphas_cumprod, torch.cumprod(1 - betas, 0))
    assert abs(ns.alphas_cumprod[0].item() - (1 - 0.00085)) < 1e-7 and ns.alphas_cumprod[-1] < 0.01
    g = torch.Generator().manual_seed(0)
    x0, eps = torch.randn(4, 4, 8, 8, generator=g), torch.randn(4, 4, 8, 8, generator=g)
    t = torch.tensor([0, 10, 500, 999])
    noisy = ns.add_noise(x0, eps, t)
    a = ns.alphas_cumprod[t].sqrt().view(-1, 1, 1, 1)
    s = (1 - ns.alphas_cumprod[t]).sqrt().view(-1, 1, 1, 1)
    assert torch.allclose(noisy, a * x0 + s * eps)
    v = ns.get_velocity(x0, eps, t)
    assert torch.allclose(a * noisy - s * v, x0, atol=1e-5)        # v-prediction identity
    assert ns.target(x0, eps, t) is eps
    assert torch.equal(T.DDPMNoiseSchedule(prediction_type="v_prediction").target(x0, eps, t), v)
    with pytest.raises(ValueError):
        T.DDPMNoiseSchedule(prediction_type="sample").target(x0, eps, t)


def test_prior_preservation_loss():
    g = torch.Generator().manual_seed(1)
    pred, tgt = torch.randn(6, 4, 8, 8, generator=g), torch.randn(6, 4, 8, 8, generator=g)
    plain = T.denoising_loss(pred, tgt)
    assert torch.allclose(plain, ((pred - tgt) ** 2).mean())
    pp = T.denoising_loss(pred, tgt, True, 0.7)
    want = ((pred[:3] - tgt[:3]) ** 2).mean([1, 2, 3]).mean() + 0.7 * ((pred[3:] - tgt[3:]) ** 2).mean()
    assert torch.allclose(pp, want)


# ------------------------------------------------------------------------------------------------
class CrossAttention(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.to_q = nn.Linear(d, d, bias=False)
        self.to_k = nn.Linear(d, d, bias=False)
        self.to_v = nn.Linear(d, d, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(d, d)])


def _model():
    import diffusion_finetuning_b200 as L
    torch.manual_seed(0)
    m = nn.Sequential(CrossAttention(16), CrossAttention(16))
    m.requires_grad_(False)
    L.inject_trainable_lora(m, {"CrossAttention"}, r=2)
    g = torch.Generator().manual_seed(3)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, L.LoraInjectedLinear):
                mod.lora_up.weight.copy_(torch.randn(16, 2, generator=g) * 0.05)
    return m


def _loss_fn(m):
    import diffusion_finetuning_b200 as L
    from oracle import lora_oracle as O

    def loss(batch):
        h = batch["x"]
        for mod in m.modules():
            if isinstance(mod, L.LoraInjectedLinear):
                h = torch.tanh(O.lora_linear_forward(h, mod.linear.weight, mod.linear.bias, mod.lora_down.weight,
                                                     mod.lora_up.weight, mod.scale))
        return ((h - batch["y"]) ** 2).mean() * 100
    return loss


def _cfg(out, **kw):
    base = dict(learning_rate=1e-2, lr_scheduler="linear", lr_warmup_steps=2, max_train_steps=8, max_grad_norm=0.05,
                gradient_accumulation_steps=2, save_steps=4, output_dir=out, output_format="both")
    base.update(kw)
    return T.TrainConfig(**base)


def _data():
    g = torch.Generator().manual_seed(7)
    return [{"x": torch.randn(8, 5, 16, generator=g), "y": torch.randn(8, 5, 16, generator=g)} for _ in range(8)]


def _lora_state(m):
    import diffusion_finetuning_b200 as L
    return [p.detach().clone() for mod in m.modules() if isinstance(mod, L.LoraInjectedLinear)
            for p in (mod.lora_up.weight, mod.lora_down.weight)]


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = _model()
        tr = T.LoraTrainer(m, _cfg(out), loss_fn=_loss_fn(m))
        data = _data()
        half = lambda b: {k: v[rank * 4:(rank + 1) * 4] for k, v in b.items()}
        steps = tr.fit(lambda epoch: (half(b) for b in data))
        assert steps == 8
        torch.save({"state": _lora_state(m), "lr": tr.lr_scheduler.get_last_lr()}, os.path.join(out, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_loop_dp2_equals_single_process(tmp_path):
    out = str(tmp_path / "dp")
    os.makedirs(out)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    r0, r1 = torch.load(os.path.join(out, "rank0.pt")), torch.load(os.path.join(out, "rank1.pt"))
    for a, b in zip(r0["state"], r1["state"]):
        assert torch.equal(a, b)                              # replicas stay identical
    # single process, whole batches
    solo = str(tmp_path / "solo")
    m = _model()
    tr = T.LoraTrainer(m, _cfg(solo), loss_fn=_loss_fn(m))
    seen = []
    tr.fit(lambda epoch: iter(_data()), on_log=lambda rec: seen.append((rec["step"], rec["lr"])))
    for a, b in zip(_lora_state(m), r0["state"]):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-7)
    assert r0["lr"] == pytest.approx(tr.lr_scheduler.get_last_lr())
    # schedule: one LambdaLR step per loop iteration (train_lora_dreambooth.py:886); its horizon is in iterations:
    # warm-up 2 x accumulation 2 = 4 of 8 x 2 = 16 (:737-742), of which the loop runs max_train_steps = 8 (:963)
    assert [round(lr / 1e-2, 6) for _, lr in seen] == [0.25, 0.5, 0.75, 1.0, round(11 / 12, 6), round(10 / 12, 6),
                                                       0.75, round(8 / 12, 6)]
    # save cadence and names (train_lora_dreambooth.py:893-958, 966-993): rank 0 only
    for d in (out, solo):
        names = sorted(n for n in os.listdir(d) if n.startswith("lora_weight"))
        assert names == ["lora_weight.pt", "lora_weight.safetensors", "lora_weight_e0_s4.pt", "lora_weight_e0_s8.pt"]
    final = torch.load(os.path.join(out, "lora_weight.pt"))
    assert len(final) == 16 and all(t.dtype == torch.float16 for t in final)
    for t, p in zip(final, r0["state"]):
        assert torch.equal(t, p.half())


def test_accumulation_equals_big_batch_and_clip_bites(tmp_path):
    """2 micro-batches of 4 with accumulation == 1 batch of 8 without (loss is a batch mean)."""
    data = _data()
    m1 = _model()
    t1 = T.LoraTrainer(m1, _cfg(str(tmp_path / "a"), gradient_accumulation_steps=2, save_steps=0, max_train_steps=8,
                                lr_scheduler="constant"), loss_fn=_loss_fn(m1))
    for b in data[:4]:
        for h in (0, 1):
            t1.train_step({k: v[h * 4:(h + 1) * 4] for k, v in b.items()})
            assert t1.sync_gradients == (h == 1)
    m2 = _model()
    t2 = T.LoraTrainer(m2, _cfg(str(tmp_path / "b"), gradient_accumulation_steps=1, save_steps=0, max_train_steps=8,
                                lr_scheduler="constant"), loss_fn=_loss_fn(m2))
    norms = []
    for b in data[:4]:
        t2.train_step(b)
        norms.append(torch.linalg.vector_norm(t2.arena.flat).item())
    for a, b in zip(_lora_state(m1), _lora_state(m2)):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-7)
    assert all(n <= 0.05 * (1 + 1e-5) for n in norms) and max(norms) > 0.049      # the clip was active
    with pytest.raises(ValueError):
        T.LoraTrainer(nn.Linear(4, 4), _cfg(str(tmp_path / "c")))                 # "No lora injected."


def test_polynomial_and_cosine_with_restarts_follow_the_diffusers_definitions():
    """diffusers.optimization: get_cosine_with_hard_restarts_schedule_with_warmup (num_cycles default 1) and
    get_polynomial_decay_schedule_with_warmup (lr_end 1e-7, power 1) — closed forms, spot values."""
    c = T._schedule_factor("cosine_with_restarts", 2, 12, num_cycles=2)
    assert c(0) == 0.0 and c(1) == 0.5 and c(2) == 1.0                       # warm-up ramp, then the first cycle starts
    assert abs(c(7) - 1.0) < 1e-12                                           # hard restart at the half-way point
    assert abs(c(4) - 0.5 * (1 + math.cos(math.pi * 0.4))) < 1e-12 and c(12) == 0.0 and c(20) == 0.0
    p = T._schedule_factor("polynomial", 2, 12, lr_init=1e-3)
    assert p(1) == 0.5 and abs(p(2) - 1.0) < 1e-12 and abs(p(7) - (0.5 * (1e-3 - 1e-7) + 1e-7) / 1e-3) < 1e-12
    assert abs(p(12) - 1e-4) < 1e-12 and p(13) == 1e-7 / 1e-3
    w = torch.nn.Parameter(torch.zeros(3))
    opt = torch.optim.SGD([w], lr=1e-3)
    sched = T.get_scheduler("polynomial", opt, num_warmup_steps=2, num_training_steps=12)
    for _ in range(12):
        sched.step()
    assert opt.param_groups[0]["lr"] == pytest.approx(1e-7)
    sched = T.get_scheduler("cosine_with_restarts", torch.optim.SGD([w], lr=1e-3), 0, 10)    # diffusers default: 1 cycle
    for _ in range(5):
        sched.step()
    assert sched.get_last_lr()[0] == pytest.approx(0.5e-3)
    with pytest.raises(ValueError):
        T.get_scheduler("polynomial", torch.optim.SGD([w], lr=1e-8), 0, 10)    # lr_end must be below the initial lr


class ResnetBlock2D(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv1 = nn.Conv2d(c, c, 3, padding=1)
        self.time_emb_proj = nn.Linear(8, c)


def test_checkpoints_of_an_extended_injection_hold_every_trained_layer(tmp_path):
    """inject_trainable_lora_extended adds layers under ResnetBlock2D; the savers must walk the set the model was
    injected under (it is recorded on the model), or the conv / time_emb_proj LoRAs are silently dropped."""
    import diffusion_finetuning_b200 as L
    torch.manual_seed(0)
    m = nn.Sequential(CrossAttention(16), ResnetBlock2D(16))
    m.requires_grad_(False)
    L.inject_trainable_lora_extended(m, r=2)
    n_injected = sum(isinstance(x, (L.LoraInjectedLinear, L.LoraInjectedConv2d)) for x in m.modules())
    assert n_injected == 6 and m._lora_injected_targets == L.UNET_EXTENDED_TARGET_REPLACE
    with torch.no_grad():
        for x in m.modules():
            if isinstance(x, (L.LoraInjectedLinear, L.LoraInjectedConv2d)):
                x.lora_up.weight.normal_(0, 0.05)
    tr = T.LoraTrainer(m, _cfg(str(tmp_path), gradient_accumulation_steps=1), loss_fn=lambda b: None)
    tr.save_final()
    saved = torch.load(tmp_path / "lora_weight.pt")
    assert len(saved) == 2 * n_injected
    assert any(t.dim() == 4 for t in saved)                   # the conv LoRA is in the file
    # round trip into a fresh model through the extended loader
    torch.manual_seed(0)
    fresh = nn.Sequential(CrossAttention(16), ResnetBlock2D(16))
    L.monkeypatch_or_replace_lora_extended(fresh, [t.clone() for t in saved],
                                           target_replace_module=L.UNET_EXTENDED_TARGET_REPLACE, r=2)
    want = [p for x in m.modules() if isinstance(x, (L.LoraInjectedLinear, L.LoraInjectedConv2d))
            for p in (x.lora_up.weight, x.lora_down.weight)]
    got = [p for x in fresh.modules() if isinstance(x, (L.LoraInjectedLinear, L.LoraInjectedConv2d))
           for p in (x.lora_up.weight, x.lora_down.weight)]
    assert len(got) == len(want) and all(torch.equal(a.half().float(), b.float()) or torch.equal(a.half(), b.half())
                                         for a, b in zip(want, got))
    from safetensors import safe_open
    f = safe_open(str(tmp_path / "lora_weight.safetensors"), "pt", "cpu")
    assert sum(k.endswith(":up") for k in f.keys()) == n_injected
    # a caller-named target set that misses layers is refused instead of writing a partial checkpoint
    tr.cfg.unet_targets = ("CrossAttention",)
    with pytest.raises(ValueError, match="not under the target classes"):
        tr.save_final()

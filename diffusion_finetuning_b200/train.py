"""Data-parallel LoRA training driver: the direct caller of the hot path.

Replaces `accelerate` + the loop of training_scripts/train_lora_dreambooth.py:811-964 (and
`cli_lora_pti.perform_tuning`, cli_lora_pti.py:408-489): one process per GPU under `torchrun`,
`dp.LoraGradArena` for the bucketed overlapped all-reduce, global-norm clip, AdamW with the reference's
parameter groups (UNet lr / text-encoder lr), the `get_scheduler` LR schedules, gradient accumulation,
prior-preservation loss, save cadence and file names of the reference, optional whole-step CUDA graph.

The order of operations inside one optimisation step is the reference's
(train_lora_dreambooth.py:876-889): backward → clip_grad_norm_ → optimizer.step() → lr_scheduler.step()
→ zero_grad.  (`perform_tuning` calls lr_scheduler.step() first, cli_lora_pti.py:434: `scheduler_first=True`.)

`diffusers` is not installed in this image, so the two pieces of it the loop needs are restated here from
their published definitions: `DDPMNoiseSchedule` (DDPMScheduler.add_noise / get_velocity with the SD-1.5
"scaled_linear" betas) and `get_scheduler` (diffusers.optimization: LambdaLR factors).

    torchrun --nproc-per-node 8 -m diffusion_finetuning_b200.train --synthetic --max-train-steps 100
"""
import argparse
import dataclasses
import math
import os
from typing import Callable, Dict, Iterable, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from . import dp
from . import lora as L


# --------------------------------------------------------------------------------------------
# LR schedules — diffusers.optimization.get_scheduler (requirements.txt: diffusers>=0.9.0), used at
# train_lora_dreambooth.py:741-747 and cli_lora_pti.py:659-664, 746-751.  All are LambdaLR factors.
# --------------------------------------------------------------------------------------------
def _schedule_factor(name: str, warmup: int, total: Optional[int], num_cycles: float = 0.5, power: float = 1.0,
                     lr_init: Optional[float] = None, lr_end: float = 1e-7):
    def ramp(step):
        return float(step) / float(max(1, warmup))

    if name == "constant":
        return lambda step: 1.0
    if name == "constant_with_warmup":
        return lambda step: ramp(step) if step < warmup else 1.0
    if total is None:
        raise ValueError(f"{name} requires `num_training_steps`, please provide that argument.")
    if name == "linear":
        return lambda step: ramp(step) if step < warmup else max(
            0.0, float(total - step) / float(max(1, total - warmup)))
    if name == "cosine":
        def f(step):
            if step < warmup:
                return ramp(step)
            progress = float(step - warmup) / float(max(1, total - warmup))
            return max(0.0, 0.5 * (1.0 + math.cos(math.pi * float(num_cycles) * 2.0 * progress)))
        return f
    if name == "cosine_with_restarts":
        # diffusers get_cosine_with_hard_restarts_schedule_with_warmup (default num_cycles = 1 there)
        def f(step):
            if step < warmup:
                return ramp(step)
            progress = float(step - warmup) / float(max(1, total - warmup))
            if progress >= 1.0:
                return 0.0
            return max(0.0, 0.5 * (1.0 + math.cos(math.pi * ((float(num_cycles) * progress) % 1.0))))
        return f
    if name == "polynomial":
        # diffusers get_polynomial_decay_schedule_with_warmup: the factor depends on the optimizer's initial lr
        if lr_init is None:
            raise ValueError("polynomial needs the initial learning rate (lr_init)")
        if not lr_init > lr_end:
            raise ValueError(f"lr_end ({lr_end}) must be be smaller than initial lr ({lr_init})")

        def f(step):
            if step < warmup:
                return ramp(step)
            if step > total:
                return lr_end / lr_init
            lr_range = lr_init - lr_end
            decay_steps = total - warmup
            pct_remaining = 1 - (step - warmup) / decay_steps
            return (lr_range * pct_remaining ** power + lr_end) / lr_init
        return f
    raise ValueError(f"{name} is not a valid SchedulerType")


class LambdaSchedule:
    """`torch.optim.lr_scheduler.LambdaLR` arithmetic for anything with `param_groups` (torch optimizers,
    `ti.PlaceholderAdamW`): lr_g = initial_lr_g · factor(step); construction applies factor(0).
    A group whose lr is a device tensor (capturable optimizers inside a CUDA graph) is updated in place."""

    def __init__(self, optimizer, factor: Callable[[int], float]):
        self.optimizer, self.factor, self.last_epoch = optimizer, factor, 0
        for g in optimizer.param_groups:
            if "initial_lr" not in g:
                lr = g["lr"]
                g["initial_lr"] = float(lr.item()) if torch.is_tensor(lr) else float(lr)
        self._apply()

    def _apply(self):
        f = self.factor(self.last_epoch)
        self._last = []
        for g in self.optimizer.param_groups:
            v = g["initial_lr"] * f
            if torch.is_tensor(g["lr"]):
                g["lr"].fill_(v)
            else:
                g["lr"] = v
            self._last.append(v)

    def step(self):
        self.last_epoch += 1
        self._apply()

    def get_last_lr(self) -> List[float]:
        return list(self._last)

    def state_dict(self):
        return {"last_epoch": self.last_epoch}

    def load_state_dict(self, sd):
        self.last_epoch = int(sd["last_epoch"])
        self._apply()


def get_scheduler(name, optimizer, num_warmup_steps: int = 0, num_training_steps: Optional[int] = None,
                  num_cycles: Optional[float] = None, power: float = 1.0):
    """`diffusers.optimization.get_scheduler`: every schedule name it accepts.  `num_cycles` defaults to diffusers'
    own per-schedule default (0.5 for cosine, 1 for cosine_with_restarts)."""
    name = str(name)
    if num_cycles is None:
        num_cycles = 1 if name == "cosine_with_restarts" else 0.5
    lr_init = None
    if name == "polynomial":        # diffusers reads optimizer.defaults["lr"]; here: the first group's initial lr
        g0 = optimizer.param_groups[0]
        lr0 = g0.get("initial_lr", g0["lr"])
        lr_init = float(lr0.item()) if torch.is_tensor(lr0) else float(lr0)
    return LambdaSchedule(optimizer, _schedule_factor(name, int(num_warmup_steps or 0), num_training_steps,
                                                      num_cycles, power, lr_init))


# --------------------------------------------------------------------------------------------
# forward diffusion — diffusers DDPMScheduler as configured for SD-1.5 (scheduler_config.json:
# beta_start 0.00085, beta_end 0.012, "scaled_linear", 1000 steps), train_lora_dreambooth.py:677-679, 833-840
# --------------------------------------------------------------------------------------------
class DDPMNoiseSchedule:
    def __init__(self, num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                 prediction_type="epsilon"):
        if beta_schedule == "scaled_linear":
            betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_train_timesteps, dtype=torch.float32) ** 2
        elif beta_schedule == "linear":
            betas = torch.linspace(beta_start, beta_end, num_train_timesteps, dtype=torch.float32)
        else:
            raise NotImplementedError(f"{beta_schedule} does is not implemented for DDPMNoiseSchedule")
        self.alphas_cumprod = torch.cumprod(1.0 - betas, dim=0)
        self._on_device = {}
        self.num_train_timesteps = num_train_timesteps
        self.prediction_type = prediction_type

    def _coefs(self, like, timesteps):
        key = (like.device, like.dtype)
        table = self._on_device.get(key)
        if table is None:       # one host->device copy per (device, dtype), not one per step
            table = self._on_device[key] = self.alphas_cumprod.to(device=like.device, dtype=like.dtype)
        ac = table[timesteps]
        shape = (-1,) + (1,) * (like.dim() - 1)
        return (ac ** 0.5).reshape(shape), ((1 - ac) ** 0.5).reshape(shape)

    def add_noise(self, original_samples, noise, timesteps):
        a, s = self._coefs(original_samples, timesteps)
        return a * original_samples + s * noise

    def get_velocity(self, sample, noise, timesteps):
        a, s = self._coefs(sample, timesteps)
        return a * noise - s * sample

    def target(self, latents, noise, timesteps):
        if self.prediction_type == "epsilon":
            return noise
        if self.prediction_type == "v_prediction":
            return self.get_velocity(latents, noise, timesteps)
        raise ValueError(f"Unknown prediction type {self.prediction_type}")


def denoising_loss(model_pred, target, with_prior_preservation=False, prior_loss_weight=1.0):
    """train_lora_dreambooth.py:853-874."""
    if not with_prior_preservation:
        return F.mse_loss(model_pred.float(), target.float(), reduction="mean")
    model_pred, model_pred_prior = torch.chunk(model_pred, 2, dim=0)
    target, target_prior = torch.chunk(target, 2, dim=0)
    loss = F.mse_loss(model_pred.float(), target.float(), reduction="none").mean([1, 2, 3]).mean()
    prior_loss = F.mse_loss(model_pred_prior.float(), target_prior.float(), reduction="mean")
    return loss + prior_loss_weight * prior_loss


class LossScaler:
    """Dynamic loss scaling for fp16 autocast over the flat gradient arena — `torch.cuda.amp.GradScaler` with its
    defaults (init 2**16, growth 2 every 2000 clean steps, backoff 0.5), which is what accelerate puts around
    `accelerator.backward` / `optimizer.step` for `--mixed_precision fp16`.  One pass over the arena unscales and
    detects overflow; the host reads the flag (GradScaler.step does the same) and skips the update on overflow."""

    def __init__(self, device, init_scale=65536.0, growth_factor=2.0, backoff_factor=0.5, growth_interval=2000):
        self.scale_value = float(init_scale)
        self.growth_factor, self.backoff_factor, self.growth_interval = growth_factor, backoff_factor, growth_interval
        self._good = 0
        self.skipped = 0

    def scale(self, loss):
        return loss * self.scale_value

    def unscale_(self, arena) -> bool:
        """Divide the (exchanged) gradients by the scale; False if any is non-finite (step must be skipped)."""
        flat = arena.flat
        flat.mul_(1.0 / self.scale_value)
        ok = bool(torch.isfinite(flat).all().item())
        if ok:
            self._good += 1
            if self._good >= self.growth_interval:
                self.scale_value *= self.growth_factor
                self._good = 0
        else:
            self.scale_value *= self.backoff_factor
            self._good = 0
            self.skipped += 1
        return ok


# --------------------------------------------------------------------------------------------
# the driver
# --------------------------------------------------------------------------------------------
@dataclasses.dataclass
class TrainConfig:
    """Names and defaults of train_lora_dreambooth.py's arguments (:60-420) that shape the loop."""
    learning_rate: float = 1e-4
    learning_rate_text: Optional[float] = None
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_weight_decay: float = 1e-2
    adam_epsilon: float = 1e-8
    max_grad_norm: float = 1.0
    lr_scheduler: str = "constant"
    lr_warmup_steps: int = 500
    max_train_steps: int = 1000
    gradient_accumulation_steps: int = 1
    with_prior_preservation: bool = False
    prior_loss_weight: float = 1.0
    mixed_precision: str = "no"                  # "no" | "fp16" | "bf16" (autocast; fp32 master LoRA weights)
    train_text_encoder: bool = False
    save_steps: int = 500
    output_dir: str = "text-inversion-model"
    output_format: str = "both"                  # "pt" | "safe" | "both"
    scheduler_first: bool = False                # cli_lora_pti.perform_tuning steps the LR schedule first (:434)
    use_cuda_graph: bool = False                 # replay fwd+bwd+exchange+clip+AdamW as one CUDA graph
    bucket_bytes: int = 1 << 20
    seed: Optional[int] = None
    learning_rate_ti: Optional[float] = None     # continue_inversion_lr / ti_lr (cli_lora_pti.py:715-722)
    clip_ti_decay: bool = False                  # perform_tuning has no norm decay; train_inversion has (:317)
    arena_optimizer: bool = False                # clip + AdamW over the flat arenas in one C-ABI call (dp.ArenaAdamW)
    unet_targets: Optional[tuple] = None         # class names the UNet was injected under (None = discover every layer)
    text_targets: Optional[tuple] = None


class LoraTrainer:
    """One replica of the data-parallel job.  `unet` (and `text_encoder`) are already injected
    (`inject_trainable_lora`), the base frozen.  `loss_fn(batch) -> loss` defaults to the reference's
    denoising loss over `batch = {"latents" | "pixel_values", "ctx" | "input_ids"}`."""

    def __init__(self, unet: nn.Module, cfg: TrainConfig, text_encoder: Optional[nn.Module] = None, vae=None,
                 noise_schedule: Optional[DDPMNoiseSchedule] = None, loss_fn: Optional[Callable] = None,
                 extra_params: Iterable[nn.Parameter] = (), process_group=None, placeholder_rows=None):
        """`placeholder_rows` (ti.PlaceholderTokenRows, installed on the text encoder's token embedding): the
        `continue_inversion` group of cli_lora_pti.py:713-728.  Its gradient rows ride in the arena (exchanged and
        clipped with the LoRA gradients); its update is `ti.PlaceholderAdamW` right after the LoRA optimizer."""
        self.cfg, self.unet, self.text_encoder, self.vae = cfg, unet, text_encoder, vae
        self.noise = noise_schedule or DDPMNoiseSchedule()
        self._own_loss = loss_fn is None
        self.loss_fn = loss_fn or self.denoising_step_loss
        self.rank = dist.get_rank(process_group) if dist.is_available() and dist.is_initialized() else 0
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        injected = (L.LoraInjectedLinear, L.LoraInjectedConv2d)

        def lora_params(model):
            out = []
            for m in model.modules():
                if isinstance(m, injected):
                    out += [m.lora_up.weight, m.lora_down.weight]      # order of inject_trainable_lora (lora.py:172-173)
            return out

        unet_params = lora_params(unet)
        if not unet_params and placeholder_rows is None:
            raise ValueError("No lora injected.")
        # no LoRA in the UNet + placeholder rows = the inversion phase (cli_lora_pti.train_inversion, :250-405):
        # only the rows train, through ti.PlaceholderAdamW (use clip_ti_decay=True, scheduler_first=True)
        groups = [{"params": unet_params, "lr": cfg.learning_rate}] if unet_params else []
        models: List[nn.Module] = [unet]
        if cfg.train_text_encoder:
            if text_encoder is None:
                raise ValueError("train_text_encoder=True needs the (LoRA-injected) text encoder")
            text_lr = cfg.learning_rate if cfg.learning_rate_text is None else cfg.learning_rate_text
            groups.append({"params": lora_params(text_encoder), "lr": text_lr})     # :659-667
            models.append(text_encoder)
        extra = list(extra_params)
        if extra:
            groups.append({"params": extra, "lr": cfg.learning_rate})
        holder = nn.ModuleList(models)          # a walk root for the arena; the models are not modified
        self.ti_opt = None
        arena_extra = list(extra)
        if placeholder_rows is not None:
            from . import ti
            arena_extra.append(placeholder_rows.rows)
            self.ti_opt = ti.PlaceholderAdamW(
                placeholder_rows, lr=cfg.learning_rate if cfg.learning_rate_ti is None else cfg.learning_rate_ti,
                betas=(cfg.adam_beta1, cfg.adam_beta2), eps=cfg.adam_epsilon, weight_decay=cfg.adam_weight_decay,
                clip_ti_decay=cfg.clip_ti_decay,
                capturable=bool(cfg.use_cuda_graph and cfg.gradient_accumulation_steps == 1
                                and cfg.mixed_precision != "fp16" and placeholder_rows.rows.is_cuda))
        self.arena = dp.LoraGradArena(holder, extra_params=arena_extra, bucket_bytes=cfg.bucket_bytes,
                                      process_group=process_group)
        self.arena.broadcast_params(0)
        self.device = self.arena.device
        on_cuda = self.device.type == "cuda"
        # fp16 runs dynamic loss scaling, whose skip-on-overflow is a host decision: not replayable
        self.graph_ok = bool(cfg.use_cuda_graph and on_cuda and cfg.gradient_accumulation_steps == 1
                             and cfg.mixed_precision != "fp16")
        opt_kw = dict(lr=cfg.learning_rate, betas=(cfg.adam_beta1, cfg.adam_beta2), weight_decay=cfg.adam_weight_decay,
                      eps=cfg.adam_epsilon)
        if on_cuda:
            opt_kw.update(fused=True, capturable=self.graph_ok)
        group_lrs = [float(g["lr"]) for g in groups]
        if self.graph_ok and not cfg.arena_optimizer:
            for g in groups:                    # device-resident lr: the schedule keeps working under replay
                g["lr"] = torch.tensor(float(g["lr"]), dtype=torch.float32, device=self.device)
            opt_kw["lr"] = torch.tensor(float(cfg.learning_rate), dtype=torch.float32, device=self.device)
        if cfg.arena_optimizer:
            # one kernel pair over the flat parameter/gradient arenas instead of clip_grad_norm_ + multi-tensor AdamW
            if len(groups) != 1 or self.ti_opt is not None or not on_cuda or any(
                    lr != float(cfg.learning_rate) for lr in group_lrs):
                raise ValueError("arena_optimizer needs ONE learning rate over all trainable parameters on a CUDA "
                                 "device (no separate text-encoder lr, no placeholder rows)")
            if self.graph_ok and cfg.lr_scheduler != "constant":
                raise ValueError("arena_optimizer takes lr as a launch argument: with use_cuda_graph the schedule must be "
                                 "'constant' (a captured launch replays the lr it was captured with)")
            self.optimizer = dp.ArenaAdamW(self.arena, lr=cfg.learning_rate, betas=(cfg.adam_beta1, cfg.adam_beta2),
                                           eps=cfg.adam_epsilon, weight_decay=cfg.adam_weight_decay,
                                           max_grad_norm=cfg.max_grad_norm if cfg.max_grad_norm and cfg.max_grad_norm > 0
                                           else None)
        else:
            self.optimizer = torch.optim.AdamW(groups, **opt_kw) if groups else None
        sched_target = self.optimizer
        if self.ti_opt is not None:             # one schedule over every group, as lr_scheduler_lora (:746-751)
            sched_target = argparse.Namespace(param_groups=(self.optimizer.param_groups if self.optimizer else [])
                                              + self.ti_opt.param_groups)
        # the schedule is stepped once per loop iteration (micro-batch), so its horizon is in iterations:
        # warm-up and total are multiplied by the accumulation factor (train_lora_dreambooth.py:737-742)
        accum = max(1, cfg.gradient_accumulation_steps)
        self.lr_scheduler = get_scheduler(cfg.lr_scheduler, sched_target, cfg.lr_warmup_steps * accum,
                                          cfg.max_train_steps * accum)
        self.global_step = 0          # counts loop iterations, as the reference's does (:891)
        self.last_save = 0
        self._micro = 0
        self._graphed = None
        self._gen = None
        if cfg.seed is not None:
            self._gen = torch.Generator(device=self.device).manual_seed(cfg.seed + self.rank)
        self.autocast_dtype = {"no": None, "fp16": torch.float16, "bf16": torch.bfloat16}[cfg.mixed_precision]
        # accelerate wraps fp16 runs in a GradScaler (accelerator.backward / optimizer.step,
        # train_lora_dreambooth.py:877-887): without it fp16 activation gradients underflow to zero silently
        self.scaler = LossScaler(self.device) if (cfg.mixed_precision == "fp16" and on_cuda) else None

    # ------------------------------------------------------------------ loss
    def prepare(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """Everything of train_lora_dreambooth.py:818-843 that happens before the UNet: latents (VAE or given),
        noise, timesteps, noisy latents, text conditioning (text encoder or given), target."""
        with torch.no_grad():
            if "latents" in batch:
                latents = batch["latents"]
            else:
                latents = self.vae.encode(batch["pixel_values"]).latent_dist.sample() * 0.18215
            noise = batch.get("noise")
            if noise is None:
                noise = torch.randn(latents.shape, device=latents.device, dtype=latents.dtype, generator=self._gen)
            t = batch.get("timesteps")
            if t is None:
                t = torch.randint(0, self.noise.num_train_timesteps, (latents.shape[0],), device=latents.device,
                                  generator=self._gen).long()
            out = {"noisy": self.noise.add_noise(latents, noise, t), "t": t, "target": self.noise.target(latents, noise, t)}
        if "ctx" in batch:
            out["ctx"] = batch["ctx"]
        else:
            out["input_ids"] = batch["input_ids"]
        return out

    def denoising_step_loss(self, b: Dict[str, torch.Tensor]) -> torch.Tensor:
        cfg = self.cfg
        with torch.autocast(self.device.type, dtype=self.autocast_dtype, enabled=self.autocast_dtype is not None):
            ctx = b["ctx"] if "ctx" in b else self.text_encoder(b["input_ids"])[0]
            pred = self.unet(b["noisy"], b["t"], ctx)
            pred = getattr(pred, "sample", pred)
        return denoising_loss(pred, b["target"], cfg.with_prior_preservation, cfg.prior_loss_weight)

    # ------------------------------------------------------------------ one iteration of the loop
    def _update(self):
        cfg = self.cfg
        self.arena.finish()
        if self.scaler is not None and not self.scaler.unscale_(self.arena):
            return                       # overflow: the step is skipped and the scale backs off (GradScaler.step)
        if cfg.max_grad_norm is not None and cfg.max_grad_norm > 0 and not cfg.arena_optimizer:
            self.arena.clip_grad_norm_(cfg.max_grad_norm)      # (ArenaAdamW clips inside its own launch)
        if self.optimizer is not None:
            self.optimizer.step()
        if self.ti_opt is not None:
            self.ti_opt.step()

    def _backward(self, loss):
        (loss if self.scaler is None else self.scaler.scale(loss)).backward()

    def _whole_step(self, b):
        self.arena.zero()
        loss = self.loss_fn(b)
        self._backward(loss)
        self._update()
        return loss

    def _capture(self, b):
        """Capture the whole step.  `dp.GraphedStep` runs the step eagerly a few times before capturing (lazy
        optimizer state, library workspaces); a training job must not see those updates, so parameters and
        optimizer state are put back IN PLACE afterwards (the graph holds their addresses)."""
        params = self.arena.params
        ti_keep = None
        if self.ti_opt is not None:
            tio = self.ti_opt
            table = tio.rows.embedding.weight
            ti_state = [tio.exp_avg, tio.exp_avg_sq, tio.step_dev, tio.rows.rows.data]
            ti_keep = ([t.clone() for t in ti_state], table.detach()[tio.rows.row_ids].clone())

        def restore_ti():
            if ti_keep is None:
                return
            with torch.no_grad():
                for t, k in zip(ti_state, ti_keep[0]):
                    t.copy_(k)
                table.data[tio.rows.row_ids] = ti_keep[1]

        if isinstance(self.optimizer, dp.ArenaAdamW):
            o = self.optimizer
            keep = [t.clone() for t in (o.flat_p, o.exp_avg, o.exp_avg_sq, o.scratch)]
            graphed = dp.GraphedStep(self._whole_step, b)
            with torch.no_grad():
                for t, k in zip((o.flat_p, o.exp_avg, o.exp_avg_sq, o.scratch), keep):
                    t.copy_(k)
            restore_ti()
            return graphed
        saved_p = [p.detach().clone() for p in params]
        opt_state = self.optimizer.state if self.optimizer is not None else {}
        saved_s = {id(p): {k: v.clone() for k, v in opt_state[p].items() if torch.is_tensor(v)}
                   for p in params if p in opt_state}
        graphed = dp.GraphedStep(self._whole_step, b)
        with torch.no_grad():
            for p, keep in zip(params, saved_p):
                p.copy_(keep)
                old = saved_s.get(id(p))
                for k, v in opt_state.get(p, {}).items():
                    if torch.is_tensor(v):
                        v.copy_(old[k]) if old is not None else v.zero_()
        restore_ti()
        return graphed

    def train_step(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        """One iteration of the reference loop body (one micro-batch).  Returns the detached loss."""
        cfg = self.cfg
        accum = max(1, cfg.gradient_accumulation_steps)
        b = self.prepare(batch) if self._own_loss else batch
        if cfg.scheduler_first:
            self.lr_scheduler.step()
        if self.graph_ok:
            if self._graphed is None:
                self._graphed = self._capture(b)
            loss = self._graphed(b)
        elif accum == 1:
            loss = self._whole_step(b)
        else:
            if self._micro == 0:
                self.arena.zero()
            last = self._micro == accum - 1
            if last:
                loss = self.loss_fn(b) / accum
                self._backward(loss)
                self._update()
            else:
                with self.arena.no_sync():
                    loss = self.loss_fn(b) / accum
                    self._backward(loss)
            self._micro = 0 if last else self._micro + 1
        if not cfg.scheduler_first:
            self.lr_scheduler.step()
        self.global_step += 1
        return loss.detach()

    @property
    def sync_gradients(self) -> bool:
        return self._micro == 0

    # ------------------------------------------------------------------ artefacts (rank 0)
    def _targets(self, model, configured, default):
        """The class names `model` was injected under: TrainConfig, else what `inject_trainable_lora*` recorded on
        the model, else the reference's default.  Every injected layer must be reachable from them — a layer the
        walk misses would be silently absent from the checkpoint."""
        targets = set(configured) if configured else set(getattr(model, "_lora_injected_targets", None) or default)
        injected = (L.LoraInjectedLinear, L.LoraInjectedConv2d)
        have = sum(1 for m in model.modules() if isinstance(m, injected))
        seen = {id(m) for _, _, m in L._find_modules(model, targets, search_class=list(injected))}
        if len(seen) != have:
            raise ValueError(f"{have - len(seen)} injected LoRA layer(s) are not under the target classes "
                             f"{sorted(targets)}: set TrainConfig.unet_targets / text_targets to the set the model "
                             "was injected with, or the checkpoint would miss them")
        return targets

    def _save(self, unet_path, text_path=None):
        if self.optimizer is None:              # inversion phase: nothing but the rows; the caller saves those
            return
        L.save_lora_weight(self.unet, unet_path,
                           target_replace_module=self._targets(self.unet, self.cfg.unet_targets, L.DEFAULT_TARGET_REPLACE))
        if self.cfg.train_text_encoder and text_path:
            L.save_lora_weight(self.text_encoder, text_path,
                               target_replace_module=self._targets(self.text_encoder, self.cfg.text_targets,
                                                                   L.TEXT_ENCODER_DEFAULT_TARGET_REPLACE))

    def maybe_save(self, epoch: int):
        """Cadence and names of train_lora_dreambooth.py:893-958."""
        cfg = self.cfg
        if not (self.sync_gradients and cfg.save_steps and self.global_step - self.last_save >= cfg.save_steps):
            return None
        name = f"{cfg.output_dir}/lora_weight_e{epoch}_s{self.global_step}.pt"
        if self.rank == 0:
            os.makedirs(cfg.output_dir, exist_ok=True)
            text = f"{cfg.output_dir}/lora_weight_e{epoch}_s{self.global_step}.text_encoder.pt"
            print(f"save weights {name}, {text}")
            self._save(name, text)
        self.last_save = self.global_step
        return name

    def save_final(self):
        """train_lora_dreambooth.py:966-993."""
        cfg = self.cfg
        if self.world > 1:
            dist.barrier()
        if self.rank != 0:
            return
        os.makedirs(cfg.output_dir, exist_ok=True)
        print("\n\nLora TRAINING DONE!\n\n")
        if cfg.output_format in ("pt", "both"):
            self._save(cfg.output_dir + "/lora_weight.pt", cfg.output_dir + "/lora_weight.text_encoder.pt")
        if cfg.output_format in ("safe", "both") and self.optimizer is not None:
            loras = {"unet": (self.unet, self._targets(self.unet, cfg.unet_targets, L.DEFAULT_TARGET_REPLACE))}
            if cfg.train_text_encoder:
                loras["text_encoder"] = (self.text_encoder, self._targets(self.text_encoder, cfg.text_targets,
                                                                          L.TEXT_ENCODER_DEFAULT_TARGET_REPLACE))
            L.save_safeloras(loras, cfg.output_dir + "/lora_weight.safetensors")

    def fit(self, batches_per_epoch: Callable[[int], Iterable[Dict[str, torch.Tensor]]], on_log: Optional[Callable] = None):
        """`batches_per_epoch(epoch)` yields this rank's shard of the epoch.  Stops at max_train_steps
        iterations (train_lora_dreambooth.py:963-964), then writes the final artefacts."""
        cfg = self.cfg
        epoch = 0
        while self.global_step < cfg.max_train_steps:
            seen = False
            for batch in batches_per_epoch(epoch):
                seen = True
                loss = self.train_step(batch)
                self.maybe_save(epoch)
                if on_log is not None:
                    on_log({"step": self.global_step, "loss": loss, "lr": self.lr_scheduler.get_last_lr()[0]})
                if self.global_step >= cfg.max_train_steps:
                    break
            if not seen:
                raise ValueError("empty epoch: batches_per_epoch yielded nothing")
            epoch += 1
        self.save_final()
        return self.global_step


# --------------------------------------------------------------------------------------------
# CLI: synthetic-data job on the SD-1.5-shaped stand-in UNet (no dataset / checkpoint access in this image)
# --------------------------------------------------------------------------------------------
def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--synthetic", action="store_true", help="random latents / text states, random-init SD-1.5-shaped UNet")
    ap.add_argument("--tiny", action="store_true", help="the tiny stand-in UNet (BASELINE config 1 shape)")
    ap.add_argument("--train-batch-size", type=int, default=8)
    ap.add_argument("--resolution", type=int, default=512)
    ap.add_argument("--lora-rank", type=int, default=4)
    ap.add_argument("--extended", action="store_true", help="inject_trainable_lora_extended (resnet convs too)")
    ap.add_argument("--fuse-siblings", action="store_true", help="q/k/v LoRA layers of a block as one launch; GEGLU projections with the activation in the GEMM epilogue")
    for f in dataclasses.fields(TrainConfig):
        flag = "--" + f.name.replace("_", "-")
        if f.type is bool or isinstance(f.default, bool):
            ap.add_argument(flag, action="store_true", default=f.default)
        else:
            typ = float if isinstance(f.default, float) or f.name == "learning_rate_text" else \
                int if isinstance(f.default, int) or f.name == "seed" else str
            ap.add_argument(flag, type=typ, default=f.default)
    a = ap.parse_args(argv)
    if not a.synthetic:
        raise SystemExit("only --synthetic jobs can run in this image (no datasets / checkpoints / diffusers); "
                         "real-data jobs build LoraTrainer around their own models and loaders")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("the LoRA layers are CUDA-only (no CPU fallback): run on a B200")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    from .fused import fuse_geglu, fuse_sibling_projections
    from .unet_sd15 import UNet2DConditionModel
    cfg = TrainConfig(**{f.name: getattr(a, f.name) for f in dataclasses.fields(TrainConfig)})
    if a.extended and cfg.mixed_precision == "no" and int(os.environ.get("RANK", "0")) == 0:
        print("note: --extended in fp32: the resnet LoRA convs run on cuDNN (the implicit-GEMM kernel is 16-bit); "
              "use --mixed-precision bf16 for the tensor-core path")
    torch.manual_seed(0)
    unet = (UNet2DConditionModel.tiny() if a.tiny else UNet2DConditionModel()).to(device)
    unet.requires_grad_(False)
    inject = L.inject_trainable_lora_extended if a.extended else L.inject_trainable_lora
    inject(unet, r=a.lora_rank)
    if a.fuse_siblings:
        fuse_sibling_projections(unet)
        fuse_geglu(unet)             # no-op unless GEGLU projections are injected (default / extended targets)
    tr = LoraTrainer(unet, cfg)
    res = a.resolution // 8
    ctx_dim = 32 if a.tiny else 768
    g = torch.Generator(device=device).manual_seed(1234 + tr.rank)

    def epoch(_):
        for _i in range(16):
            yield {"latents": torch.randn(a.train_batch_size, 4, res, res, device=device, generator=g),
                   "ctx": torch.randn(a.train_batch_size, 77, ctx_dim, device=device, generator=g)}

    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    marks = {}

    def log(rec):
        if rec["step"] == min(5, cfg.max_train_steps):
            t0.record()
            marks["s0"] = rec["step"]
        if tr.rank == 0 and (rec["step"] % 10 == 0 or rec["step"] == cfg.max_train_steps):
            print(f"step {rec['step']:5d}  loss {rec['loss'].item():.5f}  lr {rec['lr']:.3e}", flush=True)

    steps = tr.fit(epoch, on_log=log)
    t1.record()
    torch.cuda.synchronize(device)
    if tr.rank == 0 and steps > marks.get("s0", steps):
        ms = t0.elapsed_time(t1) / (steps - marks["s0"])
        print(f"{ms:.2f} ms/step, {a.train_batch_size * world / ms * 1e3:.1f} samples/s on {world} GPU(s)")
    # A captured step holds NCCL work: drop the graph FIRST, drain the device, let every rank arrive, then
    # destroy the process group and return — a communicator destroyed under a live graph blocks at exit.
    import gc
    tr._graphed = None
    tr.arena.detach()
    gc.collect()
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(device)
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())

"""Textual-inversion step beside the LoRA path.

Reference (lora_diffusion/cli_lora_pti.py): `train_inversion` (:250-405) trains the CLIP token embedding
with a DENSE AdamW over the whole `[vocab, D]` table (:651-657), then — every step — re-normalises the
placeholder rows towards norm 0.4 (:317-335) and copies EVERY other row back from a clone taken before
training (:280, :344-346).  Net effect of a step: only the placeholder rows change.  The reference pays
a 152 MB dense gradient, three dense optimizer passes and a 152 MB masked copy for it.

Here the placeholder rows are the trainable object:

  * `PlaceholderTokenRows` owns an fp32 master copy of the P rows as an `nn.Parameter` (so it joins
    `dp.LoraGradArena(extra_params=...)` and its all-reduce: §8(e) "all-reduce only the placeholder
    rows").  `install()` routes the embedding's forward through `_PlaceholderEmbedFn`: the gather is the
    stock `F.embedding` (CLIP stays stock PyTorch), the backward is ONE kernel that sums the output
    gradient of the tokens that are placeholders into `[P, D]` fp32 (`b200lora_ti_grad_rows`) — no dense
    `[vocab, D]` gradient exists.
  * `PlaceholderAdamW.step()` is ONE kernel (`b200lora_ti_step`): AdamW on the master rows, the norm
    decay with `lambda = min(1, 100·lr)`, write-through into the table in the table's dtype.  Every other
    row of the table is never written, which is what the reference's restore amounts to.

Scope note: in the tuning phase with `continue_inversion` (cli_lora_pti.py:713-728) the reference has no
restore step, so every token that occurs in a prompt drifts; this module keeps training restricted to the
placeholder rows there as well (documented deviation).

CUDA only — like the rest of the package there is no CPU fallback.
"""
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _native as nat


class _PlaceholderEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input_ids, table, rows, row_ids):
        nat.require_cuda(table, "PlaceholderTokenRows.embed")
        if input_ids.device != table.device:
            raise RuntimeError(f"input_ids are on {input_ids.device} but the embedding table is on {table.device}")
        out = F.embedding(input_ids, table)
        ctx.save_for_backward(input_ids, row_ids)
        ctx.shape = (rows.shape[0], table.shape[1])
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dy):
        input_ids, row_ids = ctx.saved_tensors
        P, D = ctx.shape
        if not ctx.needs_input_grad[2]:
            return None, None, None, None
        ids = input_ids.reshape(-1)
        if ids.dtype != torch.int64:
            ids = ids.long()
        ids = ids.contiguous()
        dy2 = dy.reshape(-1, D).contiguous()
        g = torch.zeros((P, D), dtype=torch.float32, device=dy.device)
        with torch.cuda.device(dy.device):
            rc = nat.lib().b200lora_ti_grad_rows(nat.ptr(dy2), nat.ptr(ids), ids.numel(), D, nat.ptr(row_ids), P,
                                                 nat.ptr(g), nat.dtype_code(dy2.dtype), nat.stream_ptr(dy.device))
            nat.check(rc, "b200lora_ti_grad_rows")
        return None, None, g, None


class PlaceholderTokenRows(nn.Module):
    """The trainable placeholder-token rows of a (frozen) embedding table.

    `embedding` is the text encoder's `get_input_embeddings()` (cli_lora_pti.py:280); it is NOT registered
    as a sub-module — the table stays where it is and stays frozen.  `rows` ([P, D] fp32) is the master copy
    the optimizer updates; the table holds the same values rounded to its dtype.
    """

    def __init__(self, embedding: nn.Embedding, placeholder_token_ids: Sequence[int]):
        super().__init__()
        ids = [int(i) for i in placeholder_token_ids]
        if not ids:
            raise ValueError("no placeholder token id given")
        if len(set(ids)) != len(ids):
            raise ValueError(f"duplicate placeholder token ids: {ids}")
        vocab = embedding.weight.shape[0]
        if min(ids) < 0 or max(ids) >= vocab:
            raise ValueError(f"placeholder token ids {ids} outside the embedding table (vocab {vocab})")
        if embedding.padding_idx is not None or embedding.max_norm is not None or embedding.sparse:
            raise ValueError("PlaceholderTokenRows supports a plain nn.Embedding (no padding_idx / max_norm / sparse)")
        object.__setattr__(self, "_embedding", embedding)
        dev = embedding.weight.device
        self.register_buffer("row_ids", torch.tensor(ids, dtype=torch.int64, device=dev), persistent=False)
        self.rows = nn.Parameter(embedding.weight.detach()[self.row_ids].float().clone())
        self._installed = False

    @property
    def embedding(self) -> nn.Embedding:
        return self._embedding

    def index_no_updates(self) -> torch.Tensor:
        """The reference's boolean mask of rows that must not move (cli_lora_pti.py `index_no_updates`)."""
        mask = torch.ones(self._embedding.weight.shape[0], dtype=torch.bool, device=self.row_ids.device)
        mask[self.row_ids] = False
        return mask

    def embed(self, input_ids: torch.Tensor) -> torch.Tensor:
        return _PlaceholderEmbedFn.apply(input_ids, self._embedding.weight, self.rows, self.row_ids)

    forward = embed

    def install(self):
        """Route `embedding(input_ids)` through this module (instance-level forward; class untouched)."""
        self._embedding.weight.requires_grad_(False)
        self._embedding.forward = self.embed
        self._installed = True
        return self

    def remove(self):
        if self._installed:
            del self._embedding.forward
            self._installed = False

    @torch.no_grad()
    def sync_from_table(self):
        """Re-read the master rows from the table (after `apply_learned_embed_in_clip` / a checkpoint load)."""
        self.rows.copy_(self._embedding.weight[self.row_ids].float())

    def learned_embeds(self):
        """{row id: row} in the table's dtype, as `save_all` stores them (lora.py:755-821)."""
        w = self._embedding.weight.detach()
        return {int(i): w[int(i)].clone() for i in self.row_ids.tolist()}


class PlaceholderAdamW:
    """AdamW + norm decay + table write-through for `PlaceholderTokenRows`, one kernel per step.

    Mirrors `optim.AdamW(text_encoder.get_input_embeddings().parameters(), lr=ti_lr, betas=(0.9, 0.999),
    eps=1e-8, weight_decay=weight_decay_ti)` (cli_lora_pti.py:651-657) followed by :315-346.  `param_groups`
    has the torch layout (one group) so `train.get_scheduler` and torch LR schedulers' arithmetic apply.
    """

    def __init__(self, rows: PlaceholderTokenRows, lr: float = 5e-4, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, clip_ti_decay: bool = True, target_norm: float = 0.4,
                 capturable: bool = False):
        """`capturable`: the learning rate and the step count live in device memory (`param_groups[0]["lr"]` is a
        one-element device tensor an LR schedule fills in place), so `step()` can sit in a captured CUDA graph."""
        self.rows = rows
        self.capturable = bool(capturable)
        lr_slot = torch.tensor(float(lr), dtype=torch.float32, device=rows.rows.device) if capturable else float(lr)
        self.param_groups = [{"params": [rows.rows], "lr": lr_slot, "initial_lr": float(lr), "betas": tuple(betas),
                              "eps": float(eps), "weight_decay": float(weight_decay)}]
        self.clip_ti_decay = bool(clip_ti_decay)
        self.target_norm = float(target_norm)
        p = rows.rows
        self.exp_avg = torch.zeros_like(p)
        self.exp_avg_sq = torch.zeros_like(p)
        self.norms = torch.zeros((2, p.shape[0]), dtype=torch.float32, device=p.device)   # [pre, post] per row
        self._step = 0
        self.step_dev = torch.zeros(p.shape[0], dtype=torch.float32, device=p.device) if capturable else None

    def zero_grad(self, set_to_none: bool = False):
        g = self.rows.rows.grad
        if g is not None:
            if set_to_none:
                self.rows.rows.grad = None
            else:
                g.zero_()

    @torch.no_grad()
    def step(self, lr: Optional[float] = None):
        p = self.rows.rows
        g = p.grad
        if g is None:
            return
        nat.require_cuda(p, "PlaceholderAdamW.step")
        grp = self.param_groups[0]
        lr_dev = None
        if self.capturable and lr is None:
            lr_dev = grp["lr"]                      # read by the kernel at run (replay) time
            lr = 0.0
            lam = 1.0 if self.clip_ti_decay else -1.0     # >= 0: the kernel derives min(1, 100·lr) itself
        else:
            lr = float(grp["lr"] if lr is None else lr)
            lam = min(1.0, 100.0 * lr) if self.clip_ti_decay else -1.0
        self._step += 1
        table = self.rows.embedding.weight
        if not table.is_contiguous() or not p.is_contiguous():
            raise RuntimeError("embedding table and master rows must be contiguous")
        g = g.contiguous()
        b1, b2 = grp["betas"]
        with torch.cuda.device(p.device):
            rc = nat.lib().b200lora_ti_step(nat.ptr(table), nat.dtype_code(table.dtype), p.shape[1],
                                            nat.ptr(self.rows.row_ids), p.shape[0], nat.ptr(p), nat.ptr(g),
                                            nat.ptr(self.exp_avg), nat.ptr(self.exp_avg_sq), self._step, lr,
                                            float(b1), float(b2), float(grp["eps"]), float(grp["weight_decay"]),
                                            float(lam), self.target_norm, nat.ptr(self.norms),
                                            nat.ptr(lr_dev), nat.ptr(self.step_dev) if lr_dev is not None else None,
                                            nat.stream_ptr(p.device))
            nat.check(rc, "b200lora_ti_step")
        # both tensors were written through raw pointers: tell autograd (version-keyed caches, saved tensors)
        torch.autograd.graph.increment_version(table)
        torch.autograd.graph.increment_version(p)

    def state_dict(self):
        step = int(self.step_dev[0].item()) if self.capturable else self._step
        return {"step": step, "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "lr": float(self.param_groups[0]["lr"])}

    def load_state_dict(self, sd):
        self._step = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        if self.capturable:
            self.step_dev.fill_(float(sd["step"]))
            self.param_groups[0]["lr"].fill_(float(sd["lr"]))
        else:
            self.param_groups[0]["lr"] = float(sd["lr"])

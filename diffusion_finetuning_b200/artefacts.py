"""Artefact I/O either side of the LoRA path: the on-disk formats are the reference's,
the arithmetic runs on the GPU.

  * `lora_add(...)`                 cli_lora_add.py:17-114  (`lpl` list mix; `upl` merge into a pipeline)
  * `convert(...)`                  cli_pt_to_safetensors.py:19-77  (.pt lists / TI dicts -> one .safetensors)
  * `svd_distill(...)`              cli_svd.py:29-111  (`lora_distill`: rank-r factors of tuned − base, whole model)
  * `save_lora_weight_streamed`     lora.py:201-213  — one cast launch + ONE device→host copy for all 2·L
                                    tensors instead of 2·L casts and 2·L copies
  * `save_safeloras_streamed`       lora.py:228-258  — same for the safetensors writer

`upl-ckpt-v2` additionally needs the diffusers→LDM checkpoint key remap (to_ckpt_v2.py:91-232), a table of
string renames for the WHOLE pipeline (VAE, CLIP, UNet) with no arithmetic in it: out of scope here
(DESIGN.md §8); the merge half of that mode is `lora_add(mode="upl")`.
"""
import json
import os
from typing import Callable, Dict, List, Optional, Sequence, Set, Tuple

import torch
import torch.nn as nn

from . import _native as nat
from . import lora as L


# --------------------------------------------------------------------------------------------
# the flat kernel
# --------------------------------------------------------------------------------------------
def _cuda_device(*tensors):
    for t in tensors:
        if t is not None and t.is_cuda:
            return t.device
    if not torch.cuda.is_available():
        raise nat.B200LoraError("the artefact arithmetic runs on the b200lora kernels: a CUDA device is required "
                                "(no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def flat_axpby(a: torch.Tensor, b: Optional[torch.Tensor] = None, alpha: float = 1.0, beta: float = 0.0,
               out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """round(alpha·a) (+ round(beta·b)) over flat device tensors, ATen rounding order (b200lora_flat_axpby)."""
    nat.require_cuda(a, "flat_axpby")
    a = a.contiguous().reshape(-1)
    if b is not None:
        b = b.contiguous().reshape(-1)
        if b.shape != a.shape or b.dtype != a.dtype or b.device != a.device:
            raise RuntimeError(f"flat_axpby: operands differ ({tuple(a.shape)} {a.dtype} vs {tuple(b.shape)} {b.dtype})")
    out = torch.empty(a.shape, dtype=out_dtype or a.dtype, device=a.device)
    with torch.cuda.device(a.device):
        rc = nat.lib().b200lora_flat_axpby(nat.ptr(a), nat.ptr(b), nat.ptr(out), a.numel(), float(alpha), float(beta),
                                           nat.dtype_code(a.dtype), nat.dtype_code(out.dtype), nat.stream_ptr(a.device))
    nat.check(rc, "b200lora_flat_axpby")
    return out


def _split_like(flat: torch.Tensor, like: Sequence[torch.Tensor]) -> List[torch.Tensor]:
    out, off = [], 0
    for t in like:
        n = t.numel()
        out.append(flat[off:off + n].view(t.shape).clone())     # own storage, as torch.save of the reference's list
        off += n
    return out


def mix_lora_lists(l1: Sequence[torch.Tensor], l2: Sequence[torch.Tensor], alpha: float) -> List[torch.Tensor]:
    """`alpha * x1 + (1 - alpha) * x2` for every tensor of two `.pt` LoRA lists (cli_lora_add.py:49-58), as one
    launch over the concatenated payload.  Like the reference's `zip`, stops at the shorter list."""
    n = min(len(l1), len(l2)) // 2 * 2
    l1, l2 = list(l1[:n]), list(l2[:n])
    if not l1:
        return []
    for x, y in zip(l1, l2):
        if x.shape != y.shape:
            raise RuntimeError(f"The size of tensor a {tuple(x.shape)} must match the size of tensor b {tuple(y.shape)}")
    dt = l1[0].dtype
    dev = _cuda_device(*l1)
    fa = torch.cat([t.detach().reshape(-1).to(dt) for t in l1]).to(dev, non_blocking=True)
    fb = torch.cat([t.detach().reshape(-1).to(dt) for t in l2]).to(dev, non_blocking=True)
    mixed = flat_axpby(fa, fb, alpha, 1 - alpha)
    home = l1[0].device
    return _split_like(mixed.to(home), l1)


# --------------------------------------------------------------------------------------------
# cli_lora_add
# --------------------------------------------------------------------------------------------
def _default_pipeline_loader(path):
    try:
        from diffusers import StableDiffusionPipeline
    except ImportError as exc:                  # the reference imports diffusers at module top (cli_lora_add.py:5)
        raise ImportError("mode 'upl' loads a diffusers pipeline; diffusers is not installed — pass "
                          "`load_pipeline=` returning an object with .unet / .text_encoder / .save_pretrained") from exc
    return StableDiffusionPipeline.from_pretrained(path)


def lora_add(path_1: str, path_2: str, output_path: str, alpha: float = 0.5, mode: str = "lpl",
             with_text_lora: bool = False, load_pipeline: Optional[Callable] = None):
    """cli_lora_add.add (cli_lora_add.py:17-110).

    lpl : two `.pt` LoRA lists -> their mix (and, with_text_lora, the `.text_encoder.pt` siblings when both exist).
    upl : `path_1` is a pipeline, `path_2` a `.pt` list; `W += alpha·up@down` for every target Linear on the
          merge kernel, then `save_pretrained(output_path)`.
    """
    print("Lora Add, mode " + mode)
    if mode == "lpl":
        jobs = [(path_1, path_2, "unet")]
        if with_text_lora:
            jobs.append((L._text_lora_path(path_1), L._text_lora_path(path_2), "text_encoder"))
        for p1, p2, which in jobs:
            print("Loading", p1, p2)
            if which == "text_encoder" and not (os.path.exists(p1) and os.path.exists(p2)):
                missing = p1 if not os.path.exists(p1) else p2
                print(f"No text encoder found in {missing}, skipping...")
                continue
            out_list = mix_lora_lists(torch.load(p1), torch.load(p2), alpha)
            dst = output_path if which == "unet" else L._text_lora_path(output_path)
            print(f"Saving merged {'UNET' if which == 'unet' else 'text encoder'} to", dst)
            torch.save(out_list, dst)
        return
    if mode in ("upl", "upl-ckpt-v2"):
        if mode == "upl-ckpt-v2":
            raise NotImplementedError("upl-ckpt-v2 = upl + the diffusers->LDM key remap (to_ckpt_v2.py:91-232), which "
                                      "is outside this package's scope; run mode='upl' and convert the saved pipeline")
        pipe = (load_pipeline or _default_pipeline_loader)(path_1)
        L.weight_apply_lora(pipe.unet, torch.load(path_2), alpha=alpha)
        if with_text_lora:
            L.weight_apply_lora(pipe.text_encoder, torch.load(L._text_lora_path(path_2)), alpha=alpha,
                                target_replace_module=["CLIPAttention"])
        pipe.save_pretrained(output_path)
        return pipe
    print("Unknown mode", mode)
    raise ValueError(f"Unknown mode {mode}")


# --------------------------------------------------------------------------------------------
# cli_svd (`lora_distill`)
# --------------------------------------------------------------------------------------------
SVD_UNET_TARGETS = ["CrossAttention", "Attention", "GEGLU"]       # cli_svd.py:46-48
SVD_TEXT_TARGETS = ["CLIPAttention"]                              # cli_svd.py:49


def extract_linear_weights(model, target_replace_module):
    """The `.weight` of every nn.Linear under the target classes, in discovery order (cli_svd.py:19-26)."""
    return [child.weight for _, _, child in L._find_modules(model, target_replace_module, search_class=[nn.Linear])]


def _default_fp16_pipeline_loader(path, device):
    try:
        from diffusers import StableDiffusionPipeline
    except ImportError as exc:                  # the reference imports diffusers at module top (cli_svd.py:2)
        raise ImportError("svd_distill loads two diffusers pipelines; diffusers is not installed — pass "
                          "`load_pipeline=` returning an object with .unet / .text_encoder") from exc
    return StableDiffusionPipeline.from_pretrained(path, torch_dtype=torch.float16).to(device)


def svd_distill_weights(tuned: Sequence[torch.Tensor], base: Sequence[torch.Tensor], rank: int = 4,
                        clamp_quantile: float = 0.99, device=None) -> List[torch.Tensor]:
    """`[U0, Vh0, U1, Vh1, ...]` (fp32, on `device`) for `zip(base, tuned)` — the loop of cli_svd.py:56-84 /
    :86-108.  The difference is taken in the WEIGHTS' dtype (`tuned - ori`, :59-63: fp16 pipelines, so ΔW is
    rounded to fp16 before it is widened), each ΔW goes through `ops.svd_distill_matrix` (cuSOLVER gesvd, economy
    size — the reference's full `U [N, N]` is never used — + the scale/clamp epilogue kernel)."""
    from . import ops
    dev = torch.device(device) if device is not None else _cuda_device(*tuned)
    out = []
    with torch.no_grad():
        for ori, new in zip(base, tuned):
            diff = new.detach().to(dev) - ori.detach().to(dev)        # same rounding on either device (one IEEE op)
            U, Vh = ops.svd_distill_matrix(diff, rank=rank, clamp_quantile=clamp_quantile)
            out += [U, Vh]
    return out


def svd_distill(target_model: str, base_model: str, rank: int = 4, clamp_quantile: float = 0.99,
                device: str = "cuda:0", save_path: str = "svd_distill.pt",
                load_pipeline: Optional[Callable] = None):
    """cli_svd.svd_distill (cli_svd.py:29-111): rank-`rank` LoRA factors of (tuned − base) for every Linear under
    {"CrossAttention", "Attention", "GEGLU"} of the UNet and {"CLIPAttention"} of the text encoder; saves the UNet
    list to `save_path` and the text-encoder list to `<save_path minus .pt>.text_encoder.pt` (:110-111).
    `load_pipeline(path, device)` replaces `StableDiffusionPipeline.from_pretrained(path, torch_dtype=float16)`."""
    load = load_pipeline or _default_fp16_pipeline_loader
    pipe_base = load(base_model, device)
    pipe_tuned = load(target_model, device)
    ori_unet = extract_linear_weights(pipe_base.unet, SVD_UNET_TARGETS)
    ori_clip = extract_linear_weights(pipe_base.text_encoder, SVD_TEXT_TARGETS)
    tuned_unet = extract_linear_weights(pipe_tuned.unet, SVD_UNET_TARGETS)
    tuned_clip = extract_linear_weights(pipe_tuned.text_encoder, SVD_TEXT_TARGETS)
    dev = torch.device(device) if torch.device(device).type == "cuda" else _cuda_device(*tuned_unet)
    uds_unet = svd_distill_weights(tuned_unet, ori_unet, rank, clamp_quantile, dev)
    uds_clip = svd_distill_weights(tuned_clip, ori_clip, rank, clamp_quantile, dev)
    home = torch.device(device)
    if home != dev:                              # the reference leaves the factors on `device`
        uds_unet = [t.to(home) for t in uds_unet]
        uds_clip = [t.to(home) for t in uds_clip]
    torch.save(uds_unet, save_path)
    torch.save(uds_clip, L._text_lora_path(save_path))
    return uds_unet, uds_clip


# --------------------------------------------------------------------------------------------
# cli_pt_to_safetensors
# --------------------------------------------------------------------------------------------
_TARGET_BY_NAME = {"unet": L.UNET_DEFAULT_TARGET_REPLACE, "text_encoder": L.TEXT_ENCODER_DEFAULT_TARGET_REPLACE}


def convert(*paths, outpath, overwrite=False, **settings):
    """cli_pt_to_safetensors.convert (:19-77): `.pt` LoRA lists and textual-inversion dicts -> one safetensors
    file.  The model name comes from the file name (`x.pt` -> unet, `x.text_encoder.pt` -> text_encoder);
    `**{"unet.rank": 8, "text_encoder.target_modules": [...]}` override the per-model settings."""
    if os.path.exists(outpath) and not overwrite:
        raise ValueError(f"Output path {outpath} already exists, and overwrite is not True")
    modelmap, embeds = {}, {}
    for path in paths:
        data = torch.load(path)
        if isinstance(data, dict):
            print(f"Loading textual inversion embeds {data.keys()} from {path}")
            embeds.update(data)
            continue
        parts = os.path.split(path)[1].split(".")
        name = parts[-2] if len(parts) > 2 else "unet"
        model_settings = {"target_modules": _TARGET_BY_NAME.get(name, L.DEFAULT_TARGET_REPLACE), "rank": 4}
        prefix = f"{name}."
        model_settings.update({k[len(prefix):]: v for k, v in settings.items() if k.startswith(prefix)})
        print(f"Loading Lora for {name} from {path} with settings {model_settings}")
        modelmap[name] = (path, model_settings["target_modules"], model_settings["rank"])
    L.convert_loras_to_safeloras_with_embeds(modelmap, embeds, outpath)


# --------------------------------------------------------------------------------------------
# streamed writers
# --------------------------------------------------------------------------------------------
def _gather_flat(weights: Sequence[torch.Tensor], out_dtype: torch.dtype):
    """All weights -> ONE pinned host buffer in `out_dtype`: one cast launch + one D2H copy.  When the weights
    are slices of one device buffer (dp.ArenaAdamW's parameter arena) no gather copy runs either."""
    dev = weights[0].device
    nat.require_cuda(weights[0], "streamed LoRA writer")
    dt = weights[0].dtype
    if any(w.dtype != dt or w.device != dev for w in weights):
        raise RuntimeError("streamed writer needs all LoRA weights on one device in one dtype")
    total = sum(w.numel() for w in weights)
    store = weights[0].untyped_storage().data_ptr()
    one_arena = all(w.is_contiguous() and w.untyped_storage().data_ptr() == store for w in weights)
    lo = min(w.storage_offset() for w in weights)
    hi = max(w.storage_offset() + w.numel() for w in weights)
    if one_arena and hi - lo <= 2 * total:
        # the weights are slices of ONE device buffer (dp.ArenaAdamW's parameter arena, any order): convert and
        # copy the covering span once, then cut the host buffer at the same offsets
        span = torch.empty(0, dtype=dt, device=dev).set_(weights[0].untyped_storage(), lo, (hi - lo,), (1,))
        cast = flat_axpby(span, None, 1.0, 0.0, out_dtype=out_dtype) if out_dtype != dt else span
        host = torch.empty(hi - lo, dtype=out_dtype, pin_memory=True)
        host.copy_(cast, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return [host[w.storage_offset() - lo:w.storage_offset() - lo + w.numel()].view(w.shape).clone() for w in weights]
    flat = torch.cat([w.detach().reshape(-1) for w in weights])
    cast = flat_axpby(flat.detach(), None, 1.0, 0.0, out_dtype=out_dtype) if out_dtype != dt else flat.detach()
    host = torch.empty(total, dtype=out_dtype, pin_memory=True)
    host.copy_(cast, non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()
    return _split_like(host, weights)


def save_lora_weight_streamed(model, path="./lora.pt", target_replace_module=L.DEFAULT_TARGET_REPLACE):
    """Same file as `save_lora_weight` (lora.py:201-213: `[up0, down0, up1, ...]`, fp16, CPU), produced with one
    cast launch and one device→host copy."""
    pairs = L.extract_lora_ups_down(model, target_replace_module=target_replace_module)
    weights = [w for up, down in pairs for w in (up.weight, down.weight)]
    torch.save(_gather_flat(weights, torch.float16), path)


def save_safeloras_streamed(modelmap: Dict[str, Tuple[nn.Module, Set[str]]], embeds: Optional[Dict[str, torch.Tensor]] = None,
                            outpath="./lora.safetensors"):
    """Same file as `save_safeloras_with_embeds` (lora.py:228-258; native dtype), one device→host copy per model."""
    weights, metadata = {}, {}
    for name, (model, targets) in modelmap.items():
        metadata[name] = json.dumps(list(targets))
        pairs = L.extract_lora_ups_down(model, targets)
        ws = [w for up, down in pairs for w in (up.weight, down.weight)]
        host = _gather_flat(ws, ws[0].dtype)
        for i, (up, down) in enumerate(pairs):
            rank = down.out_features if isinstance(down, nn.Linear) else down.out_channels
            metadata[f"{name}:{i}:rank"] = str(rank)
            weights[f"{name}:{i}:up"] = host[2 * i]
            weights[f"{name}:{i}:down"] = host[2 * i + 1]
    for token, tensor in (embeds or {}).items():
        metadata[token] = L.EMBED_FLAG
        weights[token] = tensor.detach().cpu()
    print(f"Saving weights to {outpath}")
    L.safe_save(weights, outpath, metadata)

"""B200-native implementation of lora_diffusion's LoRA-adapted-layer hot path.

`from diffusion_finetuning_b200 import *` gives the reference's `lora_diffusion` surface
(LoraInjectedLinear, inject_trainable_lora, monkeypatch_*, tune_lora_scale, save/load ...);
the top-level `lora_diffusion` package in this repo re-exports it under the reference's name.
"""
from .lora import *  # noqa: F401,F403
from .lora import (_find_children, _find_modules, _find_modules_old, _find_modules_v2,  # noqa: F401
                   _text_lora_path, _ti_lora_path)
from . import ops  # noqa: F401

__version__ = "0.1.0"
from .fused import fuse_sibling_projections, unfuse_sibling_projections, fuse_geglu, unfuse_geglu  # noqa: F401,E402

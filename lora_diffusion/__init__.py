"""Drop-in alias: `from lora_diffusion import *` resolves to the B200-native implementation
(the reference's package re-exports lora.py the same way, lora_diffusion/__init__.py:1-3)."""
from diffusion_finetuning_b200 import *  # noqa: F401,F403
from diffusion_finetuning_b200.lora import (_find_children, _find_modules, _find_modules_old,  # noqa: F401
                                            _find_modules_v2, _text_lora_path, _ti_lora_path)

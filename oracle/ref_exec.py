"""Execute the reference's CLI modules in the build container — TEST INFRASTRUCTURE, not product code.

`cli_svd.py`, `cli_lora_add.py` and `cli_lora_pti.py` import `diffusers`, `fire` (and a package-level
`lora_diffusion` that itself needs `diffusers`) at module top, none of which this image has.  Their ARITHMETIC needs
none of it, so `load_cli(name)` registers inert stand-ins for exactly those imports in `sys.modules`, binds the
relative import `.lora` to the reference's real `lora.py`, executes the reference file unmodified from
`/root/reference`, and restores `sys.modules` afterwards.  What the stand-ins replace is I/O only
(`StableDiffusionPipeline.from_pretrained`, `fire.Fire`, wandb logging); oracle/make_golden.py passes small
stand-in pipelines through them and commits what the reference computed.
"""
import importlib.util
import os
import sys
import types

REF = os.environ.get("B200LORA_REFERENCE", "/root/reference")


class _Inert(types.ModuleType):
    """A module whose every attribute is a do-nothing class (good enough for `from x import A, B`)."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        cls = type(name, (), {"__init__": lambda self, *a, **k: None,
                              "__call__": lambda self, *a, **k: None})
        setattr(self, name, cls)
        return cls


class _Lenient(types.ModuleType):
    """The real module, except that names it no longer exports (`HfFolder`, `Repository` ...) come back inert."""

    def __init__(self, real):
        super().__init__(real.__name__)
        self.__dict__["_real"] = real
        self.__path__ = list(getattr(real, "__path__", []))      # stays a package: submodule imports keep working

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        try:
            return getattr(self.__dict__["_real"], name)
        except AttributeError:
            return type(name, (), {"__init__": lambda self, *a, **k: None})


def _load_file(modname, path, package=None):
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    if package:
        mod.__package__ = package
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


def load_cli(name, pipeline_factory=None):
    """Import `/root/reference/lora_diffusion/<name>.py` as `lora_diffusion.<name>`.  `pipeline_factory(path, **kw)`
    stands in for `StableDiffusionPipeline.from_pretrained`."""
    from transformers import CLIPTextModel, CLIPTokenizer  # noqa: F401  (present; resolve its lazy imports now)
    import huggingface_hub
    ours = lambda k: k.split(".")[0] in ("diffusers", "fire", "wandb", "lora_diffusion") or k == "huggingface_hub"
    saved = {k: v for k, v in sys.modules.items() if ours(k)}
    for k in saved:
        del sys.modules[k]
    try:
        for missing in ("diffusers", "diffusers.optimization", "fire", "wandb"):
            sys.modules[missing] = _Inert(missing)
        sys.modules["huggingface_hub"] = _Lenient(huggingface_hub)
        if pipeline_factory is not None:
            class StableDiffusionPipeline:
                from_pretrained = staticmethod(pipeline_factory)
            sys.modules["diffusers"].StableDiffusionPipeline = StableDiffusionPipeline
        pkg = _Inert("lora_diffusion")
        pkg.__path__ = [os.path.join(REF, "lora_diffusion")]
        sys.modules["lora_diffusion"] = pkg
        ref_lora = _load_file("lora_diffusion.lora", os.path.join(REF, "lora_diffusion", "lora.py"), "lora_diffusion")
        for k, v in vars(ref_lora).items():           # `from lora_diffusion import save_all, ...`
            if not k.startswith("__"):
                setattr(pkg, k, v)
        sys.modules["lora_diffusion.to_ckpt_v2"] = _Inert("lora_diffusion.to_ckpt_v2")
        mod = _load_file(f"lora_diffusion.{name}", os.path.join(REF, "lora_diffusion", f"{name}.py"), "lora_diffusion")
        return mod, ref_lora
    finally:
        for k in [k for k in sys.modules if ours(k)]:
            del sys.modules[k]
        sys.modules.update(saved)

"""Command-line entry points of the artefact tools (the reference ships them as `fire` console scripts:
`lora_add` = cli_lora_add.py:113-114, `lora_ppim` / cli_pt_to_safetensors.py:80-85).  `fire` is not installed
here, so the same functions are exposed through argparse:

    python -m diffusion_finetuning_b200.cli lora_add PATH_1 PATH_2 OUTPUT [--alpha 0.5] [--mode lpl] [--with-text-lora]
    python -m diffusion_finetuning_b200.cli pt_to_safetensors A.pt A.text_encoder.pt A.ti.pt --outpath OUT
                                                            [--overwrite] [--set unet.rank=8 ...]
"""
import argparse
import ast
import sys

from . import artefacts


def _parse_settings(pairs):
    out = {}
    for item in pairs or []:
        if "=" not in item:
            raise SystemExit(f"--set expects name=value (e.g. unet.rank=8), got {item!r}")
        k, v = item.split("=", 1)
        try:
            out[k] = ast.literal_eval(v)
        except (ValueError, SyntaxError):
            out[k] = v
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(prog="diffusion_finetuning_b200.cli", description=__doc__.split("\n\n")[0])
    sub = ap.add_subparsers(dest="cmd", required=True)
    a = sub.add_parser("lora_add", help="cli_lora_add.add: mix two LoRA lists (lpl) or merge one into a pipeline (upl)")
    a.add_argument("path_1")
    a.add_argument("path_2")
    a.add_argument("output_path")
    a.add_argument("--alpha", type=float, default=0.5)
    a.add_argument("--mode", default="lpl", choices=["lpl", "upl", "upl-ckpt-v2"])
    a.add_argument("--with-text-lora", "--with_text_lora", action="store_true")
    c = sub.add_parser("pt_to_safetensors", help="cli_pt_to_safetensors.convert: .pt LoRA lists / TI dicts -> safetensors")
    c.add_argument("paths", nargs="+")
    c.add_argument("--outpath", required=True)
    c.add_argument("--overwrite", action="store_true")
    c.add_argument("--set", action="append", metavar="NAME=VALUE",
                   help="per-model setting, e.g. unet.rank=8 or text_encoder.target_modules=\"['CLIPAttention']\"")
    args = ap.parse_args(argv)
    if args.cmd == "lora_add":
        artefacts.lora_add(args.path_1, args.path_2, args.output_path, alpha=args.alpha, mode=args.mode,
                           with_text_lora=args.with_text_lora)
    else:
        artefacts.convert(*args.paths, outpath=args.outpath, overwrite=args.overwrite, **_parse_settings(args.set))
    return 0


if __name__ == "__main__":
    sys.exit(main())

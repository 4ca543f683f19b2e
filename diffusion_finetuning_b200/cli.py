"""Command-line entry points of the artefact tools (the reference ships them as `fire` console scripts:
`lora_add` = cli_lora_add.py:113-114, `lora_ppim` / cli_pt_to_safetensors.py:80-85).  `fire` is not installed
here, so the same functions are exposed through argparse:

    python -m diffusion_finetuning_b200.cli lora_add PATH_1 PATH_2 OUTPUT [--alpha 0.5] [--mode lpl] [--with-text-lora]
    python -m diffusion_finetuning_b200.cli pt_to_safetensors A.pt A.text_encoder.pt A.ti.pt --outpath OUT
                                                            [--overwrite] [--set unet.rank=8 ...]
    python -m diffusion_finetuning_b200.cli lora_distill TARGET_MODEL BASE_MODEL [--rank 4] [--clamp-quantile 0.99]
                                                            [--device cuda:0] [--save-path svd_distill.pt]
(`lora_distill` = cli_svd.py:114-115, setup.py:17.)
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
    d = sub.add_parser("lora_distill", help="cli_svd.svd_distill: rank-r LoRA factors of (tuned - base) for a whole pipeline")
    d.add_argument("target_model")
    d.add_argument("base_model")
    d.add_argument("--rank", type=int, default=4)
    d.add_argument("--clamp-quantile", "--clamp_quantile", type=float, default=0.99)
    d.add_argument("--device", default="cuda:0")
    d.add_argument("--save-path", "--save_path", default="svd_distill.pt")
    args = ap.parse_args(argv)
    if args.cmd == "lora_distill":
        artefacts.svd_distill(args.target_model, args.base_model, rank=args.rank, clamp_quantile=args.clamp_quantile,
                              device=args.device, save_path=args.save_path)
    elif args.cmd == "lora_add":
        artefacts.lora_add(args.path_1, args.path_2, args.output_path, alpha=args.alpha, mode=args.mode,
                           with_text_lora=args.with_text_lora)
    else:
        artefacts.convert(*args.paths, outpath=args.outpath, overwrite=args.overwrite, **_parse_settings(args.set))
    return 0


if __name__ == "__main__":
    sys.exit(main())

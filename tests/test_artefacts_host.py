"""Host side of the artefact tools, no GPU: cli_pt_to_safetensors.convert against the
reference's own packer, argument errors of lora_add, and the CUDA-only rule of the arithmetic."""
import json
import os

import pytest
import torch

from diffusion_finetuning_b200 import _native as nat
from diffusion_finetuning_b200 import artefacts as A
import diffusion_finetuning_b200 as L


def _pt_list(seed, shapes, dtype=torch.float16):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(s, generator=g) * 0.1).to(dtype) for s in shapes]


SHAPES = [(32, 4), (4, 16), (48, 4), (4, 32)]


def test_convert_matches_reference_packer(tmp_path, reference_lora):
    """cli_pt_to_safetensors.py:19-77 builds {name: (path, targets, rank)} from file names and hands it to
    convert_loras_to_safeloras_with_embeds — run the reference's packer on that map and compare files."""
    from safetensors import safe_open
    unet_pt, text_pt, ti_pt = tmp_path / "w.pt", tmp_path / "w.text_encoder.pt", tmp_path / "w.ti.pt"
    torch.save(_pt_list(0, SHAPES), unet_pt)
    torch.save(_pt_list(1, SHAPES[:2]), text_pt)
    torch.save({"<s1>": torch.randn(8), "<s2>": torch.randn(8)}, ti_pt)
    ours = tmp_path / "ours.safetensors"
    A.convert(str(unet_pt), str(text_pt), str(ti_pt), outpath=str(ours), **{"unet.rank": 8})
    ref = tmp_path / "ref.safetensors"
    reference_lora.convert_loras_to_safeloras_with_embeds(
        {"unet": (str(unet_pt), reference_lora.UNET_DEFAULT_TARGET_REPLACE, 8),
         "text_encoder": (str(text_pt), reference_lora.TEXT_ENCODER_DEFAULT_TARGET_REPLACE, 4)},
        torch.load(ti_pt), str(ref))
    fo, fr = safe_open(str(ours), "pt"), safe_open(str(ref), "pt")
    mo, mr = dict(fo.metadata()), dict(fr.metadata())
    for k in ("unet", "text_encoder"):             # target sets are serialised in set order: compare as sets
        assert set(json.loads(mo.pop(k))) == set(json.loads(mr.pop(k)))
    assert mo == mr
    assert sorted(fo.keys()) == sorted(fr.keys())
    for k in fo.keys():
        assert torch.equal(fo.get_tensor(k), fr.get_tensor(k))
    with pytest.raises(ValueError):                # exists, overwrite not set (cli_pt_to_safetensors.py:44-47)
        A.convert(str(unet_pt), outpath=str(ours))
    A.convert(str(unet_pt), outpath=str(ours), overwrite=True)
    assert dict(safe_open(str(ours), "pt").metadata())["unet:0:rank"] == "4"


def test_convert_layout_without_reference(tmp_path):
    from safetensors import safe_open
    p = tmp_path / "x.text_encoder.pt"
    torch.save(_pt_list(2, SHAPES), p)
    out = tmp_path / "o.safetensors"
    A.convert(str(p), outpath=str(out), **{"text_encoder.rank": 2})
    f = safe_open(str(out), "pt")
    assert sorted(f.keys()) == ["text_encoder:0:down", "text_encoder:0:up", "text_encoder:1:down", "text_encoder:1:up"]
    md = f.metadata()
    assert md["text_encoder:1:rank"] == "2" and set(json.loads(md["text_encoder"])) == {"CLIPAttention"}
    parsed = L.parse_safeloras(f)
    assert parsed["text_encoder"][1] == [2, 2]


def test_lora_add_errors_and_cpu_refusal(tmp_path):
    a, b = tmp_path / "a.pt", tmp_path / "b.pt"
    torch.save(_pt_list(0, SHAPES), a)
    torch.save(_pt_list(1, SHAPES), b)
    with pytest.raises(ValueError):
        A.lora_add(str(a), str(b), str(tmp_path / "o.pt"), mode="nope")
    with pytest.raises(NotImplementedError):
        A.lora_add(str(a), str(b), str(tmp_path / "o.pt"), mode="upl-ckpt-v2")
    with pytest.raises(AssertionError):            # _text_lora_path: "Only .pt files are supported"
        A.lora_add(str(tmp_path / "a.bin"), str(b), str(tmp_path / "o.pt"), with_text_lora=True)
    if not torch.cuda.is_available():
        with pytest.raises(nat.B200LoraError):     # no CPU fallback for the mix
            A.lora_add(str(a), str(b), str(tmp_path / "o.pt"))
        assert not os.path.exists(tmp_path / "o.pt")
        with pytest.raises(nat.B200LoraError):
            A.flat_axpby(torch.zeros(4))
    assert nat.lib().b200lora_flat_axpby(None, None, None, 4, 1.0, 0.0, 0, 0, None) == -6


def test_cli_pt_to_safetensors(tmp_path):
    from safetensors import safe_open
    from diffusion_finetuning_b200 import cli
    p = tmp_path / "w.pt"
    torch.save(_pt_list(4, SHAPES), p)
    out = tmp_path / "cli.safetensors"
    assert cli.main(["pt_to_safetensors", str(p), "--outpath", str(out), "--set", "unet.rank=2"]) == 0
    md = safe_open(str(out), "pt").metadata()
    assert md["unet:0:rank"] == "2" and md["unet:1:rank"] == "2"
    with pytest.raises(ValueError):
        cli.main(["pt_to_safetensors", str(p), "--outpath", str(out)])
    assert cli.main(["pt_to_safetensors", str(p), "--outpath", str(out), "--overwrite"]) == 0
    with pytest.raises(SystemExit):
        cli.main(["pt_to_safetensors", str(p), "--outpath", str(out), "--overwrite", "--set", "rank8"])

"""Textual-inversion step: the sparse placeholder-row path on the GPU against the
reference's dense AdamW + norm decay + restore, restated in oracle/lora_oracle.py (cli_lora_pti.py:280-346,
651-657).  fp32 tolerance 1e-5 relative (BASELINE.json), rows that are not placeholders bit-identical."""
import pytest
import torch
import torch.nn as nn

from conftest import rel_err
from diffusion_finetuning_b200 import _native as nat
from diffusion_finetuning_b200 import ti
from oracle import lora_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _batch(g, V, rows, B=4, T=77):
    ids = torch.randint(0, V - 8, (B, T), generator=g)
    for b in range(B):                       # every sample mentions the placeholders, one of them twice
        ids[b, 4] = rows[0]
        ids[b, 5] = rows[-1]
        if b % 2:
            ids[b, 40] = rows[0]
    return ids


@pytest.mark.parametrize("io", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("D", [768, 100])
def test_grad_rows_kernel_matches_index_add(io, D):
    g = torch.Generator().manual_seed(3)
    V, rows = 500, [497, 120, 499]
    ids = _batch(g, V, rows, B=9, T=77)          # 693 tokens: ragged against the 256-token pass
    dy = torch.randn(ids.numel(), D, generator=g).to(io)
    want = torch.zeros(V, D, dtype=torch.float64).index_add_(0, ids.reshape(-1), dy.double())[rows]
    out = torch.full((len(rows), D), 0.5, dtype=torch.float32, device=DEV)      # the kernel ADDS
    dyd, idd, rd = dy.to(DEV), ids.reshape(-1).to(DEV), torch.tensor(rows, device=DEV)
    rc = nat.lib().b200lora_ti_grad_rows(nat.ptr(dyd), nat.ptr(idd), idd.numel(), D, nat.ptr(rd), len(rows),
                                         nat.ptr(out), nat.dtype_code(io), nat.stream_ptr(torch.device(DEV)))
    nat.check(rc, "ti_grad_rows")
    assert rel_err(out - 0.5, want) < 1e-6
    again = out.clone()
    rc = nat.lib().b200lora_ti_grad_rows(nat.ptr(dyd), nat.ptr(idd), idd.numel(), D, nat.ptr(rd), len(rows),
                                         nat.ptr(again), nat.dtype_code(io), nat.stream_ptr(torch.device(DEV)))
    nat.check(rc, "ti_grad_rows")
    assert torch.equal(again - out, out - 0.5) or rel_err(again - out, out - 0.5) < 1e-6


@pytest.mark.parametrize("wd,decay", [(0.0, True), (0.01, True), (0.0, False)])
def test_inversion_steps_match_dense_reference(wd, decay):
    """6 steps of cli_lora_pti.train_inversion's update (dense AdamW -> norm decay -> restore) on the CPU
    against PlaceholderTokenRows + PlaceholderAdamW on the GPU, linear LR decay so lambda changes per step."""
    V, D, rows = 400, 768, [398, 399]
    torch.manual_seed(0)
    ref = nn.Embedding(V, D)
    with torch.no_grad():
        ref.weight.mul_(0.02)
        ref.weight[rows] *= 30.0                  # norms well away from 0.4 so the decay matters
    ours = nn.Embedding(V, D)
    ours.load_state_dict(ref.state_dict())
    ours.to(DEV)
    table0 = ours.weight.detach().clone()
    proj = torch.randn(D, 32) / D ** 0.5
    pr = ti.PlaceholderTokenRows(ours, rows).install()
    opt = ti.PlaceholderAdamW(pr, lr=5e-3, weight_decay=wd, clip_ti_decay=decay)
    ropt = O.ti_make_optimizer(ref, lr=5e-3, weight_decay=wd)
    orig = ref.weight.data.clone()                # cli_lora_pti.py:280
    no_upd = torch.ones(V, dtype=torch.bool)
    no_upd[rows] = False
    assert torch.equal(pr.index_no_updates().cpu(), no_upd)
    g = torch.Generator().manual_seed(1)
    for step in range(6):
        lr = 5e-3 * (1 - step / 6)
        for grp in ropt.param_groups:
            grp["lr"] = lr
        opt.param_groups[0]["lr"] = lr
        ids = _batch(g, V, rows)
        tgt = torch.randn(ids.shape[0], ids.shape[1], 32, generator=g)
        loss_r = ((ref(ids) @ proj - tgt) ** 2).mean() * 50
        loss_r.backward()
        ropt.step()
        ropt.zero_grad()
        pre_r, cur_r = O.ti_post_step(ref, orig, no_upd, lr, clip_ti_decay=decay)
        loss_o = ((ours(ids.to(DEV)) @ proj.to(DEV) - tgt.to(DEV)) ** 2).mean() * 50
        loss_o.backward()
        assert ours.weight.grad is None           # no dense [V, D] gradient exists
        opt.step()
        opt.zero_grad()
        assert abs(loss_o.item() - loss_r.item()) < 1e-4 * abs(loss_r.item())
        assert rel_err(ours.weight[rows], ref.weight[rows]) < 1e-5, f"step {step}"
        assert rel_err(pr.rows, ref.weight[rows]) < 1e-5
        assert rel_err(opt.norms[0], pre_r) < 1e-5 and rel_err(opt.norms[1], cur_r) < 1e-5
    keep = no_upd.to(DEV)
    assert torch.equal(ours.weight[keep], table0[keep])          # restore == never touched
    assert torch.equal(ref.weight[no_upd], orig[no_upd])
    pr.remove()
    assert "forward" not in ours.__dict__


def test_low_precision_table_is_written_through_from_fp32_master():
    V, D, rows = 300, 768, [7, 299]
    torch.manual_seed(0)
    emb = nn.Embedding(V, D).to(DEV).to(torch.bfloat16)
    before = emb.weight.detach().clone()
    pr = ti.PlaceholderTokenRows(emb, rows).install()
    opt = ti.PlaceholderAdamW(pr, lr=1e-3)
    ids = _batch(torch.Generator().manual_seed(2), V, rows).to(DEV)
    v0 = emb.weight._version
    emb(ids).float().square().mean().backward()
    assert pr.rows.grad.dtype == torch.float32 and pr.rows.grad.abs().sum() > 0
    opt.step()
    assert emb.weight._version > v0
    assert torch.equal(emb.weight[rows], pr.rows.detach().to(torch.bfloat16))
    other = torch.ones(V, dtype=torch.bool, device=DEV)
    other[rows] = False
    assert torch.equal(emb.weight[other], before[other])
    assert not torch.equal(emb.weight[rows], before[rows])
    assert set(pr.learned_embeds()) == set(rows)


def test_rows_join_the_gradient_arena():
    """§8(e): the placeholder rows ride in the flat LoRA gradient arena (extra_params)."""
    from diffusion_finetuning_b200 import dp
    emb = nn.Embedding(64, 128).to(DEV)
    pr = ti.PlaceholderTokenRows(emb, [62, 63]).install()
    arena = dp.LoraGradArena(nn.Module(), extra_params=[pr.rows])
    arena.zero()
    ids = torch.tensor([[1, 62, 3, 63, 62]], device=DEV)
    emb(ids).sum().backward()
    arena.finish()
    want = torch.tensor([[2.0], [1.0]], device=DEV).expand(2, 128)
    assert torch.equal(pr.rows.grad, want)
    assert pr.rows.grad.data_ptr() == arena.flat.data_ptr()


def test_train_inversion_against_reference_run():
    """The fixture is the REFERENCE's `cli_lora_pti.train_inversion` (cli_lora_pti.py:250-405) executed for 6 steps
    (oracle/make_golden.py): dense AdamW (:651-657), linear LR decay stepped first (:293), norm decay and restore
    (:315-346).  Here: PlaceholderTokenRows + PlaceholderAdamW on the GPU, same batches, same loss."""
    from conftest import load_golden
    blob, meta = load_golden("train_inversion_case.safetensors")
    lr, wd, steps, total = float(meta["lr"]), float(meta["weight_decay"]), int(meta["steps"]), int(meta["sched_total"])
    V, D = blob["table0"].shape
    emb = nn.Embedding(V, D).to(DEV)
    with torch.no_grad():
        emb.weight.copy_(blob["table0"].to(DEV))
    rows = blob["rows"].tolist()
    pr = ti.PlaceholderTokenRows(emb, rows).install()
    opt = ti.PlaceholderAdamW(pr, lr=lr, weight_decay=wd, clip_ti_decay=True)
    for i in range(steps):
        step_lr = lr * max(0.0, (total - (i + 1)) / total)          # scheduler stepped before the update
        opt.param_groups[0]["lr"] = step_lr
        e = emb(blob[f"b{i}.input_ids"].to(DEV))
        loss = ((e * blob[f"b{i}.w"].to(DEV)).sum(-1) - blob[f"b{i}.y"].to(DEV)).pow(2).mean()
        opt.zero_grad()
        loss.backward()
        opt.step()
    got, want = emb.weight.detach().cpu(), blob["table_final"]
    assert rel_err(got[rows], want[rows]) < 1e-5
    assert (got[rows] - want[rows]).abs().max().item() < 1e-5
    keep = torch.ones(V, dtype=torch.bool)
    keep[rows] = False
    assert torch.equal(got[keep], want[keep])


def test_capturable_step_follows_device_lr_and_step_count():
    """`capturable=True`: lr and the step count live on the device, so a captured launch keeps following an LR
    schedule and the bias correction; 4 replays of ONE captured step == 4 eager steps with the same lr sequence."""
    V, D, rows = 64, 96, [61, 63]
    torch.manual_seed(0)
    tables = [nn.Embedding(V, D).to(DEV) for _ in range(2)]
    tables[1].load_state_dict(tables[0].state_dict())
    prs = [ti.PlaceholderTokenRows(t, rows).install() for t in tables]
    eager = ti.PlaceholderAdamW(prs[0], lr=1e-2, weight_decay=1e-2)
    graphed = ti.PlaceholderAdamW(prs[1], lr=1e-2, weight_decay=1e-2, capturable=True)
    g = torch.Generator().manual_seed(1)
    grads = [torch.randn(len(rows), D, generator=g).to(DEV) for _ in range(4)]
    lrs = [1e-2, 7e-3, 4e-3, 1e-3]
    static_g = torch.zeros(len(rows), D, device=DEV)
    prs[1].rows.grad = static_g
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            graphed.step()
    torch.cuda.current_stream().wait_stream(side)
    # the capture itself launched nothing: state untouched
    assert graphed.step_dev.abs().sum().item() == 0
    for gi, lr in zip(grads, lrs):
        prs[0].rows.grad = gi.clone()
        eager.step(lr=lr)
        static_g.copy_(gi)
        graphed.param_groups[0]["lr"].fill_(lr)
        graph.replay()
    torch.cuda.synchronize()
    assert graphed.step_dev.tolist() == [4.0, 4.0]
    # (lambda = min(1, 100·lr) is formed in double on the host for the eager step and in fp32 inside the kernel for
    # the replayed one: one ulp apart at most)
    assert torch.allclose(prs[0].rows.detach(), prs[1].rows.detach(), rtol=1e-6, atol=1e-7)
    assert torch.allclose(tables[0].weight.detach(), tables[1].weight.detach(), rtol=1e-6, atol=1e-7)

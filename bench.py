"""bench.py — LoRA-UNet training throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference algorithm on the host cores

Workload (BASELINE.json configs[1]): SD-1.5-shaped UNet (random init, synthetic latents), LoRA r=4
on CrossAttention to_q/to_k/to_v/to_out (128 layers), 64x64 latents, bf16 autocast over fp32
masters (the reference's `--mixed_precision bf16` mode), batch 8 per GPU, fwd + bwd + global-norm
clip + AdamW.  A "step" is one such training step on one batch; data-parallel over N GPUs with the
bucketed LoRA-gradient all-reduce of diffusion_finetuning_b200.dp (weak scaling).

One JSON line on stdout (rank 0).  `value` = samples/s with the batch already in HBM; `e2e` = the
same through the public API with the batch in pinned HOST memory (H2D inside the timed region) and
the loss read back every step; `roofline` = the fused tcgen05 LoRA GEMM kernel (forward and
backward-dX launches), algorithmic FLOPs / CUDA-event time measured live around every launch of
the timed region; `cpu_baseline` = the oracle's restatement of the reference layer inside the same
UNet on the host cores (bounded sample: batch 1).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

METRIC = "LoRA-UNet train samples/sec"
UNIT = "samples/s"
ATTN_TARGETS = {"CrossAttention", "Attention"}
WORKLOAD = "sd15_unet_lora_r4_attn_qkvo_64x64_bf16_b8"
# BASELINE config 3 (--config 3; a parity/coverage configuration, not the headline line): extended
# injection — resnet Conv2d + time_emb_proj + attention + GEGLU/FeedForward, r = 16
EXT_TARGETS = {"ResnetBlock2D", "CrossAttention", "Attention", "GEGLU", "FeedForward"}
WORKLOAD3 = "sd15_unet_lora_r16_extended_conv_attn_ff_64x64_bf16_b8"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"tf_burst": p["bf16_tflops"], "tf_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "hbm": p["hbm_gbs"], "src": "measured"}
    return {"tf_burst": 1590.0, "tf_sustained": 1400.0, "hbm": 6650.0, "src": "fallback"}


# ------------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# live per-launch timing of the hot-path kernels (installed as ops.PROFILER)
# ------------------------------------------------------------------------------------------------
class LaunchProfiler:
    """Eager mode: a fresh event pair around every launch.  `external=True`: the pairs are recorded while
    the step is being CAPTURED (cudaEventRecordExternal -> event-record nodes of the graph), so every replay
    re-stamps them and `harvest()` reads true device-side durations of the kernels inside the replayed step —
    no host enqueue latency between the two stamps, which the eager pass cannot avoid when the host is the
    slower side."""

    def __init__(self, external=False):
        self.recs = []
        self.cur = None
        self.external = external
        self.pairs = []          # external mode: (kind, M, K, N, r, esize, a, b), stamped by each replay
        self.acc = []            # external mode: accumulated ms per pair
        self.replays = 0

    def _event(self):
        return torch.cuda.Event(enable_timing=True, external=True) if self.external \
            else torch.cuda.Event(enable_timing=True)

    def begin(self, kind, M, K, N, r, esize):
        a = self._event()
        a.record()
        self.cur = (kind, M, K, N, r, esize, a)

    def end(self):
        b = self._event()
        b.record()
        (self.pairs if self.external else self.recs).append(self.cur + (b,))

    def reset_capture(self):
        """Forget the pairs recorded by eager warm-up steps; the next pass (the capture) defines them."""
        self.pairs, self.acc, self.recs, self.replays = [], [], [], 0

    def harvest(self):
        """After a replay has finished (device synchronised): fold this replay's durations in."""
        if not self.acc:
            self.acc = [0.0] * len(self.pairs)
        for i, rec in enumerate(self.pairs):
            self.acc[i] += rec[6].elapsed_time(rec[7])
        self.replays += 1

    def _iter_ms(self):
        if self.external:
            for rec, ms in zip(self.pairs, self.acc):
                yield rec[:6], ms, self.replays
        else:
            for rec in self.recs:
                yield rec[:6], rec[6].elapsed_time(rec[7]), 1

    @staticmethod
    def work(kind, M, K, N, r, e):
        """ALGORITHMIC flops / HBM bytes per launch."""
        if kind == "fwd":      # Y = X W0^T + b + s (X A^T) B^T ; T written for the backward
            return 2.0 * M * K * N + 2.0 * M * r * (K + N), e * (M * K + N * K + M * N + r * (K + N) + M * r)
        if kind == "bwd_dx":   # dX = dY W0 + s (dY B) A ; dT written for the reductions
            return 2.0 * M * K * N + 2.0 * M * r * (K + N), e * (M * N + N * K + M * K + r * (K + N) + M * r)
        if kind == "grads":    # dA, dB token reductions: re-read X and dY once
            return 2.0 * M * r * (K + N), e * (M * K + M * N + 2 * M * r) + 4 * r * (K + N)
        if kind.startswith("conv_"):
            # implicit-GEMM LoRA conv (conv_ops._gemm): M output pixels, K = taps * C_in, N = C_out; the input
            # image is read ONCE (M * K / taps elements at stride 1), not once per tap.  Forward: base conv +
            # down conv (r more output columns) + up (K=r tail); backward dX: base + the r extra K rows per tap.
            taps = int(kind.split(":")[1])
            cin = K // taps
            if kind.startswith("conv_fwd"):
                return (2.0 * M * K * N + 2.0 * M * r * (K + N),
                        e * (M * cin + N * K + M * N + r * (K + N) + M * r))
            return (2.0 * M * K * N + 2.0 * M * r * taps * N,
                    e * (M * cin + N * K + M * N + r * taps * N + M * r))
        return 0.0, 0.0

    def summary(self):
        out = {}
        for (kind, M, K, N, r, e), ms, n in self._iter_ms():
            fl, by = self.work(kind, M, K, N, r, e)
            s = out.setdefault(kind, {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
            s["launches"] += n
            s["ms"] += ms
            s["flops"] += fl * n
            s["bytes"] += by * n
        return out

    def by_shape(self, peak_tf, peak_gbs):
        """Per (kind, M, K, N, r): launches, mean event-bracketed duration, achieved TFLOP/s and GB/s."""
        acc = {}
        for (kind, M, K, N, r, e), ms, n in self._iter_ms():
            fl, by = self.work(kind, M, K, N, r, e)
            s = acc.setdefault((kind, M, K, N, r), [0, 0.0, fl, by])
            s[0] += n
            s[1] += ms
        rows = []
        for (kind, M, K, N, r), (n, ms, fl, by) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
            us = ms * 1e3 / n
            rows.append({"kind": kind, "M": M, "K": K, "N": N, "r": r, "launches": n, "avg_us": round(us, 2),
                         "tflops": round(fl / us / 1e6, 1), "gbs": round(by / us / 1e3, 1),
                         "frac_tensor": round(fl / us / 1e6 / peak_tf, 3), "frac_hbm": round(by / us / 1e3 / peak_gbs, 3)})
        return rows


# ------------------------------------------------------------------------------------------------
# the training job
# ------------------------------------------------------------------------------------------------
def make_batches(n, batch, res, seed, device, pin=False):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        b = {"latents": torch.randn(batch, 4, res, res, generator=g),
             "t": torch.randint(0, 1000, (batch,), generator=g),
             "ctx": torch.randn(batch, 77, 768, generator=g),
             "target": torch.randn(batch, 4, res, res, generator=g)}
        if device is not None:
            b = {k: v.to(device) for k, v in b.items()}
        elif pin:
            b = {k: v.pin_memory() for k, v in b.items()}
        out.append(b)
    return out


class _StockLinear(torch.nn.Module):
    """lora.py:32-50 evaluated by stock PyTorch ops ON THE GPU (--impl stock_gpu): the number the fused
    kernels have to beat on the same B200, next to the CPU baseline.  Not a product path."""

    def __init__(self, src, r):
        super().__init__()
        self.linear = src
        self.lora_down = torch.nn.Linear(src.in_features, r, bias=False)
        self.lora_up = torch.nn.Linear(r, src.out_features, bias=False)
        self.scale = 1.0
        torch.nn.init.normal_(self.lora_down.weight, std=1 / r)

    def forward(self, x):
        return self.linear(x) + self.lora_up(self.lora_down(x)) * self.scale


class _StockConv(torch.nn.Module):
    def __init__(self, src, r):
        super().__init__()
        self.conv = src
        self.lora_down = torch.nn.Conv2d(src.in_channels, r, src.kernel_size, src.stride, src.padding, bias=False)
        self.lora_up = torch.nn.Conv2d(r, src.out_channels, 1, bias=False)
        self.scale = 1.0
        torch.nn.init.normal_(self.lora_down.weight, std=1 / r)

    def forward(self, x):
        return self.conv(x) + self.lora_up(self.lora_down(x)) * self.scale


def _stock_inject(unet, targets, r, convs):
    import diffusion_finetuning_b200 as L
    kinds = [torch.nn.Linear, torch.nn.Conv2d] if convs else [torch.nn.Linear]
    found = list(L._find_modules(unet, targets, search_class=kinds, exclude_children_of=[_StockLinear, _StockConv]))
    layers, seen = [], set()
    for parent, name, mod in found:
        if id(mod) in seen:
            continue
        seen.add(id(mod))
        new = _StockConv(mod, r) if isinstance(mod, torch.nn.Conv2d) else _StockLinear(mod, r)
        new.to(mod.weight.device)
        parent._modules[name] = new
        new.lora_up.weight.requires_grad_(True)
        new.lora_down.weight.requires_grad_(True)
        layers.append(new)
    return layers


class StockTrainer:
    """Same step as Trainer with the reference composition in stock PyTorch on the GPU; data-parallel through
    torch DistributedDataParallel when a process group is up (what accelerate wraps the UNet in,
    train_lora_dreambooth.py:751-757).  `capture()` replays the whole step as one CUDA graph, like the repo arm."""

    def __init__(self, device, rank_r, extended=False, channels_last=False, ddp=False, capturable=False):
        from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
        torch.manual_seed(0)
        self.unet = UNet2DConditionModel().to(device)
        if channels_last:
            self.unet.to(memory_format=torch.channels_last)
        self.channels_last = channels_last
        self.unet.requires_grad_(False)
        torch.manual_seed(2)      # same lora_down on every rank (DDP broadcasts rank 0's anyway)
        self.layers = _stock_inject(self.unet, EXT_TARGETS if extended else ATTN_TARGETS, rank_r, extended)
        self.n_layers = len(self.layers)
        g = torch.Generator(device="cpu").manual_seed(1)
        with torch.no_grad():
            for m in self.layers:
                m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(device))
        self.params = [p for m in self.layers for p in (m.lora_up.weight, m.lora_down.weight)]
        self.model = self.unet
        self.ddp = bool(ddp)
        self._side = torch.cuda.Stream(device=device)
        if ddp:
            # torch's recipe for whole-network capture of DDP: construct it (and warm it up) in a side-stream context,
            # or the reducer's AccumulateGrad nodes are tied to the default stream and invalidate the capture
            from torch.nn.parallel import DistributedDataParallel as DDP
            self._side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(self._side):
                self.model = DDP(self.unet, device_ids=[device.index], broadcast_buffers=False)
            torch.cuda.current_stream(device).wait_stream(self._side)
        self.opt = torch.optim.AdamW(self.params, lr=1e-4, fused=True, capturable=capturable)
        self.device = device
        self.graph = None

    def eager_step(self, b):
        self.opt.zero_grad(set_to_none=True)
        lat = b["latents"].contiguous(memory_format=torch.channels_last) if self.channels_last else b["latents"]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            pred = self.model(lat, b["t"], b["ctx"])
        loss = F.mse_loss(pred.float(), b["target"].float(), reduction="mean")
        loss.backward()
        torch.nn.utils.clip_grad_norm_(self.params, 1.0)
        self.opt.step()
        return loss

    def capture(self, example):
        """Whole-network capture as torch documents it: warm-up on a side stream (11 iterations for DDP: the
        reducer rebuilds its buckets and its logger synchronises during the first ones), then one capture."""
        self.static = {k: v.clone() for k, v in example.items()}
        side = self._side
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(11 if self.ddp else 3):
                self.eager_step(self.static)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.opt.zero_grad(set_to_none=True)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            self.loss = self.eager_step(self.static)
        self.graph = graph

    def step(self, b):
        if self.graph is None:
            return self.eager_step(b)
        for k, v in b.items():
            self.static[k].copy_(v, non_blocking=True)
        self.graph.replay()
        return self.loss

    def step_from_host(self, hb):
        if self.graph is not None:
            return self.step(hb).item()
        b = {k: v.to(self.device, non_blocking=True) for k, v in hb.items()}
        return self.step(b).item()

    def release(self):
        self.graph = None
        self.model = None


def stock_arm(args, device, dist_on, pool, host_pool, graph=True):
    """Times the stock-PyTorch composition of lora.py:49-50 on the same GPU(s), same UNet / optimizer / clip /
    batches; CUDA-graph replay like the repo arm when capture succeeds (else eager, and says so)."""
    tr = StockTrainer(device, args.rank, extended=args.config == 3, channels_last=args.memory_format == "channels_last",
                      ddp=dist_on, capturable=graph)
    launch = "python (eager)"
    if graph:
        try:
            tr.capture(pool[0])
            launch = "whole step replayed as one CUDA graph"
        except Exception as exc:
            print(f"[bench] stock arm: CUDA-graph capture failed ({type(exc).__name__}: {exc}); eager steps",
                  file=sys.stderr)
            tr.graph = None
            torch.cuda.synchronize(device)
    for i in range(max(3, args.warmup)):
        tr.step(pool[i % 4])
    tr.step_from_host(host_pool[0])
    t_dev = timed(lambda i: tr.step(pool[i % 4]), args.steps, dist_on, device)
    t_e2e = timed(lambda i: tr.step_from_host(host_pool[i % 4]), args.steps, dist_on, device)
    n_layers = tr.n_layers
    tr.release()
    del tr
    return {"t_dev": t_dev, "t_e2e": t_e2e, "launch": launch, "lora_layers": n_layers,
            "parallel": "torch DistributedDataParallel (NCCL)" if dist_on else "single process"}


class Trainer:
    """The public API a user of the reference calls, end to end: inject_trainable_lora ->
    optimizer over the returned parameter generators -> fwd/bwd/clip/step
    (train_lora_dreambooth.py:595-676, 816-888)."""

    def __init__(self, device, rank_r, autocast=True, channels_last=False, fuse=True, extended=False,
                 capturable=False, arena_optimizer=False):
        import itertools
        import diffusion_finetuning_b200 as L
        from diffusion_finetuning_b200 import dp
        from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
        torch.manual_seed(0)                      # same frozen base on every rank
        self.unet = UNet2DConditionModel().to(device)
        if channels_last:     # stock-PyTorch setting for the frozen convs/norms; the LoRA layers are 2-D GEMMs
            self.unet.to(memory_format=torch.channels_last)
        self.channels_last = channels_last
        self.unet.requires_grad_(False)
        if extended:
            params, names = L.inject_trainable_lora_extended(self.unet, EXT_TARGETS, r=rank_r)
        else:
            params, names = L.inject_trainable_lora(self.unet, ATTN_TARGETS, r=rank_r)
        self.n_layers = len(names)
        g = torch.Generator(device="cpu").manual_seed(1)
        with torch.no_grad():                     # non-zero up so every gradient is live
            for m in self.unet.modules():
                if isinstance(m, (L.LoraInjectedLinear, L.LoraInjectedConv2d)):
                    m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(device))
        self.groups = L.fuse_sibling_projections(self.unet) if fuse else []
        # GEGLU projections (only injected by the extended / default target sets: config 3): activation in the epilogue
        self.geglu = L.fuse_geglu(self.unet) if fuse and os.environ.get("B200LORA_BENCH_GEGLU", "1") != "0" else []
        self.arena = dp.LoraGradArena(self.unet, bucket_bytes=1 << 20)
        self.arena.broadcast_params(0)
        if arena_optimizer:      # clip 1.0 + AdamW over the flat arenas: one C-ABI call per step
            self.opt = dp.ArenaAdamW(self.arena, lr=1e-4, weight_decay=1e-2, max_grad_norm=1.0)
        else:
            self.opt = torch.optim.AdamW(itertools.chain(*params), lr=1e-4, fused=True, capturable=capturable)
        self.arena_optimizer = arena_optimizer
        self.device = device
        self.autocast = autocast
        self.graphed = None
        self.graph_capable = True

    def release(self):
        """Drop the captured graph (it holds NCCL work) and the arena wiring before the process group goes."""
        self.graphed = None
        self.arena.detach()

    def capture(self, example_batch):
        """Whole step as one CUDA graph (dp.GraphedStep); step()/step_from_host() replay it from then on."""
        from diffusion_finetuning_b200 import dp
        self.graphed = dp.GraphedStep(self.eager_step, example_batch)

    def step(self, b):
        return self.graphed(b) if self.graphed is not None else self.eager_step(b)

    def eager_step(self, b):
        self.arena.zero()
        lat = b["latents"].contiguous(memory_format=torch.channels_last) if self.channels_last else b["latents"]
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.autocast):
            pred = self.unet(lat, b["t"], b["ctx"])
        loss = F.mse_loss(pred.float(), b["target"].float(), reduction="mean")
        loss.backward()
        self.arena.finish()
        if not self.arena_optimizer:
            self.arena.clip_grad_norm_(1.0)
        self.opt.step()
        return loss

    def step_from_host(self, hb):
        if self.graphed is not None:            # pinned host batch -> the graph's static inputs -> replay
            return self.graphed(hb).item()
        b = {k: v.to(self.device, non_blocking=True) for k, v in hb.items()}
        return self.step(b).item()              # D2H read of the loss every step


class PtiTrainer:
    """BASELINE config 4: the tuning phase of cli_lora_pti (perform_tuning, :408-489, with continue_inversion,
    :699-751) — UNet LoRA + CLIP ViT-L/14 text-encoder LoRA (r=4, targets {"CLIPAttention"}) + the placeholder-token
    rows, one `train.LoraTrainer` (three AdamW groups, joint clip 1.0, LR schedule stepped first).  The text encoder
    is a random-init `transformers.CLIPTextModel` of the ViT-L/14 text shape; prompts are random token ids with the
    two placeholder ids at positions 4 and 5."""

    VOCAB = 49408 + 2

    def __init__(self, device, rank_r, batch, graph=True):
        from transformers import CLIPTextConfig, CLIPTextModel
        import diffusion_finetuning_b200 as L
        from diffusion_finetuning_b200 import ti, train
        from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
        torch.manual_seed(0)
        self.unet = UNet2DConditionModel().to(device)
        self.unet.requires_grad_(False)
        L.inject_trainable_lora(self.unet, ATTN_TARGETS | {"GEGLU"}, r=rank_r)
        ccfg = CLIPTextConfig(vocab_size=self.VOCAB, hidden_size=768, intermediate_size=3072, num_hidden_layers=12,
                              num_attention_heads=12, max_position_embeddings=77)
        self.clip = CLIPTextModel(ccfg).to(device)
        self.clip.requires_grad_(False)
        L.inject_trainable_lora(self.clip, {"CLIPAttention"}, r=rank_r)
        g = torch.Generator(device="cpu").manual_seed(1)
        n = 0
        with torch.no_grad():
            for model in (self.unet, self.clip):
                for m in model.modules():
                    if isinstance(m, L.LoraInjectedLinear):
                        m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(device))
                        n += 1
        self.n_layers = n
        self.groups = L.fuse_sibling_projections(self.unet) + L.fuse_sibling_projections(self.clip)
        self.rows = ti.PlaceholderTokenRows(self.clip.get_input_embeddings(), [self.VOCAB - 2, self.VOCAB - 1]).install()
        cfg = train.TrainConfig(learning_rate=1e-4, learning_rate_text=1e-5, learning_rate_ti=5e-4, lr_scheduler="constant",
                                max_train_steps=10 ** 9, max_grad_norm=1.0, train_text_encoder=True, scheduler_first=True,
                                mixed_precision="bf16", use_cuda_graph=graph, save_steps=0)
        self.tr = train.LoraTrainer(self.unet, cfg, text_encoder=self.clip, placeholder_rows=self.rows)
        self.arena = self.tr.arena
        self.device = device
        self.batch = batch
        self.graph_capable = self.tr.graph_ok
        self.graphed = None

    def make_batches(self, n, res, seed, device):
        g = torch.Generator().manual_seed(seed)
        out = []
        for _ in range(n):
            ids = torch.randint(0, self.VOCAB - 2, (self.batch, 77), generator=g)
            ids[:, 4], ids[:, 5] = self.VOCAB - 2, self.VOCAB - 1
            b = {"latents": torch.randn(self.batch, 4, res, res, generator=g), "input_ids": ids,
                 "noise": torch.randn(self.batch, 4, res, res, generator=g),
                 "timesteps": torch.randint(0, 1000, (self.batch,), generator=g)}
            out.append({k: v.to(device) for k, v in b.items()} if device is not None
                       else {k: v.pin_memory() for k, v in b.items()})
        return out

    def capture(self, example_batch):
        self.tr.train_step(example_batch)            # LoraTrainer captures on its first step
        self.graphed = self.tr._graphed

    def eager_step(self, b):
        return self.tr._whole_step(self.tr.prepare(b))

    def step(self, b):
        return self.tr.train_step(b)

    def step_from_host(self, hb):
        b = {k: v.to(self.device, non_blocking=True) for k, v in hb.items()}
        return self.step(b).item()

    def release(self):
        self.graphed = None
        self.tr._graphed = None
        self.arena.detach()


def timed(fn, steps, dist_on, device):
    """K steps between barrier+synchronize, CUDA events, MAX over ranks -> seconds."""
    import torch.distributed as dist
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize(device)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        fn(i)
    b.record()
    torch.cuda.synchronize(device)
    if dist_on:
        dist.barrier()
    ms = torch.tensor([a.elapsed_time(b)], device=device)
    if dist_on:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item() * 1e-3


# ------------------------------------------------------------------------------------------------
# CPU leg: the reference's own CPU path on the host cores.  `oracle/_ref/reference_lora.py` (the reference's
# lora.py itself, put there by oracle/build_ref.py) when it travelled with the repo — kind "reference";
# else the oracle's restatement (oracle/lora_oracle.py) — kind "port".  Config 3's extended injection does
# not exist in the reference snapshot: always the port.
# ------------------------------------------------------------------------------------------------
def cpu_leg(steps, warmup, batch, res, rank_r, config=2):
    import itertools
    from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
    from oracle import build_ref
    from oracle import lora_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    unet = UNet2DConditionModel()
    unet.requires_grad_(False)
    ref = build_ref.load() if config != 3 else None
    kind = "reference" if ref is not None else "port"
    unet_targets = EXT_TARGETS if config == 3 else (ATTN_TARGETS | {"GEGLU"} if config == 4 else ATTN_TARGETS)

    def inject(model, targets, convs=False):
        if ref is not None:                 # lora.py:137-183, executed as is
            ref.inject_trainable_lora(model, set(targets), r=rank_r)
            return [m for m in model.modules() if type(m).__name__ == "LoraInjectedLinear"]
        return O.inject(model, targets, r=rank_r, convs=convs)

    layers = inject(unet, unet_targets, convs=config == 3)
    clip = emb = None
    if config == 4:
        from transformers import CLIPTextConfig, CLIPTextModel
        clip = CLIPTextModel(CLIPTextConfig(vocab_size=PtiTrainer.VOCAB, hidden_size=768, intermediate_size=3072,
                                            num_hidden_layers=12, num_attention_heads=12, max_position_embeddings=77))
        clip.requires_grad_(False)
        layers = layers + inject(clip, {"CLIPAttention"})
        emb = clip.get_input_embeddings()
        emb.weight.requires_grad_(True)             # the reference trains the dense table (cli_lora_pti.py:713-728)
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for m in layers:
            m.lora_up.weight.copy_(torch.randn(m.lora_up.weight.shape, generator=g) * 0.02)
    params = list(itertools.chain(*[[m.lora_up.weight, m.lora_down.weight] for m in layers]))
    if emb is not None:
        params.append(emb.weight)
    opt = torch.optim.AdamW(params, lr=1e-4)
    batches = make_batches(2, batch, res, 1234, None)
    if config == 4:
        for b in batches:
            ids = torch.randint(0, PtiTrainer.VOCAB - 2, (batch, 77), generator=g)
            ids[:, 4], ids[:, 5] = PtiTrainer.VOCAB - 2, PtiTrainer.VOCAB - 1
            b["input_ids"] = ids

    def step(b):
        opt.zero_grad(set_to_none=True)
        ctx = clip(b["input_ids"])[0] if clip is not None else b["ctx"]
        pred = unet(b["latents"], b["t"], ctx)
        loss = F.mse_loss(pred.float(), b["target"].float())
        loss.backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        return loss.item()

    for i in range(warmup):
        step(batches[i % 2])
    t0 = time.perf_counter()
    for i in range(steps):
        step(batches[i % 2])
    dt = time.perf_counter() - t0
    what = ("the reference's lora.py itself (oracle/_ref/reference_lora.py) injected into" if kind == "reference"
            else "oracle/lora_oracle.py layers inside")
    return {"value": batch * steps / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
            "sample": f"{steps} steps of batch {batch} (1/8 of the per-GPU batch) at {res}x{res} latents, fp32, "
                      f"{what} the same UNet" + (" + CLIP text encoder, dense embedding training" if config == 4 else "")
                      + f", {len(layers)} LoRA layers",
            "ms_per_step": dt / steps * 1e3}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if args.config == 5:
        return run_config5_reference(args)
    cb = cpu_leg(args.steps, max(1, min(args.warmup, 2)), 1, args.res, args.rank, args.config)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config_dict(args, 1),
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _emit(line)
    return 0


WORKLOAD4 = "sd15_unet_plus_clip_vitl14_text_lora_r4_plus_placeholder_rows_64x64_bf16_b8"


def config_dict(args, batch):
    cfg = getattr(args, "config", 2)
    d = {"workload": {3: WORKLOAD3, 4: WORKLOAD4}.get(cfg, WORKLOAD),
         "unet": "SD-1.5 shape (320,640,1280,1280), 2 layers/block, ctx 77x768",
         "lora": {3: f"r={args.rank} extended: resnet convs, time_emb_proj, attention, GEGLU/FeedForward",
                  4: f"r={args.rank} UNet attention + GEGLU, r={args.rank} CLIP ViT-L/14 text encoder q/k/v/out (48 layers), "
                     "2 placeholder-token embedding rows (cli_lora_pti perform_tuning, continue_inversion)"}.get(
                      cfg, f"r={args.rank} on CrossAttention to_q/to_k/to_v/to_out"),
         "latents": f"{args.res}x{args.res}", "batch_per_gpu": batch, "precision": "bf16 autocast, fp32 masters",
         "parallelism": f"dp{args.gpus}",
         "optimizer": ("AdamW groups lr 1e-4 / 1e-5 (text) / 5e-4 (rows), joint clip 1.0" if cfg == 4 else
                       "AdamW lr 1e-4, clip 1.0" + (" (dp.ArenaAdamW)" if getattr(args, "arena_optimizer", False) else "")),
         "memory_format": getattr(args, "memory_format", "nchw"),
         "step_launch": "python" if getattr(args, "no_graph", False) else "whole step replayed as one CUDA graph",
         "sibling_fusion": "off" if getattr(args, "no_fuse", False) else "q/k/v and k/v LoRA layers of one attention block = one launch; GEGLU projections (config 3) with the activation in the GEMM epilogue",
         "l2": "inputs larger than L2: per-step activations (>10 GB) dwarf the 126 MB L2; no explicit flush"}
    return d


# ------------------------------------------------------------------------------------------------
# BASELINE config 5: whole-model merge, tune_lora_scale sweep, SVD distillation (its own metrics)
# ------------------------------------------------------------------------------------------------
METRIC5 = "whole-model LoRA merge (W0 + alpha*up@down over the 144 UNet + 48 text-encoder matrices) GB/s"


def _sd15_merge_shapes():
    """(N, K) of every nn.Linear under the default targets of the SD-1.5-shaped UNet (144) and of the ViT-L/14 text
    encoder's attention (48 x 768x768) — what cli_lora_add / weight_apply_lora walks (SURVEY 8 census)."""
    import diffusion_finetuning_b200 as L
    from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
    with torch.device("meta"):
        unet = UNet2DConditionModel()
    shapes = [tuple(m.weight.shape) for _, _, m in L._find_modules(unet, L.DEFAULT_TARGET_REPLACE,
                                                                   search_class=[torch.nn.Linear])]
    return shapes, [(768, 768)] * 48


def _time_cuda(fn, iters, device):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(device)
    a.record()
    for i in range(iters):
        fn(i)
    b.record()
    torch.cuda.synchronize(device)
    return a.elapsed_time(b) * 1e-3 / iters


def run_config5(args, device, rank):
    """One JSON line: `value` = GB/s of the fp32 whole-model merge (the reference's CLI dtype), the fp16 merge, the
    51-value tune_lora_scale sweep and the SVD distillation wall time beside it."""
    import diffusion_finetuning_b200 as L
    from diffusion_finetuning_b200 import ops
    from diffusion_finetuning_b200 import _native as nat
    if rank != 0:
        return 0
    pk = peaks()
    unet_shapes, clip_shapes = _sd15_merge_shapes()
    shapes = unet_shapes + clip_shapes
    g = torch.Generator(device=device).manual_seed(0)
    clocks = ClockSampler(device.index or 0)
    clocks.start()
    merge = {}
    launches = 0
    for tag, wdt, ldt in (("fp32", torch.float32, torch.float32), ("fp16", torch.float16, torch.float16),
                          ("bf16", torch.bfloat16, torch.bfloat16)):
        base = [(torch.randn(N, K, device=device, generator=g) * 0.03).to(wdt) for N, K in shapes]
        live = [torch.empty_like(w) for w in base]
        ups = [(torch.randn(N, args.rank, device=device, generator=g) * 0.02).to(ldt) for N, _ in shapes]
        downs = [(torch.randn(args.rank, K, device=device, generator=g) * 0.02).to(ldt) for _, K in shapes]
        plan = ops.MergePlan(base, live, ups, downs)            # Win = pristine base, Wout = live weights
        for i in range(max(3, args.warmup)):
            plan.run(0.5)
        l0 = nat.LAUNCHES
        t = _time_cuda(lambda i: plan.run(i / 50.0), args.steps, device)      # > 1 GB per pass: larger than L2
        launches += nat.LAUNCHES - l0
        merge[tag] = {"ms": t * 1e3, "gbs": plan.bytes / t / 1e9, "frac_hbm": plan.bytes / t / 1e9 / pk["hbm"],
                      "bytes": plan.bytes, "matrices": len(shapes), "launches_per_merge": 1}
        # the per-matrix kernel (one launch per layer: the round-1 path) on the same buffers
        def per_layer(i):
            for w, o, u, d in zip(base, live, ups, downs):
                rc = nat.lib().b200lora_merge(nat.ptr(w), nat.ptr(o), nat.ptr(u), nat.ptr(d), i / 50.0, w.shape[0],
                                              w.shape[1], args.rank, nat.dtype_code(wdt), nat.dtype_code(ldt),
                                              nat.stream_ptr(device))
                nat.check(rc, "b200lora_merge")
        per_layer(0)
        t1 = _time_cuda(per_layer, max(2, args.steps // 4), device)
        merge[tag]["per_matrix_launches_ms"] = t1 * 1e3
        del plan, base, live, ups, downs
        torch.cuda.empty_cache()
    # ---- tune_lora_scale sweep (lora.py:597-600; scripts/make_alpha_gifs.ipynb: alpha = i/50, 51 values):
    # scale is a RUNTIME kernel argument here, so a sweep re-packs and re-merges nothing ----
    from diffusion_finetuning_b200.unet_sd15 import UNet2DConditionModel
    torch.manual_seed(0)
    unet = UNet2DConditionModel().to(device)
    unet.requires_grad_(False)
    L.inject_trainable_lora(unet, r=args.rank)
    with torch.no_grad():
        for m in unet.modules():
            if isinstance(m, L.LoraInjectedLinear):
                m.lora_up.weight.normal_(0, 0.02, generator=g)
    L.fuse_sibling_projections(unet)
    b = make_batches(1, 2, args.res, 7, device)[0]

    def sweep_pass(_):
        for i in range(51):
            L.tune_lora_scale(unet, i / 50.0)
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                unet(b["latents"], b["t"], b["ctx"])
    sweep_pass(0)
    t_sweep = _time_cuda(sweep_pass, 2, device)
    L.tune_lora_scale(unet, 0.0)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        t_fwd0 = _time_cuda(lambda _: unet(b["latents"], b["t"], b["ctx"]), 10, device)
    del unet
    torch.cuda.empty_cache()
    sweep = {"alphas": 51, "ms_per_alpha_unet_forward_batch2": t_sweep / 51 * 1e3,
             "ms_per_forward_scale0": t_fwd0 * 1e3,
             "alternative_merge_per_alpha_ms": merge["fp16"]["ms"],
             "what": "51 x [tune_lora_scale(unet, i/50) + one UNet forward, batch 2, bf16 autocast, no grad]; the scale is "
                     "a kernel argument, nothing is re-packed or re-merged between alphas"}
    # ---- SVD distillation (cli_svd.py:29-111) on SD-1.5-shaped weight differences ----
    from diffusion_finetuning_b200 import artefacts
    uniq = {}
    for sh in shapes:
        uniq[sh] = uniq.get(sh, 0) + 1
    deltas = {}
    for (N, K) in uniq:
        ut = torch.randn(N, args.rank, device=device, generator=g) * 0.05
        dt = torch.randn(args.rank, K, device=device, generator=g) * 0.3
        deltas[(N, K)] = (ut @ dt + 1e-3 * torch.randn(N, K, device=device, generator=g)).half()
    svd_rows = []
    total_ours = total_full = 0.0
    for (N, K), cnt in sorted(uniq.items(), key=lambda kv: kv[0][0] * kv[0][1]):
        d = deltas[(N, K)]
        zero = torch.zeros_like(d)
        artefacts.svd_distill_weights([d], [zero], args.rank, 0.99, device)          # warm-up (cuSOLVER handles)
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        artefacts.svd_distill_weights([d], [zero], args.rank, 0.99, device)
        torch.cuda.synchronize(device)
        ours = time.perf_counter() - t0
        t0 = time.perf_counter()
        torch.linalg.svd(d.float())                    # the reference's call on the same GPU: full_matrices=True
        torch.cuda.synchronize(device)
        full = time.perf_counter() - t0
        svd_rows.append({"N": N, "K": K, "count": cnt, "ours_ms": ours * 1e3, "reference_call_same_gpu_ms": full * 1e3})
        total_ours += ours * cnt
        total_full += full * cnt
    svd = {"matrices": len(shapes), "wall_s_ours": total_ours, "wall_s_reference_call_same_gpu": total_full,
           "speedup": total_full / total_ours, "by_shape": svd_rows,
           "what": "per distinct shape: one timed svd_distill_weights (cuSOLVER gesvd economy + scale/clamp epilogue kernel) "
                   "vs one torch.linalg.svd(full_matrices=True) as cli_svd.py:69 calls it, on the same GPU; totals = "
                   "time x number of layers of that shape"}
    clk = clocks.stop()
    line = {"metric": METRIC5, "value": merge["fp32"]["gbs"], "unit": "GB/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": merge["fp32"]["ms"], "higher_is_better": True, "scaling": "replicas only",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"sd15_merge_144_unet_plus_48_clip_matrices_r{args.rank}_alpha_sweep_svd",
                       "l2": "inputs larger than L2: one merge pass reads and writes 1.77 GB (fp32) / 0.88 GB (16-bit)"},
            "gpu_launches": launches, "clocks": clk,
            "roofline": {"kernel": "lora_merge_stream_kernel (fp32 weights; bulk-copy ring, one launch for all 192 matrices)",
                         "bound": "hbm", "achieved": merge["fp32"]["gbs"],
                         "peak": pk["hbm"], "unit": "GB/s", "frac": merge["fp32"]["frac_hbm"], "traffic": _merge_traffic(),
                         "peak_source": pk["src"],
                         "fp16": {"achieved": merge["fp16"]["gbs"], "frac": merge["fp16"]["frac_hbm"]},
                         "bf16": {"achieved": merge["bf16"]["gbs"], "frac": merge["bf16"]["frac_hbm"]}},
            "merge": merge, "scale_sweep": sweep, "svd_distill": svd}
    # ---- e2e: the public call on HOST-resident fp32 weights (cli_lora_add merges a CPU pipeline, cli_lora_add.py:74-78):
    # weights staged host -> device, one merge launch, results copied back, inside the timed region ----
    gh = torch.Generator().manual_seed(1)
    hw = [(torch.randn(N, K, generator=gh) * 0.03).pin_memory() for N, K in shapes]
    hu = [torch.randn(N, args.rank, generator=gh) * 0.02 for N, _ in shapes]
    hd = [torch.randn(args.rank, K, generator=gh) * 0.02 for _, K in shapes]
    ops.merge_lora_weights(hw[:4], hu[:4], hd[:4], 0.5)
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    outs = ops.merge_lora_weights(hw, hu, hd, 0.5)
    torch.cuda.synchronize(device)
    t_e2e = time.perf_counter() - t0
    wbytes = sum(w.numel() * 4 for w in hw)
    line["e2e"] = {"value": merge["fp32"]["bytes"] / t_e2e / 1e9, "unit": "GB/s", "h2d_bytes_per_step": wbytes,
                   "d2h_bytes_per_step": wbytes, "ms": t_e2e * 1e3,
                   "what": "ops.merge_lora_weights on host tensors: H2D + one launch + D2H of all 192 matrices"}
    assert outs[0].device.type == "cpu"
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = config5_cpu(args)
    _emit(line)
    return 0


def _merge_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one fp32 whole-model merge launch, from the committed ncu capture."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("lora_merge_fp32_bytes_per_launch")
    except (OSError, ValueError):
        return None


def config5_cpu(args, seconds=20.0):
    """The reference's CPU merge (cli_lora_add.py:74-78: `.to("cpu")` then lora.py:423 per layer) over the 192 fp32
    matrices in walk order (stops early if it would run past `seconds`)."""
    from oracle import build_ref
    from oracle import lora_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    ref = build_ref.load()
    unet_shapes, clip_shapes = _sd15_merge_shapes()
    g = torch.Generator().manual_seed(0)
    done, nbytes, spent = 0, 0, 0.0
    for N, K in unet_shapes + clip_shapes:
        W = torch.randn(N, K, generator=g) * 0.03
        up, down = torch.randn(N, args.rank, generator=g) * 0.02, torch.randn(args.rank, K, generator=g) * 0.02
        if ref is not None:                     # lora.py:410-424 itself, on a one-Linear stand-in block
            holder = type("CrossAttention", (torch.nn.Module,), {})()
            holder.to_q = torch.nn.Linear(K, N, bias=False)
            holder.to_q.weight = torch.nn.Parameter(W)
            t1 = time.perf_counter()
            ref.weight_apply_lora(holder, [up, down], target_replace_module={"CrossAttention"}, alpha=0.5)
        else:
            t1 = time.perf_counter()
            O.merge_weight(W, up, down, 0.5)
        spent += time.perf_counter() - t1
        done += 1
        nbytes += 2 * 4 * N * K + 4 * args.rank * (N + K)
        if spent > seconds:
            break
    return {"value": nbytes / spent / 1e9, "unit": "GB/s", "cores": torch.get_num_threads(),
            "kind": "reference" if ref is not None else "port",
            "sample": f"weight_apply_lora (lora.py:410-424) on {done} of the 192 fp32 matrices, "
                      f"{nbytes / 1e9:.2f} GB of algorithmic traffic, {spent:.2f} s"}


def run_config5_reference(args):
    cb = config5_cpu(args, seconds=20.0)
    _emit({"impl": "reference", "metric": METRIC5, "value": cb["value"], "unit": "GB/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True,
           "scaling": "replicas only", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": f"sd15_merge_144_unet_plus_48_clip_matrices_r{args.rank}_alpha_sweep_svd"},
           "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "GB/s", "h2d_bytes_per_step": 0,
                                        "d2h_bytes_per_step": 0}, "gpu_launches": 0})
    return 0


def dp_replica_check(tr, device):
    """After the timed steps: the LoRA parameters of every rank must be BIT-identical (same averaged gradients,
    same optimizer) — MAX and MIN over ranks of the int32 view agree element by element."""
    import torch.distributed as dist
    flat = torch.cat([p.detach().reshape(-1).float() for p in tr.arena.params]).view(torch.int32)
    hi, lo = flat.clone(), flat.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    diff = hi != lo
    differing = int(diff.sum().item())
    out = {"params_bit_identical_across_ranks": differing == 0, "differing_elements": differing,
           "elements": flat.numel(), "checksum_i64": int(flat.to(torch.int64).sum().item())}
    if differing:          # which buckets of the gradient arena the diverged parameters belong to
        per_bucket = [0] * len(tr.arena.buckets)
        off = 0
        for p in tr.arena.params:
            n = p.numel()
            per_bucket[tr.arena._slot[id(p)][2]] += int(diff[off:off + n].sum().item())
            off += n
        out["differing_per_bucket"] = per_bucket
        out["bucket_sizes"] = [b.hi - b.lo for b in tr.arena.buckets]
        names = {id(p): n for n, p in tr.unet.named_parameters()} if hasattr(tr, "unet") else {}
        bad, off = [], 0
        for p in tr.arena.params:
            n = p.numel()
            k = int(diff[off:off + n].sum().item())
            if k:
                bad.append({"param": names.get(id(p), "?"), "shape": list(p.shape), "differing": k,
                            "bucket": tr.arena._slot[id(p)][2]})
            off += n
        out["differing_params"] = bad[:16]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "stock_gpu"])
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--res", type=int, default=64)
    ap.add_argument("--rank", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stock", action="store_true",
                    help="skip the stock-PyTorch-on-the-same-GPU arm that the b200 line carries as `stock_gpu`")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5],
                    help="BASELINE config: 2 = attention q/k/v/out r=4 (the headline), 3 = extended r=16 with resnet "
                         "convs, 4 = PTI joint UNet + CLIP text-encoder LoRA + placeholder rows, 5 = merge / scale "
                         "sweep / SVD distillation (its own metric)")
    ap.add_argument("--no-graph", action="store_true",
                    help="enqueue every step from Python instead of replaying the captured CUDA graph")
    ap.add_argument("--arena-optimizer", action="store_true",
                    help="dp.ArenaAdamW (clip + AdamW over the flat arenas, one C-ABI call) instead of "
                         "arena.clip_grad_norm_ + torch fused AdamW; measured equal within run-to-run noise")
    ap.add_argument("--no-fuse", action="store_true",
                    help="run q/k/v LoRA layers one launch each instead of fused per attention block")
    ap.add_argument("--memory-format", default="nchw", choices=["nchw", "channels_last"],
                    help="memory format of the frozen stock-PyTorch UNet (the reference scripts use nchw)")
    ap.add_argument("--eager-profile", action="store_true",
                    help="time our launches in eager steps (event pair per launch) instead of inside the replayed graph")
    ap.add_argument("--breakdown", default=None, help="write the per-kernel breakdown JSON here")
    args = ap.parse_args()
    if args.config == 3 and args.rank == 4:
        args.rank = 16
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the b200lora kernels are the only compute path")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    dist_on = world > 1
    if dist_on:
        dist.init_process_group("nccl", device_id=device)
    import __graft_entry__ as entry
    if rank == 0:
        entry.build()
    if dist_on:
        dist.barrier()
    from diffusion_finetuning_b200 import _native as nat
    from diffusion_finetuning_b200 import ops
    nat.lib()
    if args.config == 5:
        rc = run_config5(args, device, rank)
        _teardown(device, dist_on)
        return rc

    pool = make_batches(4, args.batch, args.res, 1234 + rank, device)
    host_pool = make_batches(4, args.batch, args.res, 4321 + rank, None, pin=True)
    if args.impl == "stock_gpu":
        # evidence line for profiles/: the reference composition in stock PyTorch on the same GPU(s)
        if args.config == 4:
            raise SystemExit("--impl stock_gpu covers configs 2 and 3")
        st = stock_arm(args, device, dist_on, pool, host_pool, graph=not args.no_graph)
        if rank == 0:
            samples = args.batch * world * args.steps
            _emit({"impl": "stock_gpu", "metric": METRIC, "value": samples / st["t_dev"], "unit": UNIT,
                   "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                   "ms_per_step": st["t_dev"] / args.steps * 1e3, "higher_is_better": True, "dtype": "bf16",
                   "data": "synthetic", "config": dict(config_dict(args, args.batch), step_launch=st["launch"]),
                   "e2e": {"value": samples / st["t_e2e"], "unit": UNIT}, "gpu_launches": 0,
                   "lora_layers": st["lora_layers"], "parallel": st["parallel"],
                   "note": "lora.py:49-50 as 3 F.linear + mul + add (conv: 3 convs) in stock PyTorch, "
                           "autograd backward, same UNet / optimizer / clip"})
        _teardown(device, dist_on)
        return 0

    if args.config == 4:
        tr = PtiTrainer(device, args.rank, args.batch)
        pool = tr.make_batches(4, args.res, 1234 + rank, device)
        host_pool = tr.make_batches(4, args.res, 4321 + rank, None)
    else:
        tr = Trainer(device, args.rank, channels_last=args.memory_format == "channels_last", fuse=not args.no_fuse,
                     extended=args.config == 3, capturable=not args.no_graph,
                     arena_optimizer=args.arena_optimizer)
    if not args.no_graph and tr.graph_capable:
        try:
            tr.capture(pool[0])                 # 3 eager warm-up steps + capture
        except Exception as exc:                # same kernels either way: only the launch mechanism changes
            print(f"[bench] CUDA-graph capture failed ({type(exc).__name__}: {exc}); enqueueing steps from Python",
                  file=sys.stderr)
            tr.graphed = None
            args.no_graph = True
            torch.cuda.synchronize(device)
    if tr.graphed is None:
        args.no_graph = True
    for i in range(args.warmup):
        tr.step(pool[i % 4])
    tr.step_from_host(host_pool[0])
    torch.cuda.synchronize(device)

    # ---- timed region 1: batch resident in HBM ----
    prof = LaunchProfiler()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    l0 = nat.LAUNCHES
    t_dev = timed(lambda i: tr.step(pool[i % 4]), args.steps, dist_on, device)
    launches = nat.LAUNCHES - l0
    if tr.graphed is not None:                  # replays do not pass through Python: count what was captured
        launches = tr.graphed.kernel_launches * args.steps
    # ---- timed region 2: end to end from pinned host memory, loss read back every step ----
    t_e2e = timed(lambda i: tr.step_from_host(host_pool[i % 4]), args.steps, dist_on, device)
    # ---- region 3: the same K steps with CUDA events around every launch of our kernels (roofline);
    # kept out of region 1 because an event between two launches defeats their programmatic overlap ----
    timing_mode = "eager steps, one CUDA-event pair per launch"
    t_prof = None
    exposure = []
    if tr.graphed is not None and not args.eager_profile:
        # the step captured a second time WITH external event-record nodes around our launches (and around the
        # join of the gradient exchange): each replay stamps them on the device, so the durations are those of
        # the kernels inside the replayed step, at every N
        from diffusion_finetuning_b200 import dp
        gprof = LaunchProfiler(external=True)
        try:
            ops.PROFILER = gprof
            tr.arena.exposure_probe = gprof._event if dist_on else None
            pg = dp.GraphedStep(tr.eager_step, pool[0], warmup=1,
                                before_capture=lambda: (gprof.reset_capture(), tr.arena.exposure_events.clear()))
            ops.PROFILER = None
            tr.arena.exposure_probe = None
            pg(pool[0])
            torch.cuda.synchronize(device)
            t_prof = 0.0
            for i in range(args.steps):
                t_prof += timed(lambda _: pg(pool[i % 4]), 1, dist_on, device)
                gprof.harvest()
                exposure += [a.elapsed_time(b) * 1e3 for a, b in tr.arena.exposure_events]
            prof = gprof
            timing_mode = "external CUDA-event nodes inside the replayed step graph"
            del pg
        except Exception as exc:
            ops.PROFILER = None
            tr.arena.exposure_probe = None
            t_prof = None
            print(f"[bench] in-graph launch timing unavailable ({type(exc).__name__}: {exc}); "
                  "timing eager launches instead", file=sys.stderr)
            torch.cuda.synchronize(device)
    if t_prof is None:
        ops.PROFILER = prof
        t_prof = timed(lambda i: tr.eager_step(pool[i % 4]), args.steps, dist_on, device)
        ops.PROFILER = None
    clk = clocks.stop() if rank == 0 else None

    dp_check = dp_replica_check(tr, device) if dist_on else None
    exposed_us = None
    if exposure:
        ex = torch.tensor([sum(exposure) / len(exposure)], device=device)
        if dist_on:
            dist.all_reduce(ex, op=dist.ReduceOp.MAX)
        exposed_us = ex.item()
    payload, nbuckets, n_layers = tr.arena.payload_bytes, len(tr.arena.buckets), tr.n_layers
    # every rank is done with this trainer: drop its graph (it holds NCCL work) and its memory
    tr.release()
    del tr
    import gc
    gc.collect()
    torch.cuda.synchronize(device)
    torch.cuda.empty_cache()

    # ---- the bar on the same GPU(s): lora.py:49-50 in stock PyTorch, graph-replayed like the repo arm ----
    stock = None
    if not args.no_stock and args.config in (2, 3):
        try:
            stock = stock_arm(args, device, dist_on, pool, host_pool, graph=not args.no_graph)
        except Exception as exc:
            print(f"[bench] stock arm failed ({type(exc).__name__}: {exc})", file=sys.stderr)
            torch.cuda.synchronize(device)
    _teardown(device, dist_on)
    if rank != 0:
        return 0

    pk = peaks()
    summ = prof.summary()
    gemm = {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0}
    conv = {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0}
    for k, v in summ.items():
        if k in ("fwd", "bwd_dx") or k.startswith("conv_"):
            for f in gemm:
                gemm[f] += v.get(f, 0)
                if k.startswith("conv_"):
                    conv[f] += v.get(f, 0)
    step_ms = t_prof / args.steps * 1e3      # shares below are of the profiled steps
    ach_tf = gemm["flops"] / (gemm["ms"] * 1e-3) / 1e12 if gemm["ms"] else 0.0
    roofline = {"kernel": "lora_gemm_tc2_kernel (fused LoRA linear forward + backward-dX launches"
                          + (", and the implicit-GEMM conv launches of the same kernel" if conv["launches"] else "")
                          + ", CTA-pair tcgen05)",
                "bound": "tensor", "achieved": ach_tf, "peak": pk["tf_sustained"], "unit": "TFLOP/s",
                "frac": ach_tf / pk["tf_sustained"], "peak_source": pk["src"] + " (sustained bf16)",
                "traffic": None,
                "launches_per_step": gemm["launches"] / args.steps,
                "avg_launch_us": gemm["ms"] * 1e3 / max(1, gemm["launches"]),
                "share_of_step": gemm["ms"] / args.steps / step_ms,
                "profiled_ms_per_step": step_ms, "launch_timing": timing_mode,
                "achieved_hbm_gbs": gemm["bytes"] / (gemm["ms"] * 1e-3) / 1e9 if gemm["ms"] else 0.0}
    if conv["launches"]:
        # the same kernel on its implicit-GEMM (TMA im2col) geometry: the resnet convs of the extended injection
        roofline["conv_launches"] = {"launches_per_step": conv["launches"] / args.steps,
                                     "avg_launch_us": conv["ms"] * 1e3 / conv["launches"],
                                     "achieved": conv["flops"] / (conv["ms"] * 1e-3) / 1e12,
                                     "frac": conv["flops"] / (conv["ms"] * 1e-3) / 1e12 / pk["tf_sustained"],
                                     "share_of_step": conv["ms"] / args.steps / step_ms}
        lin = {f: gemm[f] - conv[f] for f in gemm}
        roofline["linear_launches"] = {"launches_per_step": lin["launches"] / args.steps,
                                       "avg_launch_us": lin["ms"] * 1e3 / max(1, lin["launches"]),
                                       "achieved": lin["flops"] / (lin["ms"] * 1e-3) / 1e12 if lin["ms"] else 0.0,
                                       "frac": (lin["flops"] / (lin["ms"] * 1e-3) / 1e12 / pk["tf_sustained"])
                                       if lin["ms"] else 0.0,
                                       "share_of_step": lin["ms"] / args.steps / step_ms}
    if "grads" in summ:
        s = summ["grads"]
        roofline["grads_kernel"] = {"bound": "hbm", "achieved": s["bytes"] / (s["ms"] * 1e-3) / 1e9, "peak": pk["hbm"],
                                    "unit": "GB/s", "frac": s["bytes"] / (s["ms"] * 1e-3) / 1e9 / pk["hbm"],
                                    "share_of_step": s["ms"] / args.steps / step_ms}
    roofline["by_shape"] = prof.by_shape(pk["tf_sustained"], pk["hbm"])
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get("lora_gemm_bytes_per_launch")
        except (ValueError, OSError):
            pass
    step_ms = t_dev / args.steps * 1e3
    bytes_in = sum(v.numel() * v.element_size() for v in host_pool[0].values())
    samples = args.batch * world * args.steps
    line = {"metric": METRIC, "value": samples / t_dev, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": config_dict(args, args.batch),
            "e2e": {"value": samples / t_e2e, "unit": UNIT, "h2d_bytes_per_step": bytes_in, "d2h_bytes_per_step": 4},
            "gpu_launches": launches, "clocks": clk, "roofline": roofline,
            "grad_allreduce": {"payload_bytes": payload, "buckets": nbuckets,
                               "exposed_us_per_step": exposed_us,
                               "exposed_how": "event before the join of the exchange stream vs after it, inside the "
                                              "replayed step graph; max over ranks" if exposed_us is not None else None},
            "lora_layers": n_layers}
    if dp_check is not None:
        line["dp_check"] = dp_check
    if stock is not None:
        line["stock_gpu"] = {"value": samples / stock["t_dev"], "unit": UNIT,
                             "ms_per_step": stock["t_dev"] / args.steps * 1e3,
                             "e2e": samples / stock["t_e2e"], "step_launch": stock["launch"],
                             "parallel": stock["parallel"], "speedup_value": stock["t_dev"] / t_dev,
                             "speedup_e2e": stock["t_e2e"] / t_e2e,
                             "what": "lora.py:49-50 as 3 F.linear + mul + add per layer in stock PyTorch on the same "
                                     "GPU(s): same UNet, batches, AdamW, clip"}
    if args.breakdown:
        os.makedirs(os.path.dirname(os.path.abspath(args.breakdown)), exist_ok=True)
        json.dump({"per_kind": summ, "step_ms": step_ms, "steps": args.steps}, open(args.breakdown, "w"), indent=1)
    if world == 1 and not args.no_cpu_baseline:
        del pool
        cb = cpu_leg(3, 1, 1, args.res, args.rank, args.config)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
    _emit(line)
    if dp_check is not None and not dp_check["params_bit_identical_across_ranks"]:
        print("[bench] DP replicas diverged: LoRA parameters differ across ranks", file=sys.stderr)
        return 1
    return 0


def _emit(obj):
    """The ONE JSON line goes to the process's real stdout; everything else that lands on fd 1 while the
    run lasts (NCCL's version banner, library chatter) was re-pointed at stderr by _quiet_stdout()."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(obj) + "\n").encode())


_REAL_STDOUT = None


def _quiet_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def _teardown(device, dist_on):
    """Orderly end of a GPU arm: every captured graph has been dropped by the caller (a graph holds NCCL work;
    a communicator destroyed under a live graph blocks), so collect them, drain the device, let every rank
    arrive, destroy the process group, and RETURN — the interpreter exits normally and exit handlers run."""
    import gc
    import torch.distributed as dist
    gc.collect()
    torch.cuda.synchronize(device)
    if dist_on and dist.is_initialized():
        dist.barrier()
        torch.cuda.synchronize(device)
        dist.destroy_process_group()
    gc.collect()
    torch.cuda.empty_cache()


if __name__ == "__main__":
    _quiet_stdout()
    rc = main()
    sys.stdout.flush()
    sys.stderr.flush()
    sys.exit(int(rc or 0))

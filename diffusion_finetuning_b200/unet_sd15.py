"""SD-1.5-shaped conditional UNet in stock PyTorch — the CALLER of the hot path.

`diffusers` is not installed in this image and there is no network, so the
benchmark and the parity tests need a UNet of the same shape whose blocks carry the same CLASS
NAMES diffusers uses (`CrossAttention`, `GEGLU`, `ResnetBlock2D`, ...): the reference's discovery
matches ancestors by class-name string (lora.py:96), so `inject_trainable_lora` walks this model
exactly as it walks `UNet2DConditionModel` and finds the same 144 Linear layers in the same order
(16 transformer blocks x [attn1 q,k,v,out; attn2 q,k,v,out; GEGLU proj]).

Nothing in this file is a hot-path kernel: convs, norms, attention softmax·V are stock
torch (cuDNN / SDPA), as the north star prescribes ("the rest of the UNet stays in stock
PyTorch").
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class CrossAttention(nn.Module):
    def __init__(self, query_dim, context_dim=None, heads=8, dim_head=64):
        super().__init__()
        inner = heads * dim_head
        context_dim = context_dim or query_dim
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner, bias=False)
        self.to_k = nn.Linear(context_dim, inner, bias=False)
        self.to_v = nn.Linear(context_dim, inner, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(0.0)])

    def forward(self, x, context=None):
        context = x if context is None else context
        b, t, _ = x.shape
        q, k, v = self.to_q(x), self.to_k(context), self.to_v(context)
        h = self.heads
        q = q.view(b, t, h, -1).transpose(1, 2)
        k = k.view(b, context.shape[1], h, -1).transpose(1, 2)
        v = v.view(b, context.shape[1], h, -1).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v)
        o = o.transpose(1, 2).reshape(b, t, -1)
        return self.to_out[1](self.to_out[0](o))


class GEGLU(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)

    def forward(self, x):
        x, gate = self.proj(x).chunk(2, dim=-1)
        return x * F.gelu(gate)


class FeedForward(nn.Module):
    def __init__(self, dim, mult=4):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * mult), nn.Dropout(0.0), nn.Linear(dim * mult, dim)])

    def forward(self, x):
        for m in self.net:
            x = m(x)
        return x


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim, heads, dim_head, context_dim):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = CrossAttention(dim, None, heads, dim_head)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = CrossAttention(dim, context_dim, heads, dim_head)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)

    def forward(self, x, context):
        x = self.attn1(self.norm1(x)) + x
        x = self.attn2(self.norm2(x), context) + x
        return self.ff(self.norm3(x)) + x


class Transformer2DModel(nn.Module):
    def __init__(self, channels, heads, context_dim, groups=32):
        super().__init__()
        self.norm = nn.GroupNorm(groups, channels, eps=1e-6)
        self.proj_in = nn.Conv2d(channels, channels, 1)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(channels, heads, channels // heads, context_dim)])
        self.proj_out = nn.Conv2d(channels, channels, 1)

    def forward(self, x, context):
        b, c, h, w = x.shape
        res = x
        x = self.proj_in(self.norm(x))
        x = x.permute(0, 2, 3, 1).reshape(b, h * w, c)
        for blk in self.transformer_blocks:
            x = blk(x, context)
        x = x.reshape(b, h, w, c).permute(0, 3, 1, 2)
        return self.proj_out(x) + res


class ResnetBlock2D(nn.Module):
    def __init__(self, c_in, c_out, temb_dim, groups=32):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, c_in, eps=1e-5)
        self.conv1 = nn.Conv2d(c_in, c_out, 3, padding=1)
        self.time_emb_proj = nn.Linear(temb_dim, c_out)
        self.norm2 = nn.GroupNorm(groups, c_out, eps=1e-5)
        self.dropout = nn.Dropout(0.0)
        self.conv2 = nn.Conv2d(c_out, c_out, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(c_in, c_out, 1) if c_in != c_out else None

    def forward(self, x, temb):
        h = self.conv1(F.silu(self.norm1(x)))
        h = h + self.time_emb_proj(F.silu(temb))[:, :, None, None]
        h = self.conv2(self.dropout(F.silu(self.norm2(h))))
        if self.conv_shortcut is not None:
            x = self.conv_shortcut(x)
        return x + h


class Downsample2D(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, stride=2, padding=1)

    def forward(self, x):
        return self.conv(x)


class Upsample2D(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class TimestepEmbedding(nn.Module):
    def __init__(self, c_in, c_out):
        super().__init__()
        self.linear_1 = nn.Linear(c_in, c_out)
        self.linear_2 = nn.Linear(c_out, c_out)

    def forward(self, t):
        return self.linear_2(F.silu(self.linear_1(t)))


def _sinusoid(t, dim):
    half = dim // 2
    freqs = torch.exp(-math.log(10000.0) * torch.arange(half, device=t.device, dtype=torch.float32) / half)
    args = t.float()[:, None] * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


class _Stage(nn.Module):
    """One down/up stage: `resnets[i]` (+ `attentions[i]` when cross-attention) and an optional
    resampler, named like diffusers' (Cross)Attn{Down,Up}Block2D."""

    def __init__(self, res_io, temb_dim, heads, context_dim, with_attn, resample, groups):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock2D(i, o, temb_dim, groups) for i, o in res_io])
        self.attentions = nn.ModuleList(
            [Transformer2DModel(o, heads, context_dim, groups) for _, o in res_io]) if with_attn else None
        c = res_io[-1][1]
        self.resample = None if resample is None else (Downsample2D(c) if resample == "down" else Upsample2D(c))


class UNet2DConditionModel(nn.Module):
    """SD-1.5 layout by default: block_out_channels (320, 640, 1280, 1280), 2 resnets per block,
    cross-attention in the first three down stages / last three up stages, 8 heads,
    cross_attention_dim 768.  `tiny()` gives the BASELINE config-1 sized model."""

    def __init__(self, in_channels=4, out_channels=4, block_out_channels=(320, 640, 1280, 1280),
                 layers_per_block=2, cross_attention_dim=768, attention_heads=8,
                 attn_stages=(True, True, True, False), norm_groups=32):
        super().__init__()
        ch = list(block_out_channels)
        temb = ch[0] * 4
        self.time_dim = ch[0]
        self.conv_in = nn.Conv2d(in_channels, ch[0], 3, padding=1)
        self.time_embedding = TimestepEmbedding(ch[0], temb)
        nstage = len(ch)
        skips = [ch[0]]
        self.down_blocks = nn.ModuleList()
        c_prev = ch[0]
        for i, c in enumerate(ch):
            io = []
            for _ in range(layers_per_block):
                io.append((c_prev, c))
                c_prev = c
                skips.append(c)
            last = i == nstage - 1
            if not last:
                skips.append(c)
            self.down_blocks.append(_Stage(io, temb, attention_heads, cross_attention_dim,
                                           attn_stages[i], None if last else "down", norm_groups))
        self.mid_resnet1 = ResnetBlock2D(c_prev, c_prev, temb, norm_groups)
        self.mid_attn = Transformer2DModel(c_prev, attention_heads, cross_attention_dim, norm_groups)
        self.mid_resnet2 = ResnetBlock2D(c_prev, c_prev, temb, norm_groups)
        self.up_blocks = nn.ModuleList()
        for i, c in enumerate(reversed(ch)):
            io = []
            for _ in range(layers_per_block + 1):
                io.append((c_prev + skips.pop(), c))
                c_prev = c
            last = i == nstage - 1
            self.up_blocks.append(_Stage(io, temb, attention_heads, cross_attention_dim,
                                         list(reversed(attn_stages))[i], None if last else "up",
                                         norm_groups))
        self.conv_norm_out = nn.GroupNorm(norm_groups, ch[0], eps=1e-5)
        self.conv_out = nn.Conv2d(ch[0], out_channels, 3, padding=1)

    @classmethod
    def tiny(cls):
        """BASELINE config 1: block_out_channels (32, 64), 1 layer per block, ctx dim 32."""
        return cls(block_out_channels=(32, 64), layers_per_block=1, cross_attention_dim=32,
                   attention_heads=4, attn_stages=(True, False), norm_groups=8)

    def forward(self, sample, timestep, encoder_hidden_states):
        temb = self.time_embedding(_sinusoid(timestep, self.time_dim).to(sample.dtype))
        x = self.conv_in(sample)
        skips = [x]
        for st in self.down_blocks:
            for j, res in enumerate(st.resnets):
                x = res(x, temb)
                if st.attentions is not None:
                    x = st.attentions[j](x, encoder_hidden_states)
                skips.append(x)
            if st.resample is not None:
                x = st.resample(x)
                skips.append(x)
        x = self.mid_resnet1(x, temb)
        x = self.mid_attn(x, encoder_hidden_states)
        x = self.mid_resnet2(x, temb)
        for st in self.up_blocks:
            for j, res in enumerate(st.resnets):
                x = res(torch.cat([x, skips.pop()], dim=1), temb)
                if st.attentions is not None:
                    x = st.attentions[j](x, encoder_hidden_states)
            if st.resample is not None:
                x = st.resample(x)
        return self.conv_out(F.silu(self.conv_norm_out(x)))

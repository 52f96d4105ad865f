"""Random-init, plain-PyTorch stand-ins for the Stable Diffusion pieces the hot path touches.

`diffusers==0.8.0` (reference `requirements.txt:4`) is not installable offline, and there
are no checkpoints, so throughput is measured on random-init weights of the published
architecture shapes.  These modules are NOT the product: they are
"the rest of the UNet stays in stock PyTorch".  They expose exactly the attribute contract
the reference's hook relies on (`ptp_utils.py:472-573`): attention modules whose class name
is `CrossAttention` with `to_q/to_k/to_v` (bias-free), `to_out = ModuleList([Linear,
Dropout])`, `heads`, `scale`, `reshape_heads_to_batch_dim`, `reshape_batch_dim_to_heads`
and `forward(x, context=None, mask=None)`, living under children whose name contains "up".

The same classes serve the CPU oracle (the unmodified reference hook attaches to them) and
the CUDA path (our `register_attention_control` attaches to them).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# attention building blocks (diffusers-0.8.0-shaped)
# --------------------------------------------------------------------------------------
class CrossAttention(nn.Module):
    """Class name and attribute names are part of the contract (`ptp_utils.py:556,474-492`)."""

    def __init__(self, query_dim: int, context_dim: Optional[int], heads: int, dim_head: int):
        super().__init__()
        inner = heads * dim_head
        context_dim = query_dim if context_dim is None else context_dim
        self.heads = heads
        self.dim_head = dim_head
        self.scale = dim_head ** -0.5
        self.to_q = nn.Linear(query_dim, inner, bias=False)
        self.to_k = nn.Linear(context_dim, inner, bias=False)
        self.to_v = nn.Linear(context_dim, inner, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(0.0)])

    def reshape_heads_to_batch_dim(self, t: torch.Tensor) -> torch.Tensor:
        b, n, c = t.shape
        h = self.heads
        return t.reshape(b, n, h, c // h).permute(0, 2, 1, 3).reshape(b * h, n, c // h)

    def reshape_batch_dim_to_heads(self, t: torch.Tensor) -> torch.Tensor:
        bh, n, d = t.shape
        h = self.heads
        return t.reshape(bh // h, h, n, d).permute(0, 2, 1, 3).reshape(bh // h, n, h * d)

    def forward(self, x, context=None, mask=None):
        # (b, h, n, d) views of the projections, no copies: the 4-D layout is what lets
        # scaled_dot_product_attention pick its fused kernel on the GPU (the (b*h, n, d) layout of
        # reshape_heads_to_batch_dim sends it down the math path: a 2.7 GB fp32 probability tensor per
        # 64^2 self-attention, 40 ms of bmm / softmax / copies per optimisation step on B200)
        b, n, _ = x.shape
        h = self.heads
        ctx = x if context is None else context
        if context is not None and b > 1 and context.stride(0) == 0:
            ctx = context[:1]       # context.expand(B, -1, -1) (`ptp_utils.py:229`): project the one context once
        q = self.to_q(x).view(b, n, h, -1).transpose(1, 2)
        k = self.to_k(ctx).view(ctx.shape[0], ctx.shape[1], h, -1).transpose(1, 2).expand(b, -1, -1, -1)
        v = self.to_v(ctx).view(ctx.shape[0], ctx.shape[1], h, -1).transpose(1, 2).expand(b, -1, -1, -1)
        out = F.scaled_dot_product_attention(q, k, v, scale=self.scale)
        return self.to_out[0](out.transpose(1, 2).reshape(b, n, -1))


class GEGLUFeedForward(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.proj = nn.Linear(dim, dim * 8)
        self.out = nn.Linear(dim * 4, dim)

    def forward(self, x):
        a, gate = self.proj(x).chunk(2, dim=-1)
        return self.out(a * F.gelu(gate))


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim: int, heads: int, dim_head: int, context_dim: int):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = CrossAttention(dim, None, heads, dim_head)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = CrossAttention(dim, context_dim, heads, dim_head)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = GEGLUFeedForward(dim)

    def forward(self, x, context):
        x = x + self.attn1(self.norm1(x))
        x = x + self.attn2(self.norm2(x), context=context)
        x = x + self.ff(self.norm3(x))
        return x


class Transformer2D(nn.Module):
    def __init__(self, channels: int, heads: int, dim_head: int, context_dim: int,
                 depth: int = 1, linear_proj: bool = False):
        super().__init__()
        self.norm = nn.GroupNorm(32, channels, eps=1e-6)
        self.linear_proj = linear_proj
        if linear_proj:
            self.proj_in = nn.Linear(channels, channels)
            self.proj_out = nn.Linear(channels, channels)
        else:
            self.proj_in = nn.Conv2d(channels, channels, 1)
            self.proj_out = nn.Conv2d(channels, channels, 1)
        self.blocks = nn.ModuleList(
            [BasicTransformerBlock(channels, heads, dim_head, context_dim) for _ in range(depth)])

    def forward(self, x, context):
        b, c, hh, ww = x.shape
        res = x
        x = self.norm(x)
        if self.linear_proj:
            x = self.proj_in(x.permute(0, 2, 3, 1).reshape(b, hh * ww, c))
        else:
            x = self.proj_in(x).permute(0, 2, 3, 1).reshape(b, hh * ww, c)
        for blk in self.blocks:
            x = blk(x, context)
        if self.linear_proj:
            x = self.proj_out(x).reshape(b, hh, ww, c).permute(0, 3, 1, 2)
        else:
            x = self.proj_out(x.reshape(b, hh, ww, c).permute(0, 3, 1, 2))
        return x + res


class ResnetBlock(nn.Module):
    def __init__(self, cin: int, cout: int, temb_dim: Optional[int], groups: int = 32, eps: float = 1e-5):
        super().__init__()
        self.norm1 = nn.GroupNorm(groups, cin, eps=eps)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.temb = nn.Linear(temb_dim, cout) if temb_dim else None
        self.norm2 = nn.GroupNorm(groups, cout, eps=eps)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.skip = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x, temb=None):
        h = self.conv1(F.silu(self.norm1(x)))
        if self.temb is not None:
            h = h + self.temb(F.silu(temb))[:, :, None, None]
        h = self.conv2(F.silu(self.norm2(h)))
        return h + (x if self.skip is None else self.skip(x))


# --------------------------------------------------------------------------------------
# UNet
# --------------------------------------------------------------------------------------
@dataclass
class UNetConfig:
    name: str = "sd21-base"
    in_channels: int = 4
    block_channels: Tuple[int, ...] = (320, 640, 1280, 1280)
    # per resolution level: does the down block (and the mirrored up block) carry cross-attention
    attn_levels: Tuple[bool, ...] = (True, True, True, False)
    layers_per_block: int = 2
    context_dim: int = 1024
    # heads per level (dim_head = C / heads)
    heads: Tuple[int, ...] = (5, 10, 20, 20)
    transformer_depth: Tuple[int, ...] = (1, 1, 1, 1)
    linear_proj: bool = True
    latent_side: int = 64


def unet_config(name: str) -> UNetConfig:
    """Published SD shapes plus two tiny shapes for CPU tests."""
    if name in ("sd21-base", "stabilityai/stable-diffusion-2-1-base"):
        return UNetConfig()
    if name in ("sd15", "runwayml/stable-diffusion-v1-5", "CompVis/stable-diffusion-v1-4"):
        return UNetConfig(name="sd15", context_dim=768, heads=(8, 8, 8, 8), linear_proj=False)
    if name == "sdxl":
        return UNetConfig(name="sdxl", block_channels=(320, 640, 1280), attn_levels=(False, True, True),
                          context_dim=2048, heads=(5, 10, 20), transformer_depth=(1, 2, 10),
                          linear_proj=True, latent_side=128)
    if name == "tiny-sd21":   # unequal heads per level, d=16; latent 16 -> layers @4^2,8^2,16^2
        return UNetConfig(name="tiny-sd21", block_channels=(32, 64, 128, 128), context_dim=48,
                          heads=(2, 4, 8, 8), linear_proj=True, latent_side=16)
    if name == "tiny-sd15":   # equal heads (the only shape the unmodified reference collect_maps accepts)
        return UNetConfig(name="tiny-sd15", block_channels=(32, 64, 128, 128), context_dim=40,
                          heads=(4, 4, 4, 4), linear_proj=False, latent_side=16)
    raise ValueError(f"unknown UNet config {name!r}")


class DownBlock(nn.Module):
    def __init__(self, cin, cout, temb, n, attn, heads, ctx, depth, linear, downsample):
        super().__init__()
        self.resnets = nn.ModuleList([ResnetBlock(cin if i == 0 else cout, cout, temb) for i in range(n)])
        self.attentions = nn.ModuleList(
            [Transformer2D(cout, heads, cout // heads, ctx, depth, linear) for _ in range(n)]) if attn else None
        self.down = nn.Conv2d(cout, cout, 3, stride=2, padding=1) if downsample else None

    def forward(self, x, temb, context, skips):
        for i, r in enumerate(self.resnets):
            x = r(x, temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
            skips.append(x)
        if self.down is not None:
            x = self.down(x)
            skips.append(x)
        return x


class UpBlock(nn.Module):
    def __init__(self, prev, cout, skip_channels: Sequence[int], temb, attn, heads, ctx, depth, linear, upsample):
        super().__init__()
        n = len(skip_channels)
        self.resnets = nn.ModuleList(
            [ResnetBlock((prev if i == 0 else cout) + skip_channels[i], cout, temb) for i in range(n)])
        self.attentions = nn.ModuleList(
            [Transformer2D(cout, heads, cout // heads, ctx, depth, linear) for _ in range(n)]) if attn else None
        self.up = nn.Conv2d(cout, cout, 3, padding=1) if upsample else None

    def forward(self, x, temb, context, skips):
        for i, r in enumerate(self.resnets):
            x = r(torch.cat([x, skips.pop()], dim=1), temb)
            if self.attentions is not None:
                x = self.attentions[i](x, context)
        if self.up is not None:
            x = self.up(F.interpolate(x, scale_factor=2.0, mode="nearest"))
        return x


class MidBlock(nn.Module):
    def __init__(self, c, temb, heads, ctx, depth, linear):
        super().__init__()
        self.res1 = ResnetBlock(c, c, temb)
        self.attn = Transformer2D(c, heads, c // heads, ctx, depth, linear)
        self.res2 = ResnetBlock(c, c, temb)

    def forward(self, x, temb, context):
        return self.res2(self.attn(self.res1(x, temb), context), temb)


def timestep_embedding(t: torch.Tensor, dim: int) -> torch.Tensor:
    half = dim // 2
    freqs = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=torch.float32, device=t.device) / half)
    args = t.float()[:, None] * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)   # flip_sin_to_cos=True


class UNet2DConditionStandIn(nn.Module):
    """`unet(latents, t, context)["sample"]` (`ptp_utils.py:227-229`)."""

    def __init__(self, cfg: UNetConfig):
        super().__init__()
        self.cfg = cfg
        ch = cfg.block_channels
        temb = ch[0] * 4
        self.conv_in = nn.Conv2d(cfg.in_channels, ch[0], 3, padding=1)
        self.time_embedding = nn.Sequential(nn.Linear(ch[0], temb), nn.SiLU(), nn.Linear(temb, temb))
        n = cfg.layers_per_block
        self.down_blocks = nn.ModuleList()
        skip_ch = [ch[0]]
        cin = ch[0]
        for lvl, cout in enumerate(ch):
            last = lvl == len(ch) - 1
            self.down_blocks.append(DownBlock(cin, cout, temb, n, cfg.attn_levels[lvl], cfg.heads[lvl],
                                              cfg.context_dim, cfg.transformer_depth[lvl], cfg.linear_proj,
                                              downsample=not last))
            skip_ch += [cout] * n + ([] if last else [cout])
            cin = cout
        self.mid_block = MidBlock(ch[-1], temb, cfg.heads[-1], cfg.context_dim,
                                  cfg.transformer_depth[-1], cfg.linear_proj)
        self.up_blocks = nn.ModuleList()
        prev = ch[-1]
        for lvl in reversed(range(len(ch))):
            cout = ch[lvl]
            sk = [skip_ch.pop() for _ in range(n + 1)]
            self.up_blocks.append(UpBlock(prev, cout, sk, temb, cfg.attn_levels[lvl], cfg.heads[lvl],
                                          cfg.context_dim, cfg.transformer_depth[lvl], cfg.linear_proj,
                                          upsample=lvl != 0))
            prev = cout
        self.conv_norm_out = nn.GroupNorm(32, ch[0])
        self.conv_out = nn.Conv2d(ch[0], cfg.in_channels, 3, padding=1)

    def forward(self, sample, timestep, encoder_hidden_states):
        temb = self.time_embedding(timestep_embedding(timestep.reshape(-1).expand(sample.shape[0]),
                                                      self.cfg.block_channels[0]).to(sample.dtype))
        x = self.conv_in(sample)
        skips = [x]
        for blk in self.down_blocks:
            x = blk(x, temb, encoder_hidden_states, skips)
        x = self.mid_block(x, temb, encoder_hidden_states)
        for blk in self.up_blocks:
            x = blk(x, temb, encoder_hidden_states, skips)
        x = self.conv_out(F.silu(self.conv_norm_out(x)))
        return {"sample": x}


# --------------------------------------------------------------------------------------
# VAE encoder (only `.encode(img)["latent_dist"].mean` is used: `ptp_utils.py:299-302`)
# --------------------------------------------------------------------------------------
class _LatentDist:
    def __init__(self, moments):
        self.mean, self.logvar = moments.chunk(2, dim=1)


class VAEAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.norm = nn.GroupNorm(32, c, eps=1e-6)
        self.q, self.k, self.v, self.o = (nn.Linear(c, c) for _ in range(4))

    def forward(self, x):
        b, c, h, w = x.shape
        t = self.norm(x).permute(0, 2, 3, 1).reshape(b, h * w, c)
        a = F.scaled_dot_product_attention(self.q(t)[:, None], self.k(t)[:, None], self.v(t)[:, None])[:, 0]
        return x + self.o(a).reshape(b, h, w, c).permute(0, 3, 1, 2)


class VAEEncoderStandIn(nn.Module):
    def __init__(self, channels: Tuple[int, ...] = (128, 256, 512, 512), latent: int = 4):
        super().__init__()
        self.conv_in = nn.Conv2d(3, channels[0], 3, padding=1)
        blocks = []
        cin = channels[0]
        for i, c in enumerate(channels):
            blocks += [ResnetBlock(cin, c, None, eps=1e-6), ResnetBlock(c, c, None, eps=1e-6)]
            if i != len(channels) - 1:
                blocks.append(nn.Conv2d(c, c, 3, stride=2, padding=0))
            cin = c
        self.blocks = nn.ModuleList(blocks)
        self.mid = nn.ModuleList([ResnetBlock(cin, cin, None, eps=1e-6), VAEAttention(cin),
                                  ResnetBlock(cin, cin, None, eps=1e-6)])
        self.norm_out = nn.GroupNorm(32, cin, eps=1e-6)
        self.conv_out = nn.Conv2d(cin, 2 * latent, 3, padding=1)
        self.quant_conv = nn.Conv2d(2 * latent, 2 * latent, 1)

    def encode(self, x):
        x = self.conv_in(x)
        for b in self.blocks:
            if isinstance(b, nn.Conv2d):
                x = b(F.pad(x, (0, 1, 0, 1)))
            else:
                x = b(x)
        for b in self.mid:
            x = b(x)
        x = self.quant_conv(self.conv_out(F.silu(self.norm_out(x))))
        return {"latent_dist": _LatentDist(x)}

    forward = encode


# --------------------------------------------------------------------------------------
# DDIM scheduler (`optimize_token.py:25-35`)
# --------------------------------------------------------------------------------------
class DDIMSchedulerStandIn:
    def __init__(self, beta_start=0.00085, beta_end=0.012, num_train=1000, num_inference=50):
        betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_train, dtype=torch.float32) ** 2
        self.alphas_cumprod = torch.cumprod(1.0 - betas, dim=0)
        step = num_train // num_inference
        self.timesteps = (torch.arange(0, num_inference) * step).flip(0).to(torch.int64)

    def add_noise(self, original, noise, timesteps):
        ac = self.alphas_cumprod.to(original.device)
        t = timesteps.to(original.device).reshape(-1)
        a = ac[t] ** 0.5
        s = (1.0 - ac[t]) ** 0.5
        shape = (-1,) + (1,) * (original.dim() - 1)
        return a.reshape(shape) * original + s.reshape(shape) * noise


class _ModuleWrap(nn.Module):
    """Gives `.module` / `.device_ids` like `nn.DataParallel` without its per-forward replicate."""

    def __init__(self, module: nn.Module, device_ids: List[int]):
        super().__init__()
        self.module = module
        self.device_ids = device_ids

    def forward(self, *a, **k):
        return self.module(*a, **k)

    def encode(self, *a, **k):
        return self.module.encode(*a, **k)


class PipelineStandIn:
    """The `ldm` object contract of SURVEY.md §8(b): `.unet`, `.vae`, `.scheduler`, `.text_encoder`."""

    def __init__(self, cfg: UNetConfig, vae_channels=(128, 256, 512, 512), seed: int = 0):
        # weight init under a forked RNG: `torch.manual_seed` also reseeds every CUDA generator, and a pipeline
        # constructor must not leave all ranks of a job with the same CUDA seed (they would draw the same noise)
        devices = list(range(torch.cuda.device_count())) if torch.cuda.is_available() else []
        with torch.random.fork_rng(devices=devices):
            torch.manual_seed(seed)
            self.unet = UNet2DConditionStandIn(cfg)
            self.vae = VAEEncoderStandIn(vae_channels)
            self.text_encoder = nn.Linear(1, 1)   # only `.parameters()` is touched (`optimize_token.py:74`)
        self.scheduler = DDIMSchedulerStandIn()
        self.cfg = cfg

    def to(self, device):
        self.unet.to(device)
        self.vae.to(device)
        return self

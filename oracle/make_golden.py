"""Generates tests/golden/*.npz by running the UNMODIFIED reference (via oracle/ref_shim.py).

TEST INFRASTRUCTURE.  Run in the build container only (needs /root/reference):

    python -m oracle.make_golden

The reference has no golden vectors of its own; these fixtures freeze the
outputs of its own functions on seeded inputs so that the oracle and the CUDA path can be
checked on the GPU box, where /root/reference does not exist.  Every array stored under a
`ref_*` key was produced by reference code; everything else is an input.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from stablekeypoints_b200.sd_standin import CrossAttention  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def peaky_maps(n, size, seed):
    g = torch.Generator().manual_seed(seed)
    m = torch.rand(n, size, size, generator=g) * 0.2
    rows = torch.arange(size)[:, None].float()
    cols = torch.arange(size)[None, :].float()
    for i in range(n):
        r, c = torch.randint(0, size, (2,), generator=g)
        sg = 1.0 + 3 * torch.rand(1, generator=g)
        m[i] += torch.exp(-((rows - r) ** 2 + (cols - c) ** 2) / (2 * sg ** 2))
    return m


class _Holder(nn.Module):
    def __init__(self, mods):
        super().__init__()
        self.up_blocks = nn.ModuleList(mods)


def golden_attention(ref):
    """Reference hook (`ptp_utils.py:472-573`) + `collect_maps` (`optimize.py:27-79`) on four
    captured layers; stores the q/k the modules compute so the CUDA path can be fed directly."""
    torch.manual_seed(0)
    R, T, ctx_dim = 64, 100, 96
    spec = [(8, 256, 4), (8, 256, 4), (8, 256, 4), (16, 128, 4)]   # (s, C, heads)
    mods = [CrossAttention(C, ctx_dim, h, C // h) for (_, C, h) in spec]
    holder = _Holder(mods)
    ctl = ref.ptp_utils.AttentionStore()
    ref.ptp_utils.register_attention_control(holder, ctl, feature_upsample_res=R)
    g = torch.Generator().manual_seed(1)
    context = torch.randn(1, T, ctx_dim, generator=g) * 1.5
    xs = [torch.randn(1, s * s, C, generator=g) * 2.0 for (s, C, _) in spec]
    with torch.no_grad():
        for m, x in zip(mods, xs):
            m.forward(x, context=context)
    stored = [a.clone() for a in ctl.step_store["attn"]]
    assert len(stored) == 4
    out = {"R": R, "T": T, "heads": np.array([h for (_, _, h) in spec]),
           "scale": np.array([m.scale for m in mods], dtype=np.float64)}
    with torch.no_grad():
        for i, (m, x) in enumerate(zip(mods, xs)):
            out[f"q_low{i}"] = m.to_q(x)[0].numpy()          # (s*s, C)
            out[f"k{i}"] = m.to_k(context)[0].numpy()        # (T, C)
    idx = torch.tensor([5, 17, 3, 99, 42, 64, 0, 31, 77, 8])
    out["indices"] = idx.numpy()
    for name, up, ind in [("A", -1, None), ("B", 128, None), ("C", 128, idx)]:
        ctl.step_store["attn"] = [s.clone() for s in stored]
        res = ref.optimize.collect_maps(ctl, upsample_res=up, layers=[0, 1, 2, 3], indices=ind)
        out[f"ref_maps_{name}"] = (res[:8] if name == "B" else res).numpy()
    np.savez_compressed(os.path.join(OUT, "attention_maps.npz"), **out)


def golden_selection(ref):
    size = 64
    maps = peaky_maps(60, size, 7)
    out = {"maps": maps.numpy()}
    out["ref_max_pixel"] = ref.eval.find_max_pixel(maps).numpy()
    out["ref_k_max_pixels3"] = ref.eval.find_k_max_pixels(maps, 3).numpy()
    out["ref_weighted_avg"] = ref.eval.pixel_from_weighted_avg(maps.clone()).numpy()
    for ns in (1, 2):
        out[f"ref_topk_gaussian_ns{ns}"] = ref.ptp_utils.find_top_k_gaussian(maps, 25, sigma=2.0, num_subjects=ns).numpy()
    out["ref_entropy_sort"] = ref.ptp_utils.entropy_sort(maps, 25).numpy()
    cand = ref.ptp_utils.find_top_k_gaussian(maps, 25, sigma=2.0)
    out["ref_fps"] = ref.ptp_utils.furthest_point_sampling(maps, 10, cand).numpy()
    pos = ref.eval.find_k_max_pixels(maps[:6], 2) / size
    out["ref_gaussian_circles"] = ref.optimize_token.gaussian_circles(pos, size=size, sigma=2.0, device="cpu").numpy()
    # non-power-of-two plane size: fp32 location rounding matters for FPS
    maps48 = peaky_maps(30, 48, 9)
    out["maps48"] = maps48.numpy()
    out["ref_fps48"] = ref.ptp_utils.furthest_point_sampling(maps48, 10, torch.arange(25)).numpy()
    np.savez_compressed(os.path.join(OUT, "selection_readout.npz"), **out)


def golden_losses_tta(ref):
    size = 32
    T = ref.invertable_transform.RandomAffineWithInverse(degrees=15, scale=[0.8, 1.0], translate=[0.25, 0.25])
    torch.manual_seed(11)
    img = torch.rand(2, 3, size, size)
    warped = T(img)
    theta = T.last_params["theta"].clone()
    out = {"img": img.numpy(), "theta": theta.numpy(), "ref_warped": warped.numpy(),
           "ref_unwarped": T.inverse(warped).numpy()}
    a = peaky_maps(10, size, 1).requires_grad_(True)
    b = peaky_maps(10, size, 2).requires_grad_(True)
    T.last_params = {"theta": theta[:1]}
    ls = ref.optimize.sharpening_loss(a, sigma=2.0, device="cpu")
    le = ref.optimize.equivariance_loss(a, b[None], T, 0)
    ga, gb = torch.autograd.grad(ls * 100 + le * 1000, [a, b])
    out.update(a=a.detach().numpy(), b=b.detach().numpy(), ref_sharpening=ls.item(), ref_equivariance=le.item(),
               ref_grad_a=ga.numpy(), ref_grad_b=gb.numpy())
    # TTA accumulation arithmetic of eval.py:262-266, 327-330
    torch.manual_seed(5)
    num = torch.zeros(4, size, size)
    tot = torch.zeros(4, size, size)
    ms, ths = [], []
    for i in range(6):
        T(torch.rand(1, 3, size, size))
        m = peaky_maps(4, size, 20 + i)
        ms.append(m)
        ths.append(T.last_params["theta"][0].clone())
        num += T.inverse(torch.ones_like(m)[None]).sum(0)
        tot += T.inverse(m[None]).sum(0)
    ens = tot / num
    ens[ens != ens] = 0
    out.update(tta_maps=torch.stack(ms).numpy(), tta_thetas=torch.stack(ths).numpy(), ref_tta=ens.numpy(),
               ref_tta_keypoints=ref.eval.find_max_pixel(ens).numpy())
    np.savez_compressed(os.path.join(OUT, "losses_tta.npz"), **out)


def main():
    ref = ref_shim.load()
    os.makedirs(OUT, exist_ok=True)
    golden_attention(ref)
    golden_selection(ref)
    golden_losses_tta(ref)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()

"""Generates tests/golden/e2e_tiny.npz: the UNMODIFIED reference (via oracle/ref_shim.py) driving
the tiny SD-1.5-shaped stand-in pipeline end to end on CPU.

TEST INFRASTRUCTURE.  Run in the build container only (needs /root/reference):

    python -m oracle.make_golden_e2e

Pins, for one seeded 128x128 image:  (a) `run_and_find_attn` maps of the image and of its
warped copy (reference hook + collect_maps, mode A);  (b) the token ids picked by
`find_top_k_gaussian` + `furthest_point_sampling`;  (c) `sharpening_loss`, `equivariance_loss`
and the gradient of 1000*eq + 100*sharp w.r.t. the context through BOTH UNet passes
(`optimize.py:372-445`);  (d) the TTA ensemble + argmax / weighted-average keypoints, i.e. the
body of `run_image_with_context_augmented` (`eval.py:239-330`) with its hard-coded
`.to("cuda:0")` dropped.  Arrays under `ref_*` were produced by reference code.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from stablekeypoints_b200.datasets import SyntheticKeypoints  # noqa: E402
from stablekeypoints_b200.sd_standin import PipelineStandIn, unet_config  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "e2e_tiny.npz")
CFG = dict(model="tiny-sd15", image=128, R=32, T=40, top_k=5, n_cand=12, sigma=2.0, n_aug=3, upscale=128, seed=0)


def seeded_noise(seed, shape):
    return torch.randn(shape, generator=torch.Generator().manual_seed(seed))


def main():
    ref = ref_shim.load()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cfg = unet_config(CFG["model"])
    ldm = PipelineStandIn(cfg, vae_channels=(32, 32, 64, 64), seed=CFG["seed"])
    for m in (ldm.unet, ldm.vae):
        for p in m.parameters():
            p.requires_grad = False
    ctl = ref.ptp_utils.AttentionStore()
    controllers = {torch.device("cpu"): ctl}
    ref.ptp_utils.register_attention_control(ldm.unet, ctl, feature_upsample_res=CFG["R"])
    image = SyntheticKeypoints(length=4, image_size=CFG["image"], seed=3)[1]["img"][None]       # (1,3,128,128)
    context = (torch.randn(1, CFG["T"], cfg.context_dim, generator=torch.Generator().manual_seed(1))).requires_grad_(True)
    tf = ref.invertable_transform.RandomAffineWithInverse(degrees=15, scale=(0.8, 1.0), translate=(0.25, 0.25))
    torch.manual_seed(3)
    transformed = tf(image)                      # draws theta from the global generator
    theta = tf.last_params["theta"].clone()
    lat = (1, 4, cfg.latent_side, cfg.latent_side)
    out = {"image": image.numpy(), "context": context.detach().numpy(), "theta": theta.numpy()}
    out.update({"cfg_" + k: np.asarray(v) for k, v in CFG.items() if not isinstance(v, str)})

    def run(img, seed, **kw):
        # the reference draws `torch.randn_like(latent)` from the global generator (ptp_utils.py:219)
        torch.manual_seed(seed)
        noise = torch.randn(lat)
        torch.manual_seed(seed)
        maps = ref.ptp_utils.run_and_find_attn(ldm, img, context, device="cpu", layers=[0, 1, 2, 3], noise_level=-1,
                                               controllers=controllers, **kw)
        return noise, maps[0]

    noise_a, maps_a = run(image, 11, upsample_res=-1)
    noise_b, maps_b = run(transformed, 12, upsample_res=-1)
    out["noise_a"], out["noise_b"] = noise_a.numpy(), noise_b.numpy()
    out["ref_maps_a"], out["ref_maps_b"] = maps_a.detach().numpy(), maps_b.detach().numpy()
    cand = ref.ptp_utils.find_top_k_gaussian(maps_a, CFG["n_cand"], sigma=CFG["sigma"], num_subjects=1)
    idx = ref.ptp_utils.furthest_point_sampling(maps_b, CFG["top_k"], cand)
    out["ref_cand"], out["ref_idx"] = cand.numpy(), idx.numpy()
    sharp = ref.optimize.sharpening_loss(maps_a[idx], device="cpu", sigma=CFG["sigma"], num_subjects=1)
    equiv = ref.optimize.equivariance_loss(maps_a[idx], maps_b[idx][None].repeat(1, 1, 1, 1), tf, 0)
    (equiv * 1000 + sharp * 100).backward()
    out["ref_sharp"], out["ref_equiv"] = sharp.item(), equiv.item()
    out["ref_context_grad"] = context.grad.numpy().copy()

    # ---- TTA body of eval.py:239-330 ---------------------------------------------------------
    with torch.no_grad():
        up, k = CFG["upscale"], CFG["top_k"]
        num = torch.zeros(k, up, up)
        tot = torch.zeros(k, up, up)
        tta = ref.invertable_transform.RandomAffineWithInverse(degrees=15, scale=(0.8, 1.0), translate=(0.25, 0.25))
        thetas, noises = [], []
        torch.manual_seed(5)
        for i in range(CFG["n_aug"]):
            aug = tta(image)
            thetas.append(tta.last_params["theta"].clone())
            nz, m = run(aug.permute(0, 2, 3, 1).numpy(), 20 + i, upsample_res=up, indices=idx.cpu())
            noises.append(nz)
            m = m[None]
            num += tta.inverse(torch.ones_like(m)).sum(0)
            tot += tta.inverse(m).sum(0)
        ens = tot / num
        ens[ens != ens] = 0
        out["tta_thetas"] = torch.cat(thetas).numpy()
        out["tta_noise"] = torch.cat(noises).numpy()
        out["ref_tta_maps"] = ens.numpy()
        out["ref_kp_argmax"] = ref.eval.find_max_pixel(ens).numpy()
        out["ref_kp_wavg"] = ref.eval.pixel_from_weighted_avg(ens.clone()).numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()

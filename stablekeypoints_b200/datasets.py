"""Synthetic stand-in for the reference's `datasets/*.py` (out of scope: BASELINE.json measures
on synthetic 512x512 images).  Only the sample-dict contract matters (`datasets/celeba.py:93-113`,
`custom_images.py:20-25`): {"img": float (3,S,S) in [0,1], "kpts": (n,2) (row, col)/S, "visibility": (n,)}.
"""
from __future__ import annotations

import torch


class SyntheticKeypoints(torch.utils.data.Dataset):
    """Deterministic per-index images: smooth random blobs (so attention maps have structure) +
    `n_kpts` ground-truth points at the blob centres.  Image i depends only on (seed, i)."""

    def __init__(self, length: int = 64, image_size: int = 512, n_kpts: int = 5, seed: int = 0):
        self.length, self.image_size, self.n_kpts, self.seed = length, image_size, n_kpts, seed

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + int(i))
        s = self.image_size
        kp = torch.rand(self.n_kpts, 2, generator=g) * 0.8 + 0.1
        low = torch.rand(3, s // 16, s // 16, generator=g)
        img = torch.nn.functional.interpolate(low[None], size=(s, s), mode="bilinear", align_corners=False)[0] * 0.5
        ys = (torch.arange(s, dtype=torch.float32) + 0.5) / s
        for j in range(self.n_kpts):
            blob = torch.exp(-((ys[:, None] - kp[j, 0]) ** 2 + (ys[None, :] - kp[j, 1]) ** 2) / (2 * 0.03 ** 2))
            img[j % 3] = img[j % 3] + 0.5 * blob
        return {"img": img.clamp_(0, 1), "kpts": kp, "visibility": torch.ones(self.n_kpts)}

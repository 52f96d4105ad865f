"""`RandomAffineWithInverse` — host mirror of `unsupervised_keypoints/invertable_transform.py:6-92`.

Same constructor, `__call__(img_tensor, theta=None)`, `inverse(img_tensor)` and
`last_params["theta"]`; the `affine_grid` + `grid_sample` pairs (`:64-68`, `:86-90`) run as one
CUDA kernel (`skp_affine_warp`, differentiable through `skp_affine_warp_bwd`), and the 3x3
`torch.inverse` (`:77-84`) is the closed-form 2x3 inverse.
"""
from __future__ import annotations

import math

import torch

from . import ops


class RandomAffineWithInverse:
    def __init__(self, degrees=0, scale=(1.0, 1.0), translate=(0.0, 0.0)):
        self.degrees = degrees
        self.scale = scale
        self.translate = translate
        self.last_params = {"theta": torch.eye(2, 3).unsqueeze(0)}

    def create_affine_matrix(self, angle, scale, translations_percent):
        """`invertable_transform.py:22-36`: [[c*s, sn*s, tx], [-sn*s, c*s, ty]] as (1,2,3)."""
        a = math.radians(angle)
        theta = torch.tensor([[math.cos(a), math.sin(a), translations_percent[0]],
                              [-math.sin(a), math.cos(a), translations_percent[1]]], dtype=torch.float)
        theta[:, :2] = theta[:, :2] * scale
        return theta.unsqueeze(0)

    def sample_theta(self, n: int, generator=None) -> torch.Tensor:
        """Same draw order as the reference (`:44-51`): angle, scale, tx, ty per image from the
        global CPU generator, so a seeded run reproduces the reference's transforms.  `generator`
        (a CPU torch.Generator) replaces the global one: multi-rank jobs draw from a per-rank
        generator so that the ranks do not all sample the same transform."""
        def r():
            return torch.rand(1, generator=generator).item()
        out = []
        for _ in range(n):
            angle = r() * (2 * self.degrees) - self.degrees
            scale_factor = r() * (self.scale[1] - self.scale[0]) + self.scale[0]
            tr = (r() * (2 * self.translate[0]) - self.translate[0],
                  r() * (2 * self.translate[1]) - self.translate[1])
            out.append(self.create_affine_matrix(angle, scale_factor, tr))
        return torch.cat(out, dim=0)

    def __call__(self, img_tensor: torch.Tensor, theta=None, generator=None) -> torch.Tensor:
        if theta is None:
            theta = self.sample_theta(img_tensor.shape[0], generator)
        theta = theta.to(device=img_tensor.device, dtype=torch.float32)
        self.last_params = {"theta": theta}
        return ops.affine_warp(img_tensor, theta)

    def inverse_theta(self) -> torch.Tensor:
        return ops.invert_theta(self.last_params["theta"])

    def inverse(self, img_tensor: torch.Tensor) -> torch.Tensor:
        theta_inv = self.inverse_theta().to(img_tensor.device)
        return ops.affine_warp(img_tensor, theta_inv)

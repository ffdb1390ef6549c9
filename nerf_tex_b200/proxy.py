"""Proxy geometry (network/proxy.py)."""

from __future__ import annotations

import math

import torch

from ._lib import as_cuda


class AABB:
    """Axis-aligned box [b_0, b_1]; rays that miss get t = (inf, inf) (network/proxy.py:6-35)."""

    def __init__(self, b_0: list, b_1: list):
        self.b_0 = [float(v) for v in b_0]
        self.b_1 = [float(v) for v in b_1]

    def __call__(self, rays_o: torch.Tensor, rays_d: torch.Tensor) -> torch.Tensor:
        rays_o, rays_d = as_cuda(rays_o), as_cuda(rays_d)
        b0 = torch.tensor(self.b_0, dtype=torch.float32, device=rays_o.device)
        b1 = torch.tensor(self.b_1, dtype=torch.float32, device=rays_o.device)
        inv = 1.0 / rays_d                                                     # :17
        t0 = (b0 - rays_o) * inv
        t1 = (b1 - rays_o) * inv
        t0n = torch.where(t0 < t1, t0, t1)                                     # :22-24
        t1n = torch.where(t0 > t1, t0, t1)
        tn = t0n.max(dim=1).values                                             # :27-28
        tf_ = t1n.min(dim=1).values
        hit = tn < tf_                                                         # :31-33
        inf = torch.full_like(tn, math.inf)
        return torch.stack([torch.where(hit, tn, inf), torch.where(hit, tf_, inf)], -1)

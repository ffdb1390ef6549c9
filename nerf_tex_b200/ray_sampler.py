"""Camera-ray generation (network/ray_sampler.py).  Elementwise and cheap; torch ops on the GPU."""

from __future__ import annotations

from typing import Any, Tuple

import torch

from ._lib import as_cuda


def rays_from_camera(image_plane_loc, height: int, width: int, focal: float, c2w) -> Tuple[torch.Tensor, ...]:
    """network/ray_sampler.py:39-48."""
    loc, c2w = as_cuda(image_plane_loc), as_cuda(c2w)
    dirs = torch.stack([(loc[:, 1] + .5 - .5 * width) / focal, -(loc[:, 0] + .5 - .5 * height) / focal,
                        -torch.ones(loc.shape[0], dtype=torch.float32, device=loc.device)], -1)
    rays_d = (dirs[:, None, :] * c2w[:3, :3]).sum(-1)                          # :42
    rays_o = c2w[:3, -1].expand(rays_d.shape).contiguous()
    cone_scale = torch.cos(torch.atan(torch.linalg.norm(dirs[:, :2], dim=-1))) / torch.linalg.norm(dirs, dim=-1) / focal
    return rays_o, rays_d, cone_scale[:, None]


class Frustum:
    """network/ray_sampler.py:6-21: constant near / far."""

    def __init__(self, height: int, width: int, focal: float, near: float, far: float, **kwargs) -> None:
        self.height, self.width, self.focal, self.near, self.far = height, width, focal, near, far

    def __call__(self, image_plane_loc, c2w):
        rays_o, rays_d, cone_scale = rays_from_camera(image_plane_loc, self.height, self.width, self.focal, c2w)
        n = rays_o.shape[0]
        t = torch.stack([torch.full((n,), float(self.near), device=rays_o.device),
                         torch.full((n,), float(self.far), device=rays_o.device)], -1)
        return rays_o, rays_d, t, cone_scale


class Proxy:
    """network/ray_sampler.py:23-37: normalised directions, t from the proxy geometry."""

    def __init__(self, height: int, width: int, focal: float, proxy: Any, **kwargs) -> None:
        self.height, self.width, self.focal, self.proxy = height, width, focal, proxy

    def __call__(self, image_plane_loc, c2w):
        rays_o, rays_d, cone_scale = rays_from_camera(image_plane_loc, self.height, self.width, self.focal, c2w)
        rays_d = rays_d / torch.linalg.norm(rays_d, dim=-1, keepdim=True)
        t = self.proxy(rays_o, rays_d)
        return rays_o, rays_d, t, cone_scale

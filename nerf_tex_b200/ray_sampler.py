"""Camera-ray generation (network/ray_sampler.py) on the device: one kernel (ntx_rays_from_camera) produces
origins, directions, [near, far] and cone scales; a generic proxy object falls back to calling it on the rays."""

from __future__ import annotations

from typing import Any, Tuple

import torch

from . import ops
from .proxy import AABB


def rays_from_camera(image_plane_loc, height: int, width: int, focal: float, c2w) -> Tuple[torch.Tensor, ...]:
    """network/ray_sampler.py:39-48."""
    rays_o, rays_d, _, cone = ops.rays_from_camera(image_plane_loc, height, width, focal, c2w)
    return rays_o, rays_d, cone


class Frustum:
    """network/ray_sampler.py:6-21: constant near / far."""

    def __init__(self, height: int, width: int, focal: float, near: float, far: float, **kwargs) -> None:
        self.height, self.width, self.focal, self.near, self.far = height, width, focal, near, far

    def __call__(self, image_plane_loc, c2w):
        return ops.rays_from_camera(image_plane_loc, self.height, self.width, self.focal, c2w, near=self.near,
                                    far=self.far)


class Proxy:
    """network/ray_sampler.py:23-37: normalised directions, t from the proxy geometry."""

    def __init__(self, height: int, width: int, focal: float, proxy: Any, **kwargs) -> None:
        self.height, self.width, self.focal, self.proxy = height, width, focal, proxy

    def __call__(self, image_plane_loc, c2w):
        if isinstance(self.proxy, AABB):
            return ops.rays_from_camera(image_plane_loc, self.height, self.width, self.focal, c2w, normalize=True,
                                        aabb=(self.proxy.b_0, self.proxy.b_1))
        rays_o, rays_d, _, cone = ops.rays_from_camera(image_plane_loc, self.height, self.width, self.focal, c2w,
                                                       normalize=True)
        return rays_o, rays_d, self.proxy(rays_o, rays_d), cone

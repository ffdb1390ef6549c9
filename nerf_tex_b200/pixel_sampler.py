"""Pixel samplers (network/pixel_sampler.py) on the device.  Random draws come from a seeded torch CUDA
generator (TF's streams are not reproducible outside TF), so parity with the reference is distributional."""

from __future__ import annotations

import math
from typing import Any, Optional

import torch

from . import ops
from .proxy import AABB


def _device(device=None) -> torch.device:
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


class Full:
    """network/pixel_sampler.py:8-15: every pixel, row-major (row, column)."""

    def __init__(self, height: int, width: int, device=None, **kwargs) -> None:
        self.height, self.width, self.device = int(height), int(width), device

    def __call__(self, **kwargs) -> torch.Tensor:
        n = torch.arange(self.height * self.width, dtype=torch.int32, device=_device(self.device))
        return torch.stack([n // self.width, n % self.width], -1)


class Independent:
    """network/pixel_sampler.py:17-29: n_samples iid pixels."""

    def __init__(self, height: int, width: int, n_samples: int, seed: Optional[int] = None, device=None, **kwargs) -> None:
        self.height, self.width, self.n_samples, self.device = int(height), int(width), int(n_samples), device
        self._gen, self._seed = None, seed

    def _generator(self, dev):
        if self._gen is None or self._gen.device != dev:
            self._gen = torch.Generator(device=dev)
            if self._seed is not None:
                self._gen.manual_seed(self._seed)
        return self._gen

    def __call__(self, **kwargs) -> torch.Tensor:
        dev = _device(self.device)
        g = self._generator(dev)
        i = torch.randint(0, self.height, (self.n_samples,), dtype=torch.int32, device=dev, generator=g)
        j = torch.randint(0, self.width, (self.n_samples,), dtype=torch.int32, device=dev, generator=g)
        return torch.stack([i, j], -1)


class Proxy(Independent):
    """network/pixel_sampler.py:31-66: pixels whose rays hit the proxy, found on a `downsample_factor` coarser
    grid (one ntx_rays_from_camera launch), upsampled with nearest neighbour, shuffled; the first n_samples win."""

    def __init__(self, height: int, width: int, n_samples: int, proxy: Any, focal: float, downsample_factor: int = 8,
                 seed: Optional[int] = None, device=None, **kwargs) -> None:
        super().__init__(height, width, n_samples, seed, device)
        self.proxy = proxy
        self.downsample_factor = int(downsample_factor)
        self.focal = focal // downsample_factor                                  # :39 (floor division, as written)
        self.height_down = self.height // self.downsample_factor
        self.width_down = self.width // self.downsample_factor

    def hit_mask(self, c2w) -> torch.Tensor:
        """[height, width] bool: the upsampled proxy-hit mask (:46-59)."""
        dev = _device(self.device)
        if isinstance(self.proxy, AABB):
            _, _, t, _ = ops.rays_from_camera(None, self.height_down, self.width_down, self.focal, c2w,
                                              aabb=(self.proxy.b_0, self.proxy.b_1), device=dev)
        else:
            rays_o, rays_d, _, _ = ops.rays_from_camera(None, self.height_down, self.width_down, self.focal, c2w, device=dev)
            t = self.proxy(rays_o, rays_d)
        hit = (t[:, 0] != math.inf).reshape(self.height_down, self.width_down)
        if self.downsample_factor > 1:                                           # tf.image.resize(..., 'nearest')
            ys = torch.clamp(((torch.arange(self.height, device=dev) + 0.5) * (self.height_down / self.height)).floor().long(),
                             max=self.height_down - 1)
            xs = torch.clamp(((torch.arange(self.width, device=dev) + 0.5) * (self.width_down / self.width)).floor().long(),
                             max=self.width_down - 1)
            hit = hit[ys][:, xs]
        return hit

    def __call__(self, c2w, **kwargs) -> torch.Tensor:
        hit = self.hit_mask(c2w)
        idxs = torch.nonzero(hit)                                                # :62
        if idxs.shape[0] < self.n_samples:
            raise ValueError(f"pixel_sampler.Proxy: only {idxs.shape[0]} pixels hit the proxy, {self.n_samples} requested "
                             f"(the reference's gather_nd fails here too, pixel_sampler.py:66)")
        perm = torch.randperm(idxs.shape[0], device=idxs.device, generator=self._generator(idxs.device))   # :64
        return idxs[perm[:self.n_samples]].to(torch.int32)

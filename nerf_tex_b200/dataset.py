"""The two pieces of network/dataset.py that sit directly in front of the ray generator: the look-at camera
transform (:231-238) and a minimal in-memory pose / parameter list with the attributes the reference's Logger reads
from its tf.data object (height, width, focal, composite_bkgd, bkgd_color, n_parameters).  TFRecord / image-folder
loading and the data.* distributions are out of scope."""

from __future__ import annotations

import math
from typing import Iterator, Optional, Sequence

import numpy as np
import torch

from . import util


def look_at(pos, to=(0., 0., 0.), offset=(0., 0., 0.), eps: float = 1e-6) -> np.ndarray:
    """network/dataset.py:231-238: camera-to-world matrix of a camera at `pos` looking at `to` (z up)."""
    pos, to, offset = (np.asarray(v, np.float32) for v in (pos, to, offset))
    unit = lambda v: (v / np.sqrt((v * v).sum(dtype=np.float32))).astype(np.float32)
    fwd = unit(pos - to + np.float32(eps))
    right = unit(np.cross(np.array([0, 0, 1.], np.float32), fwd).astype(np.float32) + np.float32(eps))
    up = unit(np.cross(fwd, right).astype(np.float32) + np.float32(eps))
    m = np.eye(4, dtype=np.float32)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = right, up, fwd, pos + offset
    return m


class PoseDataset:
    """Iterable of renderer inputs for a list of camera poses and parameter vectors: the product of
    network/dataset.py:10-75 (`Dataset` over `GenerateData`) for full-frame evaluation, batch size 1."""

    def __init__(self, poses: Sequence, parameters: Sequence, height: int = 256, width: int = 256, angle: float = .7,
                 pixel_sampler_config: Optional[dict] = None, ray_sampler_config: Optional[dict] = None,
                 proxy_config: Optional[dict] = None, composite_bkgd: bool = False, bkgd_color=(1., 1., 1.), device=None) -> None:
        self.poses = [np.asarray(p, np.float32).reshape(4, 4) for p in poses]
        self.parameters = [np.asarray(p, np.float32).reshape(-1) for p in parameters]
        self.height, self.width = int(height), int(width)
        self.focal = width / math.tan(angle / 2) / 2                              # :229
        self.composite_bkgd, self.bkgd_color = composite_bkgd, list(bkgd_color)
        self.n_parameters = int(self.parameters[0].shape[0])
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        proxy = util.instantiate(proxy_config) if proxy_config else None          # :17
        common = {"height": self.height, "width": self.width, "focal": self.focal, "proxy": proxy}
        ps = dict(pixel_sampler_config or {"module": "network.pixel_sampler.Full"})
        ps.update(common)                                                         # :20
        self.pixel_sampler = util.instantiate(ps)
        rs = dict(ray_sampler_config or {"module": "network.ray_sampler.Proxy" if proxy is not None else "network.ray_sampler.Frustum",
                                         "near": 2., "far": 6.})
        rs.update(common)                                                         # :24
        self.ray_sampler = util.instantiate(rs)
        self.n_samples = self.height * self.width

    def __len__(self) -> int:
        return max(len(self.poses), len(self.parameters))

    def __iter__(self) -> Iterator[dict]:
        for i in range(len(self)):
            c2w = self.poses[i % len(self.poses)]
            prm = self.parameters[i % len(self.parameters)]
            loc = self.pixel_sampler(c2w=c2w)                                     # :32
            rays_o, rays_d, t, cone = self.ray_sampler(image_plane_loc=loc.to(torch.float32), c2w=c2w)   # :36
            yield {"parameters": torch.as_tensor(prm, device=self.device)[None], "rays_o": rays_o[None],
                   "rays_d": rays_d[None], "t": t[None], "cone_scale": cone[None]}

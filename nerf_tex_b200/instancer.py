"""Instancer boundary (instancer/instancer.pyx:6-54).

The reference's native instancer (Embree scene of shell patches on a mesh) cannot be built or
fed in this environment (Embree 3 is a system dependency, every mesh/texture is a git-LFS stub;
SURVEY.md D8).  ``Instancer`` is the device instancer for explicit patch transformations (no mesh);
``SyntheticInstancer`` keeps the exact constructor / ``n_instances`` /
``get_model_input`` surface and produces seeded outputs with the statistics SURVEY.md section 8d
prescribes, natively CSR-packed on the GPU (``get_model_input_csr``); ``get_model_input`` scatters
the same data into the reference's dense 10-tuple for callers written against instancer.pyx:54.
"""

from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch

from ._lib import as_cuda


class SyntheticInstancer:
    def __init__(self, b_0: list, b_1: list, cast_shadow_rays=False, textures: list = [], transformations: list = [],
                 mesh_path: str = None, patch_scale: float = 1., patch_origins_path='', min_shadow_samples: int = 4,
                 n_shadow_samples: int = 512, min_texture_samples: int = 4, n_texture_samples: int = 512,
                 jitter_amount: float = 0, instance_sampling_method: str = 'random', use_mean_distance: bool = False,
                 auxiliary_meshes=[], transformation_export_path=None, *, seed: int = 0, hit_probability: float = 0.7,
                 mean_valid: float = 120.0, sigma_valid: float = 0.6, num_instances: int = 10000) -> None:
        assert instance_sampling_method in ('random', 'nearest', 'nearest_blend')      # instancer.pyx:12
        self.b_0 = [float(v) for v in b_0]
        self.b_1 = [float(v) for v in b_1]
        self.patch_scale = float(patch_scale)
        self.seed, self.hit_probability = int(seed), float(hit_probability)
        self.mean_valid, self.sigma_valid = float(mean_valid), float(sigma_valid)
        self._n_instances = int(num_instances)
        self._calls = 0

    def n_instances(self) -> int:
        return self._n_instances

    # ------------------------------------------------------------------------------------
    def get_model_input_csr(self, rays_o, rays_d, parameters, n_samples: int, step_size: float) -> Dict[str, torch.Tensor]:
        rays_o, rays_d, parameters = as_cuda(rays_o), as_cuda(rays_d), as_cuda(parameters)
        dev = rays_o.device
        R, p = rays_o.shape[0], parameters.shape[1]
        gen = torch.Generator(device=dev).manual_seed(self.seed + 7919 * self._calls)
        self._calls += 1
        rnd = lambda *s: torch.rand(s, dtype=torch.float32, device=dev, generator=gen)
        hit = rnd(R) < self.hit_probability
        ray_index = torch.nonzero(hit).squeeze(1)
        H = ray_index.shape[0]
        logn = torch.randn(H, dtype=torch.float32, device=dev, generator=gen) * self.sigma_valid + math.log(self.mean_valid)
        n_valid = torch.clamp(torch.round(torch.exp(logn)), 1, n_samples).to(torch.int64)
        ray_ptr = torch.zeros(H + 1, dtype=torch.int64, device=dev)
        ray_ptr[1:] = torch.cumsum(n_valid, 0)
        T = int(ray_ptr[-1].item()) if H > 0 else 0
        ray_of = torch.repeat_interleave(torch.arange(H, device=dev), n_valid)
        pos_in_ray = torch.arange(T, device=dev) - ray_ptr[:-1][ray_of]
        last = pos_in_ray == (n_valid[ray_of] - 1)
        dists = torch.full((T,), float(step_size), dtype=torch.float32, device=dev)
        dists = torch.where(last, dists + rnd(T) * float(step_size), dists)
        b0 = torch.tensor(self.b_0, dtype=torch.float32, device=dev)
        b1 = torch.tensor(self.b_1, dtype=torch.float32, device=dev)
        pts = b0 + (b1 - b0) * rnd(T, 3)
        # per-sample rotation of the (normalised) ray direction about z: each sample lives in its patch frame
        d = rays_d[ray_index][ray_of]
        d = d / torch.linalg.norm(d, dim=-1, keepdim=True)
        ang = rnd(T) * (2 * math.pi)
        ca, sa = torch.cos(ang), torch.sin(ang)
        dirs = torch.stack([ca * d[:, 0] - sa * d[:, 1], sa * d[:, 0] + ca * d[:, 1], d[:, 2]], -1)
        # parameters: per-ray values, the leading (texture-modulated) ones scaled per sample, light dir kept unit
        params = parameters[ray_index][ray_of].clone()
        n_tex = max(p - 3, 0)
        if n_tex > 0:
            params[:, :n_tex] = params[:, :n_tex] * rnd(T, n_tex)
        t = pos_in_ray.to(torch.float32) * float(step_size) + 1.0
        return {
            "ray_ptr": ray_ptr, "ray_index": ray_index.to(torch.int32), "pts": pts.contiguous(), "dirs": dirs.contiguous(),
            "params": params.contiguous(), "dists": dists.contiguous(), "t": t.contiguous(),
            "alpha_weight": torch.ones(T, dtype=torch.float32, device=dev),
            "instance_id": torch.randint(0, self._n_instances, (T,), device=dev, generator=gen, dtype=torch.int32),
            "color_last": torch.zeros((H, 3), dtype=torch.float32, device=dev),
            "alpha_last": (rnd(H) < 0.5).to(torch.float32),
        }

    def get_model_input(self, rays_o, rays_d, parameters, n_samples: int, step_size: float):
        """Dense 10-tuple of instancer.pyx:54 (for small inputs / API parity)."""
        csr = self.get_model_input_csr(rays_o, rays_d, parameters, n_samples, step_size)
        return csr_to_dense(csr, as_cuda(rays_d), as_cuda(parameters), n_samples)


def csr_to_dense(csr, rays_d, parameters, n_samples: int):
    dev = rays_d.device
    R, p = rays_d.shape[0], parameters.shape[1]
    H = csr["ray_index"].shape[0]
    counts = csr["ray_ptr"][1:] - csr["ray_ptr"][:-1]
    ray_of = torch.repeat_interleave(csr["ray_index"].long(), counts)
    pos_in_ray = torch.arange(csr["pts"].shape[0], device=dev) - torch.repeat_interleave(csr["ray_ptr"][:-1], counts)
    rays_d_map = rays_d[:, None, :].repeat(1, n_samples, 1)                    # instancer.pyx:41
    pts = torch.zeros((R, n_samples, 3), device=dev)
    t = torch.zeros((R, n_samples), device=dev)
    dists = torch.zeros((R, n_samples), device=dev)
    alpha_weight = torch.ones((R, n_samples), device=dev)                      # instancer.pyx:47
    instance_id = torch.zeros((R, n_samples), dtype=torch.int32, device=dev)
    params_map = parameters[:, None, :].repeat(1, n_samples, 1)                # instancer.pyx:50
    color_last = torch.zeros((R, 1, 3), device=dev)
    alpha_last = torch.zeros((R, 1), device=dev)
    rays_d_map[ray_of, pos_in_ray] = csr["dirs"]
    pts[ray_of, pos_in_ray] = csr["pts"]
    t[ray_of, pos_in_ray] = csr["t"]
    dists[ray_of, pos_in_ray] = csr["dists"]
    alpha_weight[ray_of, pos_in_ray] = csr["alpha_weight"]
    instance_id[ray_of, pos_in_ray] = csr["instance_id"]
    params_map[ray_of, pos_in_ray] = csr["params"]
    color_last[csr["ray_index"].long(), 0] = csr["color_last"]
    alpha_last[csr["ray_index"].long(), 0] = csr["alpha_last"]
    idxs = csr["ray_index"].long()[:, None]                                    # tf.where(hit) -> [H,1]
    return rays_d_map, pts, t, dists, color_last, alpha_last, alpha_weight, instance_id, idxs, params_map


class Instancer:
    """Device shell instancer behind the reference's constructor (instancer/instancer.pyx:6-36) for explicit patch
    `transformations` (AddInstance, instancer.cpp:125-141).  What needs the Embree scene of a base mesh is refused:
    `mesh_path`, shadow rays, image parameter textures (entries other than '' / 'light' / 'point'), auxiliary meshes."""

    def __init__(self, b_0: list, b_1: list, cast_shadow_rays=False, textures: list = [], transformations: list = [],
                 mesh_path: str = None, patch_scale: float = 1., patch_origins_path='', min_shadow_samples: int = 4,
                 n_shadow_samples: int = 512, min_texture_samples: int = 4, n_texture_samples: int = 512,
                 jitter_amount: float = 0, instance_sampling_method: str = 'random', use_mean_distance: bool = False,
                 auxiliary_meshes=[], transformation_export_path=None, *, seed: int = 0, device=None) -> None:
        assert instance_sampling_method in ('random', 'nearest', 'nearest_blend')      # instancer.pyx:12
        if mesh_path:
            raise NotImplementedError("Instancer: placing patches on a mesh needs the reference's Embree scene and its "
                                      "git-LFS meshes; pass explicit `transformations` (or use SyntheticInstancer)")
        if cast_shadow_rays or auxiliary_meshes:
            raise NotImplementedError("Instancer: shadow rays / auxiliary meshes need a mesh scene")
        self.light_dir_idx = self.light_strength_idx = -1
        n_parameters = 0
        for tex in textures:                                                           # instancer.cpp:75-92
            if tex == 'light':                                                         # directional light: 3 parameters
                self.light_dir_idx = n_parameters
                n_parameters += 3
            elif tex == 'point':                                                       # point light: strength + position
                self.light_strength_idx = n_parameters
                self.light_dir_idx = n_parameters + 1
                n_parameters += 4
            elif tex in ('', None):
                n_parameters += 1
            else:
                raise NotImplementedError(f"Instancer: parameter texture {tex!r} needs a UV-mapped mesh")
        self.n_parameters = n_parameters
        import numpy as np
        T = np.asarray(transformations, dtype=np.float64).reshape(-1, 4, 4)
        if T.shape[0] == 0:
            raise ValueError("Instancer: no instances (empty `transformations`)")
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        rt = np.transpose(T[:, :3, :3], (0, 2, 1))
        as_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a.astype(np.float32))).to(dev)
        self.minv = as_dev(np.linalg.inv(T)[:, :3, :])
        self.dirt = as_dev(rt / np.linalg.norm(rt, axis=2, keepdims=True))
        self.origins = as_dev(T[:, :3, 3])
        # bounding sphere of every instanced box (centre, radius^2 inflated by 2 %): first-stage rejection in the trace
        b0, b1 = np.asarray(b_0, np.float64), np.asarray(b_1, np.float64)
        corners = np.array([[(b0, b1)[(c >> a) & 1][a] for a in range(3)] + [1.0] for c in range(8)])      # [8,4]
        world = np.einsum("iab,cb->ica", T[:, :3, :], corners)                                              # [I,8,3]
        centre = np.einsum("iab,b->ia", T[:, :3, :], np.append((b0 + b1) / 2, 1.0))
        r2 = (np.linalg.norm(world - centre[:, None, :], axis=-1).max(1) * 1.01) ** 2 * 1.0 + 1e-12
        self.spheres = as_dev(np.concatenate([centre, r2[:, None]], -1))
        self.transformations = T
        self.b_0, self.b_1 = [float(v) for v in b_0], [float(v) for v in b_1]
        self.patch_scale = float(patch_scale)
        self.method = {'random': 0, 'nearest': 1, 'nearest_blend': 2}[instance_sampling_method]   # instancer.pyx:12
        self.use_mean_distance = bool(use_mean_distance)
        self.seed = int(seed)
        self._calls = 0
        if transformation_export_path:
            np.save(transformation_export_path, T)                                    # instancer.pyx:33-36

    def n_instances(self) -> int:
        return int(self.minv.shape[0])

    def get_model_input_csr(self, rays_o, rays_d, parameters, n_samples: int, step_size: float, seed: Optional[int] = None):
        from . import ops
        if seed is None:                       # a fresh draw every call, like the reference's running mt19937
            seed = self.seed + 7919 * self._calls
            self._calls += 1
        return ops.instancer_get_model_input(rays_o, rays_d, parameters, n_samples, step_size, self.minv, self.dirt, self.origins,
                                             self.b_0, self.b_1, seed=seed, light_dir_idx=self.light_dir_idx, method=self.method,
                                             use_mean_distance=self.use_mean_distance, spheres=self.spheres,
                                             light_strength_idx=self.light_strength_idx, blend_range=float(np.float32(0.2) * np.float32(self.patch_scale)))   # fp32 product, instancer.cpp:696

    def get_model_input(self, rays_o, rays_d, parameters, n_samples: int, step_size: float, seed: Optional[int] = None):
        """Dense 10-tuple of instancer.pyx:54."""
        csr = self.get_model_input_csr(rays_o, rays_d, parameters, n_samples, step_size, seed)
        return csr_to_dense(csr, as_cuda(rays_d), as_cuda(parameters), n_samples)

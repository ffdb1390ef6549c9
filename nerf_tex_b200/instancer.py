"""Instancer boundary (instancer/instancer.pyx:6-54).

The reference's native instancer (Embree scene of shell patches on a mesh) cannot be built or
fed in this environment (Embree 3 is a system dependency, every mesh/texture is a git-LFS stub;
SURVEY.md D8).  ``Instancer`` is the device instancer (explicit patch transformations, or patches placed on a PLY
base mesh that also ends the rays; uniform-grid traversal);
``SyntheticInstancer`` keeps the exact constructor / ``n_instances`` /
``get_model_input`` surface and produces seeded outputs with the statistics SURVEY.md section 8d
prescribes, natively CSR-packed on the GPU (``get_model_input_csr``); ``get_model_input`` scatters
the same data into the reference's dense 10-tuple for callers written against instancer.pyx:54.
"""

from __future__ import annotations

import ctypes as C
import math
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib
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


def load_texture(tex):
    """loadTexture (instancer.cpp:34-52): per channel a [width, height] fp32 matrix of value / 255, columns counted from
    the image's bottom row.  `tex`: a PNG path, or an array [h, w] / [h, w, c] of uint8 (or of floats in [0, 1])."""
    if isinstance(tex, str):
        from .image import decode_png
        try:
            with open(tex, "rb") as f:
                img = decode_png(f.read())
        except OSError:
            return []                                                                  # stbi_load returns NULL: no texture
    else:
        img = np.asarray(tex)
        if img.dtype != np.uint8:
            img = np.clip(np.round(img.astype(np.float64) * 255.0), 0, 255).astype(np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    return [np.ascontiguousarray((img[::-1, :, c].T.astype(np.float32) / np.float32(255.0)).astype(np.float32))
            for c in range(img.shape[2])]


class Instancer:
    """Device shell instancer behind the reference's constructor (instancer/instancer.pyx:6-36): explicit patch
    `transformations` (AddInstance, instancer.cpp:125-141) and / or patches distributed on a base mesh
    (`mesh_path`, `patch_origins_path`, `patch_scale`, `jitter_amount`: DistributeInstancesOnMesh, :233-391), whose
    first hit ends a ray with an opaque black last sample (:804-814,:1019-1023).  Rays walk uniform grids over the patch
    boxes and the mesh triangles (scene_grid.py) where the reference walks Embree BVHs; shadow rays towards the light
    (`cast_shadow_rays`, :593-603) walk the same grids; image entries of `textures` (PNG files or arrays) modulate
    their parameter through the base mesh's UVs (getParameters, :640-667); `auxiliary_meshes` = [(mesh, texture)] end
    rays too and are shaded (AddMesh / shadeMesh, :393-418, :716-744).
    `mesh` = {'V','F','N','UV'} arrays may be given in place of `mesh_path` (procedural meshes), `patch_origins` [A,3]
    in place of `patch_origins_path`."""

    def __init__(self, b_0: list, b_1: list, cast_shadow_rays=False, textures: list = [], transformations: list = [],
                 mesh_path: str = None, patch_scale: float = 1., patch_origins_path='', min_shadow_samples: int = 4,
                 n_shadow_samples: int = 512, min_texture_samples: int = 4, n_texture_samples: int = 512,
                 jitter_amount: float = 0, instance_sampling_method: str = 'random', use_mean_distance: bool = False,
                 auxiliary_meshes=[], transformation_export_path=None, *, seed: int = 0, device=None, mesh=None,
                 patch_origins=None, use_grid: bool = True) -> None:
        assert instance_sampling_method in ('random', 'nearest', 'nearest_blend')      # instancer.pyx:12
        self.cast_shadow_rays = bool(cast_shadow_rays)
        self.min_shadow_samples, self.n_shadow_samples = int(min_shadow_samples), int(n_shadow_samples)
        self.light_dir_idx = self.light_strength_idx = -1
        n_parameters = 0
        tex_channels, tex_param_idxs = [], []                                          # flat channel list, index per file
        for tex in textures:                                                           # instancer.cpp:75-92
            name = tex if isinstance(tex, str) or tex is None else None
            if name == 'light':                                                        # directional light: 3 parameters
                self.light_dir_idx = n_parameters
                n_parameters += 3
            elif name == 'point':                                                      # point light: strength + position
                self.light_strength_idx = n_parameters
                self.light_dir_idx = n_parameters + 1
                n_parameters += 4
            elif tex is None or name == '':
                n_parameters += 1
            else:                                                                      # an image: one parameter per channel
                ch = load_texture(tex)
                if not ch:                                                             # stbi_load failed: nothing is added (:38)
                    continue
                tex_channels += ch
                tex_param_idxs.append(n_parameters)
                n_parameters += len(ch)
        self.n_parameters = n_parameters
        import numpy as np
        from . import scene_grid
        T = np.asarray(transformations, dtype=np.float64).reshape(-1, 4, 4)             # AddInstance calls come first (:17-18)
        self.patch_scale = float(patch_scale)
        self.mesh = None
        if mesh_path is not None or mesh is not None:                                   # :20-23
            from . import distribute, ply
            m = ply.read_ply(mesh_path) if mesh is None else {k: mesh.get(k) for k in ("V", "F", "N", "UV")}
            if m["F"] is None or len(m["F"]) == 0:
                raise ValueError("Instancer: the base mesh has no faces")
            if m["N"] is None or m["UV"] is None:
                raise ValueError("Instancer: the base mesh needs per-vertex normals and texture coordinates "
                                 "(the tangent frames of DistributeInstancesOnMesh come from them, instancer.cpp:248-277)")
            if patch_origins is None and patch_origins_path:
                try:
                    patch_origins = ply.read_ply(patch_origins_path)["V"]
                except (OSError, ValueError):
                    patch_origins = None                                                # readPLY fails -> mesh vertices (:350)
            placed, scale = distribute.distribute_instances_on_mesh(m["V"], m["F"], m["N"], m["UV"], patch_origins,
                                                                    float(patch_scale), float(jitter_amount), seed=0)
            self.patch_scale = scale                                                    # :245 (average edge length if <= 0)
            T = np.concatenate([T, placed.astype(np.float64)], 0)
            self.mesh = {"V": np.ascontiguousarray(m["V"], np.float32), "F": np.ascontiguousarray(m["F"], np.int32),
                         "N": np.ascontiguousarray(m["N"], np.float32), "UV": np.ascontiguousarray(m["UV"], np.float32)}
        if T.shape[0] == 0:
            raise ValueError("Instancer: no instances (empty `transformations` and no mesh)")
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        rt = np.transpose(T[:, :3, :3], (0, 2, 1))
        as_dev = lambda a, dt=np.float32: torch.from_numpy(np.ascontiguousarray(np.asarray(a).astype(dt))).to(dev)
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
        self.method = {'random': 0, 'nearest': 1, 'nearest_blend': 2}[instance_sampling_method]   # instancer.pyx:12
        self.use_mean_distance = bool(use_mean_distance)
        self.seed = int(seed)
        self._calls = 0
        self._frame = -1                       # new_frame() before the first named-ray call makes it 0
        # CommitScene (:420-424): the acceleration structures
        self._keep = []                                                                  # device arrays the scene points into
        self.scene = _lib.InstancerScene()
        self.scene.minv, self.scene.dirt, self.scene.origins = self.minv.data_ptr(), self.dirt.data_ptr(), self.origins.data_ptr()
        self.scene.spheres = self.spheres.data_ptr()
        self.scene.n_inst, self.scene.n_faces = int(T.shape[0]), 0
        if use_grid:
            self._set_grid(self.scene.inst_grid, scene_grid.build_grid(world.min(1), world.max(1)), as_dev)
        # every mesh of the scene in one vertex / face list, the base mesh first (geomID order, :374-418)
        from . import ply as _ply
        self.aux_meshes = []
        for am, atex in auxiliary_meshes:                                               # instancer.pyx:25-26
            a = _ply.read_ply(am) if isinstance(am, str) else {k: am.get(k) for k in ("V", "F", "N", "UV")}
            if a["N"] is None:
                raise ValueError("Instancer: an auxiliary mesh needs per-vertex normals (shadeMesh, instancer.cpp:722-724)")
            ch = load_texture(atex) if atex is not None and not (isinstance(atex, str) and atex == "") else []
            if ch and a["UV"] is None:
                raise ValueError("Instancer: a textured auxiliary mesh needs texture coordinates")
            self.aux_meshes.append({"V": np.ascontiguousarray(a["V"], np.float32), "F": np.ascontiguousarray(a["F"], np.int32),
                                    "N": np.ascontiguousarray(a["N"], np.float32),
                                    "UV": None if a["UV"] is None else np.ascontiguousarray(a["UV"], np.float32),
                                    "textures": ch if len(ch) in (0, 1, 3) else ch[:1]})   # :733 other counts: first channel
        if len(self.aux_meshes) > _lib.MAX_AUX_MESHES or len(tex_param_idxs) > _lib.MAX_PARAM_TEXTURES:
            raise ValueError("Instancer: at most 8 auxiliary meshes and 8 parameter textures")
        meshes = ([self.mesh] if self.mesh is not None else []) + self.aux_meshes
        self.tex_channels, self.tex_param_idxs = tex_channels, tex_param_idxs
        # getParameters searches the base mesh within the patch extent (:63, :246)
        self.patch_max_extent = float(np.linalg.norm(np.maximum(np.asarray(b_0, np.float32), np.asarray(b_1, np.float32)))) * \
            (self.patch_scale if self.mesh is not None else 1.0)
        if meshes:
            n_v = np.cumsum([0] + [m["V"].shape[0] for m in meshes])
            V = np.concatenate([m["V"] for m in meshes])
            F = np.concatenate([m["F"] + n_v[i] for i, m in enumerate(meshes)]).astype(np.int32)
            Nn = np.concatenate([m["N"] for m in meshes])
            UV = np.concatenate([m["UV"] if m.get("UV") is not None else np.zeros((m["V"].shape[0], 2), np.float32) for m in meshes])
            self.scene_mesh = (V, F)
            mv, mf = as_dev(V), as_dev(F, np.int32)
            self._keep += [mv, mf]
            self.scene.mesh_v, self.scene.mesh_f, self.scene.n_faces = mv.data_ptr(), mf.data_ptr(), int(mf.shape[0])
            if use_grid:
                self._set_grid(self.scene.tri_grid, scene_grid.build_grid(*scene_grid.triangle_boxes(V, F)), as_dev)
            param_tex = bool(tex_param_idxs) and self.mesh is not None
            if param_tex or self.aux_meshes:
                sh = _lib.InstancerShading()
                pool, chans = [], []

                def add(ch):
                    t = _lib.Texture(sum(c.size for c in pool), ch.shape[0], ch.shape[1])
                    pool.append(np.ascontiguousarray(ch, np.float32).reshape(-1))
                    return t
                sh.n_base_faces = 0 if self.mesh is None else int(self.mesh["F"].shape[0])
                if param_tex:
                    sh.n_param_tex = len(tex_param_idxs)
                    for i, idx in enumerate(tex_param_idxs):                            # file i <-> channel i of the flat list (:661)
                        sh.param_idx[i] = int(idx)
                        sh.param_tex[i] = add(tex_channels[i])
                    sh.radius, sh.min_samples, sh.n_samples = self.patch_max_extent, int(min_texture_samples), int(n_texture_samples)
                sh.n_aux = len(self.aux_meshes)
                f0 = sh.n_base_faces
                for k, am in enumerate(self.aux_meshes):
                    sh.aux_face_begin[k] = f0
                    f0 += int(am["F"].shape[0])
                    sh.aux_channels[k] = len(am["textures"])
                    for c, ch in enumerate(am["textures"]):
                        sh.aux_tex[k][c] = add(ch)
                sh.aux_face_begin[len(self.aux_meshes)] = f0
                d_uv, d_n = as_dev(UV), as_dev(Nn)
                d_pool = as_dev(np.concatenate(pool) if pool else np.zeros(1, np.float32))
                self._keep += [d_uv, d_n, d_pool]
                sh.mesh_uv, sh.mesh_n, sh.pool = d_uv.data_ptr(), d_n.data_ptr(), d_pool.data_ptr()
                self._shading = sh
                self.scene.shading = C.pointer(sh)
        self.param_textures = bool(tex_param_idxs) and self.mesh is not None
        if transformation_export_path:
            np.save(transformation_export_path, T)                                    # instancer.pyx:33-36

    def _set_grid(self, dst, g, as_dev) -> None:
        import numpy as np
        cs, it = as_dev(g["cell_start"], np.int32), as_dev(g["items"], np.int32)
        self._keep += [cs, it]
        dst.cell_start, dst.items = cs.data_ptr(), it.data_ptr()
        for a in range(3):
            dst.lo[a], dst.cell[a], dst.res[a] = float(g["lo"][a]), float(g["cell"][a]), int(g["res"][a])

    def n_instances(self) -> int:
        return int(self.minv.shape[0])

    def new_frame(self) -> None:
        """Advance the draw seed of calls that name their rays (`ray_ids`): one seed per frame, so that every chunk and
        every GPU working on the frame draws from the same stream."""
        self._frame += 1

    def get_model_input_csr(self, rays_o, rays_d, parameters, n_samples: int, step_size: float, seed: Optional[int] = None,
                            ray_ids=None, ray_offsets=None):
        """`ray_offsets` [R]: per-ray offsets in [0, 1) instead of the counter hash (reproducing a reference run)."""
        from . import ops
        if seed is None:
            if ray_ids is not None:
                seed = self.seed + 7919 * max(self._frame, 0)
            else:                              # a fresh draw every call, like the reference's running mt19937
                seed = self.seed + 7919 * self._calls
                self._calls += 1
        return ops.instancer_get_model_input(rays_o, rays_d, parameters, n_samples, step_size, self.scene, self.b_0, self.b_1,
                                             seed=seed, light_dir_idx=self.light_dir_idx, method=self.method,
                                             use_mean_distance=self.use_mean_distance, light_strength_idx=self.light_strength_idx,
                                             # fp32 product, instancer.cpp:696; the reference's patch_scale member is only
                                             # set by DistributeInstancesOnMesh (:245) and stays 1 for explicit patches
                                             blend_range=float(np.float32(0.2) * np.float32(self.patch_scale if self.mesh is not None else 1.0)),
                                             ray_ids=ray_ids, cast_shadow_rays=self.cast_shadow_rays,
                                             min_shadow_samples=self.min_shadow_samples, n_shadow_samples=self.n_shadow_samples,
                                             param_textures=self.param_textures, ray_offsets=ray_offsets)

    def get_model_input(self, rays_o, rays_d, parameters, n_samples: int, step_size: float, seed: Optional[int] = None):
        """Dense 10-tuple of instancer.pyx:54."""
        csr = self.get_model_input_csr(rays_o, rays_d, parameters, n_samples, step_size, seed)
        return csr_to_dense(csr, as_cuda(rays_d), as_cuda(parameters), n_samples)

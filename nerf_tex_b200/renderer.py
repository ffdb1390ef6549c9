"""Ray marcher with the reference's surface (network/renderer.py) on the sm_100a kernels.

``Renderer`` / ``InstanceRenderer`` accept the reference's kwargs unchanged (unknown ones are
swallowed, renderer.py:34) and return the same dict of outputs.  The per-ray work is three
kernels per network pass — K3 sampling, K1 fused encoding+MLP, K2 compositing — so the
``render_chunk`` / ``net_chunk`` knobs, which only bound TensorFlow's activation memory in the
reference, become batching hints: a chunk is one launch of each kernel.

Random draws (``z_rand``, ``u``, ``noise``) come from a seeded torch CUDA generator, or can be
passed explicitly through ``__call__`` kwargs (that is how the parity tests feed the oracle
and the kernels identical draws).
"""

from __future__ import annotations

import math
from typing import Dict, Optional

import torch

from . import ops, util
from ._lib import as_cuda


def sample_pdf(bins, weights, N_samples: int, det: bool = False, u=None) -> torch.Tensor:
    """network/renderer.py:589-618 for [..., C] bins and [..., C-1] weights."""
    bins, weights = as_cuda(bins), as_cuda(weights)
    lead = bins.shape[:-1]
    out = ops.sample_pdf(bins.reshape(-1, bins.shape[-1]), weights.reshape(-1, weights.shape[-1]), N_samples, det,
                         None if u is None else as_cuda(u).reshape(-1, N_samples))
    return out.reshape(*lead, N_samples)


class Renderer:
    """network/renderer.py:31-213."""

    def __init__(self, model, model_fine=None, n_samples: int = 64, n_importance: int = 0, perturb: bool = True,
                 raw_noise_std: float = 0, render_chunk: int = 32768, net_chunk: int = 65536,
                 downsampling_factor: int = 1, blur_idx: Optional[int] = None, map_exr: bool = False,
                 seed: Optional[int] = None, **kwargs) -> None:
        self.model = model
        self.model_fine = model_fine
        self.n_samples = n_samples
        self.n_importance = n_importance
        self.perturb = perturb
        self.raw_noise_std = raw_noise_std
        self.render_chunk = render_chunk
        self.net_chunk = net_chunk
        self.downsampling_factor = downsampling_factor
        self.blur_idx = blur_idx
        self.map_exr = map_exr
        self._gen: Optional[torch.Generator] = None
        self._seed = seed

    # -- random draws -----------------------------------------------------------------------
    def _rand(self, shape, device, normal: bool = False) -> torch.Tensor:
        if self._gen is None or self._gen.device != device:
            self._gen = torch.Generator(device=device)
            if self._seed is not None:
                self._gen.manual_seed(self._seed)
        fn = torch.randn if normal else torch.rand
        return fn(shape, dtype=torch.float32, device=device, generator=self._gen)

    # -- entry ------------------------------------------------------------------------------
    def __call__(self, rays_o, rays_d, t, parameters, cone_scale, composite_bkgd: bool = False,
                 bkgd_color=(1., 1., 1.), training: bool = True, **kwargs) -> Dict[str, torch.Tensor]:
        """network/renderer.py:47-90: flatten, cull rays with t == inf, chunk, scatter back."""
        rays_o, rays_d, t = as_cuda(rays_o), as_cuda(rays_d), as_cuda(t)
        parameters, cone_scale = as_cuda(parameters), as_cuda(cone_scale)
        B, Nr = rays_o.shape[0], rays_o.shape[1]
        dev = rays_o.device
        ro, rd, tf_ = rays_o.reshape(-1, 3), rays_d.reshape(-1, 3), t.reshape(-1, 2)
        pf = parameters.repeat_interleave(Nr, dim=0)                           # :54
        cf = cone_scale.reshape(-1, 1)
        hit = tf_[:, 0] != math.inf                                            # :58
        n_all = ro.shape[0]
        idxs = torch.nonzero(hit).squeeze(1)
        all_hit = idxs.shape[0] == n_all
        extra = {k: as_cuda(v).reshape(n_all, -1) for k, v in kwargs.items()
                 if k in ("z_rand", "u", "noise", "noise_fine") and v is not None}
        ray_ids = self._frame_ray_ids(kwargs.get("ray_ids"), n_all, dev)         # instanced renderers: names for the draws
        if ray_ids is not None:
            extra["ray_ids"] = ray_ids
        if not all_hit:
            ro, rd, tf_, pf, cf = ro[idxs], rd[idxs], tf_[idxs], pf[idxs], cf[idxs]
            extra = {k: v[idxs] for k, v in extra.items()}
        ro, rd, tf_, pf, cf = (x.contiguous() for x in (ro, rd, tf_, pf, cf))

        outs: Dict[str, list] = {}
        n = ro.shape[0]
        for i in range(0, n, self.render_chunk):                               # :72
            sl = slice(i, i + self.render_chunk)
            chunk = self.render_rays(rays_o=ro[sl], rays_d=rd[sl], t=tf_[sl], parameters=pf[sl], cone_scale=cf[sl],
                                     composite_bkgd=composite_bkgd, bkgd_color=bkgd_color, training=training,
                                     **{k: v[sl] for k, v in extra.items()})
            for k, v in chunk.items():
                outs.setdefault(k, []).append(v)

        res: Dict[str, torch.Tensor] = {}
        bk = torch.as_tensor([float(c) for c in bkgd_color], dtype=torch.float32, device=dev)
        for k, lst in outs.items():                                            # :80-88
            v = lst[0] if len(lst) == 1 else torch.cat(lst, 0)
            if all_hit:
                full = v
            else:
                full = torch.zeros((n_all,) + tuple(v.shape[1:]), dtype=v.dtype, device=dev)
                full[idxs] = v
                if composite_bkgd and "color" in k:
                    full[~hit] = bk
            res[k] = full.reshape((B, -1) + tuple(v.shape[1:]))
        return res

    def _frame_ray_ids(self, ray_ids, n_all: int, dev):
        """Hook of the instanced renderers (the plain renderer draws nothing per ray id)."""
        return None

    # -- one chunk --------------------------------------------------------------------------
    def render_rays(self, rays_o, rays_d, t, parameters, cone_scale, composite_bkgd, bkgd_color, training,
                    z_rand=None, u=None, noise=None, noise_fine=None, return_aux: bool = False, **kwargs):
        """network/renderer.py:92-143."""
        R = rays_o.shape[0]
        dev = rays_o.device
        S = self.n_samples
        if self.perturb and z_rand is None:
            z_rand = self._rand((R, S), dev)                                     # :110
        grad = torch.is_grad_enabled() and (self.model.master.requires_grad or
                                            (self.model_fine is not None and self.model_fine.master.requires_grad))
        if not return_aux and not grad and S >= 3:
            # inference: the whole chunk is one C call (ntx_render_rays)
            if self.n_importance > 0 and not self.perturb and u is None:
                u = self._rand((R, self.n_importance), dev)
            if self.raw_noise_std > 0:
                if noise is None:
                    noise = self._rand((R, S), dev, normal=True) * self.raw_noise_std
                if self.n_importance > 0 and noise_fine is None:
                    noise_fine = self._rand((R, S + self.n_importance), dev, normal=True) * self.raw_noise_std
            cfg = ops.render_config(S, self.n_importance, self.perturb, self.map_exr, composite_bkgd, self.blur_idx,
                                    bkgd_color)
            need = ops.render_rays_workspace_bytes(R, S, self.n_importance)
            ws = getattr(self, "_workspace", None)
            if ws is None or ws.numel() < need or ws.device != dev:
                ws = None                      # (re)allocated by ops.render_rays and kept for the next chunk
            out, self._workspace = ops.render_rays(self.model, self.model_fine, cfg, rays_o, rays_d, t, parameters,
                                                   cone_scale, z_rand, u, noise, noise_fine, ws)
            return out
        rays_d_n = ops.normalize_dirs(rays_d)                                    # :98
        z = ops.stratified_z(t, S, bool(self.perturb), z_rand)                   # :101-111
        color_map, alpha_map, weights, rgbs = self._eval_and_composite(
            self.model, rays_o, rays_d, rays_d_n, z, parameters, cone_scale, composite_bkgd, bkgd_color, noise,
            want_weights=self.n_importance > 0 or return_aux)
        out = {"color_pred": color_map, "alpha_pred": alpha_map}
        aux = {"z_coarse": z, "weights_coarse": weights, "rgbs_coarse": rgbs}
        if self.n_importance > 0:                                                # :125-138
            det = bool(self.perturb)                                             # :128 det=self.perturb (sic)
            if not det and u is None:
                u = self._rand((R, self.n_importance), dev)
            z_all = ops.hierarchical_z(z, weights, self.n_importance, det, u)   # :127-130
            model_imp = self.model if self.model_fine is None else self.model_fine
            c_imp, a_imp, w_imp, rgbs_imp = self._eval_and_composite(
                model_imp, rays_o, rays_d, rays_d_n, z_all, parameters, cone_scale, composite_bkgd, bkgd_color,
                noise_fine, want_weights=return_aux)
            out = {"color_pred": c_imp, "alpha_pred": a_imp, "color_pred_coarse": color_map,
                   "alpha_pred_coarse": alpha_map}
            aux.update({"z_fine": z_all, "weights_fine": w_imp, "rgbs_fine": rgbs_imp})
        return (out, aux) if return_aux else out

    def _eval_and_composite(self, model, rays_o, rays_d, rays_d_n, z, parameters, cone_scale, composite_bkgd,
                            bkgd_color, noise, want_weights):
        R, S = z.shape
        if self.raw_noise_std > 0 and noise is None:                             # :190-192
            noise = self._rand((R, S), z.device, normal=True) * self.raw_noise_std
        if torch.is_grad_enabled() and model.master.requires_grad:
            # training: same kernels, plus saved operands and a backward (train.py:62-66 tape)
            from .autograd import NetworkPass
            cmap, amap, w = NetworkPass.apply(model.master, model, rays_o, rays_d, rays_d_n.contiguous(), z, parameters,
                                              cone_scale, self.blur_idx, noise, self.map_exr, composite_bkgd,
                                              tuple(float(c) for c in bkgd_color))
            return cmap, amap, w, None
        rgbs = self.evaluate_model_packed(rays_o, rays_d, rays_d_n, parameters, cone_scale, z, model)
        cmap, amap, w, _ = ops.composite_forward(rgbs, z, rays_d, noise=noise, map_exr=self.map_exr,
                                                 composite_bkgd=composite_bkgd, bkgd_color=bkgd_color,
                                                 want_weights=want_weights)
        return cmap, amap, w, rgbs

    def evaluate_model_packed(self, rays_o, rays_d, rays_d_n, parameters, cone_scale, z, model) -> torch.Tensor:
        """network/renderer.py:145-168 without materialising points or per-sample repeats: the
        kernel forms rays_o + rays_d * z and reads dir / parameters per ray."""
        R, S = z.shape
        npar = parameters.shape[1] if parameters is not None else 0
        blur = self.blur_idx is not None
        rgbs = model.forward_rows(n_rows=R * S, rays_o=rays_o, rays_d=rays_d, z=z, dir=rays_d_n.contiguous(),
                                  params=parameters if npar > 0 else None, rows_per_group=S,
                                  row_scale=z if blur else None,
                                  group_scale=cone_scale.reshape(-1).contiguous() if blur else None,
                                  blur_idx=self.blur_idx)
        return rgbs.view(R, S, 4)

    def evaluate_model(self, pos, dirs, parameters, cone_scale, z_vals, model, training):
        """Reference-shaped variant (renderer.py:145-168): explicit points in, (color, alpha) out."""
        R, S = pos.shape[0], pos.shape[1]
        blur = self.blur_idx is not None
        rgbs = model.forward_rows(n_rows=R * S, pos=as_cuda(pos).reshape(-1, 3), dir=as_cuda(dirs),
                                  params=as_cuda(parameters) if parameters.shape[1] > 0 else None, rows_per_group=S,
                                  row_scale=as_cuda(z_vals) if blur else None,
                                  group_scale=as_cuda(cone_scale).reshape(-1) if blur else None, blur_idx=self.blur_idx)
        rgbs = rgbs.view(R, S, 4)
        return rgbs[..., :3], rgbs[..., 3]

    def map_model_output(self, color, alpha, z_vals, rays_d, composite_bkgd, bkgd_color):
        """network/renderer.py:170-213 on the reference's split (color, alpha) layout."""
        noise = None
        if self.raw_noise_std > 0:
            noise = self._rand(tuple(alpha.shape), alpha.device, normal=True) * self.raw_noise_std
        cmap, amap, w, _ = ops.composite_forward(None, z_vals, rays_d, color=color, sigma=alpha, noise=noise,
                                                 map_exr=self.map_exr, composite_bkgd=composite_bkgd,
                                                 bkgd_color=bkgd_color)
        return cmap, amap, w


class InstanceRenderer(Renderer):
    """network/renderer.py:215-354: samples come from the instancer, only in-shell points are fed
    to the network.  Here they stay CSR-packed end to end (ray_ptr + per-sample records) instead
    of the reference's dense [R, S, .] host buffers."""

    def __init__(self, instancer_config=None, step_size: float = 0.002, density_scale: float = 1,
                 density_reweighting: bool = True, false_color: bool = False, **kwargs) -> None:
        super().__init__(**kwargs)
        self.instancer = util.instantiate(instancer_config) if isinstance(instancer_config, dict) else instancer_config
        self.step_size = step_size
        self.density_scale = density_scale
        self.density_reweighting = density_reweighting
        self.false_color = false_color
        self.instance_color = None
        if false_color:
            g = torch.Generator(device="cpu").manual_seed(0)
            self.instance_color = torch.rand((self.instancer.n_instances(), 3), generator=g).cuda()   # :227
        self.patch_scale = (instancer_config["patch_scale"] if isinstance(instancer_config, dict)
                            else getattr(self.instancer, "patch_scale", 1.0))                         # :228

    def _frame_ray_ids(self, ray_ids, n_all: int, dev):
        """One call = one frame for an instancer that draws per named ray (Instancer.new_frame / ray_ids): rays are named
        by their position in the call unless the caller names them (render_sharded passes the positions in the frame)."""
        if not hasattr(self.instancer, "new_frame"):
            return None
        self.instancer.new_frame()
        if ray_ids is None:
            return torch.arange(n_all, dtype=torch.int32, device=dev).reshape(n_all, 1)
        return as_cuda(ray_ids, torch.int32, device=dev).reshape(n_all, 1)

    def render_rays(self, rays_o, rays_d, t, parameters, cone_scale, composite_bkgd, bkgd_color, training,
                    ray_ids=None, **kwargs):
        assert training is False or training == False, \
            "network.renderer.InstanceRenderer can only be used for evaluation."                      # :233
        color_map, alpha_map = self.evaluate_model(rays_o, rays_d, parameters, cone_scale, self.model, composite_bkgd,
                                                   bkgd_color, False, ray_ids=ray_ids)
        return {"color_pred": color_map, "alpha_pred": alpha_map}

    def _instancer_records(self, rays_o, rays_d, parameters, ray_ids):
        """The instancer call of renderer.py:253, CSR-packed."""
        if hasattr(self.instancer, "get_model_input_csr"):
            if ray_ids is not None and hasattr(self.instancer, "new_frame"):
                return self.instancer.get_model_input_csr(rays_o, rays_d, parameters, self.n_samples, self.step_size,
                                                          ray_ids=ray_ids.reshape(-1).contiguous())
            return self.instancer.get_model_input_csr(rays_o, rays_d, parameters, self.n_samples, self.step_size)
        return dense_to_csr(self.instancer.get_model_input(rays_o, rays_d, parameters, self.n_samples, self.step_size))

    def evaluate_model(self, rays_o, rays_d, parameters, cone_scale, model, composite_bkgd, bkgd_color, training,
                       ray_ids=None):
        """network/renderer.py:247-316."""
        n_rays = rays_o.shape[0]
        dev = rays_o.device
        csr = self._instancer_records(rays_o, rays_d, parameters, ray_ids)
        H = csr["ray_index"].shape[0]
        if H == 0:                                                                # :256-257
            return (torch.zeros((n_rays, 3), dtype=torch.float32, device=dev),
                    torch.zeros((n_rays,), dtype=torch.float32, device=dev))
        T = csr["pts"].shape[0]
        row_scale = None
        if self.blur_idx is not None:                                             # :260-263
            counts = csr["ray_ptr"][1:] - csr["ray_ptr"][:-1]
            cone_rows = cone_scale.reshape(-1)[csr["ray_index"].long()].repeat_interleave(counts)
            row_scale = ((cone_rows * csr["t"]) / self.patch_scale).contiguous()
        if T > 0:                                                                 # :284-293 only in-shell points
            rgbs = model.forward_rows(n_rows=T, pos=csr["pts"], dir=csr["dirs"], params=csr["params"],
                                      rows_per_group=1, row_scale=row_scale, blur_idx=self.blur_idx)
        else:
            rgbs = torch.zeros((0, 4), dtype=torch.float32, device=dev)
        inst_col = None
        if self.false_color:                                                      # :306-307
            inst_col = self.instance_color[csr["instance_id"].long()].contiguous()
        return ops.composite_instanced_forward(
            rgbs, csr["dists"], csr["alpha_weight"] if self.density_reweighting else None, csr["ray_ptr"],
            csr["color_last"], csr["alpha_last"], csr["ray_index"], n_rays, patch_scale=self.patch_scale,
            density_scale=self.density_scale, map_exr=self.map_exr, composite_bkgd=composite_bkgd,
            bkgd_color=bkgd_color, instance_color=inst_col)


def dense_to_csr(instancer_out) -> Dict[str, torch.Tensor]:
    """Pack the reference instancer's dense 10-tuple (instancer.pyx:54) — rays_d_map[R,S,3],
    pts[R,S,3], t[R,S], dists[R,S] (0 = padding), color_last[R,1,3], alpha_last[R,1],
    alpha_weight[R,S], instance_id[R,S], idxs[H,1], params_map[R,S,p] — into the CSR records the
    kernels consume (renderer.py:266-287: cull missed rays, then points with dists <= 0)."""
    rays_d_map, pts, t, dists, color_last, alpha_last, alpha_weight, instance_id, idxs, params_map = instancer_out
    idxs = as_cuda(idxs, torch.int64).reshape(-1)
    g = lambda a, dt=torch.float32: as_cuda(a, dt)[idxs]
    rays_d_map, pts, t, dists = g(rays_d_map), g(pts), g(t), g(dists)
    color_last, alpha_last, alpha_weight = g(color_last), g(alpha_last), g(alpha_weight)
    instance_id, params_map = g(instance_id, torch.int32), g(params_map)
    mask = dists > 0
    counts = mask.sum(1)
    ray_ptr = torch.zeros(idxs.shape[0] + 1, dtype=torch.int64, device=dists.device)
    ray_ptr[1:] = torch.cumsum(counts, 0)
    return {
        "ray_ptr": ray_ptr, "ray_index": idxs.to(torch.int32), "pts": pts[mask].contiguous(),
        "dirs": rays_d_map[mask].contiguous(), "params": params_map[mask].contiguous(), "dists": dists[mask].contiguous(),
        "t": t[mask].contiguous(), "alpha_weight": alpha_weight[mask].contiguous(),
        "instance_id": instance_id[mask].contiguous(), "color_last": color_last.reshape(-1, 3).contiguous(),
        "alpha_last": alpha_last.reshape(-1).contiguous(),
    }


class MipRenderer(Renderer):
    """network/renderer.py:356-473: cone-segment gaussians + integrated positional encoding (the model is built with
    ``n_pos=6`` and an ``IntegratedPositionalEncoding`` pos_embedding); the blur parameter is spliced out of the
    parameter vector and becomes the cone radius (:384-386)."""

    def __init__(self, blur_idx: int = None, **kwargs):
        super().__init__(**kwargs)
        self.blur_idx_mip = blur_idx                                              # :362 hidden from the base class
        if blur_idx is None:
            raise ValueError("MipRenderer needs blur_idx (the parameter that scales the cone radius, renderer.py:384)")

    def render_rays(self, rays_o, rays_d, t, parameters, cone_scale, composite_bkgd, bkgd_color, training,
                    z_rand=None, noise=None, return_aux: bool = False, **kwargs):
        if self.n_importance > 0:                                                 # :402-403
            raise NotImplementedError("Importance sampling for mip-NeRF style rendering is not implemented "
                                      "(nor is it in the reference, network/renderer.py:402-403)")
        R, dev, S = rays_o.shape[0], rays_o.device, self.n_samples
        rays_d_n = ops.normalize_dirs(rays_d)                                     # :371
        if self.perturb and z_rand is None:
            z_rand = self._rand((R, S + 1), dev)                                  # :380
        z = ops.stratified_z(t, S + 1, bool(self.perturb), z_rand)                # :374-381  S + 1 segment edges
        b = self.blur_idx_mip
        blur = (parameters[:, b] * cone_scale.reshape(-1)).contiguous()           # :384
        params = torch.cat([parameters[:, :b], parameters[:, b + 1:]], -1).contiguous()   # :385
        mean, cov = ops.mip_cone_gaussians(rays_o, rays_d, z, blur)               # :388
        if self.raw_noise_std > 0 and noise is None:                              # :456-458
            noise = self._rand((R, S), dev, normal=True) * self.raw_noise_std
        model = self.model
        if torch.is_grad_enabled() and model.master.requires_grad:
            from .autograd import MipNetworkPass
            cmap, amap, w = MipNetworkPass.apply(model.master, model, mean, cov, rays_d, rays_d_n.contiguous(), z, params,
                                                 noise, self.map_exr, composite_bkgd, tuple(float(c) for c in bkgd_color))
        else:
            rgbs = model.forward_rows(n_rows=R * S, pos=mean.view(-1, 3), pos_cov=cov.view(-1, 3), dir=rays_d_n.contiguous(),
                                      params=params if params.shape[1] > 0 else None, rows_per_group=S).view(R, S, 4)   # :392
            cmap, amap, w = ops.composite_edges_forward(rgbs, z, rays_d, noise=noise, map_exr=self.map_exr,
                                                        composite_bkgd=composite_bkgd, bkgd_color=bkgd_color,
                                                        want_weights=return_aux)   # :396
        out = {"color_pred": cmap, "alpha_pred": amap}
        return (out, {"z": z, "mean": mean, "cov": cov, "weights": w}) if return_aux else out


class MipInstanceRenderer(InstanceRenderer):
    """network/renderer.py:475-587: the instanced renderer with integrated positional encodings."""

    def __init__(self, blur_idx: int = None, **kwargs):
        super().__init__(**kwargs)
        self.blur_idx_mip = blur_idx                                              # :481
        if blur_idx is None:
            raise ValueError("MipInstanceRenderer needs blur_idx (renderer.py:511)")

    def evaluate_model(self, rays_o, rays_d, parameters, cone_scale, model, composite_bkgd, bkgd_color, training,
                       ray_ids=None):
        n_rays, dev = rays_o.shape[0], rays_o.device
        csr = self._instancer_records(rays_o, rays_d, parameters, ray_ids)
        H = csr["ray_index"].shape[0]
        if H == 0:                                                                # :494-495
            return (torch.zeros((n_rays, 3), dtype=torch.float32, device=dev),
                    torch.zeros((n_rays,), dtype=torch.float32, device=dev))
        T = csr["pts"].shape[0]
        if T > 0:
            b = self.blur_idx_mip
            counts = csr["ray_ptr"][1:] - csr["ray_ptr"][:-1]
            cone_rows = cone_scale.reshape(-1)[csr["ray_index"].long()].repeat_interleave(counts)
            blur = ((csr["params"][:, b] * cone_rows) / self.patch_scale).contiguous()           # :511
            params = torch.cat([csr["params"][:, :b], csr["params"][:, b + 1:]], -1).contiguous()   # :512
            cov = ops.mip_cone_gaussians_instanced(csr["dirs"], csr["t"], blur, csr["dists"])    # :536
            rgbs = model.forward_rows(n_rows=T, pos=csr["pts"], pos_cov=cov, dir=csr["dirs"],
                                      params=params if params.shape[1] > 0 else None, rows_per_group=1)   # :543
        else:
            rgbs = torch.zeros((0, 4), dtype=torch.float32, device=dev)
        inst_col = None
        if self.false_color:                                                      # :558-559
            inst_col = self.instance_color[csr["instance_id"].long()].contiguous()
        return ops.composite_instanced_forward(
            rgbs, csr["dists"], csr["alpha_weight"] if self.density_reweighting else None, csr["ray_ptr"],
            csr["color_last"], csr["alpha_last"], csr["ray_index"], n_rays, patch_scale=self.patch_scale,
            density_scale=self.density_scale, map_exr=self.map_exr, composite_bkgd=composite_bkgd,
            bkgd_color=bkgd_color, instance_color=inst_col)

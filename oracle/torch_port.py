"""torch-CPU fp32 port of the instanced hot path, used ONLY as the timed CPU baseline of bench.py.
TEST INFRASTRUCTURE / BASELINE ONLY (never imported by the product).  Same graph as
oracle/nerftex_oracle.py::instance_evaluate_model (network/renderer.py:247-354, network/model.py:58-125,
network/layer.py:11-23), but with threaded fp32 kernels (oneDNN / MKL GEMMs, fp32 sin / cos) instead of
NumPy's float64 transcendentals and single-threaded elementwise ops, so the CPU side is not handicapped.
Pinned against the NumPy oracle in tests/test_torch_port.py."""

from __future__ import annotations

import os

import numpy as np
import torch


def _fourier(x: torch.Tensor, n: int) -> torch.Tensor:
    feats = [x]
    for k in range(n):                                           # layer.py:16-23
        a = x * float(2 ** k)
        feats += [torch.sin(a), torch.cos(a)]
    return torch.cat(feats, -1)


class TorchPort:
    def __init__(self, oracle_model, threads: int = 0) -> None:
        torch.set_num_threads(threads or len(os.sched_getaffinity(0)))
        self.m = oracle_model
        self.W = {n: (torch.from_numpy(np.ascontiguousarray(W)), torch.from_numpy(np.ascontiguousarray(b)))
                  for n, W, b in oracle_model.layers}

    @torch.no_grad()
    def forward(self, pos, dirs, params):
        m = self.m
        g = m.n_parameters[0]
        pos_map, dir_map = _fourier(pos, m.pos_freqs), _fourier(dirs, m.dir_freqs)
        if g > 0:
            pos_map = torch.cat([pos_map, _fourier(params[:, :g], m.param_freqs)], -1)
        if m.n_parameters[1] > 0:
            dir_map = torch.cat([dir_map, _fourier(params[:, g:], m.param_freqs)], -1)
        h = pos_map
        for i in range(m.depth):                                 # model.py:104-108
            W, b = self.W[f"trunk{i}"]
            h = torch.addmm(b, h, W).relu_()
            if i in m.skips:
                h = torch.cat([pos_map, h], -1)
        alpha = torch.addmm(self.W["alpha"][1], h, self.W["alpha"][0])
        h = torch.cat([dir_map, torch.addmm(self.W["feature"][1], h, self.W["feature"][0])], -1)
        for i in range(m.color_depth):
            W, b = self.W[f"colorhid{i}"]
            h = torch.addmm(b, h, W).relu_()
        h = torch.addmm(self.W["colorhalf"][1], h, self.W["colorhalf"][0]).relu_()
        return torch.addmm(self.W["rgb"][1], h, self.W["rgb"][0]), alpha

    @torch.no_grad()
    def instance_evaluate_model(self, dense, n_rays, patch_scale=1.0, density_scale=1.0, net_chunk=32768,
                                composite_bkgd=False, bkgd_color=(1., 1., 1.)):
        """renderer.py:247-354 on the dense 10-tuple (NumPy arrays), blur_idx None, no false colour."""
        rays_d_map, pts, t, dists, color_last, alpha_last, alpha_weight, _, idxs, params_map = dense
        idxs = torch.as_tensor(np.asarray(idxs).reshape(-1), dtype=torch.int64)
        cm, am = torch.zeros((n_rays, 3)), torch.zeros((n_rays,))
        if idxs.numel() == 0:
            return cm.numpy(), am.numpy()
        g = lambda a: torch.from_numpy(np.ascontiguousarray(a))[idxs]          # :266-273
        rays_d_map, pts, dists, alpha_weight, params_map = g(rays_d_map), g(pts), g(dists), g(alpha_weight), g(params_map)
        color_last, alpha_last = g(color_last), g(alpha_last)
        H, S = dists.shape
        ip = torch.nonzero(dists.reshape(-1) > 0).squeeze(1)                   # :284
        pos, dirs, prm = pts.reshape(-1, 3)[ip], rays_d_map.reshape(-1, 3)[ip], params_map.reshape(-1, params_map.shape[-1])[ip]
        color, alpha = torch.zeros((H * S, 3)), torch.zeros((H * S, 1))
        for i in range(0, ip.shape[0], net_chunk):                             # :290-293
            c, a = self.forward(pos[i:i + net_chunk], dirs[i:i + net_chunk], prm[i:i + net_chunk])
            color[ip[i:i + net_chunk]], alpha[ip[i:i + net_chunk]] = c, a
        alpha = alpha.reshape(H, S) * alpha_weight * density_scale             # :300
        rgb = torch.cat([torch.sigmoid(color.reshape(H, S, 3)), color_last.reshape(H, 1, 3)], 1)   # :322-331
        a = 1 - torch.exp(-torch.relu(alpha) * dists / patch_scale)            # :339
        a = torch.cat([a, alpha_last.reshape(H, 1)], 1)
        trans = torch.cumprod(torch.cat([torch.ones(H, 1), 1 - a[:, :-1] + 1e-10], 1), 1)
        w = a * trans
        cmap, amap = (w[..., None] * rgb).sum(1), w.sum(1)
        if composite_bkgd:
            cmap = cmap + (1 - amap)[:, None] * torch.tensor(bkgd_color)
        cm[idxs], am[idxs] = cmap, amap                                        # :313-314
        return cm.numpy(), am.numpy()

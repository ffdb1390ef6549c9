"""Backward pass of one network pass (K1 + K2) as a torch.autograd.Function.

Forward: ``ntx_mlp_forward_train`` (fused encoding + MLP that also saves every tensor-core operand as bf16 tile
images plus ReLU sign bits) followed by ``ntx_composite_forward``.  Backward: ``ntx_composite_backward`` (warp per
ray) gives d(raw rgb, sigma); ``ntx_mlp_backward_data`` (fused tcgen05 data-gradient chain) and
``ntx_mlp_backward_weights`` (tcgen05 X^T dY on the saved images, CUDA-core heads) give d loss / d master — all
hand-written kernels of csrc/ntx_mlp_bwd.cu, no library GEMM (a layer-wise torch.mm chain on the same saved operands
lives in tests/library_backward.py as a cross-check only).  No gradient flows into the encodings, the sampled depths
or the compositing weights that drive resampling (network/renderer.py:129 stop_gradient).

The result is ``d loss / d master`` on the model's flat fp32 weight buffer, in the reference's Dense order, so the
Adam kernel (or the NCCL all-reduce of train.py) works on one tensor.
"""

from __future__ import annotations

import ctypes as C
from typing import Dict, List

import torch

from . import _lib, ops
from ._lib import MlpInputs, check, ptr, stream_ptr

ENC_COLS = 96


class MlpSchedule:
    """Host mirror of the kernel's layer schedule (ntx_mlp_step_info / ntx_mlp_enc_columns)."""

    def __init__(self, model) -> None:
        lib = _lib.load()
        h = model._h
        self.n_steps = lib.ntx_mlp_num_steps(h)
        self.steps: List[Dict[str, int]] = []
        for st in range(self.n_steps):
            v = [C.c_int32() for _ in range(6)]
            check(lib.ntx_mlp_step_info(h, st, *[C.byref(x) for x in v]))
            self.steps.append(dict(zip(("layer", "n_out", "main_row0", "enc_data", "phase", "relu"), [x.value for x in v])))
        self.enc_maps = []
        for phase in (0, 1):
            arr = (C.c_int32 * ENC_COLS)()
            check(lib.ntx_mlp_enc_columns(h, phase, arr))
            ref = torch.tensor(list(arr), dtype=torch.int64)
            self.ones_col = getattr(self, "ones_col", []) + [int(torch.nonzero(ref == -1).squeeze(1)[0])]
            cols = torch.nonzero(ref >= 0).squeeze(1)
            self.enc_maps.append((cols.to(model.device), ref[cols].to(model.device)))
        self.depth = model.cfg.depth
        self.alpha_layer = self.depth                       # Dense order: trunk0..trunk{d-1}, alpha, feature, ...
        self.rgb_layer = len(model.layers) - 1


def mlp_forward_train(model, *, n_rows, dir, params, rows_per_group, pos=None, rays_o=None, rays_d=None, z=None,
                      row_scale=None, group_scale=None, blur_idx=None, row_major_acts: bool = False, pos_cov=None):
    """Raw rgbs[N,4] plus the saved operands: (enc_pos, enc_dir, acts_img, masks[, acts])."""
    packed = model.ensure_packed()
    sched = model.schedule()
    dev = model.device
    lib = _lib.load()
    n_tiles = int(lib.ntx_mlp_train_tiles(int(n_rows)))
    rgbs = torch.empty((n_rows, 4), dtype=torch.float32, device=dev)
    # encoding tile images: 4 blocks of [128 x 32]; the kernel writes blocks 0..2, block 3 stays zero.
    # Zero-filled once, then recycled through a small per-model pool (released by the backward).
    pool = model.__dict__.setdefault("_enc_img_pool", {}).setdefault(n_tiles, [])
    enc_pos = enc_dir = None
    for ent in pool:
        if not ent[2]:
            ent[2] = True
            enc_pos, enc_dir = ent[0], ent[1]
            break
    if enc_pos is None:
        if len(pool) >= 4:
            pool.clear()
        enc_pos = torch.zeros((n_tiles, 128 * 128), dtype=torch.bfloat16, device=dev)
        enc_dir = torch.zeros((n_tiles, 128 * 128), dtype=torch.bfloat16, device=dev)
        pool.append([enc_pos, enc_dir, True])
    masks = torch.empty((sched.n_steps, n_tiles * 128, 8), dtype=torch.int32, device=dev)
    inp = MlpInputs(ptr(pos), ptr(rays_o), ptr(rays_d), ptr(z), ptr(dir), ptr(params), ptr(row_scale), ptr(group_scale),
                    -1 if blur_idx is None else int(blur_idx), int(rows_per_group), int(n_rows), ptr(pos_cov))
    f8 = fp8_state(model)
    if f8 is not None and f8["calibrated"] and not row_major_acts:
        # 8-bit saved operands: scales from the maxima of the previous step, images of 32 KB per tile
        check(lib.ntx_fp8_update_scales(ptr(f8["amax_x"]), ptr(f8["scale_x"]), ptr(f8["inv_x"]), sched.n_steps, FP8_MARGIN, stream_ptr()))
        acts_img8 = torch.empty((sched.n_steps, n_tiles, 128 * 256), dtype=torch.uint8, device=dev)
        check(lib.ntx_mlp_forward_train_fp8(model._h, ptr(packed), C.byref(inp), ptr(rgbs), ptr(enc_pos), ptr(enc_dir),
                                            ptr(acts_img8), ptr(f8["inv_x"]), ptr(f8["amax_x"]), ptr(masks), n_tiles, stream_ptr()))
        return rgbs, (enc_pos, enc_dir, acts_img8, masks, None, n_rows)
    # (the calibration pass of the 8-bit mode takes maxima over whole images: parts no kernel writes must read as zero)
    alloc = torch.zeros if (f8 is not None and not f8["calibrated"]) else torch.empty
    acts_img = alloc((sched.n_steps, n_tiles, 128 * 256), dtype=torch.bfloat16, device=dev)
    acts = torch.empty((sched.n_steps, n_rows, 256), dtype=torch.bfloat16, device=dev) if row_major_acts else None
    check(lib.ntx_mlp_forward_train(model._h, ptr(packed), C.byref(inp), ptr(rgbs), ptr(enc_pos), ptr(enc_dir),
                                    ptr(acts), ptr(acts_img), ptr(masks), n_tiles, stream_ptr()))
    return rgbs, (enc_pos, enc_dir, acts_img, masks, acts, n_rows)


FP8_MARGIN = 2.0          # head-room of the delayed scales: values up to twice last step's maximum stay finite


def fp8_state(model):
    """Per-model state of the 8-bit saved-operand mode (enabled with model.enable_fp8_saved_operands()): per-step
    scales, their inverses and the running maxima, for the activations (x) and the output gradients (y)."""
    return getattr(model, "_fp8", None)


def enable_fp8_saved_operands(model, enable: bool = True) -> None:
    """Training only: save the tile images the backward needs as fp8 e4m3 with per-step, per-tensor scales delayed by
    one step.  The first pass after enabling runs with bf16 images and measures the maxima (calibration)."""
    if not enable:
        model._fp8 = None
        return
    dev = model.device
    z32 = lambda: torch.zeros(16, dtype=torch.float32, device=dev)
    model._fp8 = {"calibrated": False, "scale_x": z32() + 1, "inv_x": z32() + 1, "amax_x": torch.zeros(16, dtype=torch.int32, device=dev),
                  "scale_y": z32() + 1, "inv_y": z32() + 1, "amax_y": torch.zeros(16, dtype=torch.int32, device=dev)}


def _calibrate(f8, key, img) -> None:
    """Maxima of a bf16 image stack [n_steps, ...] as fp32 bit patterns (what the kernels' atomicMax would have left)."""
    am = img.reshape(img.shape[0], -1).abs().amax(1).float()
    f8[key][:am.shape[0]] = torch.maximum(f8[key][:am.shape[0]].view(torch.float32), am).view(torch.int32)


def release_saved(model, saved) -> None:
    """Hand the encoding images of a finished pass back to the model's pool."""
    for pool in model.__dict__.get("_enc_img_pool", {}).values():
        for ent in pool:
            if ent[0] is saved[0]:
                ent[2] = False


def untile(img: torch.Tensor, n_rows: int, cols: int = 256) -> torch.Tensor:
    """[n_tiles, 128*W] tile images (W = 256 activations / gradients, 128 encodings) -> row-major bf16 [n_rows, cols]."""
    out = torch.empty((n_rows, cols), dtype=torch.bfloat16, device=img.device)
    check(_lib.load().ntx_untile(ptr(img), int(n_rows), int(cols), int(img.shape[1]) * 2, ptr(out), stream_ptr()))
    return out


def mlp_backward_native(model, saved, d_rgbs: torch.Tensor, return_dy: bool = False):
    """d loss / d master with the hand-written kernels: ntx_mlp_backward_data (fused dgrad chain) and
    ntx_mlp_backward_weights (X^T dY, 1^T dY, enc^T dY and the two heads on the saved tile images).  The host
    only scatters the (at most 96) encoded-input rows of the three layers that read an encoding from the
    kernel's column order to the reference's row order, and sums d_rgbs for the heads' biases."""
    enc_pos, enc_dir, acts_img, masks, _, n_rows = saved
    lib = _lib.load()
    sched = model.schedule()
    n_tiles = acts_img.shape[1]
    master = model.master.detach()
    if getattr(model, "_packed_t_version", None) != model.master._version:
        if getattr(model, "_packed_t", None) is None:
            model._packed_t = torch.empty(int(lib.ntx_mlp_packed_t_bytes(model._h)), dtype=torch.uint8, device=model.device)
        check(lib.ntx_mlp_pack_weights_t(model._h, ptr(master), ptr(model._packed_t), stream_ptr()))
        model._packed_t_version = model.master._version
    d_rgbs = d_rgbs.contiguous()
    grad = torch.zeros_like(master)
    enc_scratch = torch.zeros((sched.n_steps, 128, 256), dtype=torch.float32, device=master.device)
    f8 = fp8_state(model)
    if acts_img.dtype == torch.uint8:                                     # 8-bit saved operands
        check(lib.ntx_fp8_update_scales(ptr(f8["amax_y"]), ptr(f8["scale_y"]), ptr(f8["inv_y"]), sched.n_steps, FP8_MARGIN, stream_ptr()))
        dy_img8 = torch.empty_like(acts_img)
        # bf16 dY only for the steps that read an encoding (their weight gradient keeps bf16 operands)
        dy_img = torch.empty((sched.n_steps, n_tiles, 128 * 256), dtype=torch.bfloat16, device=master.device)
        check(lib.ntx_mlp_backward_data_fp8(model._h, ptr(model._packed_t), ptr(d_rgbs), ptr(masks), ptr(dy_img), ptr(dy_img8),
                                            ptr(f8["inv_y"]), ptr(f8["amax_y"]), n_rows, n_tiles, stream_ptr()))
        check(lib.ntx_mlp_backward_weights_fp8(model._h, ptr(acts_img), ptr(dy_img8), ptr(dy_img), ptr(f8["scale_x"]),
                                               ptr(f8["scale_y"]), ptr(enc_pos), ptr(enc_dir), ptr(d_rgbs), ptr(grad),
                                               ptr(enc_scratch), n_rows, n_tiles, stream_ptr()))
    else:
        dy_img = torch.zeros_like(acts_img) if (f8 is not None and not f8["calibrated"]) else torch.empty_like(acts_img)
        check(lib.ntx_mlp_backward_data(model._h, ptr(model._packed_t), ptr(d_rgbs), ptr(masks), ptr(dy_img), n_rows, n_tiles,
                                        stream_ptr()))
        check(lib.ntx_mlp_backward_weights(model._h, ptr(acts_img), ptr(dy_img), ptr(enc_pos), ptr(enc_dir), ptr(d_rgbs),
                                           ptr(grad), ptr(enc_scratch), n_rows, n_tiles, stream_ptr()))
        if f8 is not None and not f8["calibrated"]:                       # calibration pass of the 8-bit mode
            _calibrate(f8, "amax_x", acts_img)
            _calibrate(f8, "amax_y", dy_img)
            f8["calibrated"] = True

    def gview(layer):
        wo, bo, fi, fo = model.layers[layer]
        return grad[wo:wo + fi * fo].view(fi, fo), grad[bo:bo + fo]

    for st, s in enumerate(sched.steps):
        if not s["enc_data"]:
            continue
        gW, gb = gview(s["layer"])
        cols, rows = sched.enc_maps[s["phase"]]
        gW[rows] = enc_scratch[st][cols, :s["n_out"]]
        if st == 0:                                                       # no main operand: the bias gradient is the
            gb.copy_(enc_scratch[st][sched.ones_col[s["phase"]], :s["n_out"]])    # constant-1 column's row
    dsum = d_rgbs.sum(0)
    gview(sched.rgb_layer)[1].copy_(dsum[:3])                             # model.py:123
    gview(sched.alpha_layer)[1].copy_(dsum[3:4])                          # model.py:111
    return (grad, dy_img) if return_dy else grad


class NetworkPass(torch.autograd.Function):
    """(master, inputs...) -> (color_map, alpha_map, weights) for one network over one ray chunk."""

    @staticmethod
    def forward(ctx, master, model, rays_o, rays_d, rays_d_n, z, parameters, cone_scale, blur_idx, noise, map_exr,
                composite_bkgd, bkgd_color):
        R, S = z.shape
        npar = parameters.shape[1] if parameters is not None else 0
        blur = blur_idx is not None
        rgbs, saved = mlp_forward_train(model, n_rows=R * S, rays_o=rays_o, rays_d=rays_d, z=z, dir=rays_d_n,
                                        params=parameters if npar > 0 else None, rows_per_group=S,
                                        row_scale=z if blur else None,
                                        group_scale=cone_scale.reshape(-1).contiguous() if blur else None,
                                        blur_idx=blur_idx)
        rgbs = rgbs.view(R, S, 4)
        cmap, amap, w, _ = ops.composite_forward(rgbs, z, rays_d, noise=noise, map_exr=map_exr,
                                                 composite_bkgd=composite_bkgd, bkgd_color=bkgd_color)
        ctx.model, ctx.saved = model, saved
        ctx.misc = (rgbs, z, rays_d, noise, map_exr, composite_bkgd, bkgd_color)
        ctx.mark_non_differentiable(w)
        return cmap, amap, w

    @staticmethod
    def backward(ctx, d_cmap, d_amap, d_w):
        rgbs, z, rays_d, noise, map_exr, composite_bkgd, bkgd_color = ctx.misc
        d_rgbs = ops.composite_backward(rgbs, z, rays_d, d_cmap.contiguous(), d_amap.contiguous(), noise=noise,
                                        map_exr=map_exr, composite_bkgd=composite_bkgd, bkgd_color=bkgd_color)
        grad = mlp_backward_native(ctx.model, ctx.saved, d_rgbs.view(-1, 4))
        release_saved(ctx.model, ctx.saved)
        ctx.saved = None
        return (grad,) + (None,) * 12


class MipNetworkPass(torch.autograd.Function):
    """(master, mean, cov, ...) -> (color_map, alpha_map, weights): the mip renderer's network pass over one ray
    chunk (network/renderer.py:388-396): integrated encoding of explicit cone gaussians, S + 1 depth edges."""

    @staticmethod
    def forward(ctx, master, model, mean, cov, rays_d, rays_d_n, z_edges, parameters, noise, map_exr, composite_bkgd,
                bkgd_color):
        R, S = z_edges.shape[0], z_edges.shape[1] - 1
        npar = parameters.shape[1] if parameters is not None else 0
        rgbs, saved = mlp_forward_train(model, n_rows=R * S, pos=mean.reshape(-1, 3), pos_cov=cov.reshape(-1, 3),
                                        dir=rays_d_n, params=parameters if npar > 0 else None, rows_per_group=S)
        rgbs = rgbs.view(R, S, 4)
        cmap, amap, w = ops.composite_edges_forward(rgbs, z_edges, rays_d, noise=noise, map_exr=map_exr,
                                                    composite_bkgd=composite_bkgd, bkgd_color=bkgd_color)
        ctx.model, ctx.saved = model, saved
        ctx.misc = (rgbs, z_edges, rays_d, noise, map_exr, composite_bkgd, bkgd_color)
        ctx.mark_non_differentiable(w)
        return cmap, amap, w

    @staticmethod
    def backward(ctx, d_cmap, d_amap, d_w):
        rgbs, z_edges, rays_d, noise, map_exr, composite_bkgd, bkgd_color = ctx.misc
        d_rgbs = ops.composite_edges_backward(rgbs, z_edges, rays_d, d_cmap.contiguous(), d_amap.contiguous(), noise=noise,
                                              map_exr=map_exr, composite_bkgd=composite_bkgd, bkgd_color=bkgd_color)
        grad = mlp_backward_native(ctx.model, ctx.saved, d_rgbs.view(-1, 4))
        release_saved(ctx.model, ctx.saved)
        ctx.saved = None
        return (grad,) + (None,) * 11

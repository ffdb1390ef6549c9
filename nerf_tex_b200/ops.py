"""Tensor-level wrappers over the C ABI (one function per ``ntx_*`` entry point).

Inputs are torch CUDA tensors (or any DLPack exporter, see ``_lib.as_cuda``); outputs are freshly
allocated torch CUDA tensors.  All launches go to torch's current stream.
"""

from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import as_cuda, check, ptr, stream_ptr


def _bkgd(bkgd_color) -> "C.Array":
    vals = [float(v) for v in (bkgd_color.tolist() if hasattr(bkgd_color, "tolist") else bkgd_color)]
    return (C.c_float * 3)(*vals)


# ------------------------------------------------------------------------------------------ K3

def stratified_z(t: torch.Tensor, n_samples: int, perturb: bool, z_rand: Optional[torch.Tensor] = None) -> torch.Tensor:
    """network/renderer.py:101-111.  t[R,2] -> z[R,S]; z_rand[R,S] required iff perturb."""
    t = as_cuda(t)
    R = t.shape[0]
    z = torch.empty((R, n_samples), dtype=torch.float32, device=t.device)
    if perturb:
        if z_rand is None:
            raise _lib.NtxError("stratified_z: perturb=True needs explicit uniform draws z_rand[R,S]")
        z_rand = as_cuda(z_rand)
        assert z_rand.shape == (R, n_samples)
    check(_lib.load().ntx_stratified_z(ptr(t), ptr(z_rand) if perturb else None, R, n_samples, int(perturb), ptr(z),
                                       stream_ptr()))
    return z


def sample_pdf(bins: torch.Tensor, weights: torch.Tensor, n_samples: int, det: bool = False,
               u: Optional[torch.Tensor] = None, return_inds: bool = False):
    """network/renderer.py:589-618 on [R,C] bins / [R,C-1] weights."""
    bins, weights = as_cuda(bins), as_cuda(weights)
    R, Cb = bins.shape
    assert weights.shape == (R, Cb - 1)
    if not det:
        if u is None:
            u = torch.rand((R, n_samples), dtype=torch.float32, device=bins.device)
        u = as_cuda(u)
        assert u.shape == (R, n_samples)
    samples = torch.empty((R, n_samples), dtype=torch.float32, device=bins.device)
    inds = torch.empty((R, n_samples), dtype=torch.int32, device=bins.device) if return_inds else None
    check(_lib.load().ntx_sample_pdf(ptr(bins), ptr(weights), None if det else ptr(u), R, Cb, n_samples, int(det),
                                     ptr(samples), ptr(inds), stream_ptr()))
    return (samples, inds) if return_inds else samples


def hierarchical_z(z: torch.Tensor, weights: torch.Tensor, n_importance: int, det: bool,
                   u: Optional[torch.Tensor] = None, return_aux: bool = False):
    """network/renderer.py:127-130: merged, sorted depths [R,S+N]."""
    z, weights = as_cuda(z), as_cuda(weights)
    R, S = z.shape
    assert weights.shape == (R, S)
    if not det:
        if u is None:
            u = torch.rand((R, n_importance), dtype=torch.float32, device=z.device)
        u = as_cuda(u)
    z_out = torch.empty((R, S + n_importance), dtype=torch.float32, device=z.device)
    zs = torch.empty((R, n_importance), dtype=torch.float32, device=z.device) if return_aux else None
    inds = torch.empty((R, n_importance), dtype=torch.int32, device=z.device) if return_aux else None
    check(_lib.load().ntx_hierarchical_z(ptr(z), ptr(weights), None if det else ptr(u), R, S, n_importance, int(det),
                                         ptr(z_out), ptr(zs), ptr(inds), stream_ptr()))
    return (z_out, zs, inds) if return_aux else z_out


# ------------------------------------------------------------------------------------------ K2

def composite_forward(rgbs: Optional[torch.Tensor], z: torch.Tensor, rays_d: torch.Tensor, *,
                      color: Optional[torch.Tensor] = None, sigma: Optional[torch.Tensor] = None,
                      noise: Optional[torch.Tensor] = None, map_exr: bool = False, composite_bkgd: bool = False,
                      bkgd_color=(1., 1., 1.), want_weights: bool = True, want_depth: bool = False):
    """network/renderer.py:170-213.  Either packed rgbs[R,S,4] or split color[R,S,3] + sigma[R,S]."""
    z, rays_d = as_cuda(z), as_cuda(rays_d)
    R, S = z.shape
    packed = rgbs is not None
    if packed:
        rgbs = as_cuda(rgbs)
        assert rgbs.shape == (R, S, 4)
        cptr, sptr = ptr(rgbs), None
    else:
        color, sigma = as_cuda(color), as_cuda(sigma)
        assert color.shape == (R, S, 3) and sigma.shape == (R, S)
        cptr, sptr = ptr(color), ptr(sigma)
    if noise is not None:
        noise = as_cuda(noise)
    dev = z.device
    cmap = torch.empty((R, 3), dtype=torch.float32, device=dev)
    amap = torch.empty((R,), dtype=torch.float32, device=dev)
    w = torch.empty((R, S), dtype=torch.float32, device=dev) if want_weights else None
    d = torch.empty((R,), dtype=torch.float32, device=dev) if want_depth else None
    check(_lib.load().ntx_composite_forward(cptr, sptr, int(packed), ptr(z), ptr(rays_d), ptr(noise), R, S,
                                            int(map_exr), int(composite_bkgd), _bkgd(bkgd_color), ptr(cmap),
                                            ptr(amap), ptr(w), ptr(d), stream_ptr()))
    return cmap, amap, w, d


def composite_backward(rgbs, z, rays_d, d_color_map, d_alpha_map, *, noise=None, map_exr=False, composite_bkgd=False,
                       bkgd_color=(1., 1., 1.)) -> torch.Tensor:
    rgbs, z, rays_d = as_cuda(rgbs), as_cuda(z), as_cuda(rays_d)
    d_color_map, d_alpha_map = as_cuda(d_color_map), as_cuda(d_alpha_map)
    R, S = z.shape
    if noise is not None:
        noise = as_cuda(noise)
    d_rgbs = torch.empty_like(rgbs)
    check(_lib.load().ntx_composite_backward(ptr(rgbs), ptr(z), ptr(rays_d), ptr(noise), R, S, int(map_exr),
                                             int(composite_bkgd), _bkgd(bkgd_color), ptr(d_color_map),
                                             ptr(d_alpha_map), ptr(d_rgbs), stream_ptr()))
    return d_rgbs


def composite_edges_forward(rgbs, z_edges, rays_d, *, noise=None, map_exr=False, composite_bkgd=False,
                            bkgd_color=(1., 1., 1.), want_weights: bool = True):
    """MipRenderer.map_model_output (network/renderer.py:436-473): rgbs[R,S,4], z_edges[R,S+1]."""
    rgbs, z_edges, rays_d = as_cuda(rgbs), as_cuda(z_edges), as_cuda(rays_d)
    R, S = rgbs.shape[0], rgbs.shape[1]
    assert z_edges.shape == (R, S + 1) and rgbs.shape == (R, S, 4)
    if noise is not None:
        noise = as_cuda(noise)
    dev = rgbs.device
    cmap = torch.empty((R, 3), dtype=torch.float32, device=dev)
    amap = torch.empty((R,), dtype=torch.float32, device=dev)
    w = torch.empty((R, S), dtype=torch.float32, device=dev) if want_weights else None
    check(_lib.load().ntx_composite_edges_forward(ptr(rgbs), ptr(z_edges), ptr(rays_d), ptr(noise), R, S, int(map_exr),
                                                  int(composite_bkgd), _bkgd(bkgd_color), ptr(cmap), ptr(amap), ptr(w),
                                                  stream_ptr()))
    return cmap, amap, w


def composite_edges_backward(rgbs, z_edges, rays_d, d_color_map, d_alpha_map, *, noise=None, map_exr=False,
                             composite_bkgd=False, bkgd_color=(1., 1., 1.)) -> torch.Tensor:
    rgbs, z_edges, rays_d = as_cuda(rgbs), as_cuda(z_edges), as_cuda(rays_d)
    d_color_map, d_alpha_map = as_cuda(d_color_map), as_cuda(d_alpha_map)
    R, S = rgbs.shape[0], rgbs.shape[1]
    if noise is not None:
        noise = as_cuda(noise)
    d_rgbs = torch.empty_like(rgbs)
    check(_lib.load().ntx_composite_edges_backward(ptr(rgbs), ptr(z_edges), ptr(rays_d), ptr(noise), R, S, int(map_exr),
                                                   int(composite_bkgd), _bkgd(bkgd_color), ptr(d_color_map),
                                                   ptr(d_alpha_map), ptr(d_rgbs), stream_ptr()))
    return d_rgbs


def mip_cone_gaussians(rays_o, rays_d, z_edges, radii):
    """MipRenderer.get_cone_segment_gaussians (network/renderer.py:411-434) -> mean[R,S,3], cov[R,S,3]."""
    rays_o, rays_d, z_edges, radii = as_cuda(rays_o), as_cuda(rays_d), as_cuda(z_edges), as_cuda(radii).reshape(-1)
    R, S = z_edges.shape[0], z_edges.shape[1] - 1
    assert radii.shape[0] == R
    mean = torch.empty((R, S, 3), dtype=torch.float32, device=z_edges.device)
    cov = torch.empty_like(mean)
    check(_lib.load().ntx_mip_cone_gaussians(ptr(rays_o), ptr(rays_d), ptr(z_edges), ptr(radii), R, S, ptr(mean), ptr(cov),
                                             stream_ptr()))
    return mean, cov


def mip_cone_gaussians_instanced(dirs, t, radii, dists) -> torch.Tensor:
    """MipInstanceRenderer.get_cone_segment_gaussians (network/renderer.py:570-587) on packed samples -> cov[T,3]."""
    dirs, t, radii, dists = as_cuda(dirs), as_cuda(t), as_cuda(radii), as_cuda(dists)
    T = t.shape[0]
    cov = torch.empty((T, 3), dtype=torch.float32, device=t.device)
    check(_lib.load().ntx_mip_cone_gaussians_instanced(ptr(dirs), ptr(t), ptr(radii), ptr(dists), T, ptr(cov), stream_ptr()))
    return cov


def composite_instanced_forward(rgbs, dists, alpha_weight, ray_ptr, color_last, alpha_last, ray_index, n_out, *,
                                patch_scale: float, density_scale: float = 1.0, map_exr: bool = False,
                                composite_bkgd: bool = False, bkgd_color=(1., 1., 1.), instance_color=None):
    """network/renderer.py:300,318-354 over CSR-packed samples; rows not hit stay zero (:313-314)."""
    rgbs, dists = as_cuda(rgbs), as_cuda(dists)
    ray_ptr = as_cuda(ray_ptr, torch.int64)
    color_last, alpha_last = as_cuda(color_last), as_cuda(alpha_last)
    H = ray_ptr.shape[0] - 1
    if alpha_weight is not None:
        alpha_weight = as_cuda(alpha_weight)
    if ray_index is not None:
        ray_index = as_cuda(ray_index, torch.int32)
    if instance_color is not None:
        instance_color = as_cuda(instance_color)
    dev = rgbs.device
    cmap = torch.zeros((n_out, 3), dtype=torch.float32, device=dev)
    amap = torch.zeros((n_out,), dtype=torch.float32, device=dev)
    check(_lib.load().ntx_composite_instanced_forward(
        ptr(rgbs), ptr(dists), ptr(alpha_weight), ptr(ray_ptr), ptr(color_last), ptr(alpha_last), ptr(ray_index),
        ptr(instance_color), H, float(patch_scale), float(density_scale), int(map_exr), int(composite_bkgd),
        _bkgd(bkgd_color), ptr(cmap), ptr(amap), stream_ptr()))
    return cmap, amap


# ------------------------------------------------------------------------------------------ camera rays

def rays_from_camera(image_plane_loc, height: int, width: int, focal: float, c2w, *, normalize: bool = False,
                     aabb=None, near: float = 0.0, far: float = 0.0, n: Optional[int] = None, device=None):
    """network/ray_sampler.py:6-48 + network/proxy.py:13-35 in one launch.  image_plane_loc[n,2] or None (first n
    pixels of the frame, row-major).  aabb = (b_0, b_1) or None (constant near/far).  Returns rays_o, rays_d, t,
    cone_scale[n,1]."""
    if image_plane_loc is not None:
        image_plane_loc = as_cuda(image_plane_loc, device=device)
        n = image_plane_loc.shape[0]
        dev = image_plane_loc.device
    else:
        n = height * width if n is None else int(n)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    m = [float(v) for v in (c2w.detach().cpu().reshape(-1).tolist() if isinstance(c2w, torch.Tensor) else
                            [x for row in c2w for x in row])]
    assert len(m) == 16, "c2w must be 4x4"
    c2w_h = (C.c_float * 16)(*m)
    b0 = (C.c_float * 3)(*[float(v) for v in aabb[0]]) if aabb is not None else None
    b1 = (C.c_float * 3)(*[float(v) for v in aabb[1]]) if aabb is not None else None
    rays_o = torch.empty((n, 3), dtype=torch.float32, device=dev)
    rays_d = torch.empty((n, 3), dtype=torch.float32, device=dev)
    t = torch.empty((n, 2), dtype=torch.float32, device=dev)
    cone = torch.empty((n, 1), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.load().ntx_rays_from_camera(ptr(image_plane_loc), n, int(height), int(width), float(focal), c2w_h,
                                               int(bool(normalize)), b0, b1, float(near), float(far), ptr(rays_o),
                                               ptr(rays_d), ptr(t), ptr(cone), stream_ptr()))
    return rays_o, rays_d, t, cone


# ------------------------------------------------------------------------------------------ instancer

def instancer_get_model_input(rays_o, rays_d, parameters, n_pts: int, step_size: float, scene, b_0, b_1, *,
                              seed: int = 0, light_dir_idx: int = -1, method: int = 0, use_mean_distance: bool = False,
                              light_strength_idx: int = -1, blend_range: float = 0.0, ray_ids=None,
                              cast_shadow_rays: bool = False, min_shadow_samples: int = 4, n_shadow_samples: int = 512,
                              param_textures: bool = False, ray_offsets=None):
    """Two-pass shell instancer (instancer/src/instancer.cpp:751-1037): trace + count, prefix sum, march.
    `scene`: _lib.InstancerScene (device pointers kept alive by the caller).  `ray_ids` [R] int32: the rays' names for
    the random draws.  `ray_offsets` [R] float32 (optional): the per-ray offsets u in [0, 1) to use instead of the
    counter hash (to reproduce a run of the reference, whose std::mt19937 stream the caller then supplies).  Returns the CSR dict the instanced compositor consumes (rows = rays that touched a patch box or
    the base mesh)."""
    rays_o, rays_d, parameters = as_cuda(rays_o), as_cuda(rays_d), as_cuda(parameters)
    dev = rays_o.device
    R, p = rays_o.shape[0], parameters.shape[1]
    if ray_ids is not None:
        ray_ids = as_cuda(ray_ids, torch.int32, device=dev).reshape(-1)
        assert ray_ids.shape[0] == R
    if ray_offsets is not None:
        ray_offsets = as_cuda(ray_offsets, torch.float32, device=dev).reshape(-1)
        assert ray_offsets.shape[0] == R
    lib = _lib.load()
    b0 = (C.c_float * 3)(*[float(v) for v in b_0])
    b1 = (C.c_float * 3)(*[float(v) for v in b_1])
    n_samples = torch.empty(R, dtype=torch.int32, device=dev)
    hit = torch.empty(R, dtype=torch.int32, device=dev)
    alpha_last_all = torch.empty(R, dtype=torch.float32, device=dev)
    shadows = int(bool(cast_shadow_rays) and light_dir_idx >= 0)
    segments = int(shadows or param_textures)          # both look samples up along the ray's segments
    ws = torch.empty(max(int(lib.ntx_instancer_workspace_bytes(R, segments)), 16), dtype=torch.uint8, device=dev)
    color_last_all = torch.empty((R, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib.ntx_instancer_trace(ptr(rays_o), ptr(rays_d), ptr(ray_ids), ptr(ray_offsets), R, C.byref(scene), b0, b1,
                                      float(step_size), int(n_pts), int(seed) & 0xFFFFFFFF, int(bool(use_mean_distance)), segments,
                                      ptr(parameters) if p > 0 else None, p, int(light_dir_idx),
                                      ptr(ws), ptr(n_samples), ptr(hit), ptr(alpha_last_all), ptr(color_last_all), stream_ptr()))
        ray_ptr_all = torch.zeros(R + 1, dtype=torch.int64, device=dev)
        torch.cumsum(n_samples, 0, out=ray_ptr_all[1:])
        ray_index = torch.nonzero(hit).squeeze(1)
        T = int(ray_ptr_all[-1].item())
        f = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
        out = {"pts": f(T, 3), "dirs": f(T, 3), "params": f(T, p), "dists": f(T), "t": f(T), "alpha_weight": f(T),
               "instance_id": torch.empty(T, dtype=torch.int32, device=dev)}
        check(lib.ntx_instancer_march(ptr(rays_o), ptr(rays_d), ptr(ray_ids), ptr(parameters), p, R, C.byref(scene),
                                      float(step_size), int(n_pts), int(seed) & 0xFFFFFFFF, int(light_dir_idx), int(light_strength_idx),
                                      int(method), float(blend_range),
                                      int(bool(use_mean_distance)), shadows, int(min_shadow_samples), int(n_shadow_samples), b0, b1,
                                      ptr(ws), ptr(ray_ptr_all), T, ptr(out["pts"]), ptr(out["dirs"]),
                                      ptr(out["params"]), ptr(out["dists"]), ptr(out["t"]), ptr(out["alpha_weight"]),
                                      ptr(out["instance_id"]), stream_ptr()))
    H = ray_index.shape[0]
    # samples of non-hit rays do not exist, so the per-hit-ray offsets are a gather of the all-ray prefix sum
    out["ray_ptr"] = torch.cat([ray_ptr_all[:-1][ray_index], ray_ptr_all[-1:]]) if H > 0 else torch.zeros(1, dtype=torch.int64, device=dev)
    out["ray_index"] = ray_index.to(torch.int32)
    out["n_samples_all"] = n_samples
    out["color_last"] = color_last_all[ray_index].contiguous()                     # black on the base mesh, shaded on auxiliary ones
    out["alpha_last"] = alpha_last_all[ray_index].contiguous()                     # 1 where the ray ends on the mesh
    return out


# ------------------------------------------------------------------------------------------ whole chunk

def render_config(n_samples, n_importance, perturb, map_exr, composite_bkgd, blur_idx, bkgd_color) -> "_lib.RenderConfig":
    vals = [float(v) for v in (bkgd_color.tolist() if hasattr(bkgd_color, "tolist") else bkgd_color)]
    return _lib.RenderConfig(int(n_samples), int(n_importance), int(bool(perturb)), int(bool(map_exr)),
                             int(bool(composite_bkgd)), -1 if blur_idx is None else int(blur_idx), (C.c_float * 3)(*vals))


def normalize_dirs(rays_d: torch.Tensor) -> torch.Tensor:
    """network/renderer.py:98."""
    rays_d = as_cuda(rays_d)
    out = torch.empty_like(rays_d)
    check(_lib.load().ntx_normalize_dirs(ptr(rays_d), rays_d.shape[0], ptr(out), stream_ptr()))
    return out


def render_rays_workspace_bytes(R: int, n_samples: int, n_importance: int) -> int:
    return int(_lib.load().ntx_render_rays_workspace_bytes(int(R), int(n_samples), int(n_importance)))


def render_rays(model, model_fine, cfg, rays_o, rays_d, t, parameters, cone_scale, z_rand=None, u=None, noise=None,
                noise_fine=None, workspace: Optional[torch.Tensor] = None, entry: str = "ntx_render_rays"):
    """One chunk of Renderer.render_rays (network/renderer.py:92-143) in a single C call."""
    lib = _lib.load()
    rays_o = as_cuda(rays_o)
    dev = rays_o.device
    f32 = lambda a: None if a is None else as_cuda(a, device=dev)            # dtype / device / contiguity checked once
    rays_d, t, parameters, z_rand, u, noise, noise_fine = (f32(a) for a in (rays_d, t, parameters, z_rand, u, noise,
                                                                             noise_fine))
    # bound to a local so that a reshaped / re-laid-out copy outlives the launch (it must not be freed and its block
    # handed to the workspace allocation below)
    cone_flat = None if cone_scale is None else as_cuda(cone_scale, device=dev).reshape(-1).contiguous()
    R = rays_o.shape[0]
    fine = cfg.n_importance > 0
    color = torch.empty((R, 3), dtype=torch.float32, device=dev)
    alpha = torch.empty((R,), dtype=torch.float32, device=dev)
    color_c = torch.empty((R, 3), dtype=torch.float32, device=dev) if fine else None
    alpha_c = torch.empty((R,), dtype=torch.float32, device=dev) if fine else None
    npar = parameters.shape[1] if parameters is not None else 0
    packed = model.ensure_packed()
    packed_f = model_fine.ensure_packed() if model_fine is not None else None
    args = [model._h, ptr(packed), model_fine._h if model_fine is not None else None, ptr(packed_f), C.byref(cfg),
            ptr(rays_o), ptr(rays_d), ptr(t), ptr(parameters) if npar > 0 else None,
            ptr(cone_flat), ptr(z_rand), ptr(u), ptr(noise),
            ptr(noise_fine), R, ptr(color), ptr(alpha), ptr(color_c), ptr(alpha_c)]
    if entry == "ntx_render_rays":
        need = lib.ntx_render_rays_workspace_bytes(R, cfg.n_samples, cfg.n_importance)
        if workspace is None:
            workspace = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
        check(lib.ntx_render_rays(*args, ptr(workspace), workspace.numel(), stream_ptr()))
    else:
        check(getattr(lib, entry)(*args, stream_ptr()))
    out = {"color_pred": color, "alpha_pred": alpha}
    if fine:
        out.update({"color_pred_coarse": color_c, "alpha_pred_coarse": alpha_c})
    return out, workspace


# ------------------------------------------------------------------------------------------ self-tests

def selftest_umma(A: torch.Tensor, B: torch.Tensor, layout: int) -> torch.Tensor:
    A, B = as_cuda(A), as_cuda(B)
    assert A.shape[0] == 128 and A.shape[1] == B.shape[1]
    D = torch.empty((128, B.shape[0]), dtype=torch.float32, device=A.device)
    check(_lib.load().ntx_selftest_umma(ptr(A), ptr(B), B.shape[0], A.shape[1], layout, ptr(D), stream_ptr()))
    return D


def selftest_umma_mn(At: torch.Tensor, Bt: torch.Tensor) -> torch.Tensor:
    At, Bt = as_cuda(At), as_cuda(Bt)
    assert At.shape[1] == 128 and At.shape[0] == Bt.shape[0]
    D = torch.empty((128, Bt.shape[1]), dtype=torch.float32, device=At.device)
    check(_lib.load().ntx_selftest_umma_mn(ptr(At), ptr(Bt), Bt.shape[1], At.shape[0], ptr(D), stream_ptr()))
    return D


def selftest_umma_mn8(At: torch.Tensor, Bt: torch.Tensor) -> torch.Tensor:
    At, Bt = as_cuda(At), as_cuda(Bt)
    assert At.shape[1] == 128 and At.shape[0] == Bt.shape[0]
    D = torch.empty((128, Bt.shape[1]), dtype=torch.float32, device=At.device)
    check(_lib.load().ntx_selftest_umma_mn8(ptr(At), ptr(Bt), Bt.shape[1], At.shape[0], ptr(D), stream_ptr()))
    return D


def selftest_bulk_copy(src: torch.Tensor) -> torch.Tensor:
    dst = torch.empty_like(src)
    check(_lib.load().ntx_selftest_bulk_copy(ptr(src), ptr(dst), src.numel() * src.element_size(), stream_ptr()))
    return dst

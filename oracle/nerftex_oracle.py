"""CPU oracle for the NeRF-Tex volume-rendering hot path.  TEST INFRASTRUCTURE ONLY.

This file is a CPU restatement (NumPy fp32, every reduction in a *defined* order) of the
reference algorithm.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU
baseline legs may import it; the product path (``nerf_tex_b200``) never does.

PINNING.  The reference ships no tests, golden vectors or fixtures and its
arithmetic lives in tensorflow-gpu==2.4.1 (``/root/reference/environment.yml:12``), which cannot be
installed in this image.  What pins this file:
  * the REFERENCE'S OWN PYTHON SOURCE (network/renderer.py, model.py, layer.py, loss.py,
    ray_sampler.py, proxy.py), imported from /root/reference and run over a NumPy stand-in for the
    TensorFlow ops it calls (oracle/tf_shim): tests/test_ref_python.py, and the vectors that run
    produced (tests/golden/ref_python_small.npz, tests/golden/make_ref_python.py) for machines
    without the reference tree.  That pins the reference's LOGIC as written — which tensors are
    sliced, gathered, concatenated, scattered, in which order.
  * NOT pinned: TensorFlow's own kernels (the stand-in restates each op's documented semantics, as
    listed below), so reduction orders and transcendental round-off remain restated, not observed;
  * closed-form known answers (tests/test_oracle_kat.py), an independent C restatement
    (oracle/ntx_oracle.c) that must agree bit for bit, a torch transcription
    (tests/test_oracle_vs_torch.py).

Every function cites the reference file:line it follows (paths relative to
``/root/reference``).  TF semantics restated from documentation:
  * ``tf.linspace(a, b, n)``: endpoints exact, interior ``a + fl((b-a)/(n-1)) * i``.
  * ``tf.math.cumprod(exclusive=True)``, ``tf.cumsum``: strict left-to-right fp32.
  * ``tf.reduce_sum`` over the sample axis: strict left-to-right fp32 (TF's real order is
    unknowable here; the oracle *defines* it).
  * ``tf.searchsorted(side='right')`` == upper_bound.
  * Keras ``Dense``: ``y = x @ W[in, out] + b``, glorot_uniform / zeros.
All random draws (z_rand, u, sigma noise) are explicit inputs.
"""

from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

F32 = np.float32


def _f32(x) -> np.ndarray:
    return np.ascontiguousarray(x, dtype=np.float32)


# ----------------------------------------------------------------------------------------
# TF primitives restated
# ----------------------------------------------------------------------------------------

def tf_linspace(start: float, stop: float, num: int) -> np.ndarray:
    """tf.linspace semantics (TF 2.4 math_ops.linspace_nd): exact endpoints,
    interior = start + delta * i with delta = (stop - start) / (num - 1) in fp32."""
    start = F32(start)
    stop = F32(stop)
    if num == 1:
        return np.array([start], dtype=F32)
    delta = F32(F32(stop - start) / F32(num - 1))
    i = np.arange(num, dtype=F32)
    out = (start + delta * i).astype(F32)
    out[0] = start
    out[-1] = stop
    return out


def seq_sum(x: np.ndarray) -> np.ndarray:
    """Strict left-to-right fp32 sum over the last axis."""
    x = _f32(x)
    acc = np.zeros(x.shape[:-1], dtype=F32)
    for i in range(x.shape[-1]):
        acc = (acc + x[..., i]).astype(F32)
    return acc


def seq_cumsum(x: np.ndarray) -> np.ndarray:
    x = _f32(x)
    out = np.empty_like(x)
    acc = np.zeros(x.shape[:-1], dtype=F32)
    for i in range(x.shape[-1]):
        acc = (acc + x[..., i]).astype(F32)
        out[..., i] = acc
    return out


def seq_cumprod_exclusive(x: np.ndarray) -> np.ndarray:
    x = _f32(x)
    out = np.empty_like(x)
    acc = np.ones(x.shape[:-1], dtype=F32)
    for i in range(x.shape[-1]):
        out[..., i] = acc
        acc = (acc * x[..., i]).astype(F32)
    return out


# ----------------------------------------------------------------------------------------
# network/layer.py
# ----------------------------------------------------------------------------------------

def fourier_features(x: np.ndarray, n_freq_bands: int) -> np.ndarray:
    """network/layer.py:11-23.  concat([x] + [sin(2^k x), cos(2^k x) for k < n], -1):
    identity block first, then per frequency a d-wide sin block and a d-wide cos block."""
    x = _f32(x)
    feats = [x]
    for k in range(n_freq_bands):
        f = F32(2.0 ** k)
        arg = (f * x).astype(F32)
        # sin/cos of the fp32 argument, correctly rounded via fp64
        feats.append(np.sin(arg.astype(np.float64)).astype(F32))
        feats.append(np.cos(arg.astype(np.float64)).astype(F32))
    return np.concatenate(feats, axis=-1)


def integrated_positional_encoding(x: np.ndarray, n_freq_bands: int) -> np.ndarray:
    """network/layer.py:25-41: x = [mean(3), diagonal covariance(3)] -> [sin block, cos block] * exp(-var/2), no identity."""
    x = _f32(x)
    freq = (2.0 ** np.arange(n_freq_bands)).astype(F32)
    y = (x[..., None, :3] * freq[:, None]).reshape(-1, 3 * n_freq_bands)
    y_var = (x[..., None, 3:] * freq[:, None] ** 2).reshape(-1, 3 * n_freq_bands)
    xs = np.concatenate([y, y + F32(0.5 * math.pi)], -1)
    vs = np.concatenate([y_var, y_var], -1)
    return (np.sin(xs.astype(np.float64)) * np.exp(-0.5 * vs.astype(np.float64))).astype(F32)


# ----------------------------------------------------------------------------------------
# network/model.py
# ----------------------------------------------------------------------------------------

class ParamNerfOracle:
    """network/model.py:58-125 (ParamNerf) — and Nerf (:9-45) as the
    ``n_parameters=[0,0], color_depth=0`` case.  Weights are ``W[in, out]`` + ``b[out]``
    (Keras Dense), glorot-uniform / zeros from a NumPy generator (TF's stream is not
    reproducible outside TF)."""

    def __init__(self, n_parameters=(1, 4), n_pos: int = 3, pos_freqs: int = 10, dir_freqs: int = 4,
                 param_freqs: int = 4, depth: int = 8, width: int = 256, skips: Sequence[int] = (4,),
                 color_depth: int = 1, seed: int = 1):
        if isinstance(n_parameters, int):
            n_parameters = [n_parameters, 0]                      # model.py:63-64
        self.n_parameters = list(n_parameters)
        self.n_pos, self.depth, self.width = n_pos, depth, width
        self.skips, self.color_depth = list(skips), color_depth
        self.pos_freqs, self.dir_freqs, self.param_freqs = pos_freqs, dir_freqs, param_freqs
        g, a = self.n_parameters
        # n_pos = 6 is only ever used with IntegratedPositionalEncoding (layer.py:25-41): 2 * 3 * pos_freqs features
        self.P = (6 * pos_freqs if n_pos == 6 else n_pos * (1 + 2 * pos_freqs)) + g * (1 + 2 * param_freqs)
        self.D = 3 * (1 + 2 * dir_freqs) + a * (1 + 2 * param_freqs)
        rng = np.random.default_rng(seed)
        self.layers: List[Tuple[str, np.ndarray, np.ndarray]] = []

        def dense(name, fan_in, fan_out):
            lim = math.sqrt(6.0 / (fan_in + fan_out))
            W = rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(F32)
            b = np.zeros(fan_out, dtype=F32)
            self.layers.append((name, W, b))

        k = self.P
        for i in range(depth):                                   # model.py:104-108
            dense(f"trunk{i}", k, width)
            k = width + self.P if i in self.skips else width
        dense("alpha", k, 1)                                     # model.py:111
        dense("feature", k, width)                               # model.py:114
        k = width + self.D                                       # model.py:115
        for i in range(color_depth):                             # model.py:118-119
            dense(f"colorhid{i}", k, width)
            k = width
        dense("colorhalf", k, width // 2)                        # model.py:122
        dense("rgb", width // 2, 3)                              # model.py:123

    # -- weight access ---------------------------------------------------------------
    def named_weights(self) -> Dict[str, Tuple[np.ndarray, np.ndarray]]:
        return {n: (W, b) for n, W, b in self.layers}

    def randomize_biases(self, seed: int, scale: float = 0.1) -> None:
        rng = np.random.default_rng(seed)
        self.layers = [(n, W, rng.uniform(-scale, scale, size=b.shape).astype(F32)) for n, W, b in self.layers]

    def n_params(self) -> int:
        return sum(W.size + b.size for _, W, b in self.layers)

    def macs_per_sample(self) -> int:
        return sum(W.size for _, W, _ in self.layers)

    # -- forward ---------------------------------------------------------------------
    def encode(self, pos, dirs, params):
        g = self.n_parameters[0]
        if self.n_pos == 6:
            pos_map = integrated_positional_encoding(pos, self.pos_freqs)     # model.py:77 with layer.py:25-41
        else:
            pos_map = fourier_features(pos, self.pos_freqs)                   # model.py:77
        dir_map = fourier_features(dirs, self.dir_freqs)                      # model.py:78
        params = _f32(params)
        if g > 0:                                                             # model.py:88-93
            pos_map = np.concatenate([pos_map, fourier_features(params[:, :g], self.param_freqs)], -1)
        if self.n_parameters[1] > 0:                                          # model.py:96-101
            dir_map = np.concatenate([dir_map, fourier_features(params[:, g:], self.param_freqs)], -1)
        return pos_map, dir_map

    def forward(self, pos, dirs, params, dtype=np.float32):
        """Returns raw (color[N,3], alpha[N,1]) — activation happens in the compositor."""
        pos_map, dir_map = self.encode(pos, dirs, params)
        pos_map = pos_map.astype(dtype)
        dir_map = dir_map.astype(dtype)
        L = {n: (W.astype(dtype), b.astype(dtype)) for n, W, b in self.layers}
        h = pos_map
        for i in range(self.depth):
            W, b = L[f"trunk{i}"]
            h = np.maximum(h @ W + b, 0)
            if i in self.skips:
                h = np.concatenate([pos_map, h], -1)                          # model.py:107-108
        W, b = L["alpha"]
        alpha = h @ W + b
        W, b = L["feature"]
        h = h @ W + b
        h = np.concatenate([dir_map, h], -1)
        for i in range(self.color_depth):
            W, b = L[f"colorhid{i}"]
            h = np.maximum(h @ W + b, 0)
        W, b = L["colorhalf"]
        h = np.maximum(h @ W + b, 0)
        W, b = L["rgb"]
        color = h @ W + b
        return color.astype(F32), alpha.astype(F32)


def _bf16(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even to bfloat16, returned as fp32."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    return ((u + r) & np.uint32(0xFFFF0000)).view(np.float32)


def forward_bf16_emulated(model: "ParamNerfOracle", pos, dirs, params):
    """The reference graph (model.py:58-125) evaluated with the CUDA kernel's quantisation points:
    encodings, weights, biases of the tensor-core layers and stored activations rounded to bf16,
    fp32 accumulation, fp32 density / colour heads fed by the un-rounded fp32 activations.
    Separates logic errors (must match to ~1e-3) from bf16 rounding (looser north-star tolerance)."""
    pos_map, dir_map = model.encode(pos, dirs, params)
    pos_map, dir_map = _bf16(pos_map), _bf16(dir_map)
    L = model.named_weights()
    q = lambda n: (_bf16(L[n][0]).astype(np.float64), _bf16(L[n][1]).astype(np.float64))
    h = pos_map.astype(np.float64)
    h32 = None
    for i in range(model.depth):
        W, b = q(f"trunk{i}")
        h32 = np.maximum(h @ W + b, 0).astype(F32)
        h = _bf16(h32).astype(np.float64)
        if i in model.skips:
            h = np.concatenate([pos_map.astype(np.float64), h], -1)
    alpha = (h32.astype(np.float64) @ L["alpha"][0].astype(np.float64) + L["alpha"][1]).astype(F32)
    W, b = q("feature")
    h = _bf16((h @ W + b).astype(F32)).astype(np.float64)
    h = np.concatenate([dir_map.astype(np.float64), h], -1)
    for i in range(model.color_depth):
        W, b = q(f"colorhid{i}")
        h = _bf16(np.maximum(h @ W + b, 0).astype(F32)).astype(np.float64)
    W, b = q("colorhalf")
    h32 = np.maximum(h @ W + b, 0).astype(F32)
    color = (h32.astype(np.float64) @ L["rgb"][0].astype(np.float64) + L["rgb"][1]).astype(F32)
    return color, alpha


# ----------------------------------------------------------------------------------------
# network/renderer.py — plain renderer
# ----------------------------------------------------------------------------------------

def stratified_z(t: np.ndarray, n_samples: int, perturb: bool, z_rand: Optional[np.ndarray]) -> np.ndarray:
    """network/renderer.py:101-111."""
    t = _f32(t)
    t_vals = tf_linspace(0.0, 1.0, n_samples)
    one_minus = (F32(1.0) - t_vals).astype(F32)
    z = ((t[:, None, 0] * one_minus).astype(F32) + (t[:, None, 1] * t_vals).astype(F32)).astype(F32)
    if perturb:
        mids = (F32(0.5) * (z[..., 1:] + z[..., :-1]).astype(F32)).astype(F32)
        upper = np.concatenate([mids, z[..., -1:]], -1)
        lower = np.concatenate([z[..., :1], mids], -1)
        z_rand = _f32(z_rand)
        z = (lower + ((upper - lower).astype(F32) * z_rand).astype(F32)).astype(F32)
    return z


def ray_points(rays_o, rays_d, z_vals):
    """network/renderer.py:114."""
    return (_f32(rays_o)[..., None, :] + _f32(rays_d)[..., None, :] * _f32(z_vals)[..., :, None]).astype(F32)


def color_activation(color: np.ndarray, map_exr: bool) -> np.ndarray:
    """network/renderer.py:182-187.  sigmoid, or elu(x)+1 when map_exr."""
    c = color.astype(np.float64)
    if map_exr:
        return np.where(c > 0, c + 1.0, np.exp(np.minimum(c, 0.0))).astype(F32)
    return (1.0 / (1.0 + np.exp(-c))).astype(F32)


def composite(color, alpha, z_vals, rays_d, composite_bkgd=False, bkgd_color=(1., 1., 1.), map_exr=False,
              noise: Optional[np.ndarray] = None, edges: bool = False):
    """network/renderer.py:170-213 (Renderer.map_model_output).

    ``noise`` is the already-scaled ``N(0,1) * raw_noise_std`` tensor (explicit input).
    Returns (color_map[R,3], alpha_map[R], weights[R,S], depth_map[R]); the reference
    computes depth (:204) and discards it."""
    color, alpha, z_vals, rays_d = _f32(color), _f32(alpha), _f32(z_vals), _f32(rays_d)
    dists = (z_vals[..., 1:] - z_vals[..., :-1]).astype(F32)
    if not edges:
        dists = np.concatenate([dists, dists[..., -1:]], -1)                   # :177 last dist duplicated
    norm = np.sqrt(seq_sum(rays_d * rays_d)).astype(F32)                       # :180
    dists = (dists * norm[..., None]).astype(F32)
    rgb = color_activation(color, map_exr)
    sig = alpha if noise is None else (alpha + _f32(noise)).astype(F32)
    sig = np.maximum(sig, F32(0))
    a = (F32(1) - np.exp(-(sig * dists).astype(np.float64)).astype(F32)).astype(F32)   # :195
    trans = seq_cumprod_exclusive(((F32(1) - a).astype(F32) + F32(1e-10)).astype(F32))  # :198
    w = (a * trans).astype(F32)
    cmap = np.stack([seq_sum(w * rgb[..., c]) for c in range(3)], -1)          # :201
    depth = seq_sum(w * (z_vals[..., :-1] if edges else z_vals))               # :204
    amap = seq_sum(w)                                                          # :207
    if composite_bkgd:                                                         # :210-211
        cmap = (cmap + ((F32(1) - amap)[..., None] * _f32(bkgd_color)).astype(F32)).astype(F32)
    return cmap, amap, w, depth


def composite_instanced(color, color_last, alpha, alpha_last, dists, patch_scale, composite_bkgd=False,
                        bkgd_color=(1., 1., 1.), map_exr=False, false_color=False, noise=None):
    """network/renderer.py:318-354 (InstanceRenderer.map_model_output): the opaque-mesh
    hit (color_last, alpha_last) is appended as sample S+1."""
    color, alpha, dists = _f32(color), _f32(alpha), _f32(dists)
    rgb = color if false_color else color_activation(color, map_exr)
    rgb = np.concatenate([rgb, _f32(color_last)], axis=1)                      # :331
    sig = alpha if noise is None else (alpha + _f32(noise)).astype(F32)
    sig = np.maximum(sig, F32(0))
    x = ((sig * dists).astype(F32) / F32(patch_scale)).astype(F32)             # :339
    a = (F32(1) - np.exp(-x.astype(np.float64)).astype(F32)).astype(F32)
    a = np.concatenate([a, _f32(alpha_last)], axis=1)
    trans = seq_cumprod_exclusive(((F32(1) - a).astype(F32) + F32(1e-10)).astype(F32))
    w = (a * trans).astype(F32)
    cmap = np.stack([seq_sum(w * rgb[..., c]) for c in range(3)], -1)
    amap = seq_sum(w)
    if composite_bkgd:
        cmap = (cmap + ((F32(1) - amap)[..., None] * _f32(bkgd_color)).astype(F32)).astype(F32)
    return cmap, amap


def sample_pdf(bins, weights, n_samples: int, det: bool = False, u: Optional[np.ndarray] = None,
               return_inds: bool = False):
    """network/renderer.py:589-618.  ``u`` (random draws) is an explicit input when not det."""
    bins, weights = _f32(bins), _f32(weights)
    w = (weights + F32(1e-5)).astype(F32)                                      # :593
    total = seq_sum(w)
    pdf = (w / total[..., None]).astype(F32)                                   # :594
    cdf = seq_cumsum(pdf)                                                      # :595
    cdf = np.concatenate([np.zeros_like(cdf[..., :1]), cdf], -1)               # :596
    R, C = cdf.shape
    if det:                                                                    # :599-601
        u = np.broadcast_to(tf_linspace(0.0, 1.0, n_samples), (R, n_samples))
    u = _f32(u)
    # searchsorted(side='right') == upper_bound: number of cdf entries <= u
    inds = (cdf[:, None, :] <= u[:, :, None]).sum(-1).astype(np.int32)         # :606
    below = np.maximum(0, inds - 1)                                            # :607
    above = np.minimum(C - 1, inds)                                            # :608
    rows = np.arange(R)[:, None]
    cdf_b, cdf_a = cdf[rows, below], cdf[rows, above]
    bins_b, bins_a = bins[rows, below], bins[rows, above]
    denom = (cdf_a - cdf_b).astype(F32)                                        # :613
    denom = np.where(denom < F32(1e-5), F32(1), denom).astype(F32)             # :614
    tt = ((u - cdf_b).astype(F32) / denom).astype(F32)                         # :615
    samples = (bins_b + (tt * (bins_a - bins_b).astype(F32)).astype(F32)).astype(F32)   # :616
    if return_inds:
        return samples, inds
    return samples


def hierarchical_z(z_vals, weights, n_importance: int, perturb: bool, u: Optional[np.ndarray] = None,
                   return_inds: bool = False):
    """network/renderer.py:127-130.  NOTE det=self.perturb (flag inverted vs bmild/nerf)."""
    z_vals, weights = _f32(z_vals), _f32(weights)
    z_mid = (F32(0.5) * (z_vals[..., 1:] + z_vals[..., :-1]).astype(F32)).astype(F32)
    res = sample_pdf(z_mid, weights[..., 1:-1], n_importance, det=perturb, u=u, return_inds=return_inds)
    z_samples, inds = res if return_inds else (res, None)
    z_all = np.sort(np.concatenate([z_vals, z_samples], -1), -1)
    return (z_all, z_samples, inds) if return_inds else (z_all, z_samples)


def evaluate_model(model: ParamNerfOracle, pts, dirs_n, parameters, cone_scale, z_vals, blur_idx=None,
                   net_chunk: int = 65536):
    """network/renderer.py:145-168."""
    R, S = pts.shape[:2]
    pos_flat = pts.reshape(-1, pts.shape[-1])
    dirs_flat = np.repeat(_f32(dirs_n), S, axis=0)
    params_flat = np.repeat(_f32(parameters), S, axis=0)
    if blur_idx is not None:                                                   # :155-158
        blur = (_f32(cone_scale)[..., None, :] * _f32(z_vals)[..., :, None]).astype(F32).reshape(-1)
        params_flat = params_flat.copy()
        params_flat[:, blur_idx] = (params_flat[:, blur_idx] * blur).astype(F32)
    cs, als = [], []
    for i in range(0, pos_flat.shape[0], net_chunk):                           # :160-163
        c, a = model.forward(pos_flat[i:i + net_chunk], dirs_flat[i:i + net_chunk], params_flat[i:i + net_chunk])
        cs.append(c)
        als.append(a)
    color = np.concatenate(cs, 0).reshape(R, S, 3)
    alpha = np.concatenate(als, 0).reshape(R, S)
    return color, alpha


def render_rays(model, model_fine, rays_o, rays_d, t, parameters, cone_scale, n_samples, n_importance=0,
                perturb=True, z_rand=None, u=None, noise=None, noise_fine=None, blur_idx=None, map_exr=False,
                composite_bkgd=False, bkgd_color=(1., 1., 1.), net_chunk=65536, return_aux=False):
    """network/renderer.py:92-143."""
    rays_o, rays_d = _f32(rays_o), _f32(rays_d)
    norm = np.sqrt((rays_d * rays_d).sum(-1, keepdims=True)).astype(F32)
    rays_d_n = (rays_d / norm).astype(F32)                                     # :98
    z = stratified_z(t, n_samples, perturb, z_rand)
    pts = ray_points(rays_o, rays_d, z)
    color, alpha = evaluate_model(model, pts, rays_d_n, parameters, cone_scale, z, blur_idx, net_chunk)
    cmap, amap, w, _ = composite(color, alpha, z, rays_d, composite_bkgd, bkgd_color, map_exr, noise)
    out = {"color_pred": cmap, "alpha_pred": amap}
    aux = {"z_coarse": z, "weights_coarse": w, "raw_color_coarse": color, "raw_alpha_coarse": alpha}
    if n_importance > 0:                                                       # :125-138
        z_all, z_samples = hierarchical_z(z, w, n_importance, perturb, u)
        pts = ray_points(rays_o, rays_d, z_all)
        m = model if model_fine is None else model_fine
        color_f, alpha_f = evaluate_model(m, pts, rays_d_n, parameters, cone_scale, z_all, blur_idx, net_chunk)
        cmap_f, amap_f, w_f, _ = composite(color_f, alpha_f, z_all, rays_d, composite_bkgd, bkgd_color, map_exr,
                                           noise_fine)
        out = {"color_pred": cmap_f, "alpha_pred": amap_f, "color_pred_coarse": cmap, "alpha_pred_coarse": amap}
        aux.update({"z_fine": z_all, "z_samples": z_samples, "weights_fine": w_f})
    return (out, aux) if return_aux else out


def renderer_call(model, model_fine, rays_o, rays_d, t, parameters, cone_scale, composite_bkgd=False,
                  bkgd_color=(1., 1., 1.), render_chunk=32768, z_rand=None, u=None, **kw):
    """network/renderer.py:47-90 (Renderer.__call__): flatten [B,Nr,.], repeat parameters per
    ray, drop rays with t[:,0]==inf, chunk, scatter back, background on missed rays."""
    rays_o, rays_d, t = _f32(rays_o), _f32(rays_d), _f32(t)
    B, Nr = rays_o.shape[:2]
    ro, rd, tf_ = rays_o.reshape(-1, 3), rays_d.reshape(-1, 3), t.reshape(-1, 2)
    pf = np.repeat(_f32(parameters), Nr, axis=0)                               # :54
    cf = _f32(cone_scale).reshape(-1, 1)
    hit = tf_[:, 0] != np.inf                                                  # :58
    idx = np.nonzero(hit)[0]
    outs: Dict[str, list] = {}
    zr = None if z_rand is None else _f32(z_rand).reshape(B * Nr, -1)[idx]
    uu = None if u is None else _f32(u).reshape(B * Nr, -1)[idx]
    for i in range(0, idx.shape[0], render_chunk):                             # :72
        sl = idx[i:i + render_chunk]
        o = render_rays(model, model_fine, ro[sl], rd[sl], tf_[sl], pf[sl], cf[sl], composite_bkgd=composite_bkgd,
                        bkgd_color=bkgd_color, z_rand=None if zr is None else zr[i:i + render_chunk],
                        u=None if uu is None else uu[i:i + render_chunk], **kw)
        for k, v in o.items():
            outs.setdefault(k, []).append(v)
    res = {}
    for k, lst in outs.items():                                                # :80-88
        v = np.concatenate(lst, 0)
        full = np.zeros((B * Nr,) + v.shape[1:], dtype=F32)
        full[idx] = v
        if composite_bkgd and "color" in k:
            full[~hit] = _f32(bkgd_color)
        res[k] = full.reshape((B, -1) + v.shape[1:])
    return res


# ----------------------------------------------------------------------------------------
# network/renderer.py — instanced renderer
# ----------------------------------------------------------------------------------------

def instance_evaluate_model(model, instancer_out, n_rays, cone_scale=None, blur_idx=None, patch_scale=1.0,
                            density_scale=1.0, density_reweighting=True, composite_bkgd=False,
                            bkgd_color=(1., 1., 1.), map_exr=False, net_chunk=32768, instance_color=None):
    """network/renderer.py:247-316.  ``instancer_out`` is the 10-tuple of
    instancer.pyx:54 as dense NumPy arrays; ``idxs`` is an int array of hit rays."""
    rays_d_map, pts, t, dists, color_last, alpha_last, alpha_weight, instance_id, idxs, params_map = instancer_out
    idxs = np.asarray(idxs).reshape(-1)
    if idxs.shape[0] == 0:                                                     # :256-257
        return np.zeros((n_rays, 3), F32), np.zeros((n_rays,), F32)
    params_map = _f32(params_map)
    if blur_idx is not None:                                                   # :260-263
        blur = ((_f32(cone_scale)[..., None, :] * _f32(t)[..., :, None]).astype(F32) / F32(patch_scale)).astype(F32)
        params_map = params_map.copy()
        params_map[..., blur_idx] = (params_map[..., blur_idx] * blur[..., 0]).astype(F32)
    g = lambda a: np.asarray(a)[idxs]                                          # :266-273
    rays_d_map, pts, dists = g(rays_d_map), g(pts), g(dists)
    color_last, alpha_last, alpha_weight = g(color_last), g(alpha_last), g(alpha_weight)
    instance_id, params_map = g(instance_id), g(params_map)
    H, S = dists.shape
    pos_flat, dirs_flat = pts.reshape(-1, 3), rays_d_map.reshape(-1, 3)
    params_flat = params_map.reshape(-1, params_map.shape[-1])
    ip = np.nonzero(dists.reshape(-1) > 0)[0]                                  # :284
    color = np.zeros((H * S, 3), F32)
    alpha = np.zeros((H * S, 1), F32)
    if ip.shape[0] > 0:
        for i in range(0, ip.shape[0], net_chunk):                             # :290-293
            sl = ip[i:i + net_chunk]
            c, a = model.forward(pos_flat[sl], dirs_flat[sl], params_flat[sl])
            color[sl], alpha[sl] = c, a
        alpha = alpha.reshape(H, S)
        mult = (_f32(alpha_weight) if density_reweighting else F32(1)) * F32(density_scale)
        alpha = (alpha * mult).astype(F32)                                     # :300
    else:
        alpha = alpha.reshape(H, S)
    color = color.reshape(H, S, 3)
    false_color = instance_color is not None
    if false_color:                                                            # :306-307
        color = _f32(instance_color)[instance_id.reshape(-1)].reshape(H, S, 3)
    cmap, amap = composite_instanced(color, color_last, alpha, alpha_last, dists, patch_scale, composite_bkgd,
                                     bkgd_color, map_exr, false_color)
    cm = np.zeros((n_rays, 3), F32)
    am = np.zeros((n_rays,), F32)
    cm[idxs], am[idxs] = cmap, amap                                            # :313-314
    return cm, am


# ----------------------------------------------------------------------------------------
# network/renderer.py — mip variants (:356-587)
# ----------------------------------------------------------------------------------------

def mip_cone_gaussians(rays_o, rays_d, t_vals, radii):
    """MipRenderer.get_cone_segment_gaussians (renderer.py:411-434): t_vals[R,S+1] edges, radii[R,1]."""
    rays_o, rays_d, t_vals, radii = _f32(rays_o), _f32(rays_d), _f32(t_vals), _f32(radii)
    t0, t1 = t_vals[..., :-1], t_vals[..., 1:]
    mu = ((t0 + t1) / F32(2)).astype(F32)
    hw = ((t1 - t0) / F32(2)).astype(F32)
    den = (F32(3) * mu ** 2 + hw ** 2).astype(F32)
    t_mean = (mu + (F32(2) * mu * hw ** 2) / den).astype(F32)
    t_var = ((hw ** 2) / F32(3) - F32(4 / 15) * ((hw ** 4 * (F32(12) * mu ** 2 - hw ** 2)) / den ** 2)).astype(F32)
    r_var = (radii ** 2 * ((mu ** 2) / F32(4) + F32(5 / 12) * hw ** 2 - F32(4 / 15) * (hw ** 4) / den)).astype(F32)
    mean = (rays_o[..., None, :] + rays_d[..., None, :] * t_mean[..., None]).astype(F32)
    d_mag_sq = np.maximum(F32(1e-10), (rays_d ** 2).sum(-1, keepdims=True)).astype(F32)
    d_outer = (rays_d ** 2).astype(F32)
    null_outer = (F32(1) - d_outer / d_mag_sq).astype(F32)
    cov = (t_var[..., None] * d_outer[..., None, :] + r_var[..., None] * null_outer[..., None, :]).astype(F32)
    return mean, cov


def mip_cone_gaussians_instanced(rays_d, t_vals, radii, dists):
    """MipInstanceRenderer.get_cone_segment_gaussians (renderer.py:570-587): flat per-sample arrays; mu = t, hw = dists."""
    rays_d, mu, radii, hw = _f32(rays_d), _f32(t_vals), _f32(radii), _f32(dists)
    den = (F32(3) * mu ** 2 + hw ** 2).astype(F32)
    t_var = ((hw ** 2) / F32(3) - F32(4 / 15) * ((hw ** 4 * (F32(12) * mu ** 2 - hw ** 2)) / den ** 2)).astype(F32)
    r_var = (radii ** 2 * ((mu ** 2) / F32(4) + F32(5 / 12) * hw ** 2 - F32(4 / 15) * (hw ** 4) / den)).astype(F32)
    d_mag_sq = np.maximum(F32(1e-10), (rays_d ** 2).sum(-1, keepdims=True)).astype(F32)
    d_outer = (rays_d ** 2).astype(F32)
    null_outer = (F32(1) - d_outer / d_mag_sq).astype(F32)
    return (t_var[:, None] * d_outer + r_var[:, None] * null_outer).astype(F32)


def composite_edges(color, alpha, z_edges, rays_d, composite_bkgd=False, bkgd_color=(1., 1., 1.), map_exr=False, noise=None):
    """MipRenderer.map_model_output (renderer.py:436-473): S + 1 depth edges, no duplicated last distance."""
    cmap, amap, w, _ = composite(color, alpha, z_edges, rays_d, composite_bkgd, bkgd_color, map_exr, noise, edges=True)
    return cmap, amap, w


def mip_render_rays(model, rays_o, rays_d, t, parameters, cone_scale, n_samples, blur_idx, perturb=True, z_rand=None,
                    noise=None, map_exr=False, composite_bkgd=False, bkgd_color=(1., 1., 1.), net_chunk=65536):
    """MipRenderer.render_rays (renderer.py:365-409); `model` has n_pos = 6, and one parameter fewer than
    `parameters` has columns (the blur parameter is spliced out, :385-386)."""
    rays_o, rays_d, parameters = _f32(rays_o), _f32(rays_d), _f32(parameters)
    norm = np.sqrt((rays_d * rays_d).sum(-1, keepdims=True)).astype(F32)
    rays_d_n = (rays_d / norm).astype(F32)
    z = stratified_z(t, n_samples + 1, perturb, z_rand)                        # :371-381 (S + 1 edges)
    blur = (parameters[..., blur_idx, None] * _f32(cone_scale)).astype(F32)    # :384
    params = np.concatenate([parameters[..., :blur_idx], parameters[..., blur_idx + 1:]], -1)
    mean, cov = mip_cone_gaussians(rays_o, rays_d, z, blur)
    pts = np.concatenate([mean, cov], -1)
    color, alpha = evaluate_model(model, pts, rays_d_n, params, None, None, None, net_chunk)
    cmap, amap, w = composite_edges(color, alpha, z, rays_d, composite_bkgd, bkgd_color, map_exr, noise)
    return {"color_pred": cmap, "alpha_pred": amap}, {"z": z, "mean": mean, "cov": cov, "weights": w}


def mip_instance_evaluate_model(model, instancer_out, n_rays, cone_scale, blur_idx, patch_scale=1.0, density_scale=1.0,
                                density_reweighting=True, composite_bkgd=False, bkgd_color=(1., 1., 1.), map_exr=False,
                                net_chunk=32768, instance_color=None):
    """MipInstanceRenderer.evaluate_model (renderer.py:484-568)."""
    rays_d_map, pts, t, dists, color_last, alpha_last, alpha_weight, instance_id, idxs, params_map = instancer_out
    idxs = np.asarray(idxs).reshape(-1)
    if idxs.shape[0] == 0:
        return np.zeros((n_rays, 3), F32), np.zeros((n_rays,), F32)
    g = lambda a: np.asarray(a)[idxs]
    rays_d_map, pts, t, dists = g(rays_d_map), g(pts), g(t), g(dists)
    color_last, alpha_last, alpha_weight = g(color_last), g(alpha_last), g(alpha_weight)
    instance_id, params_map, cone = g(instance_id), _f32(g(params_map)), _f32(g(cone_scale))
    blur = ((params_map[..., blur_idx] * cone[..., None, 0]).astype(F32) / F32(patch_scale)).astype(F32)   # :511
    params_map = np.concatenate([params_map[..., :blur_idx], params_map[..., blur_idx + 1:]], -1)
    H, S = dists.shape
    pos_flat, dirs_flat = _f32(pts).reshape(-1, 3), _f32(rays_d_map).reshape(-1, 3)
    params_flat = params_map.reshape(-1, params_map.shape[-1])
    ip = np.nonzero(dists.reshape(-1) > 0)[0]                                  # :526
    color = np.zeros((H * S, 3), F32)
    alpha = np.zeros((H * S, 1), F32)
    if ip.shape[0] > 0:
        cov = mip_cone_gaussians_instanced(dirs_flat[ip], _f32(t).reshape(-1)[ip], blur.reshape(-1)[ip],
                                           _f32(dists).reshape(-1)[ip])
        pos6 = np.concatenate([pos_flat[ip], cov], -1)
        for i in range(0, ip.shape[0], net_chunk):
            c, a = model.forward(pos6[i:i + net_chunk], dirs_flat[ip[i:i + net_chunk]], params_flat[ip[i:i + net_chunk]])
            color[ip[i:i + net_chunk]], alpha[ip[i:i + net_chunk]] = c, a
        alpha = alpha.reshape(H, S)
        mult = (_f32(alpha_weight) if density_reweighting else F32(1)) * F32(density_scale)
        alpha = (alpha * mult).astype(F32)
    else:
        alpha = alpha.reshape(H, S)
    color = color.reshape(H, S, 3)
    false_color = instance_color is not None
    if false_color:
        color = _f32(instance_color)[instance_id.reshape(-1)].reshape(H, S, 3)
    cmap, amap = composite_instanced(color, color_last, alpha, alpha_last, dists, patch_scale, composite_bkgd,
                                     bkgd_color, map_exr, false_color)
    cm = np.zeros((n_rays, 3), F32)
    am = np.zeros((n_rays,), F32)
    cm[idxs], am[idxs] = cmap, amap
    return cm, am


# ----------------------------------------------------------------------------------------
# network/ray_sampler.py, network/proxy.py
# ----------------------------------------------------------------------------------------

def rays_from_camera(image_plane_loc, height, width, focal, c2w):
    """network/ray_sampler.py:39-48."""
    loc, c2w = _f32(image_plane_loc), _f32(c2w)
    focal = F32(focal)
    dx = ((loc[:, 1] + F32(.5) - F32(.5 * width)) / focal).astype(F32)
    dy = (-(loc[:, 0] + F32(.5) - F32(.5 * height)) / focal).astype(F32)
    dirs = np.stack([dx, dy, -np.ones_like(dx)], -1)
    rays_d = seq_sum(dirs[:, None, :] * c2w[:3, :3])                            # :42
    rays_o = np.broadcast_to(c2w[:3, -1], rays_d.shape).copy()
    nxy = np.sqrt(seq_sum(dirs[:, :2] ** 2)).astype(F32)
    nall = np.sqrt(seq_sum(dirs ** 2)).astype(F32)
    cone = (np.cos(np.arctan(nxy.astype(np.float64))).astype(F32) / nall / focal).astype(F32)   # :46
    return rays_o, rays_d, cone[:, None]


def aabb_intersect(b_0, b_1, rays_o, rays_d):
    """network/proxy.py:13-35.  Misses give (inf, inf)."""
    b_0, b_1, rays_o, rays_d = _f32(b_0), _f32(b_1), _f32(rays_o), _f32(rays_d)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = (F32(1) / rays_d).astype(F32)
        t0 = ((b_0 - rays_o) * inv).astype(F32)
        t1 = ((b_1 - rays_o) * inv).astype(F32)
    t0n = np.where(t0 < t1, t0, t1)
    t1n = np.where(t0 > t1, t0, t1)
    tn = t0n.max(axis=1)
    tf_ = t1n.min(axis=1)
    hit = tn < tf_
    return np.stack([np.where(hit, tn, np.inf), np.where(hit, tf_, np.inf)], -1).astype(F32)


def proxy_rays(image_plane_loc, height, width, focal, c2w, b_0, b_1):
    """network/ray_sampler.py:23-37 (Proxy): normalised directions, t from the AABB."""
    o, d, cone = rays_from_camera(image_plane_loc, height, width, focal, c2w)
    d = (d / np.sqrt(seq_sum(d * d))[:, None]).astype(F32)
    return o, d, aabb_intersect(b_0, b_1, o, d), cone


# ----------------------------------------------------------------------------------------
# network/loss.py, optimiser of network/train.py:49-52
# ----------------------------------------------------------------------------------------

def mse(y_true, y_pred):
    """network/loss.py:51-54."""
    d = _f32(y_true).astype(np.float64) - _f32(y_pred).astype(np.float64)
    return F32(np.mean(d * d))


def smape(y_true, y_pred, eps=1e-2):
    """network/loss.py:56-58."""
    a, b = _f32(y_true).astype(np.float64), _f32(y_pred).astype(np.float64)
    return F32(np.mean(np.abs(a - b) / (a + b + eps)))


def alpha_loss(color_true, alpha_true, color_pred, alpha_pred, color_pred_coarse=None, alpha_pred_coarse=None,
               loss_fn=smape, alpha_loss_fn=mse, gamma=1.0, filter_color_loss=True, use_hard_mask=True):
    """network/loss.py:20-49 (AlphaLoss.__call__)."""
    color_true, alpha_true = _f32(color_true), _f32(alpha_true)
    mask = None
    if filter_color_loss:
        mask = (alpha_true[..., None] > 0).astype(F32) if use_hard_mask else alpha_true[..., None]
        color_true = color_true * mask
        color_pred = _f32(color_pred) * mask
    loss = loss_fn(color_true, color_pred) + F32(gamma) * alpha_loss_fn(alpha_true, alpha_pred)
    if color_pred_coarse is not None:
        cpc = _f32(color_pred_coarse) * mask if filter_color_loss else _f32(color_pred_coarse)
        loss = loss + loss_fn(color_true, cpc) + F32(gamma) * alpha_loss_fn(alpha_true, alpha_pred_coarse)
    return F32(loss)


def exponential_decay_lr(lrate: float, lrate_decay: float, step: int) -> float:
    """network/train.py:49-50: ExponentialDecay(lrate, decay_steps=lrate_decay*1e3, 0.1), non-staircase."""
    if lrate_decay <= 0:
        return lrate
    return lrate * 0.1 ** (step / (lrate_decay * 1e3))


def adam_step(p, g, m, v, step, lr, b1=0.9, b2=0.999, eps=1e-7):
    """tf.keras.optimizers.Adam defaults (epsilon=1e-7), non-amsgrad; ``step`` counts from 1."""
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    lr_t = lr * math.sqrt(1 - b2 ** step) / (1 - b1 ** step)
    p = p - lr_t * m / (np.sqrt(v) + eps)
    return p.astype(F32), m.astype(F32), v.astype(F32)


# ----------------------------------------------------------------------------------------
# network/logger.py:121-144 + util/interpolate.py:67-82 — image post-processing
# ----------------------------------------------------------------------------------------

def gaussian_kernel_2d(size: int, std: float) -> np.ndarray:
    """util/interpolate.py:67-75."""
    x = np.linspace(-(size - 1) / 2, (size - 1) / 2, size).astype(F32) + F32(.5 if size % 2 == 0 else 0)
    k1 = np.exp(-.5 * (x / F32(std)) ** 2).astype(F32)
    k2 = np.tensordot(k1, k1, axes=0).astype(F32)
    return (k2 / k2.sum()).astype(F32)


def filtered_downsample(img: np.ndarray, factor: int, std: float = .5) -> np.ndarray:
    """util/interpolate.py:78-82: tf.nn.depthwise_conv2d(img, gaussian, strides=factor, padding='SAME')."""
    img = _f32(img)
    H, W, Cn = img.shape
    k = int(factor * std * 6)
    K = gaussian_kernel_2d(k, factor * std).astype(np.float64)
    h, w = -(-H // factor), -(-W // factor)
    pt = max((h - 1) * factor + k - H, 0) // 2
    pl = max((w - 1) * factor + k - W, 0) // 2
    pad = np.zeros((H + 2 * k, W + 2 * k, Cn), np.float64)
    pad[k:k + H, k:k + W] = img
    out = np.zeros((h, w, Cn), np.float64)
    for i in range(k):
        for j in range(k):
            ys = k - pt + i + factor * np.arange(h)
            xs = k - pl + j + factor * np.arange(w)
            out += K[i, j] * pad[np.ix_(ys, xs)]
    return out.astype(F32)


def render_image_post(color_pred, alpha_pred, height, width, downsampling_factor=1, write_exr=False):
    """Logger.render_image after the renderer call (logger.py:126-137)."""
    img = np.concatenate([_f32(color_pred).reshape(-1, 3), _f32(alpha_pred).reshape(-1, 1)], -1).reshape(height, width, 4)
    if downsampling_factor > 1:
        img = filtered_downsample(img, downsampling_factor)
    if not write_exr:
        img = np.concatenate([img[..., :3] / (img[..., 3:] + F32(1e-5)), img[..., 3:]], -1).astype(F32)
    return img


def to_uint8(img):
    """tf.image.convert_image_dtype(img, tf.uint8) for in-range values: floor(x * 255.5) (logger.py:144)."""
    return (np.clip(_f32(img), 0, 1) * F32(255.5)).astype(np.uint8)

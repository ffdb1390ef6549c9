"""Known-answer tests that pin the oracle to closed forms (the reference has
no tests or golden vectors of its own, so these are authored from the formulas) and check that
the two independent restatements (NumPy and C) agree bit for bit."""

import math

import numpy as np
import pytest

from oracle import nerftex_oracle as O
from oracle import cbind as OC


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_linspace_endpoints_and_interior():
    v = O.tf_linspace(0., 1., 64)
    assert v[0] == 0 and v[-1] == 1
    d = np.float32(1) / np.float32(63)
    assert v[7] == np.float32(d * np.float32(7))
    assert O.tf_linspace(0., 1., 128)[-1] == 1.0


def test_fourier_features_layout_and_zero():
    # layer.py:11-23: identity block, then per frequency (sin block, cos block)
    x = np.zeros((2, 3), np.float32)
    f = O.fourier_features(x, 4)
    assert f.shape == (2, 27)
    expect = np.concatenate([np.zeros(3)] + [np.r_[np.zeros(3), np.ones(3)] for _ in range(4)])
    assert np.array_equal(f[0], expect.astype(np.float32))
    x = np.array([[0.5, -1.0, 2.0]], np.float32)
    f = O.fourier_features(x, 10)
    assert f.shape == (1, 63)
    assert np.allclose(f[0, 3:6], np.sin(x[0]), atol=1e-7)
    assert np.allclose(f[0, 6:9], np.cos(x[0]), atol=1e-7)
    assert np.allclose(f[0, 3 + 6 * 9:6 + 6 * 9], np.sin(512 * x[0].astype(np.float64)), atol=1e-6)


def test_model_shapes_and_param_counts():
    # params/model and MACs/sample for the three shipped variants
    for npar, P, D, macs, params in [((1, 4), 72, 63, 676224, 678916), ((2, 3), 81, 54, 678528, 681220),
                                     ((1, 6), 72, 81, 680832, 683524)]:
        m = O.ParamNerfOracle(n_parameters=npar)
        assert (m.P, m.D) == (P, D)
        assert m.macs_per_sample() == macs
        assert m.n_params() == params
    m = O.ParamNerfOracle(n_parameters=(1, 4), seed=3)
    c, a = m.forward(np.zeros((5, 3), np.float32), np.ones((5, 3), np.float32), np.zeros((5, 5), np.float32))
    assert c.shape == (5, 3) and a.shape == (5, 1)
    nerf = O.ParamNerfOracle(n_parameters=(0, 0), color_depth=0)
    assert (nerf.P, nerf.D) == (63, 27)


def test_composite_constant_sigma_closed_form():
    # constant sigma, unit direction, equal spacing: w_i = (1 - e^{-s d}) e^{-s d i}
    R, S, s, d = 3, 64, 2.5, 0.05
    z = np.tile(np.arange(S, dtype=np.float32) * np.float32(d), (R, 1))
    color = np.zeros((R, S, 3), np.float32)
    sigma = np.full((R, S), s, np.float32)
    rays_d = np.tile(np.array([[0, 0, 1.]], np.float32), (R, 1))
    cmap, amap, w, depth = O.composite(color, sigma, z, rays_d)
    i = np.arange(S)
    expect = (1 - math.exp(-s * d)) * np.exp(-s * d * i)
    assert np.allclose(w[0], expect, rtol=2e-5, atol=1e-7)
    assert np.allclose(amap, 1 - math.exp(-s * d * S), rtol=1e-5)
    assert np.allclose(cmap, 0.5 * amap[:, None], rtol=1e-5)        # sigmoid(0) = .5
    # |rays_d| scales the distances (renderer.py:180)
    _, amap2, _, _ = O.composite(color, sigma, z, 2 * rays_d)
    assert np.allclose(amap2, 1 - math.exp(-2 * s * d * S), rtol=1e-5)
    # background: C += (1 - A) * bkgd
    cm3, am3, _, _ = O.composite(color, sigma, z, rays_d, composite_bkgd=True, bkgd_color=(1, 0, .25))
    assert np.allclose(cm3, cmap + (1 - am3)[:, None] * np.array([1, 0, .25]), rtol=1e-6)


def test_composite_last_dist_duplicated_and_exr():
    z = np.array([[0., 1., 3.]], np.float32)
    sigma = np.array([[0., 0., 1.]], np.float32)
    color = np.array([[[0, 0, 0], [0, 0, 0], [-1., 0., 2.]]], np.float32)
    rays_d = np.array([[1., 0, 0]], np.float32)
    cmap, amap, w, _ = O.composite(color, sigma, z, rays_d, map_exr=True)
    assert np.allclose(w[0], [0, 0, 1 - math.exp(-2.0)], rtol=1e-6)  # last dist = 3-1 = 2 (renderer.py:177)
    assert np.allclose(cmap[0], w[0, 2] * np.array([math.exp(-1.), 1., 3.]), rtol=1e-6)   # elu(x)+1


def test_instanced_alpha_last_absorbs_remaining_transmittance():
    H, S = 2, 5
    rng = np.random.default_rng(0)
    color = rng.normal(size=(H, S, 3)).astype(np.float32)
    sigma = np.abs(rng.normal(size=(H, S))).astype(np.float32) * 30
    dists = np.full((H, S), 0.002, np.float32)
    color_last = np.tile(np.array([[[0.2, 0.4, 0.6]]], np.float32), (H, 1, 1))
    alpha_last = np.ones((H, 1), np.float32)
    cmap, amap = O.composite_instanced(color, color_last, sigma, alpha_last, dists, patch_scale=0.09)
    assert np.allclose(amap, 1.0, atol=1e-6)
    # zero-length padding samples are exact no-ops (1 - 0 + 1e-10 == 1 in fp32)
    color_p = np.concatenate([color, np.zeros((H, 3, 3), np.float32)], 1)
    sigma_p = np.concatenate([sigma, np.zeros((H, 3), np.float32)], 1)
    dists_p = np.concatenate([dists, np.zeros((H, 3), np.float32)], 1)
    cmap2, amap2 = O.composite_instanced(color_p, color_last, sigma_p, alpha_last, dists_p, patch_scale=0.09)
    assert np.array_equal(bits(cmap), bits(cmap2)) and np.array_equal(bits(amap), bits(amap2))


def test_sample_pdf_uniform_weights_is_identity_lerp():
    C = 63
    bins = np.tile(np.linspace(2, 6, C, dtype=np.float32), (2, 1))
    w = np.ones((2, C - 1), np.float32)
    s, inds = O.sample_pdf(bins, w, 128, det=True, return_inds=True)
    u = O.tf_linspace(0, 1, 128)
    assert np.allclose(s[0], 2 + 4 * u, atol=2e-5)
    assert inds.min() >= 1 and inds.max() <= C
    # u == 1.0: inds == len(cdf) iff cdf[-1] <= 1 -> below == above -> denom -> 1 -> bins[-1]
    assert inds[0, -1] in (C - 1, C)
    assert abs(s[0, -1] - 6.0) < 1e-5


def test_sample_pdf_peaky_and_bounds():
    rng = np.random.default_rng(3)
    R, C, N = 64, 63, 128
    bins = np.sort(rng.uniform(2, 6, size=(R, C)).astype(np.float32), -1)
    w = (rng.uniform(size=(R, C - 1)) ** 8).astype(np.float32)
    u = rng.uniform(size=(R, N)).astype(np.float32)
    s = O.sample_pdf(bins, w, N, det=False, u=u)
    assert np.all(s >= bins[:, :1] - 1e-6) and np.all(s <= bins[:, -1:] + 1e-6)
    z_all, zs = O.hierarchical_z(bins, np.pad(w, ((0, 0), (1, 0)))[:, :C], 16, perturb=True)
    assert np.all(np.diff(z_all, axis=-1) >= 0)


def test_numpy_and_c_restatements_agree_bitwise():
    rng = np.random.default_rng(7)
    R, S, N = 37, 64, 128
    t = np.stack([rng.uniform(2, 3, R), rng.uniform(4, 6, R)], -1).astype(np.float32)
    zr = rng.uniform(size=(R, S)).astype(np.float32)
    for perturb in (False, True):
        a = O.stratified_z(t, S, perturb, zr)
        b = OC.stratified_z(t, S, perturb, zr)
        assert np.array_equal(bits(a), bits(b))
    z = O.stratified_z(t, S, True, zr)
    color = rng.normal(size=(R, S, 3)).astype(np.float32)
    sigma = (rng.normal(size=(R, S)) * 20).astype(np.float32)
    rays_d = rng.normal(size=(R, 3)).astype(np.float32)
    noise = (rng.normal(size=(R, S)) * 0.1).astype(np.float32)
    for exr in (False, True):
        a = O.composite(color, sigma, z, rays_d, True, (1., .5, 0.), exr, noise)
        b = OC.composite(color, sigma, z, rays_d, noise, exr, True, (1., .5, 0.))
        for x, y in zip(a, b):
            assert np.array_equal(bits(x), bits(y))
    w = a[2]
    u = rng.uniform(size=(R, N)).astype(np.float32)
    for det in (True, False):
        za, sa, ia = O.hierarchical_z(z, w, N, det, u, return_inds=True)
        zb, sb, ib = OC.hierarchical_z(z, w, N, det, u)
        assert np.array_equal(ia, ib)
        assert np.array_equal(bits(sa), bits(sb))
        assert np.array_equal(bits(za), bits(zb))
    bins = np.sort(rng.uniform(0, 1, (R, 40)).astype(np.float32), -1)
    ww = rng.uniform(0, 1, (R, 39)).astype(np.float32) ** 8
    sa, ia = O.sample_pdf(bins, ww, 50, False, u[:, :50], return_inds=True)
    sb, ib = OC.sample_pdf(bins, ww, 50, False, u[:, :50])
    assert np.array_equal(ia, ib) and np.array_equal(bits(sa), bits(sb))


def test_rays_from_camera_and_aabb():
    H = W = 8
    focal = 0.5 * W / math.tan(0.5 * 0.63)
    ii, jj = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    loc = np.stack([ii.ravel(), jj.ravel()], -1).astype(np.float32)
    c2w = np.eye(4, dtype=np.float32)
    c2w[:3, 3] = [0, 0, 5]
    o, d, cone = O.rays_from_camera(loc, H, W, focal, c2w)
    assert np.allclose(o, [0, 0, 5]) and np.allclose(d[:, 2], -1)
    assert np.allclose(d[0, :2], [(0.5 - 4) / focal, -(0.5 - 4) / focal], rtol=1e-6)
    t = O.aabb_intersect([-1, -1, -1], [1, 1, 1], o, d / np.linalg.norm(d, axis=-1, keepdims=True))
    assert np.all(np.isfinite(t[:, 0]) == np.isfinite(t[:, 1]))
    ctr = (H // 2) * W + W // 2
    assert abs(t[ctr, 0] - 4.0) < 0.05 and abs(t[ctr, 1] - 6.0) < 0.1
    miss = O.aabb_intersect([-1, -1, -1], [1, 1, 1], np.array([[0, 0, 5.]], np.float32), np.array([[1., 0, 0]], np.float32))
    assert np.all(np.isinf(miss))


def test_losses_and_adam():
    y, p = np.array([0., 1., 2.], np.float32), np.array([1., 1., 0.], np.float32)
    assert abs(O.mse(y, p) - 5 / 3) < 1e-6
    assert abs(O.smape(y, p) - np.mean([1 / 1.01, 0, 2 / 2.01])) < 1e-6
    assert abs(O.exponential_decay_lr(5e-4, 500, 500000) - 5e-5) < 1e-12
    pw, m, v = O.adam_step(np.ones(3, np.float32), np.full(3, 0.5, np.float32), np.zeros(3), np.zeros(3), 1, 1e-3)
    assert np.allclose(pw, 1 - 1e-3, atol=1e-6)     # first Adam step moves by ~lr

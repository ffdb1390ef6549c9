"""K2 parity: compositing within 1e-5 relative (fp32) of the oracle; tolerance per north_star."""

import numpy as np
import pytest
import torch

from oracle import nerftex_oracle as O

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 1e-6     # "within 1e-5 relative in fp32"; atol covers |x| << 1 outputs in [0,1]


def dev(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a)).to("cuda", dtype)


def make(R, S, seed, sigma_scale=20.0):
    rng = np.random.default_rng(seed)
    t = np.stack([rng.uniform(2, 3, R), rng.uniform(4, 6, R)], -1).astype(np.float32)
    z = O.stratified_z(t, S, True, rng.uniform(size=(R, S)).astype(np.float32)) if S > 1 else t[:, :1]
    color = rng.normal(size=(R, S, 3)).astype(np.float32) * 2
    sigma = (rng.normal(size=(R, S)) * sigma_scale).astype(np.float32)
    rays_d = rng.normal(size=(R, 3)).astype(np.float32)
    noise = (rng.normal(size=(R, S)) * 0.1).astype(np.float32)
    return z, color, sigma, rays_d, noise


@pytest.mark.parametrize("R,S", [(4096, 64), (4096, 192), (300, 256), (3, 2), (7, 33), (2, 1024)])
@pytest.mark.parametrize("packed", [True, False])
@pytest.mark.parametrize("variant", ["plain", "noise_bkgd", "exr"])
def test_composite_forward(R, S, packed, variant):
    from nerf_tex_b200 import ops
    z, color, sigma, rays_d, noise = make(R, S, R + S)
    kw = dict(map_exr=variant == "exr", composite_bkgd=variant == "noise_bkgd", bkgd_color=(1., .5, .25))
    nz = noise if variant == "noise_bkgd" else None
    wc, wa, ww, wd = O.composite(color, sigma, z, rays_d, kw["composite_bkgd"], kw["bkgd_color"], kw["map_exr"], nz)
    if packed:
        rgbs = dev(np.concatenate([color, sigma[..., None]], -1))
        gc, ga, gw, gd = ops.composite_forward(rgbs, dev(z), dev(rays_d), noise=None if nz is None else dev(nz),
                                               want_depth=True, **kw)
    else:
        gc, ga, gw, gd = ops.composite_forward(None, dev(z), dev(rays_d), color=dev(color), sigma=dev(sigma),
                                               noise=None if nz is None else dev(nz), want_depth=True, **kw)
    np.testing.assert_allclose(gw.cpu().numpy(), ww, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(ga.cpu().numpy(), wa, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(gc.cpu().numpy(), wc, rtol=RTOL, atol=ATOL * 4)
    np.testing.assert_allclose(gd.cpu().numpy(), wd, rtol=RTOL, atol=ATOL * 8)


def test_composite_properties_full_size():
    """Size-independent properties at the full C1 size: weights >= 0, sum <= 1, linear in colour."""
    from nerf_tex_b200 import ops
    R, S = 65536, 192
    g = torch.Generator(device="cuda").manual_seed(0)
    rgbs = torch.randn((R, S, 4), device="cuda", generator=g) * 5
    z = torch.sort(torch.rand((R, S), device="cuda", generator=g) * 4 + 2, -1).values
    rays_d = torch.randn((R, 3), device="cuda", generator=g)
    c, a, w, _ = ops.composite_forward(rgbs, z, rays_d)
    assert (w >= 0).all() and (a <= 1 + 1e-5).all() and (c >= -1e-6).all() and (c <= 1 + 1e-5).all()
    torch.testing.assert_close(w.sum(-1), a, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("exr,bk", [(False, False), (True, True)])
def test_composite_backward_matches_autograd_of_oracle_formula(exr, bk):
    from nerf_tex_b200 import ops
    R, S = 257, 96
    z, color, sigma, rays_d, noise = make(R, S, 11, sigma_scale=8.0)
    rgbs = torch.tensor(np.concatenate([color, sigma[..., None]], -1), dtype=torch.float64, requires_grad=True)
    zt, dt, nt = torch.tensor(z, dtype=torch.float64), torch.tensor(rays_d, dtype=torch.float64), torch.tensor(noise, dtype=torch.float64)
    bkc = torch.tensor([1., .5, .25], dtype=torch.float64)
    # float64 torch restatement of renderer.py:170-213 (differentiable twin of oracle.composite)
    dists = zt[:, 1:] - zt[:, :-1]
    dists = torch.cat([dists, dists[:, -1:]], -1) * dt.norm(dim=-1, keepdim=True)
    col = torch.nn.functional.elu(rgbs[..., :3]) + 1 if exr else torch.sigmoid(rgbs[..., :3])
    a = 1 - torch.exp(-torch.relu(rgbs[..., 3] + nt) * dists)
    T = torch.cumprod(torch.cat([torch.ones_like(a[:, :1]), 1 - a + 1e-10], -1), -1)[:, :-1]
    w = a * T
    C, A = (w[..., None] * col).sum(-2), w.sum(-1)
    if bk:
        C = C + (1 - A)[:, None] * bkc
    gC = torch.randn(R, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(1))
    gA = torch.randn(R, dtype=torch.float64, generator=torch.Generator().manual_seed(2))
    ((C * gC).sum() + (A * gA).sum()).backward()
    got = ops.composite_backward(dev(rgbs.detach().numpy()), dev(z), dev(rays_d), dev(gC.numpy()), dev(gA.numpy()),
                                 noise=dev(noise), map_exr=exr, composite_bkgd=bk, bkgd_color=(1., .5, .25))
    want = rgbs.grad.numpy()
    scale = np.abs(want).max()
    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=2e-4, atol=2e-5 * scale)


@pytest.mark.parametrize("false_color", [False, True])
def test_composite_instanced_matches_dense_oracle(false_color):
    from nerf_tex_b200 import ops
    rng = np.random.default_rng(5)
    n_rays, S, ps = 300, 40, 0.09
    hit = rng.uniform(size=n_rays) < 0.7
    idxs = np.nonzero(hit)[0]
    H = idxs.shape[0]
    n_valid = rng.integers(0, S + 1, size=H)
    n_valid[0] = 0                                            # a hit ray with no in-shell samples
    dists = np.zeros((H, S), np.float32)
    for h in range(H):
        dists[h, :n_valid[h]] = 0.002
    color = rng.normal(size=(H, S, 3)).astype(np.float32)
    sigma = (rng.normal(size=(H, S)) * 40).astype(np.float32)
    aw = rng.uniform(0.3, 1, size=(H, S)).astype(np.float32)
    color_last = rng.uniform(size=(H, 1, 3)).astype(np.float32)
    alpha_last = (rng.uniform(size=(H, 1)) < 0.5).astype(np.float32)
    inst_col = rng.uniform(size=(H, S, 3)).astype(np.float32)
    mask = dists > 0
    # dense oracle path: renderer.py:300 then :318-354
    sig_dense = np.where(mask, sigma, 0).astype(np.float32) * (aw * np.float32(1.5))
    col_dense = inst_col if false_color else np.where(mask[..., None], color, 0).astype(np.float32)
    wc, wa = O.composite_instanced(col_dense, color_last, sig_dense, alpha_last, dists, ps, True, (1., 1., 1.),
                                   false_color=false_color)
    want_c = np.zeros((n_rays, 3), np.float32); want_a = np.zeros(n_rays, np.float32)
    want_c[idxs], want_a[idxs] = wc, wa
    # CSR-packed GPU path
    ray_ptr = np.concatenate([[0], np.cumsum(mask.sum(1))]).astype(np.int64)
    rgbs = np.concatenate([color, sigma[..., None]], -1)[mask]
    gc, ga = ops.composite_instanced_forward(dev(rgbs), dev(dists[mask]), dev(aw[mask]), dev(ray_ptr, torch.int64),
                                             dev(color_last.reshape(H, 3)), dev(alpha_last.reshape(H)),
                                             dev(idxs, torch.int32), n_rays, patch_scale=ps, density_scale=1.5,
                                             composite_bkgd=True, bkgd_color=(1., 1., 1.),
                                             instance_color=dev(inst_col[mask]) if false_color else None)
    np.testing.assert_allclose(ga.cpu().numpy(), want_a, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(gc.cpu().numpy(), want_c, rtol=RTOL, atol=ATOL * 4)

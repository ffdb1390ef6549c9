"""Device ray generation (ntx_rays_from_camera) against the oracle's rays_from_camera / AABB / Proxy."""

import math

import numpy as np
import pytest
import torch

from oracle import nerftex_oracle as O

pytestmark = pytest.mark.gpu


def pose(seed):
    rng = np.random.default_rng(seed)
    a, b = rng.uniform(0, 2 * math.pi), rng.uniform(-0.4, 0.4)
    rz = np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]])
    rx = np.array([[1, 0, 0], [0, math.cos(b), -math.sin(b)], [0, math.sin(b), math.cos(b)]])
    c2w = np.eye(4, dtype=np.float32)
    c2w[:3, :3] = (rz @ rx).astype(np.float32)
    c2w[:3, 3] = (c2w[:3, :3] @ np.array([0, 0, 4.0], np.float32))
    return c2w


@pytest.mark.parametrize("H,W", [(8, 8), (64, 48), (800, 800)])
def test_full_frame_proxy_rays(H, W):
    from nerf_tex_b200.proxy import AABB
    from nerf_tex_b200.ray_sampler import Proxy
    focal = 0.5 * W / math.tan(0.5 * 0.63)
    c2w = pose(H)
    b0, b1 = [-0.6, -0.5, -0.7], [0.5, 0.6, 0.4]
    ii, jj = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    loc = np.stack([ii.ravel(), jj.ravel()], -1).astype(np.float32)
    wo, wd, wt, wc = O.proxy_rays(loc, H, W, focal, c2w, b0, b1)
    for given_loc in (torch.as_tensor(loc).cuda(), None):
        go, gd, gt, gc = Proxy(H, W, focal, AABB(b0, b1))(given_loc, torch.as_tensor(c2w))
        torch.cuda.synchronize()
        assert np.array_equal(go.cpu().numpy(), wo)
        assert np.array_equal(gd.cpu().numpy().view(np.uint32), wd.view(np.uint32))        # same op order: bit-exact
        gtn = gt.cpu().numpy()
        assert np.array_equal(np.isinf(gtn), np.isinf(wt))
        fin = np.isfinite(wt)
        assert np.array_equal(gtn[fin].view(np.uint32), wt[fin].view(np.uint32))
        np.testing.assert_allclose(gc.cpu().numpy(), wc, rtol=1e-6, atol=0)
    assert np.isinf(wt).any() and np.isfinite(wt).any()


def test_frustum_and_pixel_list():
    from nerf_tex_b200.ray_sampler import Frustum, rays_from_camera
    H, W = 32, 40
    focal = 55.0
    c2w = pose(3)
    loc = np.random.default_rng(0).uniform(0, 31, (500, 2)).astype(np.float32)
    wo, wd, wc = O.rays_from_camera(loc, H, W, focal, c2w)
    go, gd, gt, gc = Frustum(H, W, focal, 2.0, 6.0)(torch.as_tensor(loc).cuda(), torch.as_tensor(c2w))
    assert np.array_equal(gd.cpu().numpy().view(np.uint32), wd.view(np.uint32))
    assert torch.equal(gt, torch.tensor([[2.0, 6.0]], device="cuda").repeat(500, 1))
    np.testing.assert_allclose(gc.cpu().numpy(), wc, rtol=1e-6)
    o2, d2, c2 = rays_from_camera(torch.as_tensor(loc).cuda(), H, W, focal, c2w.tolist())
    assert torch.equal(d2, gd) and torch.equal(c2, gc)


def test_generic_proxy_object_still_works():
    from nerf_tex_b200.ray_sampler import Proxy

    class Sphere:
        def __call__(self, o, d):
            return torch.stack([torch.full_like(o[:, 0], 1.0), torch.full_like(o[:, 0], 2.0)], -1)

    o, d, t, c = Proxy(8, 8, 10.0, Sphere())(None, torch.eye(4))
    assert t.shape == (64, 2) and torch.allclose(d.norm(dim=-1), torch.ones(64, device="cuda"), atol=1e-6)

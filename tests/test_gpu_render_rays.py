"""ntx_render_rays (one C call per ray chunk, scratch supplied by the caller) against the staged
K3 / K1 / K2 calls: the same kernels run in the same order, so the outputs must be identical."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def make(npar=(1, 4), fine=True, seed=0):
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import Renderer
    m = FusedNerfTexMLP(n_geo=npar[0], n_app=npar[1], seed=seed)
    mf = FusedNerfTexMLP(n_geo=npar[0], n_app=npar[1], seed=seed + 1) if fine else None
    return m, mf, Renderer


def rays(R, p, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    ro = torch.tensor([[0., 0., 5.]], device="cuda").repeat(R, 1)
    rd = torch.rand((R, 3), device="cuda", generator=g) * 2 - 1 - ro
    rd = rd / rd.norm(dim=-1, keepdim=True) * (0.8 + 0.4 * torch.rand((R, 1), device="cuda", generator=g))
    t = torch.stack([3 + torch.rand(R, device="cuda", generator=g), 6 + torch.rand(R, device="cuda", generator=g)], -1)
    return ro, rd, t, torch.rand((R, p), device="cuda", generator=g), torch.rand((R, 1), device="cuda", generator=g) * .01, g


@pytest.mark.parametrize("S,N,perturb,blur,noise,R", [(64, 128, True, None, 0.0, 1000), (64, 128, False, None, 0.0, 333),
                                                      (96, 0, True, 0, 0.1, 257), (32, 16, True, 0, 0.1, 5)])
def test_fused_entry_equals_staged_calls(S, N, perturb, blur, noise, R):
    m, mf, Renderer = make((2, 3), fine=N > 0)
    r = Renderer(model=m, model_fine=mf, n_samples=S, n_importance=N, perturb=perturb, blur_idx=blur, raw_noise_std=noise)
    ro, rd, t, p, cs, g = rays(R, 5, R)
    z_rand = torch.rand((R, S), device="cuda", generator=g)
    u = torch.rand((R, max(N, 1)), device="cuda", generator=g)
    nz = torch.randn((R, S), device="cuda", generator=g) * noise if noise > 0 else None
    nzf = torch.randn((R, S + N), device="cuda", generator=g) * noise if noise > 0 and N > 0 else None
    kw = dict(z_rand=z_rand, u=u if N > 0 else None, noise=nz, noise_fine=nzf)
    staged, _ = r.render_rays(ro, rd, t, p, cs, True, (1., .5, .25), False, return_aux=True, **kw)
    fused = r.render_rays(ro, rd, t, p, cs, True, (1., .5, .25), False, **kw)
    assert set(staged) == set(fused)
    for k in staged:
        assert torch.equal(staged[k], fused[k]), k


def test_workspace_query_and_errors():
    from nerf_tex_b200 import _lib, ops
    lib = _lib.load()
    assert lib.ntx_render_rays_workspace_bytes(0, 64, 128) == 0
    b1 = lib.ntx_render_rays_workspace_bytes(4096, 64, 128)
    assert b1 >= 4096 * (12 + 64 * 24 + 192 * 20) and b1 % 256 == 0
    m, mf, _ = make()
    ro, rd, t, p, cs, g = rays(64, 5, 1)
    cfg = ops.render_config(64, 128, True, False, False, None, (1., 1., 1.))
    with pytest.raises(_lib.NtxError, match="workspace"):
        ops.render_rays(m, mf, cfg, ro, rd, t, p, cs, z_rand=torch.rand((64, 64), device="cuda"),
                        workspace=torch.empty(1024, dtype=torch.uint8, device="cuda"))

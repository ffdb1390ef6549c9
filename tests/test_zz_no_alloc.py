"""ntx_render_rays_static — the entry point for hosts that cannot allocate device memory (for
example this library loaded next to TensorFlow, whose allocator already owns the GPU's memory, or
a frame loop that cannot afford the stall of an allocation).  Its scratch is a fixed array inside
the library image instead of a caller workspace.

Checked here, through the C ABI: two calls in a row (the second with a different, smaller chunk, so
stale contents of the shared array would show) give the same colours / alphas as ntx_render_rays
with a caller workspace, within the compositing tolerance used across the GPU tests; a chunk beyond
the built-in capacity is refused with a clear message.  Collected last (file name) so that a failure
here cannot hide any other test when the run stops at the first failure."""

import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 1e-6


def _inputs(R, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    ro = torch.tensor([[0., 0.3, 5.]], device="cuda").repeat(R, 1)
    rd = torch.rand((R, 3), device="cuda", generator=g) * 2 - 1 - ro
    rd = rd / rd.norm(dim=-1, keepdim=True)
    t = torch.stack([3 + torch.rand(R, device="cuda", generator=g), 6 + torch.rand(R, device="cuda", generator=g)], -1)
    params = torch.rand((R, 5), device="cuda", generator=g)
    cone = torch.rand((R, 1), device="cuda", generator=g) * 0.01
    z_rand = torch.rand((R, 64), device="cuda", generator=g)
    return ro, rd, t, params, cone, z_rand


def test_static_scratch_entry_matches_workspace_entry_twice_in_a_row():
    from nerf_tex_b200 import ops
    from nerf_tex_b200.model import FusedNerfTexMLP
    coarse = FusedNerfTexMLP(n_geo=1, n_app=4, seed=11)
    fine = FusedNerfTexMLP(n_geo=1, n_app=4, seed=12)
    cfg = ops.render_config(64, 128, True, False, True, None, (1., 1., 1.))
    for R, seed in ((700, 1), (300, 2)):
        ro, rd, t, params, cone, z_rand = _inputs(R, seed)
        want, _ = ops.render_rays(coarse, fine, cfg, ro, rd, t, params, cone, z_rand=z_rand)
        got, _ = ops.render_rays(coarse, fine, cfg, ro, rd, t, params, cone, z_rand=z_rand,
                                 entry="ntx_render_rays_static")
        torch.cuda.synchronize()
        assert set(got) == set(want) == {"color_pred", "alpha_pred", "color_pred_coarse", "alpha_pred_coarse"}
        for k in want:
            torch.testing.assert_close(got[k], want[k], rtol=RTOL, atol=ATOL, msg=lambda m, k=k: f"{k}: {m}")
        assert float(want["alpha_pred"].max()) > 0.05          # the comparison is about non-trivial renders


def test_static_scratch_entry_refuses_chunks_beyond_its_capacity():
    from nerf_tex_b200 import _lib, ops
    from nerf_tex_b200.model import FusedNerfTexMLP
    m = FusedNerfTexMLP(n_geo=1, n_app=4, seed=3)
    cfg = ops.render_config(64, 128, True, False, False, None, (1., 1., 1.))
    R = 8192                                                   # 44 MB of scratch > the 24 MiB built in
    assert ops.render_rays_workspace_bytes(R, 64, 128) > (24 << 20)
    ro, rd, t, params, cone, _ = _inputs(R, 5)
    z_rand = torch.rand((R, 64), device="cuda")
    with pytest.raises(_lib.NtxError, match="built-in buffer"):
        ops.render_rays(m, None, cfg, ro, rd, t, params, cone, z_rand=z_rand, entry="ntx_render_rays_static")
    # the capacity covers the 4096-ray, 64 + 128 sample batch of BASELINE.json configs[0]
    assert ops.render_rays_workspace_bytes(4096, 64, 128) <= (24 << 20)

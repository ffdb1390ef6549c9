"""ntx_render_rays_static — the entry point for hosts that cannot allocate device memory (for
example this library loaded next to TensorFlow, whose allocator already owns the GPU's memory, or
a frame loop that cannot afford the stall of an allocation).  Its scratch is a fixed array inside
the library image instead of a caller workspace.

Checked here, through the C ABI: two calls in a row (the second with a different, smaller chunk, so
stale contents of the shared array would show) give bit-for-bit the colours / alphas of
ntx_render_rays with a caller workspace (both entries launch the same kernels in the same order,
only the scratch address differs); the free device memory is the same before and after a call made
with caller-owned outputs; a blur parameter without cone scales and a chunk beyond the built-in
capacity are refused with clear messages.  Collected last (file name) so that a failure
here cannot hide any other test when the run stops at the first failure."""

import pytest
import torch

pytestmark = pytest.mark.gpu


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
            assert torch.equal(got[k], want[k]), f"{k}: max abs diff {float((got[k] - want[k]).abs().max())}"
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


def test_static_scratch_entry_leaves_device_memory_as_it_found_it():
    """The raw C call with outputs the caller already owns: cudaMemGetInfo reports the same free bytes before and
    after (the first call on the device, which makes CUDA load the library's image with the array in it, is made
    beforehand — that is the set-up step the header asks for)."""
    import ctypes as C
    from nerf_tex_b200 import _lib, ops
    from nerf_tex_b200._lib import check, ptr, stream_ptr
    from nerf_tex_b200.model import FusedNerfTexMLP
    lib = _lib.load()
    m = FusedNerfTexMLP(n_geo=1, n_app=4, seed=21)
    cfg = ops.render_config(64, 0, True, False, False, 0, (1., 1., 1.))          # blur_idx 0: cone scales are used
    R = 512
    ro, rd, t, params, cone, z_rand = _inputs(R, 9)
    cone = cone.reshape(-1).contiguous()
    color = torch.empty((R, 3), device="cuda")
    alpha = torch.empty((R,), device="cuda")
    packed = m.ensure_packed()

    def call(cone_ptr):
        return lib.ntx_render_rays_static(m._h, ptr(packed), None, None, C.byref(cfg), ptr(ro), ptr(rd), ptr(t), ptr(params),
                                          cone_ptr, ptr(z_rand), None, None, None, R, ptr(color), ptr(alpha), None, None,
                                          stream_ptr())
    check(call(ptr(cone)))                                     # set-up call
    torch.cuda.synchronize()
    first = color.clone()
    free_before, _ = torch.cuda.mem_get_info()
    check(call(ptr(cone)))
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, (free_before, free_after)
    assert torch.equal(color, first)
    assert call(None) != 0                                     # blur_idx >= 0 without cone scales is refused, not guessed
    assert b"cone_scale" in lib.ntx_last_error()

"""Committed vectors of the widened rows (tests/golden/widened_small.npz, made by tests/golden/make_golden_wide.py from
the oracles): the oracles must keep reproducing them (CPU), the CUDA paths must match them (GPU)."""

from pathlib import Path

import numpy as np
import pytest

from oracle import instancer_oracle as IO
from oracle import nerftex_oracle as O

G = np.load(Path(__file__).parent / "golden" / "widened_small.npz")
B0, B1 = [-0.5, -0.5, -0.1], [0.5, 0.5, 0.1]
REC = ("n_samples", "hit", "pts", "dirs", "params", "dists", "t", "alpha_weight", "instance_id")


def same_bits(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype == np.float32:
        return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))
    return np.array_equal(a, b)


def test_oracles_reproduce_golden_vectors():
    for method in ("random", "nearest", "nearest_blend"):
        rec = IO.get_model_input(G["inst_o"], G["inst_d"], G["inst_prm"], 24, 0.02, G["inst_T"], B0, B1, seed=9, light_dir_idx=2,
                                 method=method, patch_scale=1.0)
        for k in REC:
            assert same_bits(rec[k], G[f"inst_{method}_{k}"]), (method, k)
    mean, cov = O.mip_cone_gaussians(G["mip_o"], G["mip_d"], G["mip_z"], G["mip_radii"])
    np.testing.assert_allclose(mean, G["mip_mean"], rtol=1e-6)
    np.testing.assert_allclose(cov, G["mip_cov"], rtol=1e-5, atol=1e-12)
    ec, ea, ew = O.composite_edges(G["mip_rgbs"][..., :3], G["mip_rgbs"][..., 3], G["mip_z"], G["mip_d"], True, (1., .5, .25))
    assert same_bits(ew, G["mip_w"]) and same_bits(ec, G["mip_cmap"]) and same_bits(ea, G["mip_amap"])
    ipe = O.integrated_positional_encoding(np.concatenate([G["mip_mean"], G["mip_cov"]], -1).reshape(-1, 6)[:64], 10)
    np.testing.assert_allclose(ipe, G["mip_ipe"], atol=1e-6)
    H, W = 12, 16
    ii, jj = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    loc = np.stack([ii.ravel(), jj.ravel()], -1).astype(np.float32)
    ro, rd, t, cone = O.proxy_rays(loc, H, W, 20.0, G["ray_c2w"], [-1, -1, -1], [1, 1, 1])
    assert same_bits(rd, G["ray_d"]) and np.array_equal(np.isinf(t), np.isinf(G["ray_t"]))
    np.testing.assert_allclose(O.render_image_post(G["img_color"], G["img_alpha"], 17, 22, 3), G["img_out"], rtol=1e-6, atol=1e-7)


@pytest.mark.gpu
def test_cuda_paths_match_golden_vectors():
    import torch
    from nerf_tex_b200 import ops
    from nerf_tex_b200.image import finish_image
    from nerf_tex_b200.instancer import Instancer
    from nerf_tex_b200.layer import IntegratedPositionalEncoding
    cu = lambda k: torch.as_tensor(G[k]).cuda()
    for method in ("random", "nearest", "nearest_blend"):
        inst = Instancer(B0, B1, textures=["", "", "light"], transformations=G["inst_T"].tolist(), instance_sampling_method=method,
                         patch_scale=1.2)
        got = inst.get_model_input_csr(cu("inst_o"), cu("inst_d"), cu("inst_prm"), 24, 0.02, seed=9)
        assert np.array_equal(got["n_samples_all"].cpu().numpy(), G[f"inst_{method}_n_samples"])
        for k in ("pts", "dirs", "params", "dists", "t", "alpha_weight", "instance_id"):
            assert same_bits(got[k].cpu().numpy(), G[f"inst_{method}_{k}"]), (method, k)
    mean, cov = ops.mip_cone_gaussians(cu("mip_o"), cu("mip_d"), cu("mip_z"), cu("mip_radii"))
    np.testing.assert_allclose(mean.cpu().numpy(), G["mip_mean"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(cov.cpu().numpy(), G["mip_cov"], rtol=2e-4, atol=1e-9)
    c, a, w = ops.composite_edges_forward(cu("mip_rgbs"), cu("mip_z"), cu("mip_d"), composite_bkgd=True, bkgd_color=(1., .5, .25))
    np.testing.assert_allclose(w.cpu().numpy(), G["mip_w"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(c.cpu().numpy(), G["mip_cmap"], rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(a.cpu().numpy(), G["mip_amap"], rtol=1e-5, atol=2e-6)
    x6 = torch.cat([cu("mip_mean"), cu("mip_cov")], -1).reshape(-1, 6)[:64]
    np.testing.assert_allclose(IntegratedPositionalEncoding(10)(x6).cpu().numpy(), G["mip_ipe"], atol=3e-4)   # fp32 sin(512 x)
    ro, rd, t, cone = ops.rays_from_camera(None, 12, 16, 20.0, G["ray_c2w"], normalize=True, aabb=([-1, -1, -1], [1, 1, 1]))
    assert same_bits(rd.cpu().numpy(), G["ray_d"])
    fin = np.isfinite(G["ray_t"])
    assert np.array_equal(np.isfinite(t.cpu().numpy()), fin) and same_bits(t.cpu().numpy()[fin], G["ray_t"][fin])
    np.testing.assert_allclose(cone.cpu().numpy(), G["ray_cone"], rtol=1e-6)
    img, _ = finish_image(cu("img_color"), cu("img_alpha"), 17, 22, 3, straight_alpha=True)
    np.testing.assert_allclose(img.cpu().numpy(), G["img_out"], rtol=2e-5, atol=2e-6)

"""Image finishing kernel, pixel samplers, and the Render entry end to end (network/render.py, network/logger.py)."""

import math
import os

import numpy as np
import pytest
import torch

from oracle import instancer_oracle as IO
from oracle import nerftex_oracle as O

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()


@pytest.mark.parametrize("H,W,f", [(8, 8, 1), (32, 48, 2), (17, 22, 3), (64, 64, 4), (50, 31, 8)])
def test_finish_image_matches_oracle(H, W, f):
    from nerf_tex_b200.image import finish_image
    rng = np.random.default_rng(H * W + f)
    alpha = rng.uniform(0, 1, (H * W,)).astype(np.float32)
    alpha[rng.uniform(size=H * W) < 0.2] = 0.0                                     # fully transparent pixels: 0 / 1e-5
    color = (rng.uniform(0, 1, (H * W, 3)).astype(np.float32) * alpha[:, None]).astype(np.float32)
    for straight in (True, False):
        want = O.render_image_post(color, alpha, H, W, f, write_exr=not straight)
        img, u8 = finish_image(dev(color), dev(alpha), H, W, f, straight_alpha=straight, want_u8=True)
        img, u8 = img.cpu().numpy(), u8.cpu().numpy()
        assert img.shape == want.shape == (-(-H // f), -(-W // f), 4)
        np.testing.assert_allclose(img, want, rtol=2e-5, atol=2e-6)
        assert np.abs(u8.astype(np.int32) - O.to_uint8(want).astype(np.int32)).max() <= 1     # truncation at a bin edge


def test_pixel_samplers():
    from nerf_tex_b200.pixel_sampler import Full, Independent, Proxy
    from nerf_tex_b200.proxy import AABB
    from nerf_tex_b200.dataset import look_at
    H, W = 64, 96
    full = Full(H, W)().cpu().numpy()
    n = np.arange(H * W)
    np.testing.assert_array_equal(full, np.stack([n // W, n % W], -1))
    ind = Independent(H, W, 5000, seed=1)().cpu().numpy()
    assert ind.shape == (5000, 2) and ind[:, 0].max() < H and ind[:, 1].max() < W and ind.min() >= 0
    assert abs(ind[:, 0].mean() - (H - 1) / 2) < 2 and abs(ind[:, 1].mean() - (W - 1) / 2) < 3
    # proxy sampler: every returned pixel's (coarse) ray hits the box, no duplicates, right count
    focal = W / math.tan(0.35) / 2
    box = AABB([-0.5, -0.5, -0.5], [0.5, 0.5, 0.5])
    c2w = look_at([4., 1., 2.])
    ps = Proxy(H, W, 300, box, focal, downsample_factor=8, seed=3)
    mask = ps.hit_mask(c2w).cpu().numpy()
    # oracle: coarse grid rays against the box, nearest-neighbour upsampling
    Hd, Wd = H // 8, W // 8
    ii, jj = np.meshgrid(np.arange(Hd), np.arange(Wd), indexing="ij")
    loc = np.stack([ii.ravel(), jj.ravel()], -1).astype(np.float32)
    ro, rd, _ = O.rays_from_camera(loc, Hd, Wd, focal // 8, c2w)
    t = O.aabb_intersect(box.b_0, box.b_1, ro, rd)
    want = np.isfinite(t[:, 0]).reshape(Hd, Wd).repeat(8, 0).repeat(8, 1)
    np.testing.assert_array_equal(mask, want)
    assert 0 < mask.sum() < H * W
    got = ps(c2w).cpu().numpy()
    assert got.shape == (300, 2) and mask[got[:, 0], got[:, 1]].all()
    assert len({(a, b) for a, b in got}) == 300
    with pytest.raises(ValueError):
        Proxy(H, W, H * W, box, focal, seed=0)(c2w)


def test_render_entry_end_to_end(tmp_path):
    """network.render.Render on a config dict: pose dataset -> Proxy rays -> device Instancer -> InstanceRenderer ->
    Logger writes PNGs; the same frames through the oracles (rays, instancer, network, compositor, post-process)."""
    from nerf_tex_b200 import util
    from nerf_tex_b200.dataset import look_at
    from nerf_tex_b200.image import decode_png
    from test_gpu_instancer import B0, B1, patch_grid
    Ts = patch_grid(3, seed=2, tilt=0.2)
    Ts[:, :3, 3] -= [0.9, 0.9, 0.0]
    H = W = 40
    poses = [look_at([2.5, 0.5, 3.0]), look_at([-1.0, 2.0, 3.5])]
    params = [[1, 1, 1, .1, .3, -.2, 1], [0.5, 0.2, 1, .4, -.3, .2, 1]]
    om = O.ParamNerfOracle(n_parameters=(1, 6), seed=5)
    om.randomize_biases(9, 0.05)
    cfg = {
        "module": "network.render.Render", "target_path": str(tmp_path / "logs"),
        "test_dataset_config": {"module": "network.dataset.PoseDataset", "poses": [p.tolist() for p in poses],
                                "parameters": params, "height": H, "width": W, "angle": 0.6, "composite_bkgd": True,
                                "pixel_sampler_config": {"module": "network.pixel_sampler.Full"},
                                "ray_sampler_config": {"module": "network.ray_sampler.Proxy"},
                                "proxy_config": {"module": "network.proxy.AABB", "b_0": [-1.6, -1.6, -0.4], "b_1": [1.6, 1.6, 0.4]}},
        "model_config": {"module": "network.model.ParamNerf",
                         "pos_embedding": {"module": "network.model.FourierFeatures", "n_freq_bands": 10},
                         "dir_embedding": {"module": "network.model.FourierFeatures", "n_freq_bands": 4},
                         "param_embedding": {"module": "network.model.FourierFeatures", "n_freq_bands": 4},
                         "n_parameters": [1, 6]},
        "renderer_config": {"module": "network.renderer.InstanceRenderer", "n_samples": 32, "step_size": 0.01, "density_scale": 100.0,
                            "instancer_config": {"module": "instancer.instancer.Instancer", "b_0": B0, "b_1": B1,
                                                 "textures": ["", "", "", "", "light"], "transformations": Ts.tolist(),
                                                 "patch_scale": 1.0, "seed": 4}},
        "logger_config": {"module": "network.logger.Logger", "downsampling_factor": 2},
    }
    # a checkpoint for the logger to restore: the oracle's weights in the flat layout
    from nerf_tex_b200.model import FusedNerfTexMLP
    tmp = FusedNerfTexMLP(1, 6)
    tmp.load_named_weights(om.named_weights())
    os.makedirs(tmp_path / "logs" / "checkpoints")
    np.savez(tmp_path / "logs" / "checkpoints" / "ckpt-7.npz", step=np.int64(7), model=tmp.master.cpu().numpy())
    logger = util.instantiate(cfg)
    assert logger.step == 7
    files = sorted(os.listdir(tmp_path / "logs" / "media" / "test"))
    assert files == ["0.png", "1.png"]
    focal = W / math.tan(0.3) / 2
    ii, jj = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    loc = np.stack([ii.ravel(), jj.ravel()], -1).astype(np.float32)
    for i, (pose, prm) in enumerate(zip(poses, params)):
        ro, rd, t, cone = O.proxy_rays(loc, H, W, focal, pose, [-1.6, -1.6, -0.4], [1.6, 1.6, 0.4])
        hitp = np.isfinite(t[:, 0])
        R = int(hitp.sum())
        pf = np.tile(np.asarray(prm, np.float32), (R, 1))
        # the renderer culls proxy misses first (renderer.py:58-66), so the instancer sees the surviving rays, in order;
        # it names every ray by its position in the frame (the draws do not depend on chunking) and seeds per frame
        rec = IO.get_model_input(ro[hitp], rd[hitp], pf, 32, 0.01, Ts, B0, B1, seed=4 + 7919 * i, light_dir_idx=4,
                                 ray_ids=np.nonzero(hitp)[0])
        csr = {k: torch.as_tensor(rec[k]) for k in ("pts", "dirs", "params", "dists", "t", "alpha_weight", "instance_id")}
        idx = np.nonzero(rec["hit"])[0]
        ptr_all = np.concatenate([[0], np.cumsum(rec["n_samples"])])
        csr["ray_ptr"] = torch.as_tensor(np.concatenate([ptr_all[:-1][idx], ptr_all[-1:]]))
        csr["ray_index"] = torch.as_tensor(idx.astype(np.int32))
        csr["color_last"], csr["alpha_last"] = torch.zeros(len(idx), 3), torch.zeros(len(idx))
        from nerf_tex_b200.instancer import csr_to_dense
        dense = [x.numpy() for x in csr_to_dense(csr, torch.as_tensor(rd[hitp]), torch.as_tensor(pf), 32)]
        wc, wa = O.instance_evaluate_model(om, dense, R, cone_scale=cone[hitp], blur_idx=None, patch_scale=1.0,
                                           density_scale=100.0, composite_bkgd=True)
        color = np.ones((H * W, 3), np.float32)                                     # missed rays: background (renderer.py:86-87)
        alpha = np.zeros((H * W,), np.float32)
        color[hitp], alpha[hitp] = wc, wa
        want = O.to_uint8(O.render_image_post(color, alpha, H, W, 2))
        got = decode_png(open(tmp_path / "logs" / "media" / "test" / files[i], "rb").read())
        assert got.shape == want.shape == (H // 2, W // 2, 4)
        assert (want[..., 3] > 12).mean() > 0.05                                    # the patches are visible
        # straight colour = premultiplied / alpha amplifies the bf16 error where alpha is small: compare alpha
        # everywhere, colour where the pixel is reasonably opaque
        assert np.abs(got[..., 3].astype(int) - want[..., 3].astype(int)).max() <= 3
        solid = want[..., 3] > 64
        assert np.abs(got[..., :3].astype(int) - want[..., :3].astype(int))[solid].max() <= 6

"""mip variants (network/renderer.py:356-587, network/layer.py:25-41): integrated positional encoding inside K1,
cone-segment gaussians, compositing over S + 1 edges, MipRenderer / MipInstanceRenderer against the oracle."""

import numpy as np
import pytest
import torch

from oracle import nerftex_oracle as O

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()


def psnr(a, b):
    return -10 * np.log10(np.mean((a.astype(np.float64) - b) ** 2) + 1e-20)


def make_model(npar, seed, via_config=False):
    om = O.ParamNerfOracle(n_parameters=npar, n_pos=6, seed=seed)
    om.randomize_biases(seed + 100, 0.1)
    if via_config:
        from nerf_tex_b200 import util
        gm = util.instantiate({"module": "network.model.ParamNerf", "n_pos": 6, "n_parameters": list(npar),
                               "pos_embedding": {"module": "network.layer.IntegratedPositionalEncoding", "n_freq_bands": 10},
                               "dir_embedding": {"module": "network.layer.FourierFeatures", "n_freq_bands": 4},
                               "param_embedding": {"module": "network.layer.FourierFeatures", "n_freq_bands": 4}})["model"]
    else:
        from nerf_tex_b200.model import FusedNerfTexMLP
        gm = FusedNerfTexMLP(n_geo=npar[0], n_app=npar[1], n_pos=6)
    gm.load_named_weights(om.named_weights())
    return om, gm


@pytest.mark.parametrize("npar", [(1, 4), (2, 3), (3, 7)])
@pytest.mark.parametrize("n", [1, 1000, 33000])
def test_ipe_mlp_matches_oracle(npar, n):
    om, gm = make_model(npar, seed=3, via_config=(n == 1000))
    assert om.P == 60 + 9 * npar[0] and gm.layers[0][2] == om.P
    rng = np.random.default_rng(5)
    mean = rng.uniform(-1.4, 1.5, (n, 3)).astype(np.float32)
    # variances over many decades: from "no attenuation at any frequency" to "only the lowest band survives"
    cov = (10.0 ** rng.uniform(-8, 0, (n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    params = rng.uniform(0, 1, (n, sum(npar))).astype(np.float32)
    pos6 = np.concatenate([mean, cov], -1)
    color, alpha = gm((dev(pos6), dev(d), dev(params)))
    torch.cuda.synchronize()
    color, alpha = color.cpu().numpy(), alpha.cpu().numpy()
    ec, ea = O.forward_bf16_emulated(om, pos6, d, params)
    assert np.abs(color - ec).max() < 5e-3 * max(1.0, np.abs(ec).max()), np.abs(color - ec).max()
    assert np.abs(alpha - ea).max() < 5e-3 * max(1.0, np.abs(ea).max()), np.abs(alpha - ea).max()
    fc, fa = om.forward(pos6, d, params)
    sig = lambda x: 1 / (1 + np.exp(-x))
    err = np.abs(sig(color) - sig(fc))
    assert err.max() < 4e-3 and psnr(sig(color), sig(fc)) > 50, (err.max(), psnr(sig(color), sig(fc)))


def test_ipe_layer_and_loud_errors():
    from nerf_tex_b200 import _lib
    from nerf_tex_b200.layer import IntegratedPositionalEncoding
    from nerf_tex_b200.model import FusedNerfTexMLP, ParamNerf
    x = np.random.default_rng(0).uniform(0.0, 1.0, (50, 6)).astype(np.float32)
    got = IntegratedPositionalEncoding(10)(dev(x)).cpu().numpy()
    np.testing.assert_allclose(got, O.integrated_positional_encoding(x, 10), atol=2e-5)
    gm = FusedNerfTexMLP(n_geo=1, n_app=4, n_pos=6)
    with pytest.raises(_lib.NtxError):                     # variances missing
        gm.forward_rows(n_rows=4, pos=torch.zeros(4, 3).cuda(), dir=torch.zeros(4, 3).cuda(),
                        params=torch.zeros(4, 5).cuda(), rows_per_group=1)
    with pytest.raises(NotImplementedError):               # IPE config with a 3-wide position
        ParamNerf({"module": "network.layer.IntegratedPositionalEncoding", "n_freq_bands": 10},
                  {"module": "network.layer.FourierFeatures", "n_freq_bands": 4},
                  {"module": "network.layer.FourierFeatures", "n_freq_bands": 4}, [1, 4])


@pytest.mark.parametrize("R,S", [(1, 1), (37, 64), (500, 100)])
def test_cone_gaussians(R, S):
    from nerf_tex_b200 import ops
    rng = np.random.default_rng(R)
    o = rng.normal(size=(R, 3)).astype(np.float32)
    d = rng.normal(size=(R, 3)).astype(np.float32)
    z = np.sort(rng.uniform(2.0, 6.0, (R, S + 1)).astype(np.float32), -1)
    radii = rng.uniform(1e-3, 5e-2, (R, 1)).astype(np.float32)
    wm, wc = O.mip_cone_gaussians(o, d, z, radii)
    gm, gc = ops.mip_cone_gaussians(dev(o), dev(d), dev(z), dev(radii))
    np.testing.assert_allclose(gm.cpu().numpy(), wm, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gc.cpu().numpy(), wc, rtol=2e-4, atol=1e-9)
    T = R * S
    dirs = rng.normal(size=(T, 3)).astype(np.float32)
    t = rng.uniform(1.0, 4.0, T).astype(np.float32)
    rad = rng.uniform(1e-3, 5e-2, T).astype(np.float32)
    dists = rng.uniform(1e-3, 5e-3, T).astype(np.float32)
    want = O.mip_cone_gaussians_instanced(dirs, t, rad, dists)
    got = ops.mip_cone_gaussians_instanced(dev(dirs), dev(t), dev(rad), dev(dists)).cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=2e-4, atol=1e-12)


@pytest.mark.parametrize("S,exr,bk,with_noise", [(1, False, False, False), (64, False, True, True), (97, True, True, False)])
def test_composite_edges_forward_backward(S, exr, bk, with_noise):
    from nerf_tex_b200 import ops
    import torch_ref as T
    rng = np.random.default_rng(S)
    R = 300
    rgbs = rng.normal(size=(R, S, 4)).astype(np.float32)
    rgbs[..., 3] = rgbs[..., 3] * 3 + 1
    z = np.sort(rng.uniform(2, 6, (R, S + 1)).astype(np.float32), -1)
    d = rng.normal(size=(R, 3)).astype(np.float32)
    noise = rng.normal(size=(R, S)).astype(np.float32) * 0.3 if with_noise else None
    wc, wa, ww = O.composite_edges(rgbs[..., :3], rgbs[..., 3], z, d, bk, (1., .5, .25), exr, noise)
    gc, ga, gw = ops.composite_edges_forward(dev(rgbs), dev(z), dev(d), noise=None if noise is None else dev(noise),
                                             map_exr=exr, composite_bkgd=bk, bkgd_color=(1., .5, .25))
    np.testing.assert_allclose(gw.cpu().numpy(), ww, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(gc.cpu().numpy(), wc, rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(ga.cpu().numpy(), wa, rtol=1e-5, atol=2e-6)
    # backward against autograd of the same math in fp64
    dc = rng.normal(size=(R, 3)).astype(np.float32)
    da = rng.normal(size=(R,)).astype(np.float32)
    got = ops.composite_edges_backward(dev(rgbs), dev(z), dev(d), dev(dc), dev(da), noise=None if noise is None else dev(noise),
                                       map_exr=exr, composite_bkgd=bk, bkgd_color=(1., .5, .25)).cpu().numpy()
    x = torch.tensor(rgbs, dtype=torch.float64, requires_grad=True)
    zt, dt = torch.tensor(z, dtype=torch.float64), torch.tensor(d, dtype=torch.float64)
    dists = (zt[:, 1:] - zt[:, :-1]) * dt.norm(dim=-1, keepdim=True)
    col = torch.where(x[..., :3] > 0, x[..., :3] + 1, torch.exp(torch.clamp(x[..., :3], max=0))) if exr else torch.sigmoid(x[..., :3])
    sig = x[..., 3] + (torch.tensor(noise, dtype=torch.float64) if noise is not None else 0)
    a = 1 - torch.exp(-torch.relu(sig) * dists)
    om = 1 - a + 1e-10
    trans = torch.cumprod(torch.cat([torch.ones(R, 1, dtype=torch.float64), om[:, :-1]], 1), 1)
    w = a * trans
    C = (w[..., None] * col).sum(1)
    A = w.sum(1)
    if bk:
        C = C + (1 - A)[:, None] * torch.tensor([1., .5, .25], dtype=torch.float64)
    ((C * torch.tensor(dc, dtype=torch.float64)).sum() + (A * torch.tensor(da, dtype=torch.float64)).sum()).backward()
    want = x.grad.numpy()
    assert np.abs(got - want).max() < 1e-4 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("perturb,bk", [(True, True), (False, False)])
def test_mip_renderer_matches_oracle(perturb, bk):
    from nerf_tex_b200.renderer import MipRenderer
    npar_model = (1, 4)                                   # the renderer gets 6 parameters, index 1 is the blur radius
    om, gm = make_model(npar_model, seed=11)
    R, S = 600, 48
    rng = np.random.default_rng(2)
    o = np.tile(np.array([[0., 0.3, 4.]], np.float32), (R, 1))
    d = (rng.uniform(-1, 1, (R, 3)).astype(np.float32) - o)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    d *= rng.uniform(0.8, 1.2, (R, 1)).astype(np.float32)                 # not unit: exercises the ||d|| factors
    t = np.stack([rng.uniform(2.5, 3, R), rng.uniform(5, 6, R)], -1).astype(np.float32)
    params = np.array([[0.4, 0.7, 0.2, 0.6, -0.3, 0.74]], np.float32)
    cone = rng.uniform(1e-3, 2e-2, (1, R, 1)).astype(np.float32)
    z_rand = rng.uniform(size=(R, S + 1)).astype(np.float32)
    r = MipRenderer(model=gm, n_samples=S, perturb=perturb, blur_idx=1, seed=0)
    got = r(dev(o[None]), dev(d[None]), dev(t[None]), dev(params), dev(cone), composite_bkgd=bk, bkgd_color=(1., 1., 1.),
            training=False, z_rand=dev(z_rand))
    want, aux = O.mip_render_rays(om, o, d, t, np.repeat(params, R, 0), cone[0], S, blur_idx=1, perturb=perturb,
                                  z_rand=z_rand, composite_bkgd=bk)
    gc, ga = got["color_pred"].cpu().numpy()[0], got["alpha_pred"].cpu().numpy()[0]
    assert (want["alpha_pred"] > 0.05).mean() > 0.3
    assert np.abs(ga - want["alpha_pred"]).max() < 2e-3 and np.abs(gc - want["color_pred"]).max() < 2e-3, \
        (np.abs(ga - want["alpha_pred"]).max(), np.abs(gc - want["color_pred"]).max())
    assert psnr(gc, want["color_pred"]) > 50
    with pytest.raises(NotImplementedError):
        MipRenderer(model=gm, n_samples=S, n_importance=8, blur_idx=1)(dev(o[None]), dev(d[None]), dev(t[None]), dev(params),
                                                                        dev(cone), training=False)


def test_mip_instance_renderer_matches_oracle():
    from nerf_tex_b200.instancer import SyntheticInstancer
    from nerf_tex_b200.renderer import MipInstanceRenderer
    om, gm = make_model((1, 5), seed=7)                   # instancer parameters: 7 columns, column 3 is the blur radius
    kw = dict(b_0=[-1.4, -1.2, -.1], b_1=[1.2, 1.2, 1.8], patch_scale=0.09, seed=0, mean_valid=40)
    r = MipInstanceRenderer(instancer_config=SyntheticInstancer(**kw), model=gm, n_samples=64, step_size=0.002,
                            density_scale=10.0, blur_idx=3)
    R = 400
    rng = np.random.default_rng(0)
    o = rng.normal(size=(1, R, 3)).astype(np.float32)
    d = rng.normal(size=(1, R, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    t = np.tile(np.array([[[1., 3.]]], np.float32), (1, R, 1))
    params = np.array([[1, 1, 1, .1, 0, 0, 1]], np.float32)
    cone = rng.uniform(0.01, 0.05, (1, R, 1)).astype(np.float32)
    got = r(dev(o), dev(d), dev(t), dev(params), dev(cone), composite_bkgd=True, training=False)
    pf = np.repeat(params, R, 0)
    dense = [x.cpu().numpy() for x in SyntheticInstancer(**kw).get_model_input(dev(o[0]), dev(d[0]), dev(pf), 64, 0.002)]
    wc, wa = O.mip_instance_evaluate_model(om, dense, R, cone[0], blur_idx=3, patch_scale=0.09, density_scale=10.0,
                                           composite_bkgd=True)
    gc, ga = got["color_pred"].cpu().numpy()[0], got["alpha_pred"].cpu().numpy()[0]
    assert (wa > 0.05).mean() > 0.3
    assert np.abs(ga - wa).max() < 2e-3 and np.abs(gc - wc).max() < 2e-3, (np.abs(ga - wa).max(), np.abs(gc - wc).max())
    assert psnr(gc, wc) > 50


def test_mip_training_gradients():
    """One MipRenderer training pass: loss and weight gradients against fp64 autograd of the oracle's graph."""
    import torch_ref as T
    from nerf_tex_b200.renderer import MipRenderer
    om, gm = make_model((1, 4), seed=21)
    gm.requires_grad_(True)
    R, S = 256, 32
    rng = np.random.default_rng(4)
    o = np.tile(np.array([[0., 0.3, 4.]], np.float32), (R, 1))
    d = rng.uniform(-1, 1, (R, 3)).astype(np.float32) - o
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    t = np.stack([rng.uniform(2.5, 3, R), rng.uniform(5, 6, R)], -1).astype(np.float32)
    params = np.array([[0.4, 0.7, 0.2, 0.6, -0.3, 0.74]], np.float32)
    cone = rng.uniform(1e-3, 2e-2, (1, R, 1)).astype(np.float32)
    z_rand = rng.uniform(size=(R, S + 1)).astype(np.float32)
    tc = rng.uniform(size=(R, 3)).astype(np.float32)
    ta = rng.uniform(size=(R,)).astype(np.float32)
    r = MipRenderer(model=gm, n_samples=S, perturb=True, blur_idx=1)
    out = r(dev(o[None]), dev(d[None]), dev(t[None]), dev(params), dev(cone), composite_bkgd=False, training=True,
            z_rand=dev(z_rand))
    loss = ((out["color_pred"][0] - dev(tc)) ** 2).mean() + ((out["alpha_pred"][0] - dev(ta)) ** 2).mean()
    loss.backward()
    g = gm.master.grad.detach().cpu().numpy().astype(np.float64)
    # reference: the oracle's forward in torch fp64 on the oracle's own cone gaussians
    _, aux = O.mip_render_rays(om, o, d, t, np.repeat(params, R, 0), cone[0], S, blur_idx=1, perturb=True, z_rand=z_rand)
    pts6 = np.concatenate([aux["mean"], aux["cov"]], -1).reshape(-1, 6)
    dn = np.repeat(d / np.linalg.norm(d, axis=-1, keepdims=True), S, 0)
    pr = np.repeat(np.delete(np.repeat(params, R, 0), 1, axis=1), S, 0)
    pos_map, dir_map = om.encode(pts6, dn, pr)
    ws = T.make_params(om)
    color, alpha = T.mlp_from_maps(om, ws, torch.tensor(pos_map, dtype=torch.float64), torch.tensor(dir_map, dtype=torch.float64))
    zt, dt = torch.tensor(aux["z"], dtype=torch.float64), torch.tensor(d, dtype=torch.float64)
    dists = (zt[:, 1:] - zt[:, :-1]) * dt.norm(dim=-1, keepdim=True)
    a = 1 - torch.exp(-torch.relu(alpha.reshape(R, S)) * dists)
    trans = torch.cumprod(torch.cat([torch.ones(R, 1, dtype=torch.float64), (1 - a + 1e-10)[:, :-1]], 1), 1)
    w = a * trans
    C = (w[..., None] * torch.sigmoid(color.reshape(R, S, 3))).sum(1)
    A = w.sum(1)
    ref_loss = ((C - torch.tensor(tc, dtype=torch.float64)) ** 2).mean() + ((A - torch.tensor(ta, dtype=torch.float64)) ** 2).mean()
    ref_loss.backward()
    assert abs(float(loss.detach()) - float(ref_loss.detach())) < 1e-3
    ref = T.flat_grads(om, ws).numpy()
    cos = float(g @ ref / (np.linalg.norm(g) * np.linalg.norm(ref)))
    rel = float(np.linalg.norm(g - ref) / np.linalg.norm(ref))
    assert cos > 0.999 and rel < 3e-2, (cos, rel)

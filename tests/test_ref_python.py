"""The oracle of the TensorFlow path (oracle/nerftex_oracle.py) against the REFERENCE'S OWN Python source — network/
renderer.py, model.py, layer.py, loss.py, ray_sampler.py, proxy.py imported from /root/reference and executed over a
NumPy stand-in for the TensorFlow ops they call (oracle/tf_shim: our code; it restates the documented semantics of each
op in fp32).  What this pins is the reference's LOGIC as written — which tensors are sliced, concatenated, gathered,
scattered, in which order; chunking; culling of rays with t = inf; the `det=self.perturb` inversion; `weights[..., 1:-1]`;
the blur scaling; the instanced renderer's masking and last-sample handling; the mip cone gaussians; the training
loop of network/train.py (tape, AlphaLoss, Adam with its decayed rate); every shipped config through the reference's own
util.instantiate — not TensorFlow's kernels (TensorFlow 2.4 cannot run here).  Random draws are taken from the
stand-in's log and handed to the oracle.

CPU only; needs /root/reference (skipped elsewhere; tests/golden/ref_python_small.npz carries vectors made by the same
runs for machines without it, see tests/golden/make_ref_python.py)."""

import numpy as np
import pytest

import ref_python as RP
from oracle import nerftex_oracle as O

pytestmark = pytest.mark.skipif(not RP.available(), reason="the reference tree is not on this machine")
F = np.float32


@pytest.fixture(scope="module")
def ref():
    return RP.load()


def rays(R, seed, B=1):
    rng = np.random.default_rng(seed)
    o = np.tile(np.asarray([[0.0, 0.3, 5.0]], F), (B * R, 1))
    d = rng.uniform(-1, 1, (B * R, 3)).astype(F) - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True) * rng.uniform(0.8, 1.3, (B * R, 1))).astype(F)     # not unit length
    t = np.stack([rng.uniform(3.0, 4.0, B * R), rng.uniform(5.5, 7.0, B * R)], -1).astype(F)
    cone = rng.uniform(0.001, 0.01, (B * R, 1)).astype(F)
    return o.reshape(B, R, 3), d.reshape(B, R, 3), t.reshape(B, R, 2), cone.reshape(B, R, 1)


@pytest.mark.parametrize("npar,color_depth", [((1, 4), 1), ((2, 3), 1), ((1, 6), 2), ((0, 0), 1)])
def test_param_nerf_graph(ref, npar, color_depth):
    """network/model.py:58-125 + layer.py:8-23: the reference's functional Keras graph, weights copied in."""
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    om = O.ParamNerfOracle(n_parameters=npar, color_depth=color_depth, seed=3)
    om.randomize_biases(4, 0.1)
    km = RP.build_param_nerf(model, util, om)
    rng = np.random.default_rng(5)
    n = 300
    pos, dirs = rng.uniform(-1, 1, (n, 3)).astype(F), rng.normal(size=(n, 3)).astype(F)
    prm = rng.uniform(0, 1, (n, sum(npar))).astype(F)
    c_ref, a_ref = km((tf.constant(pos), tf.constant(dirs), tf.constant(prm)), training=False)
    c, a = om.forward(pos, dirs, prm, dtype=np.float64)
    scale = max(np.abs(c).max(), np.abs(a).max())
    assert np.abs(c_ref.numpy() - c).max() <= 2e-5 * scale and np.abs(a_ref.numpy() - a).max() <= 2e-5 * scale


def test_nerf_graph_is_the_parameterless_case(ref):
    """network/model.py:9-45 (Nerf): ParamNerf with n_parameters [0, 0] and color_depth 0."""
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    om = O.ParamNerfOracle(n_parameters=(0, 0), color_depth=0, seed=6)
    km = model.Nerf(RP.ff(util, 10), RP.ff(util, 4))["model"]
    RP.set_weights(km, om)
    rng = np.random.default_rng(7)
    pos, dirs = rng.uniform(-1, 1, (64, 3)).astype(F), rng.normal(size=(64, 3)).astype(F)
    c_ref, a_ref = km((tf.constant(pos), tf.constant(dirs), tf.constant(np.zeros((64, 0), F))), training=False)
    c, a = om.forward(pos, dirs, np.zeros((64, 0), F), dtype=np.float64)
    assert np.abs(c_ref.numpy() - c).max() <= 2e-5 * np.abs(c).max() and np.abs(a_ref.numpy() - a).max() <= 2e-5 * np.abs(a).max()


@pytest.mark.parametrize("det", [True, False])
def test_sample_pdf(ref, det):
    """network/renderer.py:589-618: the oracle's restatement is bit-equal to the reference's source over the stand-in."""
    tf, renderer, *_ = ref
    rng = np.random.default_rng(8)
    R, C, N = 200, 63, 128
    bins = np.sort(rng.uniform(2, 6, (R, C)), -1).astype(F)
    w = (rng.uniform(0, 1, (R, C - 1)) ** 6).astype(F)
    w[::7] = 0                                                       # all-zero rows: only the 1e-5 floor
    tf.random.seed(9)
    got = renderer.sample_pdf(tf.constant(bins), tf.constant(w.copy()), N, det=det).numpy()
    u = None if det else tf.random.log[-1][1]
    want = O.sample_pdf(bins, w, N, det=det, u=u)
    want = want[0] if isinstance(want, tuple) else want
    assert np.array_equal(got.view(np.uint32), np.asarray(want, F).view(np.uint32))


CASES = [
    dict(n_samples=64, n_importance=0, perturb=True),
    dict(n_samples=64, n_importance=0, perturb=False),
    dict(n_samples=48, n_importance=96, perturb=True),                  # det resampling (det=self.perturb)
    dict(n_samples=48, n_importance=96, perturb=False),                 # random resampling
    dict(n_samples=32, n_importance=0, perturb=True, blur_idx=0, raw_noise_std=0.5),
    dict(n_samples=32, n_importance=0, perturb=True, map_exr=True),
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("composite_bkgd", [False, True])
def test_renderer_call(ref, case, composite_bkgd):
    """Renderer.__call__ / render_rays / evaluate_model / map_model_output (renderer.py:34-213) on a [B, Nr] batch with
    missed rays, several render chunks and net chunks, coarse + fine models."""
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    npar = (2, 3)
    om, omf = O.ParamNerfOracle(n_parameters=npar, seed=11), O.ParamNerfOracle(n_parameters=npar, seed=12)
    for m_, s_ in ((om, 13), (omf, 14)):
        m_.randomize_biases(s_, 0.2)
    km, kmf = RP.build_param_nerf(model, util, om), RP.build_param_nerf(model, util, omf, name="model_fine")
    B, R = 2, 70
    o, d, t, cone = rays(R, 15, B)
    t[0, ::9, 0] = np.inf                                            # rays that miss the proxy
    t[1, 3::11] = np.inf
    prm = np.random.default_rng(16).uniform(0.2, 1.0, (B, sum(npar))).astype(F)
    fine = case["n_importance"] > 0
    r = renderer.Renderer(model=km, model_fine=kmf if fine else None, render_chunk=50, net_chunk=1000, **case)
    tf.random.seed(17)
    got = r(tf.constant(o), tf.constant(d), tf.constant(t), tf.constant(prm), tf.constant(cone), composite_bkgd=composite_bkgd,
            bkgd_color=[1.0, 0.5, 0.25], training=True)
    # the draws, chunk by chunk, in the order the reference made them -> per-ray arrays for the oracle
    hit = (t.reshape(-1, 2)[:, 0] != np.inf)
    n_hit, S, N = int(hit.sum()), case["n_samples"], case["n_importance"]
    z_rand = np.zeros((B * R, S), F)
    noise = np.zeros((B * R, S), F)
    noise_f = np.zeros((B * R, S + N), F)
    u = np.zeros((B * R, max(N, 1)), F)
    log = list(tf.random.log)
    idx = np.nonzero(hit)[0]
    for i in range(0, n_hit, 50):
        sl = idx[i:i + 50]
        if case["perturb"]:
            z_rand[sl] = log.pop(0)[1]
        if case.get("raw_noise_std", 0) > 0:
            noise[sl] = log.pop(0)[1]
        if fine and not case["perturb"]:
            u[sl] = log.pop(0)[1]
        if fine and case.get("raw_noise_std", 0) > 0:
            noise_f[sl] = log.pop(0)[1]
    assert not log
    std = F(case.get("raw_noise_std", 0))
    kw = dict(n_samples=S, n_importance=N, perturb=case["perturb"], blur_idx=case.get("blur_idx"), map_exr=case.get("map_exr", False),
              net_chunk=1000)
    if std > 0:
        kw.update(noise=None, noise_fine=None)
    # the oracle takes per-chunk draws through renderer_call's z_rand / u (noise per chunk needs its own loop)
    if std > 0:
        outs = {}
        for i in range(0, n_hit, 50):
            sl = idx[i:i + 50]
            o_ = O.render_rays(om, omf if fine else None, o.reshape(-1, 3)[sl], d.reshape(-1, 3)[sl], t.reshape(-1, 2)[sl],
                               np.repeat(prm, R, 0)[sl], cone.reshape(-1, 1)[sl], S, N, case["perturb"], z_rand=z_rand[sl],
                               u=u[sl] if fine and not case["perturb"] else None, noise=noise[sl] * std, noise_fine=noise_f[sl] * std,
                               blur_idx=case.get("blur_idx"), map_exr=case.get("map_exr", False), composite_bkgd=composite_bkgd,
                               bkgd_color=(1.0, 0.5, 0.25), net_chunk=1000)
            for k, v in o_.items():
                outs.setdefault(k, []).append(v)
        want = {}
        for k, lst in outs.items():
            v = np.concatenate(lst, 0)
            full = np.zeros((B * R,) + v.shape[1:], F)
            full[idx] = v
            if composite_bkgd and "color" in k:
                full[~hit] = np.asarray([1.0, 0.5, 0.25], F)
            want[k] = full.reshape((B, -1) + v.shape[1:])
    else:
        want = O.renderer_call(om, omf if fine else None, o, d, t, prm, cone, composite_bkgd=composite_bkgd, bkgd_color=(1.0, 0.5, 0.25),
                               render_chunk=50, z_rand=z_rand if case["perturb"] else None,
                               u=u if fine and not case["perturb"] else None, **kw)
    assert set(got) == set(want)
    for k in want:
        g = got[k].numpy()
        assert g.shape == want[k].shape, (k, g.shape, want[k].shape)
        # coarse outputs: fp32 round-off of two evaluations of the same expressions; fine outputs sit behind the
        # resampling, which turns a last-bit difference of a coarse weight into a moved sample
        tol = 2e-5 if ("coarse" in k or not fine) else 5e-3
        assert np.abs(g - want[k]).max() <= tol, (k, float(np.abs(g - want[k]).max()))
    miss = ~hit.reshape(B, R)
    assert (got["alpha_pred"].numpy()[miss] == 0).all()


# ------------------------------------------------------------------------------------------------ rays, proxy, losses
def test_rays_from_camera_frustum_and_proxy(ref):
    """network/ray_sampler.py:6-48, network/proxy.py:6-35: directions, origins, cone scale, Frustum t, AABB t (misses = inf)."""
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    H, W, focal = 37, 53, 60.0
    ii, jj = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    loc = np.stack([ii.reshape(-1), jj.reshape(-1)], -1).astype(F)
    th = 0.4
    c2w = np.asarray([[np.cos(th), 0, np.sin(th), 2.0], [0, 1, 0, 0.3], [-np.sin(th), 0, np.cos(th), 4.0], [0, 0, 0, 1]], F)
    o_r, d_r, c_r = ray_sampler.rays_from_camera(tf.constant(loc), H, W, focal, tf.constant(c2w))
    o, d, c = O.rays_from_camera(loc, H, W, focal, c2w)
    assert np.array_equal(o_r.numpy(), o) and np.array_equal(d_r.numpy(), d) and np.abs(c_r.numpy() - c).max() <= 4e-9   # (1 ulp: atan rounded before cos)
    fo, fd, ft, fc = ray_sampler.Frustum(H, W, focal, 2.0, 6.0)(tf.constant(loc), tf.constant(c2w))
    assert np.array_equal(ft.numpy(), np.tile(np.asarray([[2.0, 6.0]], F), (H * W, 1))) and np.array_equal(fd.numpy(), d)
    b0, b1 = [-1.0, -1.2, -0.5], [1.0, 0.8, 0.7]
    po, pd, pt, pc = ray_sampler.Proxy(H, W, focal, proxy.AABB(b0, b1))(tf.constant(loc), tf.constant(c2w))
    wo, wd, wt, wc = O.proxy_rays(loc, H, W, focal, c2w, b0, b1)
    assert np.array_equal(pd.numpy(), wd) and np.array_equal(pt.numpy(), wt)
    assert 0.05 < np.isinf(wt[:, 0]).mean() < 0.95


@pytest.mark.parametrize("coarse", [False, True])
@pytest.mark.parametrize("hard", [True, False])
def test_losses(ref, coarse, hard):
    """network/loss.py: mse, smape, NerfLoss, AlphaLoss (hard / soft mask, with and without the coarse pair)."""
    tf, renderer, model, loss, *_ = ref
    rng = np.random.default_rng(21)
    ct, cp, cc = (rng.uniform(0, 1, (2, 50, 3)).astype(F) for _ in range(3))
    at, ap, ac = (rng.uniform(0, 1, (2, 50)).astype(F) for _ in range(3))
    at[0, ::3] = 0
    c = tf.constant
    got = loss.AlphaLoss("network.loss.smape", "network.loss.mse", gamma=0.7, use_hard_mask=hard)(
        color_true=c(ct), alpha_true=c(at), color_pred=c(cp), alpha_pred=c(ap), color_pred_coarse=c(cc) if coarse else None,
        alpha_pred_coarse=c(ac) if coarse else None)
    want = O.alpha_loss(ct, at, cp, ap, cc if coarse else None, ac if coarse else None, gamma=0.7, use_hard_mask=hard)
    assert abs(float(got) - float(want)) <= 2e-6 * abs(float(want))
    assert abs(float(loss.mse(c(ct), c(cp))) - float(O.mse(ct, cp))) <= 1e-7
    got_n = loss.NerfLoss()(color_true=c(ct), color_pred=c(cp), color_pred_coarse=c(cc) if coarse else None)
    assert abs(float(got_n) - float(O.mse(ct, cp) + (O.mse(ct, cc) if coarse else 0))) <= 1e-6


# ------------------------------------------------------------------------------------------------ mip renderer
@pytest.mark.parametrize("perturb,noise_std", [(True, 0.0), (False, 0.3)])
def test_mip_renderer(ref, perturb, noise_std):
    """MipRenderer (renderer.py:356-473) with IntegratedPositionalEncoding (layer.py:25-41): S + 1 edges, the blur parameter
    spliced out, cone-segment gaussians, edge compositing."""
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    npar = (2, 2)                                                      # after the blur parameter is spliced out
    om = O.ParamNerfOracle(n_parameters=npar, n_pos=6, seed=31)
    om.randomize_biases(32, 0.2)
    ipe = util.EasyDict({"module": "network.layer.IntegratedPositionalEncoding", "n_freq_bands": 10})
    km = model.ParamNerf(ipe, RP.ff(util, 4), RP.ff(util, 4), n_parameters=list(npar), n_pos=6)["model"]
    RP.set_weights(km, om)
    R, S = 60, 24
    o, d, t, cone = rays(R, 33)
    prm = np.random.default_rng(34).uniform(0.2, 1.0, (1, 5)).astype(F)
    r = renderer.MipRenderer(model=km, n_samples=S, perturb=perturb, raw_noise_std=noise_std, blur_idx=1, render_chunk=25)
    tf.random.seed(35)
    got = r(tf.constant(o), tf.constant(d), tf.constant(t), tf.constant(prm), tf.constant(cone), composite_bkgd=True, training=True)
    log = list(tf.random.log)
    outs = {"color_pred": [], "alpha_pred": []}
    for i in range(0, R, 25):
        zr = log.pop(0)[1] if perturb else None
        nz = log.pop(0)[1] * F(noise_std) if noise_std > 0 else None
        w, _ = O.mip_render_rays(om, o[0, i:i + 25], d[0, i:i + 25], t[0, i:i + 25], np.repeat(prm, R, 0)[i:i + 25], cone[0, i:i + 25], S, 1,
                                 perturb=perturb, z_rand=zr, noise=nz, composite_bkgd=True)
        for k in outs:
            outs[k].append(w[k])
    assert not log
    for k in outs:
        assert np.abs(got[k].numpy()[0] - np.concatenate(outs[k], 0)).max() <= 2e-5, k


# ------------------------------------------------------------------------------------------------ instanced renderers
def instanced_setup(ref, tmp_path, mip):
    """The reference's InstanceRenderer over the reference's own C++ instancer (oracle/tf_shim/instancer wraps
    oracle/ref_instancer like instancer.pyx wraps the class) on a mesh scene with a base mesh that ends rays."""
    import test_ref_instancer as TI
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    mesh, T, ref_kw = TI.mesh_scene(tmp_path, 41, n_anchor=120)
    npar = (1, 3) if not mip else (1, 2)
    om = O.ParamNerfOracle(n_parameters=npar, n_pos=6 if mip else 3, seed=42)
    om.randomize_biases(43, 0.2)
    if mip:
        ipe = util.EasyDict({"module": "network.layer.IntegratedPositionalEncoding", "n_freq_bands": 10})
        km = model.ParamNerf(ipe, RP.ff(util, 4), RP.ff(util, 4), n_parameters=list(npar), n_pos=6)["model"]
        RP.set_weights(km, om)
    else:
        km = RP.build_param_nerf(model, util, om)
    icfg = util.EasyDict(dict(module="instancer.instancer.Instancer", b_0=TI.B0, b_1=TI.B1, textures=["", "light"], **ref_kw))
    R = 160
    o, d = TI.rays_from_above(R, 44)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(F)
    t = np.tile(np.asarray([[0.1, 5.0]], F), (R, 1))
    t[::13, 0] = np.inf
    cone = np.random.default_rng(45).uniform(0.001, 0.01, (R, 1)).astype(F)
    prm = np.asarray([[0.7, 0.2, 0.3, 0.9]], F)
    return tf, renderer, km, om, icfg, o, d, t, cone, prm, R


@pytest.mark.parametrize("variant", [dict(), dict(blur_idx=0), dict(density_reweighting=False, density_scale=2.5), dict(false_color=True),
                                     dict(map_exr=True)])
def test_instance_renderer(ref, variant, tmp_path):
    """InstanceRenderer (renderer.py:215-354) end to end on the reference's side: its Python over its C++ instancer.  The
    oracle gets the very records the C++ returned."""
    tf, renderer, km, om, icfg, o, d, t, cone, prm, R = instanced_setup(ref, tmp_path, mip=False)
    S, step = 64, 0.004
    tf.random.seed(46)
    r = renderer.InstanceRenderer(instancer_config=icfg, model=km, n_samples=S, step_size=step, render_chunk=70, net_chunk=900, **variant)
    got = r(tf.constant(o[None]), tf.constant(d[None]), tf.constant(t[None]), tf.constant(prm), tf.constant(cone[None]),
            composite_bkgd=True, bkgd_color=[1.0, 0.5, 0.25], training=False)
    inst_color = tf.random.log[0][1] if variant.get("false_color") else None
    # replay chunk by chunk with a second instancer of the same seed: the same draws in the same order
    inst2 = icfg_instancer(icfg)
    hit = t[:, 0] != np.inf
    idx = np.nonzero(hit)[0]
    cm, am = np.zeros((R, 3), F), np.zeros(R, F)
    cm[~hit] = np.asarray([1.0, 0.5, 0.25], F)
    for i in range(0, idx.shape[0], 70):
        sl = idx[i:i + 70]
        rec = inst2.get_model_input(o[sl], d[sl], np.repeat(prm, sl.shape[0], 0), S, step)
        tup = (rec["rays_d"], rec["pts"], rec["t"], rec["dist"], rec["color"], rec["density"], rec["density_weight"], rec["instance_id"],
               np.nonzero(rec["hit"])[0], rec["parameters"])
        c, a = O.instance_evaluate_model(om, tup, sl.shape[0], cone_scale=cone[sl], blur_idx=variant.get("blur_idx"), patch_scale=0.09,
                                         density_scale=variant.get("density_scale", 1.0), density_reweighting=variant.get("density_reweighting", True),
                                         composite_bkgd=True, bkgd_color=(1.0, 0.5, 0.25), map_exr=variant.get("map_exr", False),
                                         net_chunk=900, instance_color=inst_color)
        cm[sl], am[sl] = c, a
    assert np.abs(got["color_pred"].numpy()[0] - cm).max() <= 3e-5 and np.abs(got["alpha_pred"].numpy()[0] - am).max() <= 3e-5
    assert (am[hit] > 0.5).any() and (am[hit] == 0).any()


def icfg_instancer(icfg):
    from oracle import ref_instancer as RI
    kw = {k: v for k, v in icfg.items() if k != "module"}
    return RI.RefInstancer(kw.pop("b_0"), kw.pop("b_1"), **kw)


def test_mip_instance_renderer(ref, tmp_path):
    """MipInstanceRenderer (renderer.py:475-587)."""
    tf, renderer, km, om, icfg, o, d, t, cone, prm, R = instanced_setup(ref, tmp_path, mip=True)
    S, step = 64, 0.004
    r = renderer.MipInstanceRenderer(instancer_config=icfg, model=km, n_samples=S, step_size=step, render_chunk=70, net_chunk=900, blur_idx=1)
    got = r(tf.constant(o[None]), tf.constant(d[None]), tf.constant(t[None]), tf.constant(prm), tf.constant(cone[None]),
            composite_bkgd=False, training=False)
    inst2 = icfg_instancer(icfg)
    hit = t[:, 0] != np.inf
    idx = np.nonzero(hit)[0]
    cm, am = np.zeros((R, 3), F), np.zeros(R, F)
    for i in range(0, idx.shape[0], 70):
        sl = idx[i:i + 70]
        rec = inst2.get_model_input(o[sl], d[sl], np.repeat(prm, sl.shape[0], 0), S, step)
        tup = (rec["rays_d"], rec["pts"], rec["t"], rec["dist"], rec["color"], rec["density"], rec["density_weight"], rec["instance_id"],
               np.nonzero(rec["hit"])[0], rec["parameters"])
        c, a = O.mip_instance_evaluate_model(om, tup, sl.shape[0], cone[sl], 1, patch_scale=0.09, net_chunk=900)
        cm[sl], am[sl] = c, a
    assert np.abs(got["color_pred"].numpy()[0] - cm).max() <= 3e-5 and np.abs(got["alpha_pred"].numpy()[0] - am).max() <= 3e-5
    assert (am[hit] > 0.5).any()


@pytest.mark.parametrize("factor,H,W", [(2, 16, 20), (3, 17, 23), (4, 32, 32)])
def test_filtered_downsample(ref, factor, H, W):
    """util/interpolate.py:67-82 (what Logger.render_image applies, logger.py:126-131): gaussian kernel size and
    normalisation, strided 'SAME' depthwise convolution."""
    tf, *_ = ref
    from util import interpolate
    img = np.random.default_rng(51).uniform(0, 1, (H, W, 4)).astype(F)
    got = interpolate.filtered_downsample(tf.constant(img), factor).numpy()
    want = O.filtered_downsample(img, factor)
    assert got.shape == want.shape and np.abs(got - want).max() <= 2e-6
    k = int(factor * .5 * 6)
    ker = interpolate.gaussian_kernel(k, factor * .5, 4).numpy()
    assert ker.shape == (k, k, 4, 1) and np.abs(ker[:, :, 0, 0] - O.gaussian_kernel_2d(k, factor * .5)).max() <= 1e-7


def test_look_at(ref):
    """network/dataset.py:231-238 against the product's dataset.look_at (the bench's camera poses)."""
    tf, *_ = ref
    from network import dataset as ref_dataset
    from nerf_tex_b200.dataset import look_at
    for pos in ([5.0, 0.3, 2.0], [0.1, -4.0, 3.0], [-2.0, 2.0, 0.5], [0.0, 0.0, 6.0]):
        got = ref_dataset.look_at(tf.constant(pos)).numpy()
        assert np.abs(got - look_at(pos)).max() <= 1e-6, pos


# ------------------------------------------------------------------------------------------------ the shipped configs
def _config_names():
    return sorted(p.stem for p in (RP.REF / "configs").glob("config_*_*.py")) if RP.available() else []


@pytest.mark.parametrize("name", _config_names())
def test_shipped_config_through_the_references_own_classes(ref, name, tmp_path, monkeypatch):
    """Every configs/config_*_{train,render}.py, imported as the reference's main.py imports it, its `model_config` and
    `renderer_config` instantiated by the REFERENCE'S util.instantiate into the reference's own classes exactly as
    network/render.py:21-25 and network/train.py:30-43 do (render configs: the reference's C++ instancer on generated
    assets under the configs' file names) — against the oracle with the same keyword values."""
    import copy
    import importlib
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    cfg = copy.deepcopy(importlib.import_module("configs." + name).config)
    mc, rc = util.EasyDict(cfg["model_config"]), util.EasyDict(cfg["renderer_config"])
    npar = tuple(mc["n_parameters"]) if "n_parameters" in mc else tuple(mc["model_config"]["n_parameters"])
    is_render = "instancer_config" in rc
    if is_render:
        from test_reference_configs import procedural_assets
        procedural_assets(str(tmp_path), rc["instancer_config"])
        monkeypatch.chdir(tmp_path)                                   # the configs name their assets relative to the run directory
    models = util.instantiate(mc)                                    # the reference's ParamNerf / CoarseFine
    oracles = {}
    for i, (mname, km) in enumerate(models.items()):
        om = O.ParamNerfOracle(n_parameters=npar, seed=70 + i)
        om.randomize_biases(80 + i, 0.1)
        RP.set_weights(km, om)
        oracles[mname] = om
    rc.update(models)
    r = util.instantiate(rc)                                         # the reference's Renderer / InstanceRenderer
    p = sum(npar)
    rng = np.random.default_rng(90)
    if not is_render:
        R = 24
        o, d, t, cone = rays(R, 91)
        t[0, ::5, 0] = np.inf
        prm = rng.uniform(0.2, 1.0, (1, p)).astype(F)
        tf.random.seed(92)
        got = r(tf.constant(o), tf.constant(d), tf.constant(t), tf.constant(prm), tf.constant(cone), composite_bkgd=True, training=True)
        hit = t[0, :, 0] != np.inf
        S, N = rc["n_samples"], rc.get("n_importance", 0)
        log = list(tf.random.log)
        z_rand = np.zeros((R, S), F)
        if rc.get("perturb", True):
            z_rand[hit] = log.pop(0)[1]
        noise = None
        if rc.get("raw_noise_std", 0) > 0:
            noise = np.zeros((R, S), F)
            noise[hit] = log.pop(0)[1] * F(rc["raw_noise_std"])
        assert not log and N == 0
        out = O.render_rays(oracles["model"], None, o[0][hit], d[0][hit], t[0][hit], np.repeat(prm, R, 0)[hit], cone[0][hit], S, 0,
                            rc.get("perturb", True), z_rand=z_rand[hit], noise=None if noise is None else noise[hit],
                            blur_idx=rc.get("blur_idx"), composite_bkgd=True)
        assert np.abs(got["color_pred"].numpy()[0][hit] - out["color_pred"]).max() <= 3e-5
        assert np.abs(got["alpha_pred"].numpy()[0][hit] - out["alpha_pred"]).max() <= 3e-5
        assert (got["color_pred"].numpy()[0][~hit] == 1).all()
        return
    ic = rc["instancer_config"]
    R = 40
    o = np.stack([rng.uniform(-1.3, 1.3, R), rng.uniform(-1.3, 1.3, R), rng.uniform(0.8, 1.6, R)], 1).astype(F)
    d = np.stack([rng.uniform(-.6, .6, R), rng.uniform(-.6, .6, R), np.full(R, -1.0)], 1).astype(F)
    d = (d / np.linalg.norm(d, axis=-1, keepdims=True)).astype(F)
    light = [3.0, 1.0, 0.4, 0.8] if "point" in ic["textures"] else [0.9, 0.1, 0.12]
    prm = np.asarray([list(rng.uniform(0.3, 1.0, p - len(light))) + light], F)
    t = np.tile(np.asarray([[0.1, 5.0]], F), (R, 1))
    cone = rng.uniform(0.01, 0.05, (R, 1)).astype(F)
    r.density_scale = 3.0
    got = r(tf.constant(o[None]), tf.constant(d[None]), tf.constant(t[None]), tf.constant(prm), tf.constant(cone[None]),
            composite_bkgd=True, training=False)
    rec = r.instancer.last                                            # the C++ instancer's records of the one chunk
    tup = (rec["rays_d"], rec["pts"], rec["t"], rec["dist"], rec["color"], rec["density"], rec["density_weight"], rec["instance_id"],
           np.nonzero(rec["hit"])[0], rec["parameters"])
    c, a = O.instance_evaluate_model(oracles["model"], tup, R, cone_scale=cone, blur_idx=rc.get("blur_idx"), patch_scale=ic["patch_scale"],
                                     density_scale=3.0, density_reweighting=rc.get("density_reweighting", True), composite_bkgd=True,
                                     net_chunk=rc["net_chunk"])
    assert np.abs(got["color_pred"].numpy()[0] - c).max() <= 3e-5 and np.abs(got["alpha_pred"].numpy()[0] - a).max() <= 3e-5
    assert rec["hit"].sum() > R // 2 and (a > 0.2).any()


# ------------------------------------------------------------------------------------------------ the training step
@pytest.mark.parametrize("cfg", ["grass_filtered", "coarse_fine"])
def test_train_loop(ref, cfg, tmp_path, monkeypatch):
    """network/train.py::Train, UNMODIFIED, for three iterations — its GradientTape over the renderer and AlphaLoss, its
    Adam with its ExponentialDecay — with stand-ins only for the tf.data dataset and the logger (tests/ref_train_stubs.py).
    Replica on our side: float64 autograd of the oracle graph (tests/torch_ref.py) + oracle.adam_step with
    oracle.exponential_decay_lr.  (The stand-in differentiates in fp32 with rules of its own, oracle/tf_shim/
    tensorflow/_autodiff.py, and implements Keras' documented Adam: what is pinned is the reference's use of them.)"""
    import torch
    import torch_ref as T
    import ref_train_stubs as stubs
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    from network import train as ref_train
    npar = (2, 3)
    fine = cfg == "coarse_fine"
    names = ["model", "model_fine"] if fine else ["model"]
    oms = {n: O.ParamNerfOracle(n_parameters=npar, seed=201 + i) for i, n in enumerate(names)}
    for i, om in enumerate(oms.values()):
        om.randomize_biases(211 + i, 0.1)
    R, S, N, iters, lrate, decay = 20, 16, (24 if fine else 0), 3, 5e-4, 0.002      # decay_steps = 2: the schedule moves in 3 steps
    rng = np.random.default_rng(221)
    batches = []
    for _ in range(iters):
        o, d, t, cone = rays(R, int(rng.integers(1 << 30)))
        batches.append(dict(rays_o=o, rays_d=d, t=t, cone_scale=cone, parameters=rng.uniform(0.2, 1.0, (1, 5)).astype(F),
                            color=rng.uniform(0, 1, (1, R, 3)).astype(F), alpha=(rng.uniform(0, 1, (1, R)) * (rng.uniform(size=(1, R)) > .3)).astype(F)))
    ffc = lambda n: {"module": "network.layer.FourierFeatures", "n_freq_bands": n}
    pn = {"module": "network.model.ParamNerf", "pos_embedding": ffc(10), "dir_embedding": ffc(4), "param_embedding": ffc(4),
          "n_parameters": list(npar)}
    model_config = {"module": "network.model.CoarseFine", "model_config": pn} if fine else pn
    renderer_config = {"module": "network.renderer.Renderer", "n_samples": S, "n_importance": N, "perturb": True,
                       "raw_noise_std": 0.0 if fine else 0.1, "blur_idx": None if fine else 0}
    tf.random.seed(231)
    tf._autodiff.DROPPED.clear()
    # listen at the optimizer (the stand-in's class, not the reference's code): gradients and weights of every step
    steps = []
    Adam = tf.keras.optimizers.Adam
    orig_apply = Adam.apply_gradients

    def spy(self, grads_and_vars):
        gv = list(grads_and_vars)
        steps.append({"lr": self._lr(), "grads": [None if g is None else g.numpy().copy() for g, _ in gv],
                      "before": [v.numpy().copy() for _, v in gv]})
        orig_apply(self, gv)
        steps[-1]["after"] = [v.numpy().copy() for _, v in gv]
    monkeypatch.setattr(Adam, "apply_gradients", spy)
    ref_train.Train(target_path=str(tmp_path),
                    train_dataset_config=util.EasyDict({"module": "ref_train_stubs.Batches", "n_parameters": list(npar), "composite_bkgd": True,
                                                        "bkgd_color": [1., 1., 1.],
                                                        "batches": [{k: tf.constant(v) for k, v in b.items()} for b in batches]}),
                    val_dataset_config=util.EasyDict({"module": "ref_train_stubs.Batches"}),
                    model_config=util.EasyDict(model_config),
                    loss_config=util.EasyDict({"module": "network.loss.AlphaLoss", "loss_fn": "network.loss.smape", "alpha_loss_fn": "network.loss.mse"}),
                    n_iters=iters, lrate=lrate, lrate_decay=decay, renderer_config=util.EasyDict(renderer_config),
                    logger_config=util.EasyDict({"module": "ref_train_stubs.Logger",
                                                 "initial": {n: [(W, b) for _, W, b in om.layers] for n, om in oms.items()}}))
    logger = stubs.LAST["logger"]
    assert len(logger.losses) == iters == len(steps)
    # operations that saw a tracked tensor without a differentiation rule: only what lies behind tf.stop_gradient
    assert tf._autodiff.DROPPED <= {"cumsum", "stack", "gather", "searchsorted", "where", "maximum", "minimum", "sort"}, tf._autodiff.DROPPED
    # one optimizer over the variables of all models, in model order, kernel before bias (train.py:33-34)
    n_vars = sum(2 * len(om.layers) for om in oms.values())
    assert all(len(st["grads"]) == n_vars for st in steps)
    log = list(tf.random.log)
    t64 = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.float64)
    mom = [(np.zeros_like(w), np.zeros_like(w)) for w in steps[0]["before"]]
    for step, (b, st) in enumerate(zip(batches, steps)):
        # ---- the reference's weights of this step into the float64 twin, its draws, the same graph
        for n, om in oms.items():
            k0 = sum(2 * len(oms[m].layers) for m in names[:names.index(n)])
            om.layers = [(name, st["before"][k0 + 2 * j], st["before"][k0 + 2 * j + 1]) for j, (name, _, _) in enumerate(om.layers)]
        P = {n: T.make_params(om) for n, om in oms.items()}
        z_rand = log.pop(0)[1]
        noise = None if fine else log.pop(0)[1] * F(0.1)
        o, d, t, cone = b["rays_o"][0], b["rays_d"][0], b["t"][0], b["cone_scale"][0]
        prm = np.repeat(b["parameters"], R, 0)
        z = t64(O.stratified_z(t, S, True, z_rand))
        C, A, W_c = T.network_pass(oms["model"], P["model"], t64(o), t64(d), z, t64(prm), t64(cone), None if fine else 0,
                                   composite_bkgd=True, noise=None if noise is None else t64(noise))
        if fine:                                                      # det=self.perturb=True: deterministic resampling of the coarse weights
            z_all, _ = O.hierarchical_z(z.numpy().astype(F), W_c.detach().numpy().astype(F), N, True)
            Cf, Af, _ = T.network_pass(oms["model_fine"], P["model_fine"], t64(o), t64(d), t64(z_all), t64(prm), t64(cone), None, composite_bkgd=True)
            L = T.alpha_loss(t64(b["color"][0]), t64(b["alpha"][0]), Cf, Af, C, A)
        else:
            L = T.alpha_loss(t64(b["color"][0]), t64(b["alpha"][0]), C, A)
        assert abs(logger.losses[step] - float(L.detach())) <= 2e-4 * max(abs(float(L.detach())), 1e-3)
        leaves = [x for n in names for pair in P[n].values() for x in pair]
        want = torch.autograd.grad(L, leaves, allow_unused=True)
        # ---- the tape's gradients (fp32, the stand-in's rules) against float64 autograd of the oracle graph
        flat_g = np.concatenate([np.zeros(w.size) if g is None else g.astype(np.float64).ravel() for g, w in zip(st["grads"], st["before"])])
        flat_w = np.concatenate([np.zeros(x.numel()) if g is None else g.numpy().ravel() for g, x in zip(want, leaves)])
        cos = float(flat_g @ flat_w / (np.linalg.norm(flat_g) * np.linalg.norm(flat_w)))
        rel = float(np.linalg.norm(flat_g - flat_w) / np.linalg.norm(flat_w))
        # (coarse + fine: the fine samples sit behind the resampling of fp32 against float64 coarse weights, which moves
        # a few of them: the bars of the GPU gradient tests apply there)
        assert (cos >= 0.9999 and rel <= 2e-2) if fine else (cos >= 0.99999 and rel <= 2e-3), (step, cos, rel)
        # ---- the update: Adam as Keras documents it with the decayed rate, from the gradients the tape produced
        assert abs(st["lr"] - O.exponential_decay_lr(lrate, decay, step)) <= 1e-9
        for j, (g, w0, w1) in enumerate(zip(st["grads"], st["before"], st["after"])):
            p_new, m_new, v_new = O.adam_step(w0, g, mom[j][0], mom[j][1], step + 1, st["lr"])
            mom[j] = (m_new, v_new)
            assert np.abs(p_new - w1).max() <= 2e-8 + 1e-6 * lrate, (step, j, np.abs(p_new - w1).max())
            if step == 0:                                             # Adam's first step moves every weight by ~lr
                assert np.abs(w1 - w0).max() > 0.5 * st["lr"], j
    assert not log


def test_pixel_samplers(ref):
    """network/pixel_sampler.py: Full (every pixel, row-major), Independent (iid, in range), Proxy (pixels whose rays hit
    the proxy on a grid `downsample_factor` coarser — focal // factor as written — upsampled with 'nearest', shuffled)."""
    tf, renderer, model, loss, ray_sampler, proxy, util = ref
    from network import pixel_sampler as ps
    H, W, focal, factor = 48, 64, 70.0, 8
    full = ps.Full(H, W)().numpy()
    assert np.array_equal(full, np.stack(np.meshgrid(np.arange(H), np.arange(W), indexing="ij"), -1).reshape(-1, 2))
    ind = ps.Independent(H, W, 500)().numpy()
    assert ind.shape == (500, 2) and ind[:, 0].max() < H and ind[:, 1].max() < W and ind.min() >= 0
    b0, b1 = [-0.6, -0.5, -0.4], [0.5, 0.6, 0.4]
    c2w = np.asarray([[np.cos(.4), 0, np.sin(.4), 1.2], [0, 1, 0, 0.1], [-np.sin(.4), 0, np.cos(.4), 2.6], [0, 0, 0, 1]], F)
    # the oracle's mask: coarse grid through rays_from_camera + AABB, half-pixel-centre nearest upsampling
    hd, wd = H // factor, W // factor
    ii, jj = np.meshgrid(np.arange(hd), np.arange(wd), indexing="ij")
    o, d, _ = O.rays_from_camera(np.stack([ii.reshape(-1), jj.reshape(-1)], -1).astype(F), hd, wd, focal // factor, c2w)
    hit = (O.aabb_intersect(b0, b1, o, d)[:, 0] != np.inf).reshape(hd, wd)
    ys = np.minimum(np.floor((np.arange(H) + 0.5) * (hd / H)).astype(int), hd - 1)
    xs = np.minimum(np.floor((np.arange(W) + 0.5) * (wd / W)).astype(int), wd - 1)
    want = np.argwhere(hit[ys][:, xs])
    assert 0.05 < hit.mean() < 0.95
    tf.random.seed(61)
    got = ps.Proxy(H, W, want.shape[0], proxy.AABB(b0, b1), focal, downsample_factor=factor)(tf.constant(c2w)).numpy()
    perm = tf.random.log[-1][1]
    assert np.array_equal(got, want[perm])                            # all hit pixels, in the drawn order

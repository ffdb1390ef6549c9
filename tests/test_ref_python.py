"""The oracle of the TensorFlow path (oracle/nerftex_oracle.py) against the REFERENCE'S OWN Python source — network/
renderer.py, model.py, layer.py, loss.py, ray_sampler.py, proxy.py imported from /root/reference and executed over a
NumPy stand-in for the TensorFlow ops they call (oracle/tf_shim: our code; it restates the documented semantics of each
op in fp32).  What this pins is the reference's LOGIC as written — which tensors are sliced, concatenated, gathered,
scattered, in which order; chunking; culling of rays with t = inf; the `det=self.perturb` inversion; `weights[..., 1:-1]`;
the blur scaling; the instanced renderer's masking and last-sample handling; the mip cone gaussians — not TensorFlow's
kernels (TensorFlow 2.4 cannot run here).  Random draws are taken from the stand-in's log and handed to the oracle.

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

"""A third, library-backed restatement of the sampling / compositing semantics, held against the oracle (CPU only).

Parity stays UNPINNED (no TensorFlow here, the reference holds no vectors); what this adds is an implementation of the
same TF ops by an independent library.  Each function below follows the reference line by line with the torch op of
the same meaning substituted for the tf op:

    tf.linspace -> torch.linspace          tf.cumsum -> torch.cumsum          tf.math.cumprod(exclusive) -> torch.cumprod
    tf.searchsorted(side='right') -> torch.searchsorted(right=True)           tf.sort -> torch.sort
    tf.gather(batch_dims) -> torch.gather  tf.reduce_sum -> torch.sum         tf.math.sigmoid / exp -> torch.sigmoid / exp

and is asserted against oracle/nerftex_oracle.py: integer results (search indices) equal, floating results within a
stated number of fp32 ulps.  The places where the two libraries legitimately differ are enumerated, counted and
bounded in the tests instead of being hidden behind a loose tolerance:

  E1  torch.linspace fills the upper half from the end (`stop - step * (n-1-i)`), tf.linspace (and the oracle) use
      `start + step * i` throughout: interior points of the upper half may differ by 1 ulp, the rest is identical.
  E2  torch.sum reduces in vectorised chunks, the oracle strictly left to right: a few ulps on sums of 62..256 terms.
  E3  torch.cumsum on CPU accumulates float32 inputs in double and rounds each prefix: <= 1 ulp per entry against the
      oracle's fp32 running sum (1.5 ulp in total after E2 moved the normaliser).
  E4  torch.exp / torch.sigmoid are float32 vector routines (<= 1 ulp); the oracle rounds a float64 evaluation.
  E5  a search index may differ only where u lies within a few ulps of a CDF entry (consequence of E2/E3; mostly the
      last sample, u = 1 against cdf[-1] = 1 -+ 1 ulp); the inverse CDF is continuous there, so the sample still
      agrees to ~1e-5 of the bin range.
  E6  `denom < 1e-5 -> 1` (renderer.py:614) is a genuine discontinuity of the reference function: a bin whose CDF
      increment sits within rounding of 1e-5 lands on either side, and the sample moves by up to the bin width.  Such
      entries are counted (a few per 100 000) and excluded from the value comparison.  The same switch makes the
      inverse CDF jump across a (near-)zero-width bin, so an E5 index difference next to such a bin moves the sample
      by that bin's width in depth; entries whose index differs are therefore bounded by the local bin width instead.
"""

import numpy as np
import torch

from oracle import nerftex_oracle as O

torch.set_num_threads(4)


def ulps(a, b):
    """|a - b| in units of the fp32 spacing at max(|a|, |b|) (elementwise)."""
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    sp = np.spacing(np.maximum(np.abs(a), np.abs(b)).astype(np.float32))
    return np.abs(a.astype(np.float64) - b.astype(np.float64)) / np.maximum(sp, np.float32(1e-45))


# ------------------------------------------------------------------------------------------------ torch restatements
def t_stratified_z(t, n_samples, perturb, z_rand, t_vals=None):
    """network/renderer.py:101-111."""
    t_vals = torch.linspace(0., 1., n_samples) if t_vals is None else t_vals
    z = t[:, None, 0] * (1 - t_vals) + t[:, None, 1] * t_vals
    z = torch.broadcast_to(z, (t.shape[0], n_samples))
    if perturb:
        mids = .5 * (z[..., 1:] + z[..., :-1])
        upper = torch.cat([mids, z[..., -1:]], -1)
        lower = torch.cat([z[..., :1], mids], -1)
        z = lower + (upper - lower) * z_rand
    return z


def t_map_model_output(color, alpha, z, rays_d, composite_bkgd, bkgd):
    """network/renderer.py:170-213."""
    dists = z[..., 1:] - z[..., :-1]
    dists = torch.cat([dists, dists[..., -1:]], -1)
    dists = dists * torch.linalg.norm(rays_d[..., None, :], dim=-1)
    color_map = torch.sigmoid(color)
    alpha_map = 1 - torch.exp(-torch.relu(alpha) * dists)
    x = 1. - alpha_map + 1e-10
    excl = torch.cat([torch.ones_like(x[..., :1]), torch.cumprod(x, -1)[..., :-1]], -1)     # cumprod(exclusive=True)
    weights = alpha_map * excl
    color_map = torch.sum(weights[..., None] * color_map, -2)
    alpha_map = torch.sum(weights, -1)
    if composite_bkgd:
        color_map = color_map + (1. - alpha_map[..., None]) * bkgd
    return color_map, alpha_map, weights


def t_sample_pdf(bins, weights, n, det, u=None):
    """network/renderer.py:589-618."""
    weights = weights + 1e-5
    pdf = weights / torch.sum(weights, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    if det:
        u = torch.broadcast_to(torch.linspace(0., 1., n), (cdf.shape[0], n))
    u = u.contiguous()
    inds = torch.searchsorted(cdf.contiguous(), u, right=True)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_b, cdf_a = torch.gather(cdf, -1, below), torch.gather(cdf, -1, above)
    bins_b, bins_a = torch.gather(bins, -1, below), torch.gather(bins, -1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    tt = (u - cdf_b) / denom
    return bins_b + tt * (bins_a - bins_b), inds, cdf, u


def value_tolerance(cdf, inds, bins):
    """Conditioning-aware bound for one sample: t = (u - cdf_b) / denom amplifies a CDF rounding difference of a few
    ulps of 1 (E2/E3; 32 allowed) by 1 / denom, and the sample is bins_b + t * (bin width)."""
    C = cdf.shape[-1]
    below, above = np.maximum(inds - 1, 0), np.minimum(inds, C - 1)
    denom = np.take_along_axis(cdf, above, -1) - np.take_along_axis(cdf, below, -1)
    denom = np.where(denom < 1e-5, 1.0, denom)
    width = np.take_along_axis(bins, above, -1) - np.take_along_axis(bins, below, -1)
    span = (bins[:, -1] - bins[:, 0])[:, None]
    return (32 * 5.96e-8 / denom) * width + 4e-6 * span


def on_denom_threshold(cdf, inds, width=2e-8):
    """Entries whose CDF increment is within rounding of the 1e-5 switch of renderer.py:614 (E6)."""
    C = cdf.shape[-1]
    below, above = np.maximum(inds - 1, 0), np.minimum(inds, C - 1)
    denom = np.take_along_axis(cdf, above, -1) - np.take_along_axis(cdf, below, -1)
    return np.abs(denom - np.float32(1e-5)) < width


# ------------------------------------------------------------------------------------------------ tests
def test_linspace_E1():
    for n in (2, 3, 64, 128, 192, 256, 1024):
        a, b = torch.linspace(0., 1., n).numpy(), O.tf_linspace(0., 1., n)
        assert a[0] == b[0] == 0 and a[-1] == b[-1] == 1
        diff = np.nonzero(a != b)[0]
        assert ulps(a, b).max() <= 1
        assert np.all(diff >= n // 2), (n, diff)                       # E1: only the half torch fills from the end


def test_stratified_z_matches_bitwise_given_the_same_t_vals():
    rng = np.random.default_rng(0)
    R, S = 257, 64
    t = np.stack([rng.uniform(2, 4, R), rng.uniform(5, 7, R)], -1).astype(np.float32)
    zr = rng.uniform(size=(R, S)).astype(np.float32)
    for perturb in (False, True):
        want = O.stratified_z(t, S, perturb, zr)
        got = t_stratified_z(torch.from_numpy(t), S, perturb, torch.from_numpy(zr), torch.from_numpy(O.tf_linspace(0., 1., S)))
        assert np.array_equal(got.numpy().view(np.uint32), want.view(np.uint32))        # same op order: bit-equal
        got2 = t_stratified_z(torch.from_numpy(t), S, perturb, torch.from_numpy(zr))    # torch's own linspace: E1
        assert ulps(got2.numpy(), want).max() <= 2


def test_searchsorted_right_is_the_oracles_upper_bound():
    """Indices on IDENTICAL cdf / u (no float path in between): must be equal everywhere, ties included."""
    rng = np.random.default_rng(1)
    R, C, N = 300, 63, 128
    cdf = np.sort(rng.uniform(size=(R, C)).astype(np.float32), -1)
    cdf[:, 0], cdf[:, -1] = 0, 1
    cdf[::7, 10:14] = cdf[::7, 10:11]                                  # runs of equal entries (zero-width bins)
    u = rng.uniform(size=(R, N)).astype(np.float32)
    u[:, 0], u[:, -1] = 0, 1                                           # the two ends of linspace
    u[:, 5:25] = cdf[:, 20:40]                                         # exact ties with CDF entries
    want = (cdf[:, None, :] <= u[:, :, None]).sum(-1)
    got = torch.searchsorted(torch.from_numpy(cdf), torch.from_numpy(u), right=True).numpy()
    assert np.array_equal(got, want)
    assert got.min() >= 1 and got.max() == C                           # u >= cdf[0] = 0 always; u = 1 -> C


def test_sample_pdf_E2_E3_E5():
    rng = np.random.default_rng(2)
    R, C, N = 2000, 63, 128
    bins = np.sort(rng.uniform(2, 6, (R, C)).astype(np.float32), -1)
    weights = (rng.uniform(size=(R, C - 1)) ** 8).astype(np.float32)   # peaky, like real compositing weights
    weights[::50] = 0                                                   # all-zero rows (the 1e-5 floor carries them)
    for det in (True, False):
        u_in = None if det else rng.uniform(size=(R, N)).astype(np.float32)
        want, want_inds = O.sample_pdf(bins, weights, N, det, u_in, return_inds=True)
        got, inds, cdf, u = t_sample_pdf(torch.from_numpy(bins), torch.from_numpy(weights), N, det,
                                         None if det else torch.from_numpy(u_in))
        got, inds, cdf, u = got.numpy(), inds.numpy(), cdf.numpy(), u.numpy()
        mism = inds != want_inds
        assert mism.mean() < 1e-2, mism.mean()                           # at most ~ the u = 1 column
        if mism.any():                                                  # E5: only next to a CDF entry, and by one bin
            r, k = np.nonzero(mism)
            assert np.abs(inds[r, k] - want_inds[r, k]).max() == 1
            near = np.abs(cdf[r] - u[r, k][:, None]).min(-1)
            assert near.max() <= 8 * np.spacing(np.float32(1)), near.max()
        span = (bins[:, -1] - bins[:, 0])[:, None]
        rel = np.abs(got - want) / span
        edge = on_denom_threshold(cdf, inds)                            # E6
        assert edge.mean() < 1e-3, edge.mean()
        ok = ~edge & ~mism
        assert np.all((np.abs(got - want) <= value_tolerance(cdf, inds, bins))[ok])
        assert np.percentile(rel[ok], 99.9) < 1e-5                      # and almost everywhere it is rounding noise
        if mism.any():                                                  # E5 next to E6: at most the neighbouring bins
            lo = np.clip(np.minimum(inds, want_inds) - 1, 0, C - 2)[mism]
            hi = np.clip(np.maximum(inds, want_inds), 1, C - 1)[mism]
            rows = np.nonzero(mism)[0]
            assert np.all(np.abs(got - want)[mism] <= (bins[rows, hi] - bins[rows, lo]) * (1 + 1e-5) + 1e-6)
        assert np.median(ulps(got, want)) <= 1


def test_map_model_output_E2_E4():
    rng = np.random.default_rng(3)
    R, S = 500, 192
    z = np.sort(rng.uniform(2, 6, (R, S)).astype(np.float32), -1)
    color = rng.normal(size=(R, S, 3)).astype(np.float32)
    alpha = (rng.normal(size=(R, S)) * 20).astype(np.float32)
    rays_d = rng.normal(size=(R, 3)).astype(np.float32)
    for bk in (False, True):
        wc, wa, ww, _ = O.composite(color, alpha, z, rays_d, bk)
        gc, ga, gw = t_map_model_output(*(torch.from_numpy(a) for a in (color, alpha, z, rays_d)), bk, torch.ones(3))
        # weights: products of exp results (E4) through a sequential cumprod: tiny relative error, no cancellation
        np.testing.assert_allclose(gw.numpy(), ww, rtol=2e-5, atol=1e-7)
        np.testing.assert_allclose(ga.numpy(), wa, rtol=1e-5, atol=1e-6)       # E2 on 192 terms
        np.testing.assert_allclose(gc.numpy(), wc, rtol=1e-5, atol=1e-6)


def test_hierarchical_merge_is_a_plain_sort():
    """network/renderer.py:125-130: z_mid, sample_pdf on weights[1:-1] with det = perturb, sort(concat)."""
    rng = np.random.default_rng(4)
    R, S, N = 400, 64, 128
    z = np.sort(rng.uniform(2, 6, (R, S)).astype(np.float32), -1)
    w = (rng.uniform(size=(R, S)) ** 6).astype(np.float32)
    z_all, z_samples = O.hierarchical_z(z, w, N, True)
    zt, wt = torch.from_numpy(z), torch.from_numpy(w)
    z_mid = .5 * (zt[..., 1:] + zt[..., :-1])
    assert np.array_equal(z_mid.numpy().view(np.uint32),
                          (np.float32(0.5) * (z[..., 1:] + z[..., :-1])).astype(np.float32).view(np.uint32))
    # feed the oracle's own samples so that only the merge is compared: torch.sort == np.sort, values only
    merged = torch.sort(torch.cat([zt, torch.from_numpy(z_samples)], -1), -1).values.numpy()
    assert np.array_equal(merged.view(np.uint32), z_all.view(np.uint32))
    _, _, want_inds = O.hierarchical_z(z, w, N, True, return_inds=True)
    got, inds, cdf, _ = t_sample_pdf(z_mid, wt[..., 1:-1], N, True)
    rel = np.abs(got.numpy() - z_samples) / (z[:, -1:] - z[:, :1])
    ok = ~on_denom_threshold(cdf.numpy(), inds.numpy()) & (inds.numpy() == want_inds)      # E5 / E6
    assert ok.mean() > 0.99
    assert np.all((np.abs(got.numpy() - z_samples) <= value_tolerance(cdf.numpy(), inds.numpy(), z_mid.numpy()))[ok])
    assert np.percentile(rel[ok], 99.9) < 1e-5

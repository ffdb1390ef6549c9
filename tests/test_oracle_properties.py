"""Property tests of the oracles (hypothesis), the invariants SURVEY.md section 4 lists: compositing weights are a
sub-probability, the CDF is monotone, resampled depths stay inside the bin range and come out sorted, the merged
depths are the sorted multiset union; instancer records stay inside the patch box and tile the in-shell length."""

import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import instancer_oracle as IO
from oracle import nerftex_oracle as O

SET = settings(max_examples=30, deadline=None)


@SET
@given(st.integers(0, 2 ** 31 - 1), st.integers(1, 9), st.integers(2, 70), st.booleans(), st.booleans())
def test_composite_weights_are_a_sub_probability(seed, R, S, exr, bk):
    rng = np.random.default_rng(seed)
    color = (rng.normal(size=(R, S, 3)) * 3).astype(np.float32)
    sigma = (rng.normal(size=(R, S)) * rng.choice([0.1, 5, 200])).astype(np.float32)
    z = np.sort(rng.uniform(1, 7, (R, S)).astype(np.float32), -1)
    d = rng.normal(size=(R, 3)).astype(np.float32)
    cmap, amap, w, _ = O.composite(color, sigma, z, d, bk, (1., 1., 1.), exr)
    assert (w >= 0).all() and (amap <= 1 + 1e-5).all() and (amap >= 0).all()
    assert np.isfinite(cmap).all()
    if not exr:                                   # sigmoid colours: a convex combination (plus the background's share)
        assert (cmap >= -1e-6).all() and (cmap <= 1 + 1e-5).all()
    dead = np.maximum(sigma, 0).sum(-1) == 0      # no density anywhere on the ray: nothing is accumulated
    assert (amap[dead] == 0).all()


@SET
@given(st.integers(0, 2 ** 31 - 1), st.integers(1, 6), st.integers(3, 40), st.integers(1, 50), st.booleans())
def test_sample_pdf_and_merge_invariants(seed, R, S, N, det):
    rng = np.random.default_rng(seed)
    z = np.sort(rng.uniform(2, 6, (R, S)).astype(np.float32), -1)
    w = (rng.uniform(size=(R, S)) ** rng.choice([1, 8])).astype(np.float32)
    w[rng.uniform(size=R) < 0.3] = 0                        # all-zero rows: the 1e-5 floor makes the pdf uniform
    u = None if det else rng.uniform(size=(R, N)).astype(np.float32)
    z_all, zs, inds = O.hierarchical_z(z, w, N, det, u, return_inds=True)
    mids = 0.5 * (z[:, 1:] + z[:, :-1])
    assert (zs >= mids[:, :1] - 1e-6).all() and (zs <= mids[:, -1:] + 1e-6).all()
    assert inds.min() >= 1 and inds.max() <= S - 1           # searchsorted(right) of u in [0, 1] on a cdf starting at 0
    assert (np.diff(z_all, axis=-1) >= 0).all()
    assert np.array_equal(z_all, np.sort(np.concatenate([z, zs], -1), -1))
    if det:                                                  # increasing u -> non-decreasing samples
        assert (np.diff(zs, axis=-1) >= 0).all()


@SET
@given(st.integers(0, 2 ** 31 - 1), st.integers(2, 40), st.booleans())
def test_stratified_depths_stay_in_their_strata(seed, S, perturb):
    rng = np.random.default_rng(seed)
    R = 5
    t = np.stack([rng.uniform(1, 3, R), rng.uniform(4, 9, R)], -1).astype(np.float32)
    zr = rng.uniform(size=(R, S)).astype(np.float32) if perturb else None
    z = O.stratified_z(t, S, perturb, zr)
    assert (z >= t[:, :1] - 1e-6).all() and (z <= t[:, 1:] + 1e-6).all() and (np.diff(z, axis=-1) >= -1e-6).all()


@settings(max_examples=12, deadline=None)
@given(st.integers(0, 2 ** 31 - 1), st.sampled_from(["random", "nearest", "nearest_blend"]), st.booleans())
def test_instancer_records_invariants(seed, method, mean_dist):
    rng = np.random.default_rng(seed)
    B0, B1 = [-0.5, -0.5, -0.1], [0.5, 0.5, 0.1]
    Ts = []
    for _ in range(6):
        a = rng.uniform(0, 6.28)
        T = np.eye(4)
        T[:3, :3] = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]]) * rng.uniform(0.7, 1.4)
        T[:3, 3] = [rng.uniform(0, 2), rng.uniform(0, 2), rng.uniform(0, 0.1)]
        Ts.append(T)
    R = 25
    o = np.stack([rng.uniform(-1, 3, R), rng.uniform(-1, 3, R), np.full(R, 3.0)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-.4, .4, R), rng.uniform(-.4, .4, R), np.full(R, -1.0)], 1).astype(np.float32)
    step, n_pts = 0.02, 40
    rec = IO.get_model_input(o, d, np.zeros((R, 1), np.float32), n_pts, step, np.array(Ts), B0, B1, seed=seed % 1000,
                             method=method, use_mean_distance=mean_dist)
    ns = rec["n_samples"]
    assert (ns <= n_pts).all() and (ns[~rec["hit"]] == 0).all() and ns.sum() == rec["pts"].shape[0]
    tol = 2e-2 if mean_dist else 1e-4                       # the mean-distance shift moves a point by < one step
    assert (np.abs(rec["pts"]) <= np.array([0.5, 0.5, 0.1]) + tol).all()          # patch-local points lie in the patch box
    np.testing.assert_allclose(np.linalg.norm(rec["dirs"], axis=1), 1, atol=1e-5)
    assert (rec["alpha_weight"] >= 1).all() and (rec["instance_id"] >= 0).all() and (rec["instance_id"] < 6).all()
    k = 0
    for r in range(R):                                      # per ray: increasing t, steps tile the in-shell length
        n = ns[r]
        if n:
            t, dist = rec["t"][k:k + n], rec["dists"][k:k + n]
            assert (np.diff(t) > 0).all() and (dist[:-1] == np.float32(step)).all() and dist[-1] > 0
        k += n

"""CPU checks of oracle/instancer_oracle.py against cases with a closed-form answer (the reference instancer has
no tests and cannot be built here, so these analytic cases are what pins the restatement)."""

import numpy as np

from oracle import instancer_oracle as IO

B0, B1 = [-0.5, -0.5, -0.1], [0.5, 0.5, 0.1]
KNOWN_HASHES = [129708002, 2831084092, 4099845390, 3861530882]       # pcg_hash of 0, 1, 12345, 2^32-1


def translate(x, y, z=0.0):
    T = np.eye(4)
    T[:3, 3] = [x, y, z]
    return T


def test_single_patch_straight_down():
    """Ray through the middle of one identity patch: enters at t = 2.9, leaves at 3.1, ten steps of 0.02."""
    o = np.array([[0.1, -0.2, 3.0]], np.float32)
    d = np.array([[0.0, 0.0, -1.0]], np.float32)
    prm = np.array([[0.5, 0.0, 0.0, 2.0]], np.float32)
    out = IO.get_model_input(o, d, prm, 64, 0.02, [np.eye(4)], B0, B1, seed=1, light_dir_idx=1)
    n = int(out["n_samples"][0])
    assert out["hit"][0] and n in (9, 10)                    # fp32 rounding of 0.2 / 0.02 may fall either side
    u = float(IO.ray_uniform(1, 0))
    assert 0.0 <= u < 1.0
    np.testing.assert_allclose(out["t"], 2.9 + u * 0.02 + 0.02 * np.arange(n), rtol=0, atol=2e-6)
    np.testing.assert_allclose(out["pts"][:, 2], 3.0 - out["t"], atol=1e-6)
    np.testing.assert_allclose(out["pts"][:, :2], np.tile([[0.1, -0.2]], (n, 1)), atol=1e-7)
    np.testing.assert_allclose(out["dirs"], np.tile([[0, 0, -1.0]], (n, 1)), atol=0)
    np.testing.assert_allclose(out["dists"][:-1], 0.02, rtol=1e-7)
    np.testing.assert_allclose(out["dists"].sum(), 0.2, atol=2e-6)          # the steps tile the in-shell length
    np.testing.assert_allclose(out["params"][:, 1:], np.tile([[0, 0, 1.0]], (n, 1)), atol=0)   # light dir, normalised
    assert np.all(out["alpha_weight"] == 1) and np.all(out["instance_id"] == 0)


def test_two_overlapping_patches_union_and_choice():
    """Two patches offset by 0.1 in z: the union of [2.9, 3.1] and [2.8, 3.0] is one segment of length 0.3; inside the
    overlap 'random' reports weight 2 and 'nearest' picks the closer origin with weight 1."""
    o = np.array([[0.0, 0.0, 3.0]], np.float32)
    d = np.array([[0.0, 0.0, -1.0]], np.float32)
    Ts = [np.eye(4), translate(0.05, 0.0, 0.1)]
    prm = np.zeros((1, 1), np.float32)
    rnd = IO.get_model_input(o, d, prm, 64, 0.01, Ts, B0, B1, seed=3, method="random")
    near = IO.get_model_input(o, d, prm, 64, 0.01, Ts, B0, B1, seed=3, method="nearest")
    np.testing.assert_allclose(rnd["dists"].sum(), 0.3, atol=3e-6)
    np.testing.assert_array_equal(rnd["t"], near["t"])
    t = rnd["t"]
    in_overlap = (t > 2.9) & (t < 3.0)
    assert np.array_equal(rnd["alpha_weight"] == 2, in_overlap) and np.all(near["alpha_weight"] == 1)
    assert set(rnd["instance_id"][in_overlap]) == {0, 1}                    # both get picked somewhere in the overlap
    assert np.all(rnd["instance_id"][t < 2.9] == 1) and np.all(rnd["instance_id"][t > 3.0] == 0)
    z = 3.0 - t
    assert np.array_equal(near["instance_id"][in_overlap], (np.hypot(0.05, z - 0.1) < np.abs(z)).astype(np.int32)[in_overlap])
    # patch-local points: the second patch sees z shifted by its origin
    sel1 = rnd["instance_id"] == 1
    np.testing.assert_allclose(rnd["pts"][sel1, 2], z[sel1] - 0.1, atol=1e-6)
    np.testing.assert_allclose(rnd["pts"][sel1, 0], -0.05, atol=1e-7)


def test_disjoint_segments_are_concatenated():
    """Two patches far apart along the ray: samples continue in the second segment where the first one stopped,
    and n_pts clips the march."""
    o = np.array([[0.0, 0.0, 3.0]], np.float32)
    d = np.array([[0.0, 0.0, -1.0]], np.float32)
    Ts = [np.eye(4), translate(0, 0, -1.0)]
    prm = np.zeros((1, 0), np.float32)
    out = IO.get_model_input(o, d, prm, 64, 0.02, Ts, B0, B1, seed=0)
    t = out["t"]
    first, second = t < 3.5, t > 3.5
    assert first.sum() + second.sum() == len(t) and abs(int(first.sum()) - 10) <= 1 and abs(int(second.sum()) - 10) <= 1
    assert np.all((t[first] >= 2.9) & (t[first] <= 3.1)) and np.all((t[second] >= 3.9) & (t[second] <= 4.1))
    np.testing.assert_allclose(np.diff(t[first]), 0.02, atol=1e-6)
    # spacing is kept across the gap: t jumps by the empty length 0.8 plus one step
    np.testing.assert_allclose(t[second][0] - t[first][-1], 0.8 + 0.02, atol=2e-6)
    np.testing.assert_allclose(out["dists"].sum(), 0.4, atol=4e-6)
    clipped = IO.get_model_input(o, d, prm, 7, 0.02, Ts, B0, B1, seed=0)
    assert clipped["n_samples"][0] == 7 and np.array_equal(clipped["t"], t[:7])


def test_thin_crossing_and_rotation():
    """A crossing shorter than one step yields one sample whose dist is the whole crossing; a rotated, scaled
    patch maps the direction with the row-normalised transposed rotation."""
    a = 0.7
    T = np.eye(4)
    T[:3, :3] = 2.0 * np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    o = np.array([[0.0, 0.0, 3.0]], np.float32)
    d = np.array([[0.1, 0.0, -1.0]], np.float32)
    out = IO.get_model_input(o, d, np.zeros((1, 0), np.float32), 16, 1.0, [T], B0, B1, seed=2)
    assert out["n_samples"][0] == 1
    np.testing.assert_allclose(out["dists"][0], 0.4, atol=1e-6)                 # box is 0.4 thick after scaling by 2
    dn = d[0] / np.linalg.norm(d[0])
    np.testing.assert_allclose(out["dirs"][0], (T[:3, :3].T / 2.0) @ dn, atol=1e-6)
    assert abs(np.linalg.norm(out["dirs"][0]) - 1) < 1e-6
    assert np.all(np.abs(out["pts"][0]) <= np.array([0.5, 0.5, 0.1]) + 1e-6)


def test_hash_is_stable():
    """Known answers of the counter hash (the GPU kernel carries the same constants)."""
    assert [IO.hash_u32(x) for x in (0, 1, 12345, 0xFFFFFFFF)] == KNOWN_HASHES
    us = np.array([IO.ray_uniform(7, r) for r in range(4000)])
    assert 0.47 < us.mean() < 0.53 and us.min() >= 0 and us.max() < 1
    picks = np.bincount([IO.pick(7, r, s, 3) for r in range(300) for s in range(10)], minlength=3)
    assert picks.min() > 850

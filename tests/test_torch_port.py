"""The torch-CPU port timed as bench.py's CPU baseline must compute what the NumPy oracle computes."""

import numpy as np
import torch

from oracle import nerftex_oracle as O
from oracle.torch_port import TorchPort


def test_port_matches_numpy_oracle():
    rng = np.random.default_rng(0)
    om = O.ParamNerfOracle(n_parameters=(1, 6), seed=1)
    om.randomize_biases(5, 0.05)
    port = TorchPort(om)
    n = 3000
    pos = rng.uniform(-1.4, 1.4, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    prm = rng.uniform(size=(n, 7)).astype(np.float32)
    c, a = port.forward(torch.from_numpy(pos), torch.from_numpy(d), torch.from_numpy(prm))
    wc, wa = om.forward(pos, d, prm)
    np.testing.assert_allclose(c.numpy(), wc, rtol=1e-3, atol=2e-4)      # fp32 sin(512 x) vs float64: ~3e-5 in the angle
    np.testing.assert_allclose(a.numpy(), wa, rtol=1e-3, atol=2e-4)
    R, S = 64, 48
    hit = rng.uniform(size=R) < 0.7
    nv = np.where(hit, rng.integers(1, S, R), 0)
    dists = np.where(np.arange(S)[None] < nv[:, None], np.float32(0.01), np.float32(0)).astype(np.float32)
    dense = (rng.normal(size=(R, S, 3)).astype(np.float32), rng.uniform(-1, 1, (R, S, 3)).astype(np.float32),
             np.zeros((R, S), np.float32), dists, np.zeros((R, 1, 3), np.float32),
             (rng.uniform(size=(R, 1)) < .5).astype(np.float32), np.ones((R, S), np.float32), np.zeros((R, S), np.int32),
             np.nonzero(hit)[0], rng.uniform(size=(R, S, 7)).astype(np.float32))
    gc, ga = port.instance_evaluate_model(dense, R, patch_scale=0.09, density_scale=10.0, composite_bkgd=True)
    wc, wa = O.instance_evaluate_model(om, dense, R, patch_scale=0.09, density_scale=10.0, composite_bkgd=True)
    np.testing.assert_allclose(ga, wa, atol=2e-5)
    np.testing.assert_allclose(gc, wc, atol=2e-5)
    assert (wa > 0.05).any()

"""Generates tests/golden/hotpath_small.npz from the oracle (NumPy, defined op order).

The reference cannot run in this image (TensorFlow 2.4.1 not installable) and ships no vectors of
its own, so these are ORACLE-generated regression vectors, not reference outputs: they freeze the
oracle's behaviour (guarding it against drift) and give the GPU tests a committed input/output
set.  Re-run:  python tests/golden/make_golden.py
"""

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import nerftex_oracle as O  # noqa: E402


def main():
    rng = np.random.default_rng(20261106)
    R, S, N = 48, 64, 128
    t = np.stack([rng.uniform(2, 3, R), rng.uniform(4, 6, R)], -1).astype(np.float32)
    z_rand = rng.uniform(size=(R, S)).astype(np.float32)
    z = O.stratified_z(t, S, True, z_rand)
    color = (rng.normal(size=(R, S, 3)) * 2).astype(np.float32)
    sigma = (rng.normal(size=(R, S)) * 20).astype(np.float32)
    rays_d = rng.normal(size=(R, 3)).astype(np.float32)
    cmap, amap, w, depth = O.composite(color, sigma, z, rays_d, True, (1., .5, .25))
    u = rng.uniform(size=(R, N)).astype(np.float32)
    z_det, zs_det, i_det = O.hierarchical_z(z, w, N, True, return_inds=True)
    z_rnd, zs_rnd, i_rnd = O.hierarchical_z(z, w, N, False, u, return_inds=True)
    npar = (1, 4)
    om = O.ParamNerfOracle(n_parameters=npar, seed=7)
    om.randomize_biases(8, 0.05)
    n = 96
    pos = rng.uniform(-1.4, 1.5, (n, 3)).astype(np.float32)
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=-1, keepdims=True)
    params = rng.uniform(0, 1, (n, 5)).astype(np.float32)
    mc, ma = om.forward(pos, dirs, params)
    out = dict(t=t, z_rand=z_rand, z=z, color=color, sigma=sigma, rays_d=rays_d, cmap=cmap, amap=amap, weights=w,
               depth=depth, u=u, z_det=z_det, zs_det=zs_det, i_det=i_det, z_rnd=z_rnd, zs_rnd=zs_rnd, i_rnd=i_rnd,
               pos=pos, dirs=dirs, params=params, mlp_color=mc, mlp_alpha=ma, mlp_seed=np.int64(7), bias_seed=np.int64(8))
    np.savez_compressed(Path(__file__).with_name("hotpath_small.npz"), **out)
    print("wrote", Path(__file__).with_name("hotpath_small.npz"))


if __name__ == "__main__":
    main()

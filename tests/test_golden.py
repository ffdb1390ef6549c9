"""Committed vectors (tests/golden/hotpath_small.npz, made by tests/golden/make_golden.py from the
oracle — the reference itself cannot run here): the oracle must keep reproducing them bit for bit
(CPU), and the CUDA kernels must match them at the stated bars (GPU)."""

from pathlib import Path

import numpy as np
import pytest

from oracle import nerftex_oracle as O
from oracle import cbind as OC

G = np.load(Path(__file__).parent / "golden" / "hotpath_small.npz")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_oracle_reproduces_golden_vectors():
    z = O.stratified_z(G["t"], 64, True, G["z_rand"])
    assert np.array_equal(bits(z), bits(G["z"]))
    cmap, amap, w, depth = O.composite(G["color"], G["sigma"], G["z"], G["rays_d"], True, (1., .5, .25))
    for a, b in ((cmap, G["cmap"]), (amap, G["amap"]), (w, G["weights"]), (depth, G["depth"])):
        assert np.array_equal(bits(a), bits(b))
    zd, zsd, idd = O.hierarchical_z(G["z"], G["weights"], 128, True, return_inds=True)
    zr, zsr, ir = O.hierarchical_z(G["z"], G["weights"], 128, False, G["u"], return_inds=True)
    assert np.array_equal(idd, G["i_det"]) and np.array_equal(ir, G["i_rnd"])
    assert np.array_equal(bits(zd), bits(G["z_det"])) and np.array_equal(bits(zr), bits(G["z_rnd"]))
    # the C restatement agrees with the same vectors
    cz, czs, ci = OC.hierarchical_z(G["z"], G["weights"], 128, False, G["u"])
    assert np.array_equal(ci, G["i_rnd"]) and np.array_equal(bits(cz), bits(G["z_rnd"]))
    om = O.ParamNerfOracle(n_parameters=(1, 4), seed=int(G["mlp_seed"]))
    om.randomize_biases(int(G["bias_seed"]), 0.05)
    mc, ma = om.forward(G["pos"], G["dirs"], G["params"])
    np.testing.assert_allclose(mc, G["mlp_color"], rtol=1e-5, atol=1e-6)      # BLAS order may differ across hosts
    np.testing.assert_allclose(ma, G["mlp_alpha"], rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
def test_kernels_match_golden_vectors():
    import torch
    from nerf_tex_b200 import ops
    from nerf_tex_b200.model import FusedNerfTexMLP
    cu = lambda k: torch.as_tensor(G[k]).cuda()
    z = ops.stratified_z(cu("t"), 64, True, cu("z_rand"))
    assert np.array_equal(bits(z.cpu().numpy()), bits(G["z"]))
    zd, zsd, idd = ops.hierarchical_z(cu("z"), cu("weights"), 128, True, return_aux=True)
    zr, zsr, ir = ops.hierarchical_z(cu("z"), cu("weights"), 128, False, cu("u"), return_aux=True)
    assert np.array_equal(idd.cpu().numpy(), G["i_det"]) and np.array_equal(ir.cpu().numpy(), G["i_rnd"])
    assert np.array_equal(bits(zd.cpu().numpy()), bits(G["z_det"])) and np.array_equal(bits(zr.cpu().numpy()), bits(G["z_rnd"]))
    c, a, w, d = ops.composite_forward(None, cu("z"), cu("rays_d"), color=cu("color"), sigma=cu("sigma"),
                                       composite_bkgd=True, bkgd_color=(1., .5, .25), want_depth=True)
    np.testing.assert_allclose(w.cpu().numpy(), G["weights"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(c.cpu().numpy(), G["cmap"], rtol=1e-5, atol=4e-6)
    np.testing.assert_allclose(a.cpu().numpy(), G["amap"], rtol=1e-5, atol=1e-6)
    om = O.ParamNerfOracle(n_parameters=(1, 4), seed=int(G["mlp_seed"]))
    om.randomize_biases(int(G["bias_seed"]), 0.05)
    gm = FusedNerfTexMLP(n_geo=1, n_app=4)
    gm.load_named_weights(om.named_weights())
    col, alp = gm((cu("pos"), cu("dirs"), cu("params")))
    sig = lambda x: 1 / (1 + np.exp(-x))
    assert np.abs(sig(col.cpu().numpy()) - sig(G["mlp_color"])).max() < 4e-3
    assert np.abs(alp.cpu().numpy() - G["mlp_alpha"]).max() < 2e-2

"""The oracle against vectors the REFERENCE'S OWN source produced (tests/golden/ref_python_small.npz, written by
tests/golden/make_ref_python.py in the container that holds /root/reference: network/*.py over the NumPy stand-in for
TensorFlow, the instanced case over the reference's own C++ instancer).  CPU only; runs anywhere."""

from pathlib import Path

import numpy as np

from oracle import instancer_oracle as IO
from oracle import nerftex_oracle as O

F = np.float32
G = np.load(Path(__file__).resolve().parent / "golden" / "ref_python_small.npz")


def oracle_model(npar, seed, **kw):
    om = O.ParamNerfOracle(n_parameters=npar, seed=seed, **kw)
    om.randomize_biases(seed + 1, 0.2)
    return om


def test_model_forward():
    om = oracle_model((1, 6), 101)
    c, a = om.forward(G["a_pos"], G["a_dirs"], G["a_params"], dtype=np.float64)
    s = max(np.abs(G["a_color"]).max(), np.abs(G["a_alpha"]).max())
    assert np.abs(c - G["a_color"]).max() <= 2e-5 * s and np.abs(a - G["a_alpha"]).max() <= 2e-5 * s


def test_renderer_call_coarse_fine():
    omc, omf = oracle_model((2, 3), 111), oracle_model((2, 3), 113)
    want = O.renderer_call(omc, omf, G["b_o"][None], G["b_d"][None], G["b_t"][None], G["b_params"], G["b_cone"][None], composite_bkgd=True,
                           bkgd_color=(1.0, 0.5, 0.25), render_chunk=20, z_rand=G["b_z_rand"], n_samples=32, n_importance=64,
                           perturb=True, blur_idx=0, net_chunk=700)
    for k, tol in (("color_pred_coarse", 2e-5), ("alpha_pred_coarse", 2e-5), ("color_pred", 5e-3), ("alpha_pred", 5e-3)):
        assert np.abs(want[k][0] - G["b_" + k]).max() <= tol, k
    miss = G["b_t"][:, 0] == np.inf
    assert miss.any() and np.array_equal(G["b_color_pred"][miss], np.tile(np.asarray([[1.0, 0.5, 0.25]], F), (int(miss.sum()), 1)))


def records():
    return (G["c_rec_dirs"], G["c_rec_pts"], G["c_rec_t"], G["c_rec_dist"], G["c_rec_color"], G["c_rec_density"], G["c_rec_weight"],
            G["c_rec_id"], np.nonzero(G["c_rec_hit"])[0], G["c_rec_params"])


def test_instance_renderer_on_the_reference_instancers_records():
    om = oracle_model((1, 4), 131)
    hit = G["c_t"][:, 0] != np.inf
    c, a = O.instance_evaluate_model(om, records(), int(hit.sum()), cone_scale=G["c_cone"][hit], blur_idx=1, patch_scale=0.2, density_scale=3.0,
                                     composite_bkgd=True)
    assert np.abs(c - G["c_color_pred"][hit]).max() <= 3e-5 and np.abs(a - G["c_alpha_pred"][hit]).max() <= 3e-5
    assert not G["c_alpha_pred"][~hit].any() and (G["c_color_pred"][~hit] == 1).all()           # culled rays: background
    assert (a > 0.3).any()


def test_instancer_oracle_reproduces_the_reference_instancers_records():
    hit = G["c_t"][:, 0] != np.inf
    o, d = G["c_o"][hit], G["c_d"][hit]
    prm = np.repeat(G["c_params"], o.shape[0], 0)
    ora = IO.get_model_input(o, d, prm, 48, 0.01, G["c_T"], [-1.0, -1.0, -0.1], [1.0, 1.0, 1.8], light_dir_idx=2, method="nearest",
                             ray_u=G["c_ray_u"])
    assert np.array_equal(ora["hit"], G["c_rec_hit"])
    n_ref = (G["c_rec_dist"] > 0).sum(1)
    assert (ora["n_samples"] == n_ref).mean() >= 0.99
    ptr = np.concatenate([[0], np.cumsum(ora["n_samples"])])
    for i in np.nonzero((ora["n_samples"] == n_ref) & (n_ref > 0))[0]:
        a, b = ptr[i], ptr[i + 1]
        assert np.abs(G["c_rec_t"][i, :b - a] - ora["t"][a:b]).max() <= 2e-5
        same = G["c_rec_id"][i, :b - a] == ora["instance_id"][a:b]
        assert same.mean() > 0.9 and np.abs(G["c_rec_pts"][i, :b - a] - ora["pts"][a:b])[same].max() <= 2e-4


def test_proxy_rays():
    o, d, t, cone = O.proxy_rays(G["d_loc"], 13, 17, 21.0, G["d_c2w"], [-1.0, -1.2, -0.5], [1.0, 0.8, 0.7])
    assert np.array_equal(o, G["d_o"]) and np.array_equal(d, G["d_d"]) and np.array_equal(t, G["d_t"])
    assert np.abs(cone - G["d_cone"]).max() <= 4e-9


def test_mip_renderer():
    om = oracle_model((2, 2), 141, n_pos=6)
    out, _ = O.mip_render_rays(om, G["e_o"], G["e_d"], G["e_t"], np.repeat(G["e_params"], 40, 0), G["e_cone"], 24, 1, perturb=True,
                               z_rand=G["e_z_rand"], composite_bkgd=True)
    assert np.abs(out["color_pred"] - G["e_color_pred"]).max() <= 2e-5 and np.abs(out["alpha_pred"] - G["e_alpha_pred"]).max() <= 2e-5


def test_training_renderer_with_blur_and_noise():
    om = oracle_model((2, 3), 151)
    out = O.render_rays(om, None, G["f_o"], G["f_d"], G["f_t"], np.repeat(G["f_params"], 24, 0), G["f_cone"], 256, 0, True,
                        z_rand=G["f_z_rand"], noise=G["f_noise"], blur_idx=0, composite_bkgd=True)
    assert np.abs(out["color_pred"] - G["f_color_pred"]).max() <= 3e-5 and np.abs(out["alpha_pred"] - G["f_alpha_pred"]).max() <= 3e-5

"""Writes tests/golden/ref_python_small.npz: inputs and the outputs the REFERENCE'S OWN source produced for them here —
network/*.py imported from /root/reference and run over the NumPy stand-in for TensorFlow (oracle/tf_shim), the
instanced case over the reference's own C++ instancer (oracle/ref_instancer.py).  Run in the container that has
/root/reference:   python tests/golden/make_ref_python.py
The fixtures let machines without the reference tree (the GPU box) hold the oracle and the CUDA path to the same
vectors (tests/test_golden_ref.py, tests/test_gpu_golden_ref.py).  Model weights are not stored: they are
oracle.ParamNerfOracle(seed=...) + randomize_biases(seed + 1, 0.2), reproducible from the seeds below."""

import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent))
sys.path.insert(0, str(HERE.parent.parent))
import ref_python as RP  # noqa: E402
from oracle import nerftex_oracle as O  # noqa: E402

F = np.float32


def oracle_model(npar, seed, **kw):
    om = O.ParamNerfOracle(n_parameters=npar, seed=seed, **kw)
    om.randomize_biases(seed + 1, 0.2)
    return om


def main():
    tf, renderer, model, loss, ray_sampler, proxy, util = RP.load()
    out = {}
    rng = np.random.default_rng(100)

    # ---- (a) ParamNerf [1,6] forward
    om = oracle_model((1, 6), 101)
    km = RP.build_param_nerf(model, util, om)
    pos, dirs = rng.uniform(-1, 1, (96, 3)).astype(F), rng.normal(size=(96, 3)).astype(F)
    prm = rng.uniform(0, 1, (96, 7)).astype(F)
    c, a = km((tf.constant(pos), tf.constant(dirs), tf.constant(prm)), training=False)
    out.update(a_pos=pos, a_dirs=dirs, a_params=prm, a_color=c.numpy(), a_alpha=a.numpy())

    # ---- (b) Renderer.__call__, coarse + fine [2,3], perturb (deterministic resampling), blur, background, missed rays
    omc, omf = oracle_model((2, 3), 111), oracle_model((2, 3), 113)
    kc, kf = RP.build_param_nerf(model, util, omc), RP.build_param_nerf(model, util, omf, name="model_fine")
    B, R, S, N = 1, 48, 32, 64
    o = np.tile(np.asarray([[0.0, 0.3, 5.0]], F), (R, 1))
    d = rng.uniform(-1, 1, (R, 3)).astype(F) - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(F)
    t = np.stack([rng.uniform(3.0, 4.0, R), rng.uniform(5.5, 7.0, R)], -1).astype(F)
    t[::7] = np.inf
    cone = rng.uniform(0.001, 0.01, (R, 1)).astype(F)
    prm_b = rng.uniform(0.2, 1.0, (1, 5)).astype(F)
    r = renderer.Renderer(model=kc, model_fine=kf, n_samples=S, n_importance=N, perturb=True, blur_idx=0, render_chunk=20, net_chunk=700)
    tf.random.seed(120)
    got = r(tf.constant(o[None]), tf.constant(d[None]), tf.constant(t[None]), tf.constant(prm_b), tf.constant(cone[None]),
            composite_bkgd=True, bkgd_color=[1.0, 0.5, 0.25], training=True)
    hit = t[:, 0] != np.inf
    z_rand = np.zeros((R, S), F)
    z_rand[hit] = np.concatenate([a_ for _, a_ in tf.random.log], 0)
    out.update(b_o=o, b_d=d, b_t=t, b_cone=cone, b_params=prm_b, b_z_rand=z_rand,
               **{"b_" + k: v.numpy()[0] for k, v in got.items()})

    # ---- (c) InstanceRenderer [1,4] over the reference's C++ instancer, explicit patches (no files), 'nearest', light
    import test_ref_instancer as TI
    Tm = TI.patches(60, 1)
    omi = oracle_model((1, 4), 131)
    ki = RP.build_param_nerf(model, util, omi)
    Ri, Si, step = 96, 48, 0.01
    oi, di = TI.camera_rays(Ri, 2)
    ti = np.tile(np.asarray([[0.1, 9.0]], F), (Ri, 1))
    ti[::11, 0] = np.inf
    conei = rng.uniform(0.001, 0.01, (Ri, 1)).astype(F)
    prm_c = np.asarray([[0.5, 0.7, 0.3, 0.2, 1.0]], F)
    icfg = util.EasyDict(dict(module="instancer.instancer.Instancer", b_0=TI.B0, b_1=TI.B1, textures=["", "", "light"],
                              transformations=Tm.tolist(), instance_sampling_method="nearest", patch_scale=0.2))
    ri = renderer.InstanceRenderer(instancer_config=icfg, model=ki, n_samples=Si, step_size=step, density_scale=3.0, blur_idx=1)
    goti = ri(tf.constant(oi[None]), tf.constant(di[None]), tf.constant(ti[None]), tf.constant(prm_c), tf.constant(conei[None]),
              composite_bkgd=True, training=False)
    rec = ri.instancer.last                                          # the C++ instancer's records of the one chunk
    hiti = np.nonzero(ti[:, 0] != np.inf)[0]
    drew = rec["dist"][:, 0] > 0
    u = np.zeros(hiti.shape[0], F)
    u[drew] = TI.mt_stream(0, int(drew.sum()))
    out.update(c_T=Tm, c_o=oi, c_d=di, c_t=ti, c_cone=conei, c_params=prm_c, c_ray_u=u,
               c_rec_dirs=rec["rays_d"], c_rec_pts=rec["pts"], c_rec_t=rec["t"], c_rec_dist=rec["dist"], c_rec_color=rec["color"],
               c_rec_density=rec["density"], c_rec_weight=rec["density_weight"], c_rec_id=rec["instance_id"], c_rec_hit=rec["hit"],
               c_rec_params=rec["parameters"], c_color_pred=goti["color_pred"].numpy()[0], c_alpha_pred=goti["alpha_pred"].numpy()[0])

    # ---- (d) rays: rays_from_camera + Proxy(AABB)
    H, W, focal = 13, 17, 21.0
    ii, jj = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    loc = np.stack([ii.reshape(-1), jj.reshape(-1)], -1).astype(F)
    c2w = np.asarray([[np.cos(.4), 0, np.sin(.4), 2.0], [0, 1, 0, 0.3], [-np.sin(.4), 0, np.cos(.4), 4.0], [0, 0, 0, 1]], F)
    po, pd, pt, pc = ray_sampler.Proxy(H, W, focal, proxy.AABB([-1.0, -1.2, -0.5], [1.0, 0.8, 0.7]))(tf.constant(loc), tf.constant(c2w))
    out.update(d_loc=loc, d_c2w=c2w, d_o=po.numpy(), d_d=pd.numpy(), d_t=pt.numpy(), d_cone=pc.numpy())

    # ---- (e) MipRenderer [2,2] (n_pos = 6, IntegratedPositionalEncoding), perturb, blur parameter 1, background
    ome = oracle_model((2, 2), 141, n_pos=6)
    ipe = util.EasyDict({"module": "network.layer.IntegratedPositionalEncoding", "n_freq_bands": 10})
    ke = model.ParamNerf(ipe, RP.ff(util, 4), RP.ff(util, 4), n_parameters=[2, 2], n_pos=6)["model"]
    RP.set_weights(ke, ome)
    Re, Se = 40, 24
    oe = np.tile(np.asarray([[0.0, 0.3, 5.0]], F), (Re, 1))
    de = rng.uniform(-1, 1, (Re, 3)).astype(F) - oe
    de = (de / np.linalg.norm(de, axis=1, keepdims=True)).astype(F)
    te = np.stack([rng.uniform(3.0, 4.0, Re), rng.uniform(5.5, 7.0, Re)], -1).astype(F)
    conee = rng.uniform(0.001, 0.01, (Re, 1)).astype(F)
    prm_e = rng.uniform(0.2, 1.0, (1, 5)).astype(F)
    re_ = renderer.MipRenderer(model=ke, n_samples=Se, perturb=True, blur_idx=1)
    tf.random.seed(142)
    gote = re_(tf.constant(oe[None]), tf.constant(de[None]), tf.constant(te[None]), tf.constant(prm_e), tf.constant(conee[None]),
               composite_bkgd=True, training=True)
    out.update(e_o=oe, e_d=de, e_t=te, e_cone=conee, e_params=prm_e, e_z_rand=tf.random.log[0][1],
               e_color_pred=gote["color_pred"].numpy()[0], e_alpha_pred=gote["alpha_pred"].numpy()[0])

    # ---- (f) the grass_filtered training renderer (config values: S = 256, blur_idx 0, raw_noise_std 0.1), single net [2,3]
    omg = oracle_model((2, 3), 151)
    kg = RP.build_param_nerf(model, util, omg)
    Rg, Sg = 24, 256
    og = np.tile(np.asarray([[0.0, 0.3, 5.0]], F), (Rg, 1))
    dg = rng.uniform(-1, 1, (Rg, 3)).astype(F) - og
    dg = (dg / np.linalg.norm(dg, axis=1, keepdims=True)).astype(F)
    tg = np.stack([rng.uniform(3.0, 4.0, Rg), rng.uniform(5.5, 7.0, Rg)], -1).astype(F)
    coneg = rng.uniform(0.001, 0.01, (Rg, 1)).astype(F)
    prm_g = rng.uniform(0.2, 1.0, (1, 5)).astype(F)
    rg = renderer.Renderer(model=kg, n_samples=Sg, perturb=True, raw_noise_std=0.1, blur_idx=0, render_chunk=32768, net_chunk=65536)
    tf.random.seed(152)
    gotg = rg(tf.constant(og[None]), tf.constant(dg[None]), tf.constant(tg[None]), tf.constant(prm_g), tf.constant(coneg[None]),
              composite_bkgd=True, training=True)
    out.update(f_o=og, f_d=dg, f_t=tg, f_cone=coneg, f_params=prm_g, f_z_rand=tf.random.log[0][1], f_noise=tf.random.log[1][1] * F(0.1),
               f_color_pred=gotg["color_pred"].numpy()[0], f_alpha_pred=gotg["alpha_pred"].numpy()[0])

    path = HERE / "ref_python_small.npz"
    np.savez_compressed(path, **out)
    print(path, path.stat().st_size, "bytes;", ", ".join(f"{k}{tuple(v.shape)}" for k, v in sorted(out.items())))


if __name__ == "__main__":
    main()

"""smoke(): one small pass of the whole hot path (K3 -> K1 -> K2, coarse + fine) on cuda:0,
checked against the CPU oracle.  Called by __graft_entry__.smoke(); the oracle import is the
checker, not the thing run."""

from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch


def run() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError("smoke() needs a CUDA device: the hot path has no CPU fallback")
    root = Path(__file__).resolve().parent.parent
    if str(root) not in sys.path:
        sys.path.insert(0, str(root))
    from oracle import nerftex_oracle as O
    from nerf_tex_b200 import util

    torch.cuda.set_device(0)
    npar = (1, 4)
    mcfg = {"module": "network.model.CoarseFine",
            "model_config": {"module": "network.model.ParamNerf",
                             "pos_embedding": {"module": "network.model.FourierFeatures", "n_freq_bands": 10},
                             "dir_embedding": {"module": "network.model.FourierFeatures", "n_freq_bands": 4},
                             "param_embedding": {"module": "network.model.FourierFeatures", "n_freq_bands": 4},
                             "n_parameters": list(npar)}}
    models = util.instantiate(mcfg)
    oracles = {}
    for i, (name, m) in enumerate(models.items()):
        om = O.ParamNerfOracle(n_parameters=npar, seed=1 + i)
        m.load_named_weights(om.named_weights())
        oracles[name] = om
    rcfg = {"module": "network.renderer.Renderer", "n_samples": 64, "n_importance": 128, "perturb": True}
    rcfg.update(models)
    renderer = util.instantiate(rcfg)

    rng = np.random.default_rng(0)
    R = 256
    rays_o = np.tile(np.array([[0., 0., 5.]], np.float32), (R, 1))
    rays_d = rng.uniform(-1, 1, (R, 3)).astype(np.float32) - rays_o
    rays_d /= np.linalg.norm(rays_d, axis=-1, keepdims=True)
    t = O.aabb_intersect([-1.4, -1.1, -.3], [1.4, 1.5, 3], rays_o, rays_d)
    hit = np.isfinite(t[:, 0])
    rays_o, rays_d, t = rays_o[hit], rays_d[hit], t[hit]
    R = rays_o.shape[0]
    params = np.repeat(rng.uniform(0, 1, (1, 5)).astype(np.float32), R, 0)
    cone = np.full((R, 1), 2e-3, np.float32)
    z_rand = rng.uniform(size=(R, 64)).astype(np.float32)
    cu = lambda a: torch.as_tensor(a).cuda()
    out, aux = renderer.render_rays(cu(rays_o), cu(rays_d), cu(t), cu(params), cu(cone), False, (1., 1., 1.), False,
                                    z_rand=cu(z_rand), return_aux=True)
    torch.cuda.synchronize()

    # stage-wise check against the oracle (see tests/test_gpu_renderer.py::test_coarse_fine_stagewise)
    dn = rays_d / np.linalg.norm(rays_d, axis=-1, keepdims=True)
    z = O.stratified_z(t, 64, True, z_rand)
    col, alp = O.evaluate_model(oracles["model"], O.ray_points(rays_o, rays_d, z), dn, params, cone, z)
    wc, wa, _, _ = O.composite(col, alp, z, rays_d)
    e_coarse = float(np.abs(out["color_pred_coarse"].cpu().numpy() - wc).max())
    z_all, _ = O.hierarchical_z(z, aux["weights_coarse"].cpu().numpy(), 128, True)
    gz = aux["z_fine"].cpu().numpy()
    if not np.array_equal(gz.view(np.uint32), z_all.view(np.uint32)):
        raise AssertionError("smoke: hierarchical depths are not bit-exact against the oracle")
    colf, alpf = O.evaluate_model(oracles["model_fine"], O.ray_points(rays_o, rays_d, gz), dn, params, cone, gz)
    fc, fa, _, _ = O.composite(colf, alpf, gz, rays_d)
    e_fine = float(np.abs(out["color_pred"].cpu().numpy() - fc).max())
    e_alpha = float(np.abs(out["alpha_pred"].cpu().numpy() - fa).max())
    if not (e_coarse < 2e-3 and e_fine < 2e-3 and e_alpha < 2e-3):
        raise AssertionError(f"smoke: rgb/alpha differ from the oracle: coarse {e_coarse}, fine {e_fine}, alpha {e_alpha}")
    print(f"smoke ok: {R} rays x (64 + 192) samples; sampling bit-exact; max |rgb - oracle| coarse {e_coarse:.2e}, "
          f"fine {e_fine:.2e}")

"""The reference's shipped configs as fixtures: every configs/config_*_{train,render}.py hands
its `model_config` / `renderer_config` / `loss_config` sub-dicts UNCHANGED to util.instantiate (the reference's own
mechanism, util/util.py:44-54; main.py:56, train.py:31-46, render.py:21-25) and the result is held to the oracle.

The dicts come from tests/golden/reference_configs.json (extracted verbatim by tests/golden/make_reference_configs.py;
the GPU box has no /root/reference).  Only one thing is substituted: the instancer class of the render configs — their
`instancer_config` names meshes and textures that are git-LFS stubs in the reference tree, so the module path is pointed
at SyntheticInstancer, which takes the very same kwargs (instancer/instancer.pyx:9) and seeded records.

Bars: rendered colour / alpha <= 2e-3 max abs and >= 50 dB against the fp32 oracle (north_star); training gradients
cos >= 0.999 / rel <= 2e-2 against float64 autograd of the oracle graph; one Adam step against the oracle's Keras form."""

import copy
import json
import math
import os
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import nerftex_oracle as O

FIXTURE = Path(__file__).resolve().parent / "golden" / "reference_configs.json"
CONFIGS = json.loads(FIXTURE.read_text())
RENDER = sorted(k for k, c in CONFIGS.items() if c["module"] == "network.render.Render")
TRAIN = sorted(k for k, c in CONFIGS.items() if c["module"] == "network.train.Train")


# ----------------------------------------------------------------------------------------------- CPU side
def test_fixture_lists_every_shipped_config():
    assert len(RENDER) == 4 and len(TRAIN) == 5
    for name, c in CONFIGS.items():
        assert {"module", "model_config", "renderer_config"} <= set(c), name


@pytest.mark.skipif(not os.path.isdir("/root/reference/configs"), reason="reference tree not present on this box")
def test_fixture_is_what_the_reference_ships():
    import importlib.util
    spec = importlib.util.spec_from_file_location("_mk", FIXTURE.parent / "make_reference_configs.py")
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    assert json.loads(json.dumps(mk.extract("/root/reference"))) == CONFIGS


def test_every_module_path_in_the_configs_resolves():
    """util.get_attr_from_path maps network.* / instancer.instancer.* onto this package (no GPU needed to import)."""
    from nerf_tex_b200 import util

    def walk(d):
        for k, v in d.items():
            if isinstance(v, dict):
                yield from walk(v)
            elif k == "module" or (k.endswith("loss_fn") and isinstance(v, str)):
                yield v
    for name, c in CONFIGS.items():
        for sub in ("model_config", "renderer_config", "loss_config"):
            for path in walk(c.get(sub, {})):
                assert callable(util.get_attr_from_path(path)), (name, path)


# ----------------------------------------------------------------------------------------------- GPU side
def dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()


def psnr(a, b):
    return -10 * math.log10(float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)) + 1e-30)


def build(name, seed):
    """model + renderer exactly as Render / Train build them (render.py:21-25, train.py:30-43)."""
    from nerf_tex_b200 import util
    c = copy.deepcopy(CONFIGS[name])
    model_config = util.EasyDict(c["model_config"])
    models = util.instantiate(model_config)
    oracles = {}
    for i, (mname, m) in enumerate(models.items()):
        om = O.ParamNerfOracle(n_parameters=tuple(c["model_config"]["n_parameters"]), seed=seed + i)
        om.randomize_biases(90 + i, 0.05)
        m.load_named_weights(om.named_weights())
        oracles[mname] = om
    renderer_config = util.EasyDict(c["renderer_config"])
    if "instancer_config" in renderer_config:                       # the one substitution (see the module docstring)
        assert renderer_config.instancer_config.module == "instancer.instancer.Instancer"
        renderer_config.instancer_config.module = "nerf_tex_b200.instancer.SyntheticInstancer"
    renderer_config.update(models)
    return c, models, oracles, util.instantiate(renderer_config)


@pytest.mark.gpu
@pytest.mark.parametrize("name", RENDER)
def test_render_config_matches_oracle(name):
    from nerf_tex_b200.instancer import SyntheticInstancer
    c, models, oracles, r = build(name, seed=3)
    rc = c["renderer_config"]
    ic = rc["instancer_config"]
    assert r.n_samples == rc["n_samples"] and r.step_size == rc["step_size"] and r.patch_scale == ic["patch_scale"]
    assert r.blur_idx == rc.get("blur_idx") and r.density_reweighting == rc.get("density_reweighting", True)
    p = sum(c["model_config"]["n_parameters"])
    R = 256
    rng = np.random.default_rng(11)
    rays_o = rng.normal(size=(1, R, 3)).astype(np.float32)
    rays_d = rng.normal(size=(1, R, 3)).astype(np.float32)
    rays_d /= np.linalg.norm(rays_d, axis=-1, keepdims=True)
    t = np.tile(np.array([[[1., 3.]]], np.float32), (1, R, 1))
    params = np.concatenate([rng.uniform(0.2, 1, (1, p - 3)), [[0, 0, 1.]]], -1).astype(np.float32)
    cone = rng.uniform(0.01, 0.05, (1, R, 1)).astype(np.float32)
    r.density_scale = 4.0                                           # untrained weights: keep the opacity non-trivial
    got = r(dev(rays_o), dev(rays_d), dev(t), dev(params), dev(cone), composite_bkgd=True, training=False)
    replay = SyntheticInstancer(**{k: v for k, v in ic.items() if k != "module"})   # same seed, same records
    dense = replay.get_model_input(dev(rays_o[0]), dev(rays_d[0]), dev(np.repeat(params, R, 0)), rc["n_samples"], rc["step_size"])
    dense = [x.cpu().numpy() for x in dense]
    wc, wa = O.instance_evaluate_model(oracles["model"], dense, R, cone_scale=cone[0], blur_idx=rc.get("blur_idx"),
                                       patch_scale=ic["patch_scale"], density_scale=4.0, composite_bkgd=True)
    gc, ga = got["color_pred"].cpu().numpy()[0], got["alpha_pred"].cpu().numpy()[0]
    assert (wa > 0.05).mean() > 0.3
    assert np.abs(ga - wa).max() < 2e-3 and np.abs(gc - wc).max() < 2e-3, (np.abs(ga - wa).max(), np.abs(gc - wc).max())
    assert psnr(gc, wc) > 50


@pytest.mark.gpu
@pytest.mark.parametrize("name", TRAIN)
def test_train_config_forward_gradients_and_step(name):
    import torch_ref as T
    from nerf_tex_b200 import util
    from nerf_tex_b200.train import Trainer
    from test_gpu_train import check_grads, t64
    c, models, oracles, r = build(name, seed=5)
    rc = c["renderer_config"]
    gm, om = models["model"], oracles["model"]
    S, blur, noise_std = rc["n_samples"], rc.get("blur_idx"), rc.get("raw_noise_std", 0)
    assert r.n_samples == S and r.perturb is True and r.blur_idx == blur and r.raw_noise_std == noise_std
    p = sum(c["model_config"]["n_parameters"])
    R = 256                                                        # (the C3 shape at 512 rays: test_gpu_config_shapes.py)
    rng = np.random.default_rng(17)
    rays_o = np.tile(np.array([[0., 0.5, 5.]], np.float32), (R, 1))
    rays_d = rng.uniform(-1, 1, (R, 3)).astype(np.float32) - rays_o
    rays_d /= np.linalg.norm(rays_d, axis=-1, keepdims=True)
    t = np.stack([rng.uniform(3, 3.5, R), rng.uniform(5.5, 6.5, R)], -1).astype(np.float32)
    params = rng.uniform(0, 1, (1, p)).astype(np.float32)
    cone = rng.uniform(0.05, 0.2, (R, 1)).astype(np.float32)
    ct = rng.uniform(0, 1, (1, R, 3)).astype(np.float32)
    at = np.where(rng.uniform(size=(1, R)) < 0.3, 0, rng.uniform(0, 1, (1, R))).astype(np.float32)
    z_rand = rng.uniform(size=(R, S)).astype(np.float32)
    noise = (rng.normal(size=(R, S)) * noise_std).astype(np.float32) if noise_std > 0 else None
    draws = {"z_rand": dev(z_rand)}
    if noise is not None:
        draws["noise"] = dev(noise)

    # inference pass against the fp32 oracle (renderer.py:47-213)
    with torch.no_grad():
        got = r(dev(rays_o[None]), dev(rays_d[None]), dev(t[None]), dev(params), dev(cone[None]), composite_bkgd=True,
                training=False, **draws)
    want = O.renderer_call(om, None, rays_o[None], rays_d[None], t[None], params, cone[None], composite_bkgd=True,
                           z_rand=z_rand[None], n_samples=S, perturb=True, blur_idx=blur, noise=noise)
    for k in want:
        g, w = got[k].cpu().numpy(), want[k]
        assert np.abs(g - w).max() < 2e-3 and psnr(g, w) > 50, (k, np.abs(g - w).max(), psnr(g, w))

    # training step: loss and optimiser built from the config (train.py:44-52)
    loss_fn = util.instantiate(c["loss_config"])
    tr = Trainer(models, r, loss_fn, lrate=c["lrate"], lrate_decay=c["lrate_decay"])
    data = {"rays_o": dev(rays_o[None]), "rays_d": dev(rays_d[None]), "t": dev(t[None]), "parameters": dev(params),
            "cone_scale": dev(cone[None]), "color": dev(ct), "alpha": dev(at)}
    before = gm.master.detach().clone()
    loss = float(tr.train_step(data, composite_bkgd=True, **draws))
    P = T.make_params(om)
    z = t64(O.stratified_z(t, S, True, z_rand))
    C, A, _ = T.network_pass(om, P, t64(rays_o), t64(rays_d), z, t64(np.repeat(params, R, 0)), t64(cone), blur,
                             composite_bkgd=True, noise=None if noise is None else t64(noise))
    ref = T.alpha_loss(t64(ct[0]), t64(at[0]), C, A)
    ref.backward()
    assert abs(loss - float(ref)) < 1e-3
    g = gm.master.grad.detach().cpu()
    check_grads(g, T.flat_grads(om, P), om, name)
    stepped, _, _ = O.adam_step(before.cpu().numpy(), g.numpy(), np.zeros(g.numel(), np.float32),
                                np.zeros(g.numel(), np.float32), 1, c["lrate"])
    np.testing.assert_allclose(gm.master.detach().cpu().numpy(), stepped, rtol=1e-5, atol=1e-7)
    assert abs(tr.current_lr() - O.exponential_decay_lr(c["lrate"], c["lrate_decay"], 1)) < 1e-12


# ----------------------------------------------------------------------------------------------- the device instancer
def procedural_assets(root, ic):
    """Files with the names an instancer_config asks for, generated: a hilly plane with normals and UVs for the mesh,
    points on it for the patch origins, noisy checkerboards for the textures (the reference's assets are git-LFS stubs)."""
    from nerf_tex_b200 import ply
    from nerf_tex_b200.image import encode_png
    from test_gpu_instancer import checker_texture, plane_mesh
    rng = np.random.default_rng(21)
    mesh = plane_mesh(20, 1.5, wobble=0.12)
    os.makedirs(os.path.join(root, os.path.dirname(ic["mesh_path"])), exist_ok=True)
    ply.write_ply(os.path.join(root, ic["mesh_path"]), mesh["V"], mesh["F"], mesh["N"], mesh["UV"])
    anchors = None
    if ic.get("patch_origins_path"):
        xy = rng.uniform(-1.3, 1.3, (260, 2))
        anchors = np.concatenate([xy, 0.12 * np.sin(2.1 * xy[:, :1]) * np.cos(1.7 * xy[:, 1:])], 1).astype(np.float32)
        ply.write_ply(os.path.join(root, ic["patch_origins_path"]), anchors)
    images = {}
    for k, tex in enumerate(ic.get("textures", [])):
        if tex not in ("", "light", "point"):
            images[tex] = checker_texture(24, 3, seed=k)
            with open(os.path.join(root, tex), "wb") as f:
                f.write(encode_png(images[tex][:, :, None]))
    return mesh, anchors, images


@pytest.mark.gpu
@pytest.mark.parametrize("name", RENDER)
def test_render_config_on_the_device_instancer(name, tmp_path, monkeypatch):
    """Every shipped render config, instancer_config included, UNCHANGED through util.instantiate: mesh + anchor PLYs,
    texture PNGs, shadow rays, point / directional lights, 'nearest' / 'nearest_blend' — on generated assets of the same
    file names.  Records bit-exact against the instancer oracle, rendered colour / alpha against the oracle chain."""
    from nerf_tex_b200 import util
    from oracle import instancer_oracle as IO
    c = copy.deepcopy(CONFIGS[name])
    rc, ic = c["renderer_config"], c["renderer_config"]["instancer_config"]
    mesh, anchors, images = procedural_assets(str(tmp_path), ic)
    monkeypatch.chdir(tmp_path)                                      # the configs name their assets relative to the run directory
    models = util.instantiate(util.EasyDict(c["model_config"]))
    om = O.ParamNerfOracle(n_parameters=tuple(c["model_config"]["n_parameters"]), seed=3)
    om.randomize_biases(91, 0.05)
    models["model"].load_named_weights(om.named_weights())
    renderer_config = util.EasyDict(rc)
    renderer_config.update(models)
    r = util.instantiate(renderer_config)
    inst = r.instancer
    p = sum(c["model_config"]["n_parameters"])
    assert inst.n_parameters == p and inst.cast_shadow_rays == ic.get("cast_shadow_rays", False)
    assert inst.n_instances() == (260 if anchors is not None else len(np.unique(mesh["V"], axis=0)))
    S, step = rc["n_samples"], rc["step_size"]
    R = 96
    rng = np.random.default_rng(22)
    o = np.stack([rng.uniform(-1.3, 1.3, R), rng.uniform(-1.3, 1.3, R), rng.uniform(0.8, 1.6, R)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-.6, .6, R), rng.uniform(-.6, .6, R), np.full(R, -1.0)], 1).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    if "point" in ic["textures"]:
        light = [3.0, 1.0, 0.4, 0.8]                                 # strength, position
    else:
        light = [0.9, 0.1, 0.12]                                     # low sun
    params = np.array([list(rng.uniform(0.3, 1.0, p - len(light))) + light], np.float32)
    t = np.tile(np.array([[[0.1, 5.0]]], np.float32), (1, R, 1))
    cone = rng.uniform(0.01, 0.05, (1, R, 1)).astype(np.float32)
    r.density_scale = 3.0
    got = r(dev(o[None]), dev(d[None]), dev(t), dev(params), dev(cone), composite_bkgd=True, training=False)
    # the oracle chain on the same rays (ray names = positions in the call, frame 0)
    tex_idx, n_par, chans = [], 0, []
    for tex in ic["textures"]:
        if tex == "light":
            n_par += 3
        elif tex == "point":
            n_par += 4
        elif tex == "":
            n_par += 1
        else:
            tex_idx.append(n_par)
            chans += IO.texture_channels(images[tex])
            n_par += 1
    texture = None
    if tex_idx:
        texture = dict(UV=mesh["UV"], textures=chans, idxs=tex_idx, radius=inst.patch_max_extent,
                       min_samples=ic.get("min_texture_samples", 4), n_samples=ic.get("n_texture_samples", 512))
    pf = np.repeat(params, R, 0)
    want = IO.get_model_input(o, d, pf, S, step, inst.transformations, ic["b_0"], ic["b_1"], seed=0,
                              light_dir_idx=inst.light_dir_idx, light_strength_idx=inst.light_strength_idx,
                              method=ic["instance_sampling_method"], patch_scale=inst.patch_scale, mesh=(mesh["V"], mesh["F"]),
                              cast_shadow_rays=ic.get("cast_shadow_rays", False), min_shadow_samples=ic.get("min_shadow_samples", 4),
                              n_shadow_samples=ic.get("n_shadow_samples", 512), texture=texture)
    rec = inst.get_model_input_csr(dev(o), dev(d), dev(pf), S, step, seed=0, ray_ids=torch.arange(R, dtype=torch.int32).cuda())
    from test_gpu_instancer import check_records
    check_records(rec, want, R)
    assert np.array_equal(rec["alpha_last"].cpu().numpy(), want["alpha_last"][want["hit"]])
    from nerf_tex_b200.instancer import csr_to_dense
    dense = [x.cpu().numpy() for x in csr_to_dense(rec, dev(d), dev(pf), S)]
    wc, wa = O.instance_evaluate_model(om, dense, R, cone_scale=cone[0], blur_idx=rc.get("blur_idx"), patch_scale=inst.patch_scale,
                                       density_scale=3.0, density_reweighting=rc.get("density_reweighting", True), composite_bkgd=True)
    gc, ga = got["color_pred"].cpu().numpy()[0], got["alpha_pred"].cpu().numpy()[0]
    assert want["n_samples"].max() > 50 and (want["alpha_last"] == 1).mean() > 0.5
    assert np.abs(ga - wa).max() < 2e-3 and np.abs(gc - wc).max() < 2e-3, (np.abs(ga - wa).max(), np.abs(gc - wc).max())
    assert psnr(gc, wc) > 50

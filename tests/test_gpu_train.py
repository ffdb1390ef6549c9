"""Training path (C3/C4 of SURVEY.md section 8d): gradients of the fused forward + hand-written
compositing backward + layer-wise MLP backward against float64 torch autograd of the oracle graph.
Tolerances: relative L2 error <= 2e-2 and cosine >= 0.999 on the whole flat gradient;
per layer cosine >= 0.995 and relative error <= 0.1 (with a bf16 forward the disagreement grows
from 0.3 % at the heads to 4-9 % at trunk0, whose small gradient is a sum of cancelling terms; an
fp32 gradient chain gives the same numbers, i.e. the residual is the bf16 forward); loss within 1e-3."""

import os

import numpy as np
import pytest
import torch

from oracle import nerftex_oracle as O
import torch_ref as T

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()


def t64(a):
    return torch.tensor(np.asarray(a), dtype=torch.float64)


def setup(npar, coarse_fine, R, seed=0):
    from nerf_tex_b200.model import FusedNerfTexMLP
    rng = np.random.default_rng(seed)
    oms, gms = {}, {}
    for i, name in enumerate(["model", "model_fine"] if coarse_fine else ["model"]):
        om = O.ParamNerfOracle(n_parameters=npar, seed=1 + i)
        om.randomize_biases(20 + i, 0.05)
        gm = FusedNerfTexMLP(n_geo=npar[0], n_app=npar[1], name=name)
        gm.load_named_weights(om.named_weights())
        oms[name], gms[name] = om, gm
    rays_o = np.tile(np.array([[0., 0.5, 5.]], np.float32), (R, 1))
    rays_d = rng.uniform(-1, 1, (R, 3)).astype(np.float32) - rays_o
    rays_d /= np.linalg.norm(rays_d, axis=-1, keepdims=True)
    rays_d *= rng.uniform(0.8, 1.2, (R, 1)).astype(np.float32)          # |d| != 1 exercises renderer.py:180
    t = np.stack([rng.uniform(3, 3.5, R), rng.uniform(5.5, 6.5, R)], -1).astype(np.float32)
    params = rng.uniform(0, 1, (1, sum(npar))).astype(np.float32)
    cone = rng.uniform(0.05, 0.2, (R, 1)).astype(np.float32)
    color_true = rng.uniform(0, 1, (1, R, 3)).astype(np.float32)
    alpha_true = np.where(rng.uniform(size=(1, R)) < 0.3, 0, rng.uniform(0, 1, (1, R))).astype(np.float32)
    return oms, gms, rays_o, rays_d, t, params, cone, color_true, alpha_true


def check_grads(got, want, om, tag):
    off = 0
    for name, W, b in om.layers:
        for part, n in (("W", W.size), ("b", b.size)):
            g, w = got[off:off + n].double(), want[off:off + n]
            off += n
            if w.norm() < 1e-12:
                assert g.norm() < 1e-6, (tag, name, part)
                continue
            cos = float((g * w).sum() / (g.norm() * w.norm()))
            rel = float((g - w).norm() / w.norm())
            assert cos >= 0.995 and rel <= 0.1, (tag, name, part, cos, rel)
    assert off == got.numel()
    total_rel = float((got.double() - want).norm() / want.norm())
    total_cos = float((got.double() * want).sum() / (got.double().norm() * want.norm()))
    assert total_rel <= 2e-2 and total_cos >= 0.999, (tag, total_rel, total_cos)


@pytest.mark.parametrize("npar,blur,noise_std,exr", [((1, 4), None, 0.0, False), ((2, 3), 0, 0.1, False), ((1, 6), None, 0.0, True)])
def test_single_network_gradients(npar, blur, noise_std, exr):
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.loss import AlphaLoss
    R, S = 300, 64
    oms, gms, rays_o, rays_d, t, params, cone, ct, at = setup(npar, False, R)
    gm, om = gms["model"].requires_grad_(True), oms["model"]
    rng = np.random.default_rng(5)
    z_rand = rng.uniform(size=(R, S)).astype(np.float32)
    noise = (rng.normal(size=(R, S)) * noise_std).astype(np.float32) if noise_std > 0 else None
    r = Renderer(model=gm, n_samples=S, perturb=True, blur_idx=blur, raw_noise_std=noise_std, map_exr=exr)
    pred = r(dev(rays_o[None]), dev(rays_d[None]), dev(t[None]), dev(params), dev(cone[None]), composite_bkgd=True,
             bkgd_color=(1., 1., 1.), training=True, z_rand=dev(z_rand), noise=None if noise is None else dev(noise))
    loss = AlphaLoss("network.loss.smape", "network.loss.mse")(color_true=dev(ct), alpha_true=dev(at), **pred)
    loss.backward()
    got = gm.master.grad.detach().cpu()

    P = T.make_params(om)
    z = t64(O.stratified_z(t, S, True, z_rand))
    C, A, _ = T.network_pass(om, P, t64(rays_o), t64(rays_d), z, t64(np.repeat(params, R, 0)), t64(cone), blur,
                             composite_bkgd=True, map_exr=exr, noise=None if noise is None else t64(noise))
    ref_loss = T.alpha_loss(t64(ct[0]), t64(at[0]), C, A)
    ref_loss.backward()
    assert abs(float(loss) - float(ref_loss)) < 1e-3
    check_grads(got, T.flat_grads(om, P), om, "single")


def test_coarse_fine_gradients_and_stop_gradient():
    """CoarseFine 64 + 128 (BASELINE config 3 variant ii): both networks receive gradients; the
    fine pass is evaluated at the kernel's own resampled depths (no gradient through them)."""
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.loss import AlphaLoss
    npar, R = (2, 3), 200
    oms, gms, rays_o, rays_d, t, params, cone, ct, at = setup(npar, True, R, seed=3)
    for m in gms.values():
        m.requires_grad_(True)
    r = Renderer(model=gms["model"], model_fine=gms["model_fine"], n_samples=64, n_importance=128, perturb=True)
    z_rand = np.random.default_rng(1).uniform(size=(R, 64)).astype(np.float32)
    pf = dev(np.repeat(params, R, 0))
    out, aux = r.render_rays(dev(rays_o), dev(rays_d), dev(t), pf, dev(cone), False, (1., 1., 1.), True,
                             z_rand=dev(z_rand), return_aux=True)
    loss = AlphaLoss("network.loss.smape", "network.loss.mse")(color_true=dev(ct[0]), alpha_true=dev(at[0]), **out)
    loss.backward()
    z_c, z_f = t64(aux["z_coarse"].cpu().numpy()), t64(aux["z_fine"].cpu().numpy())
    Pc, Pf = T.make_params(oms["model"]), T.make_params(oms["model_fine"])
    args = (t64(rays_o), t64(rays_d))
    Cc, Ac, _ = T.network_pass(oms["model"], Pc, *args, z_c, t64(np.repeat(params, R, 0)))
    Cf, Af, _ = T.network_pass(oms["model_fine"], Pf, *args, z_f, t64(np.repeat(params, R, 0)))
    ref = T.alpha_loss(t64(ct[0]), t64(at[0]), Cf, Af, Cc, Ac)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 1e-3
    check_grads(gms["model"].master.grad.cpu(), T.flat_grads(oms["model"], Pc), oms["model"], "coarse")
    check_grads(gms["model_fine"].master.grad.cpu(), T.flat_grads(oms["model_fine"], Pf), oms["model_fine"], "fine")


def test_trainer_adam_step_matches_oracle_and_learns():
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.loss import AlphaLoss
    from nerf_tex_b200.train import Trainer
    npar, R = (1, 4), 256
    oms, gms, rays_o, rays_d, t, params, cone, ct, at = setup(npar, False, R, seed=7)
    r = Renderer(model=gms["model"], n_samples=64, perturb=True, seed=0)
    tr = Trainer(gms, r, AlphaLoss("network.loss.smape", "network.loss.mse"), lrate=5e-4, lrate_decay=500)
    data = {"rays_o": dev(rays_o[None]), "rays_d": dev(rays_d[None]), "t": dev(t[None]), "parameters": dev(params),
            "cone_scale": dev(cone[None]), "color": dev(ct), "alpha": dev(at)}
    before = gms["model"].master.detach().clone()
    z_rand = dev(np.random.default_rng(0).uniform(size=(R, 64)).astype(np.float32))
    l0 = float(tr.train_step(data, z_rand=z_rand))
    g = gms["model"].master.grad.detach().cpu().numpy()
    want, _, _ = O.adam_step(before.cpu().numpy(), g, np.zeros_like(g), np.zeros_like(g), 1, 5e-4)
    np.testing.assert_allclose(gms["model"].master.detach().cpu().numpy(), want, rtol=1e-5, atol=1e-7)
    assert tr.step == 1 and abs(tr.current_lr() - 5e-4 * 0.1 ** (1 / 5e5)) < 1e-12
    losses = [l0] + [float(tr.train_step(data, z_rand=z_rand)) for _ in range(25)]
    assert losses[-1] < 0.8 * losses[0], losses                    # it learns on a fixed batch


@pytest.mark.parametrize("n", [100, 512, 1300])
def test_training_forward_saves_tile_images_and_masks(n):
    """The bulk-stored tile images (the kernels' SWIZZLE_64B shared-memory operand layout) hold the
    same activations as the row-major dump, and the mask words are their sign bits."""
    from nerf_tex_b200.autograd import mlp_forward_train, untile
    from nerf_tex_b200.model import FusedNerfTexMLP
    m = FusedNerfTexMLP(n_geo=1, n_app=6, seed=5)
    g = torch.Generator(device="cuda").manual_seed(n)
    pos = torch.rand((n, 3), device="cuda", generator=g) * 2 - 1
    d = torch.randn((n, 3), device="cuda", generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    prm = torch.rand((n, 7), device="cuda", generator=g)
    rgbs, (enc_pos, enc_dir, acts_img, masks, acts, _) = mlp_forward_train(m, n_rows=n, pos=pos, dir=d, params=prm,
                                                                        rows_per_group=1, row_major_acts=True)
    ref = m.forward_rows(n_rows=n, pos=pos, dir=d, params=prm, rows_per_group=1)
    torch.cuda.synchronize()
    assert torch.equal(rgbs, ref)                                   # the training kernel computes the same outputs
    sched = m.schedule()
    for st, s in enumerate(sched.steps):
        cols = s["n_out"]
        got = untile(acts_img[st], n, cols)
        assert torch.equal(got, acts[st][:, :cols]), st
        if not s["relu"]:
            continue                                                # the linear feature layer needs no mask
        bits = (masks[st][:n].view(n, 8, 1) >> torch.arange(32, device="cuda", dtype=torch.int32).view(1, 1, 32)) & 1
        want = (acts[st][:, :cols] > 0).view(n, cols // 32, 32).to(torch.int32)
        assert torch.equal(bits[:, :cols // 32], want), st


@pytest.mark.parametrize("npar,n", [((1, 4), 300), ((1, 6), 1500), ((0, 0), 700)])
def test_native_backward_kernels_match_library_gemm_backward(npar, n):
    """ntx_mlp_backward_data / _weights against the layer-wise torch.mm backward on the same saved
    operands: both use bf16 operands with fp32 accumulation, so they agree far tighter than either
    agrees with float64."""
    from library_backward import mlp_backward
    from nerf_tex_b200.autograd import mlp_backward_native, mlp_forward_train, untile
    from nerf_tex_b200.model import FusedNerfTexMLP
    cd = 0 if npar == (0, 0) else 1
    m = FusedNerfTexMLP(n_geo=npar[0], n_app=npar[1], color_depth=cd, seed=9)
    g = torch.Generator(device="cuda").manual_seed(n)
    pos = torch.rand((n, 3), device="cuda", generator=g) * 2 - 1
    d = torch.randn((n, 3), device="cuda", generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    prm = torch.rand((n, sum(npar)), device="cuda", generator=g) if sum(npar) else None
    rgbs, saved = mlp_forward_train(m, n_rows=n, pos=pos, dir=d, params=prm, rows_per_group=1)
    d_rgbs = torch.randn((n, 4), device="cuda", generator=g) * 0.01
    want = mlp_backward(m, saved, d_rgbs)
    got, dy_img = mlp_backward_native(m, saved, d_rgbs, return_dy=True)
    torch.cuda.synchronize()
    rel = float((got - want).norm() / want.norm())
    assert rel < 5e-3, rel
    for (wo, bo, fi, fo), name in zip(m.layers, m.layer_names):
        for a, b_ in ((wo, wo + fi * fo), (bo, bo + fo)):
            w, g_ = want[a:b_], got[a:b_]
            if float(w.norm()) > 0:
                assert float((g_ - w).norm() / w.norm()) < 2e-2, (name, float((g_ - w).norm() / w.norm()))
    # rows beyond n inside the last tile carry exactly zero gradient
    tail = untile(dy_img[0], saved[2].shape[1] * 128)[n:]
    assert float(tail.abs().max()) == 0.0 if tail.numel() else True


@pytest.mark.parametrize("kinds,coarse,hard,filt", [(("smape", "mse"), True, True, True), (("mse", "mse"), False, True, True),
                                                     (("smape", "smape"), True, False, True), (("mse", "smape"), True, True, False)])
def test_fused_alpha_loss_matches_torch_expression(kinds, coarse, hard, filt):
    """ntx_alpha_loss (value + gradients in one launch) against the generic torch form of network/loss.py:30-58."""
    from nerf_tex_b200 import loss as L
    g = torch.Generator(device="cuda").manual_seed(3)
    B, Nr = 4, 777
    ct = torch.rand((B, Nr, 3), device="cuda", generator=g)
    at = torch.rand((B, Nr), device="cuda", generator=g)
    at[at < 0.3] = 0.0
    mk = lambda *s: torch.rand(s, device="cuda", generator=g).requires_grad_(True)
    cp, ap = mk(B, Nr, 3), mk(B, Nr)
    cc, ac = (mk(B, Nr, 3), mk(B, Nr)) if coarse else (None, None)
    fn = L.AlphaLoss("network.loss." + kinds[0], "network.loss." + kinds[1], gamma=0.7, filter_color_loss=filt, use_hard_mask=hard)
    assert fn._kinds is not None
    fused = fn(ct, at, cp, ap, cc, ac)
    (fused * 3.0).backward()
    got = [x.grad.clone() for x in (cp, ap, cc, ac) if x is not None]
    for x in (cp, ap, cc, ac):
        if x is not None:
            x.grad = None
    fn._kinds = None                                                   # generic path
    ref = fn(ct, at, cp, ap, cc, ac)
    (ref * 3.0).backward()
    want = [x.grad for x in (cp, ap, cc, ac) if x is not None]
    assert abs(float(fused.detach()) - float(ref.detach())) < 1e-5 * max(1.0, abs(float(ref.detach())))
    for a, b in zip(got, want):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-9), float((a - b).abs().max())


def test_adam_kernel_matches_keras_form():
    """Trainer.adam_update through ntx_adam_step against the oracle's tf.keras-form Adam (train.py:52)."""
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.train import Trainer
    m = FusedNerfTexMLP(1, 4, seed=1)
    tr = Trainer({"model": m}, renderer=None, loss_fn=None, lrate=5e-4, lrate_decay=500)
    rng = np.random.default_rng(0)
    p = m.master.detach().cpu().numpy().copy()
    mm, vv = np.zeros_like(p), np.zeros_like(p)
    for step in range(1, 4):
        gnp = rng.normal(size=p.shape).astype(np.float32) * 1e-2
        m.master.grad = torch.as_tensor(gnp).cuda()
        lr = tr.current_lr()
        tr.step += 1
        tr.adam_update(lr)
        p, mm, vv = O.adam_step(p, gnp, mm, vv, step, lr)
        assert m._packed_version == -1
    np.testing.assert_allclose(m.master.detach().cpu().numpy(), p, rtol=2e-6, atol=5e-8)      # steps are ~5e-4
    np.testing.assert_allclose(tr.m["model"].cpu().numpy(), mm, rtol=2e-6, atol=5e-9)          # |m| ~ 1e-3, fused multiply-adds
    # beta2 = 0.999 is not an fp32 number: the kernel (like TF, which keeps hyper-parameters in the variable dtype) uses
    # 1 - fl32(0.999), the oracle 1 - 0.999 in double: 1.3e-5 apart
    np.testing.assert_allclose(tr.v["model"].cpu().numpy(), vv, rtol=3e-5, atol=1e-11)


def test_tensorflow_checkpoint_restores_into_trainable_models(tmp_path):
    """Train() builds the Trainer (which makes every master buffer a trainable leaf) before the Logger, whose
    TensorFlow-checkpoint path (logger.py:38-40 restore) then copies weights in place: must not trip autograd."""
    from nerf_tex_b200 import tf_checkpoint as TF
    from nerf_tex_b200.logger import Logger
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.train import Trainer
    from test_tf_checkpoint import keras_keys
    om = O.ParamNerfOracle((1, 4), seed=5)
    om.randomize_biases(6, 0.1)
    creation = [n for n, _, _ in om.layers]
    tensors = {"step" + TF.VALUE_SUFFIX: np.array(77, np.int64)}
    tensors.update(keras_keys(om, "model", creation))
    os.makedirs(tmp_path / "checkpoints")
    TF.write_bundle(str(tmp_path / "checkpoints" / "ckpt-1"), tensors)
    with open(tmp_path / "checkpoints" / "checkpoint", "w") as f:
        f.write('model_checkpoint_path: "ckpt-1"\n')
    m = FusedNerfTexMLP(n_geo=1, n_app=4, seed=0)
    tr = Trainer({"model": m}, renderer=None, loss_fn=None)
    assert m.master.requires_grad
    Logger(str(tmp_path), {"model": m, "step": 0, "optimizer": tr}, is_training=True)
    assert tr.step == 77
    flat = np.concatenate([np.concatenate([W.reshape(-1), b]) for _, W, b in om.layers])
    assert np.array_equal(m.master.detach().cpu().numpy(), flat)
    m.reset_parameters(seed=3)                                   # the other in-place writer
    assert m.master.requires_grad and m.master.is_leaf


@pytest.mark.parametrize("npar,blur,noise_std,R,S", [((1, 4), None, 0.0, 300, 64), ((2, 3), 0, 0.1, 512, 256)])
def test_fp8_saved_operands_keep_the_gradient_bars(npar, blur, noise_std, R, S):
    """Training with the saved tile images in fp8 e4m3 (per-step scales delayed by one step; the first pass calibrates
    with bf16 images): the gradient of the second pass meets the very bars of the bf16 path against float64 autograd."""
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.loss import AlphaLoss
    oms, gms, rays_o, rays_d, t, params, cone, ct, at = setup(npar, False, R)
    gm, om = gms["model"].requires_grad_(True), oms["model"]
    gm.enable_fp8_saved_operands()
    junk = torch.full((96 << 20,), float("inf"), dtype=torch.bfloat16, device="cuda")   # what a recycled block may hold:
    del junk                                                       # the calibration must not read unwritten image parts
    rng = np.random.default_rng(5)
    z_rand = rng.uniform(size=(R, S)).astype(np.float32)
    noise = (rng.normal(size=(R, S)) * noise_std).astype(np.float32) if noise_std > 0 else None
    r = Renderer(model=gm, n_samples=S, perturb=True, blur_idx=blur, raw_noise_std=noise_std)
    loss_fn = AlphaLoss("network.loss.smape", "network.loss.mse")
    grads = []
    for step in range(2):
        gm.master.grad = None
        pred = r(dev(rays_o[None]), dev(rays_d[None]), dev(t[None]), dev(params), dev(cone[None]), composite_bkgd=True,
                 bkgd_color=(1., 1., 1.), training=True, z_rand=dev(z_rand), noise=None if noise is None else dev(noise))
        loss = loss_fn(color_true=dev(ct), alpha_true=dev(at), **pred)
        loss.backward()
        grads.append(gm.master.grad.detach().cpu().clone())
        assert gm._fp8["calibrated"]
    assert not torch.equal(grads[0], grads[1])                     # the second pass really took the 8-bit path
    sx, sy = gm._fp8["scale_x"].cpu(), gm._fp8["scale_y"].cpu()
    n_steps = gm.schedule().n_steps
    assert bool((sx[:n_steps] > 0).all()) and bool((sy[:n_steps] > 0).all())
    # the kernels' maxima replaced the initial 1.0
    assert int((sx[:n_steps] < 1.0).sum()) >= n_steps - 1 and int((sy[:n_steps] != 1.0).sum()) >= n_steps - 1
    P = T.make_params(om)
    z = t64(O.stratified_z(t, S, True, z_rand))
    C, A, _ = T.network_pass(om, P, t64(rays_o), t64(rays_d), z, t64(np.repeat(params, R, 0)), t64(cone), blur,
                             composite_bkgd=True, noise=None if noise is None else t64(noise))
    T.alpha_loss(t64(ct[0]), t64(at[0]), C, A).backward()
    want = T.flat_grads(om, P)
    check_grads(grads[1], want, om, "fp8 saved operands")
    rel16 = float((grads[0].double() - want).norm() / want.norm())
    rel8 = float((grads[1].double() - want).norm() / want.norm())
    print(f"flat gradient rel error vs float64: bf16 images {rel16:.3e}, e4m3 images {rel8:.3e}")
    assert rel8 < 2e-2

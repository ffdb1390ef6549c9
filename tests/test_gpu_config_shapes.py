"""Parity at the shapes the benchmark and the shipped configs actually run, not at toy sizes:

* C2 carpet render: model [1,6], S = 1024, step 0.002, patch_scale 0.09, one 16 384-ray chunk, ~120 in-shell samples
  per hit ray (configs/config_carpet_render.py:76-99);
* C5 plush render: model [1,4], S = 1280, step 0.0005, patch_scale 0.04, ~200 samples per hit ray, 32 768-ray chunk
  (configs/config_plush_render.py:87-112);
* C3 grass_filtered train step: model [2,3], S = 256, blur_idx 0, raw_noise_std 0.1 on 512 rays
  (configs/config_grass_filtered_train.py:95-103).

The instanced chunks are rendered whole on the GPU; the oracle (network/renderer.py:247-354 restated in
oracle/nerftex_oracle.py) replays a 2 048-hit-ray subset of the very same records.  Bars: 2e-3 max abs and 50 dB for
rendered colour / alpha; gradients cos >= 0.999, rel <= 2e-2 vs float64 autograd."""

import math

import numpy as np
import pytest
import torch

from oracle import nerftex_oracle as O

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()


def psnr(a, b):
    return -10 * math.log10(float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)) + 1e-30)


def sub_csr(csr, sel):
    """CSR records of the hit rays `sel` (indices into the hit list), re-indexed as rays 0..len(sel)-1."""
    ptr = csr["ray_ptr"]
    counts = (ptr[1:] - ptr[:-1])[sel]
    starts = ptr[:-1][sel]
    new_ptr = torch.zeros(sel.shape[0] + 1, dtype=torch.int64, device=ptr.device)
    new_ptr[1:] = torch.cumsum(counts, 0)
    rows = torch.repeat_interleave(starts - new_ptr[:-1], counts) + torch.arange(int(new_ptr[-1]), device=ptr.device)
    out = {k: csr[k][rows].contiguous() for k in ("pts", "dirs", "params", "dists", "t", "alpha_weight", "instance_id")}
    out.update({"ray_ptr": new_ptr, "ray_index": torch.arange(sel.shape[0], device=ptr.device, dtype=torch.int32),
                "color_last": csr["color_last"][sel].contiguous(), "alpha_last": csr["alpha_last"][sel].contiguous()})
    return out


@pytest.mark.parametrize("tag,npar,S,step,patch_scale,chunk,mean_valid,b0,b1,blur", [
    ("C2 carpet", (1, 6), 1024, 0.002, 0.09, 16384, 120.0, [-1.4, -1.2, -.1], [1.2, 1.2, 1.8], None),
    ("C5 plush", (1, 4), 1280, 0.0005, 0.04, 32768, 200.0, [-1.1, -1.1, -.2], [1.1, 1.1, 1.1], None),
    ("grass_filtered render", (2, 3), 1024, 0.001, 0.1, 16384, 120.0, [-2, -2, -.5], [2, 2, 2.5], 0),
])
def test_instanced_chunk_at_config_shape(tag, npar, S, step, patch_scale, chunk, mean_valid, b0, b1, blur):
    from nerf_tex_b200.instancer import SyntheticInstancer, csr_to_dense
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import InstanceRenderer
    om = O.ParamNerfOracle(n_parameters=npar, seed=31)
    om.randomize_biases(32, 0.05)
    gm = FusedNerfTexMLP(n_geo=npar[0], n_app=npar[1])
    gm.load_named_weights(om.named_weights())
    mk = lambda: SyntheticInstancer(b_0=b0, b_1=b1, patch_scale=patch_scale, seed=0, mean_valid=mean_valid)
    density_scale = 1.5                                            # untrained weights: opacity neither 0 nor 1
    r = InstanceRenderer(instancer_config=mk(), model=gm, n_samples=S, render_chunk=chunk, step_size=step,
                         density_scale=density_scale, blur_idx=blur)
    R, p = chunk, sum(npar)
    g = torch.Generator(device="cuda").manual_seed(5)
    rays_o = torch.randn((1, R, 3), device="cuda", generator=g)
    rays_d = torch.nn.functional.normalize(torch.randn((1, R, 3), device="cuda", generator=g), dim=-1)
    t = torch.tensor([1., 3.], device="cuda").expand(1, R, 2).contiguous()
    params = torch.cat([torch.rand((1, p - 3), device="cuda", generator=g) * 0.8 + 0.2,
                        torch.tensor([[0., 0., 1.]], device="cuda")], -1)
    cone = torch.rand((1, R, 1), device="cuda", generator=g) * 0.04 + 0.01
    got = r(rays_o, rays_d, t, params, cone, composite_bkgd=True, training=False)
    gc, ga = got["color_pred"][0], got["alpha_pred"][0]

    csr = mk().get_model_input_csr(rays_o[0], rays_d[0], params.expand(R, p).contiguous(), S, step)
    counts = csr["ray_ptr"][1:] - csr["ray_ptr"][:-1]
    assert abs(float(counts.float().mean()) / mean_valid - 1.2) < 0.25            # LogNormal(ln m, 0.6): mean ~ 1.2 m
    assert int(counts.max()) > 0.45 * S                                            # long rays are in the chunk
    H = csr["ray_index"].shape[0]
    sel = torch.randperm(H, device="cuda", generator=g)[:2048].sort().values
    sel[-1] = int(counts.argmax())                                                 # the longest ray is always checked
    sub = sub_csr(csr, sel)
    ray_ids = csr["ray_index"][sel].long()
    n_sub = sel.shape[0]
    dense = csr_to_dense(sub, rays_d[0][ray_ids].contiguous(), params.expand(n_sub, p).contiguous(), S)
    dense = [x.cpu().numpy() for x in dense]
    wc, wa = O.instance_evaluate_model(om, dense, n_sub, cone_scale=cone[0][ray_ids].cpu().numpy(), blur_idx=blur,
                                       patch_scale=patch_scale, density_scale=density_scale, composite_bkgd=True)
    gcs, gas = gc[ray_ids].cpu().numpy(), ga[ray_ids].cpu().numpy()
    assert 0.1 < (wa > 0.05).mean() and (wa < 0.95).mean() > 0.1, (tag, float((wa > 0.05).mean()), float((wa < 0.95).mean()))
    assert np.abs(gas - wa).max() < 2e-3 and np.abs(gcs - wc).max() < 2e-3, (tag, np.abs(gas - wa).max(), np.abs(gcs - wc).max())
    assert psnr(gcs, wc) > 50 and psnr(gas, wa) > 50, (tag, psnr(gcs, wc), psnr(gas, wa))
    miss = torch.ones(R, dtype=torch.bool, device="cuda")
    miss[csr["ray_index"].long()] = False
    assert bool((gc[miss] == 0.0).all()) and bool((ga[miss] == 0.0).all())         # renderer.py:313-314: culled rays are 0


def test_c3_shape_gradients():
    """C3 (i): single network, n_samples = 256, blur_idx 0, noise 0.1, AlphaLoss(smape, mse), 512 rays = 131 072 rows."""
    import torch_ref as T
    from nerf_tex_b200.loss import AlphaLoss
    from nerf_tex_b200.renderer import Renderer
    from test_gpu_train import check_grads, setup, t64
    npar, R, S = (2, 3), 512, 256
    oms, gms, rays_o, rays_d, t, params, cone, ct, at = setup(npar, False, R, seed=13)
    gm, om = gms["model"].requires_grad_(True), oms["model"]
    rng = np.random.default_rng(14)
    z_rand = rng.uniform(size=(R, S)).astype(np.float32)
    noise = (rng.normal(size=(R, S)) * 0.1).astype(np.float32)
    r = Renderer(model=gm, n_samples=S, perturb=True, blur_idx=0, raw_noise_std=0.1, render_chunk=32768, net_chunk=65536)
    pred = r(dev(rays_o[None]), dev(rays_d[None]), dev(t[None]), dev(params), dev(cone[None]), composite_bkgd=True,
             bkgd_color=(1., 1., 1.), training=True, z_rand=dev(z_rand), noise=dev(noise))
    loss = AlphaLoss("network.loss.smape", "network.loss.mse")(color_true=dev(ct), alpha_true=dev(at), **pred)
    loss.backward()
    P = T.make_params(om)
    z = t64(O.stratified_z(t, S, True, z_rand))
    C, A, _ = T.network_pass(om, P, t64(rays_o), t64(rays_d), z, t64(np.repeat(params, R, 0)), t64(cone), 0,
                             composite_bkgd=True, noise=t64(noise))
    ref = T.alpha_loss(t64(ct[0]), t64(at[0]), C, A)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 1e-3
    check_grads(gm.master.grad.detach().cpu(), T.flat_grads(om, P), om, "C3 shape")

"""Would 8-bit saved operands keep the gradient bars?  (CPU experiment, float64 reference.)

The training step is bound by the bytes of the saved activation / gradient images.  Storing them
as fp8 e4m3 for the weight-gradient GEMM would halve that traffic.  This script takes the float64 graph of the C3 shape
(tests/torch_ref.py), rounds the two operands of every layer's dW = X^T dY to bf16 and to e4m3 (per-tensor scale to the
format's maximum, the usual fp8 recipe), and reports the error of the flat weight gradient against the unrounded one:
the bars are rel <= 2e-2 and cos >= 0.999 (tests/test_gpu_train.py)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch_ref as T  # noqa: E402
from oracle import nerftex_oracle as O  # noqa: E402


def q(x, kind):
    if kind == "f64":
        return x
    if kind == "bf16":
        return x.to(torch.bfloat16).to(torch.float64)
    s = x.abs().max().clamp_min(1e-30) / 448.0                       # e4m3 max normal
    return (x / s).to(torch.float32).to(torch.float8_e4m3fn).to(torch.float64) * s


def main(R=256, S=64):
    torch.manual_seed(0)
    rng = np.random.default_rng(0)
    om = O.ParamNerfOracle(n_parameters=(2, 3), seed=1)
    om.randomize_biases(20, 0.05)
    P = T.make_params(om)
    rays_o = np.tile(np.array([[0., 0.5, 5.]], np.float32), (R, 1))
    rays_d = rng.uniform(-1, 1, (R, 3)).astype(np.float32) - rays_o
    rays_d /= np.linalg.norm(rays_d, axis=-1, keepdims=True)
    t = np.stack([rng.uniform(3, 3.5, R), rng.uniform(5.5, 6.5, R)], -1).astype(np.float32)
    z = torch.tensor(O.stratified_z(t, S, True, rng.uniform(size=(R, S)).astype(np.float32)), dtype=torch.float64)
    params = torch.tensor(rng.uniform(0, 1, (1, 5)), dtype=torch.float64).repeat(R, 1)
    ct = torch.tensor(rng.uniform(0, 1, (R, 3)))
    at = torch.tensor(np.where(rng.uniform(size=R) < 0.3, 0, rng.uniform(0, 1, R)))
    # forward with hooks on every Dense input / output
    acts = {}
    pts = (torch.tensor(rays_o, dtype=torch.float64)[:, None] + torch.tensor(rays_d, dtype=torch.float64)[:, None] * z[:, :, None]).reshape(-1, 3)
    dn = torch.tensor(rays_d, dtype=torch.float64).repeat_interleave(S, 0)
    prm = params.repeat_interleave(S, 0)
    g = om.n_parameters[0]
    pos_map = torch.cat([T.fourier(pts, om.pos_freqs), T.fourier(prm[:, :g], om.param_freqs)], -1)
    dir_map = torch.cat([T.fourier(dn, om.dir_freqs), T.fourier(prm[:, g:], om.param_freqs)], -1)

    def dense(name, x, relu):
        W, b = P[name]
        y = x @ W + b
        y.retain_grad()
        acts[name] = (x.detach(), y)
        return torch.relu(y) if relu else y
    h = pos_map
    for i in range(om.depth):
        h = dense(f"trunk{i}", h, True)
        if i in om.skips:
            h = torch.cat([pos_map, h], -1)
    alpha = dense("alpha", h, False)
    h = torch.cat([dir_map, dense("feature", h, False)], -1)
    for i in range(om.color_depth):
        h = dense(f"colorhid{i}", h, True)
    h = dense("colorhalf", h, True)
    color = dense("rgb", h, False)
    C, A, _ = T.composite(color.reshape(R, S, 3), alpha.reshape(R, S), z, torch.tensor(rays_d, dtype=torch.float64), composite_bkgd=True)
    T.alpha_loss(ct, at, C, A).backward()
    names = [n for n, _, _ in om.layers]
    ref = torch.cat([P[n][0].grad.reshape(-1) for n in names])
    for kind in ("bf16", "fp8"):
        got = torch.cat([(q(acts[n][0], kind).T @ q(acts[n][1].grad, kind)).reshape(-1) for n in names])
        rel = float((got - ref).norm() / ref.norm())
        cos = float((got * ref).sum() / (got.norm() * ref.norm()))
        worst = max(float(((q(acts[n][0], kind).T @ q(acts[n][1].grad, kind)) - P[n][0].grad).norm() / P[n][0].grad.norm()) for n in names)
        print(f"{kind}: flat weight-gradient rel error {rel:.3e}, cos {cos:.6f}, worst layer rel {worst:.3e}")


if __name__ == "__main__":
    main()

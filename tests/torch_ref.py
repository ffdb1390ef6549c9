"""Differentiable float64 torch-CPU twin of the oracle graph (test infrastructure).

Same ops as oracle/nerftex_oracle.py (layer.py:11-23, model.py:58-125, renderer.py:170-213),
written in torch so that autograd supplies reference gradients for the training-path tests."""

import torch


def fourier(x, n):
    feats = [x]
    for k in range(n):
        feats += [torch.sin((2.0 ** k) * x), torch.cos((2.0 ** k) * x)]
    return torch.cat(feats, -1)


def make_params(oracle_model):
    """{name: (W, b)} float64 leaf tensors from a ParamNerfOracle."""
    out = {}
    for n, W, b in oracle_model.layers:
        out[n] = (torch.tensor(W, dtype=torch.float64, requires_grad=True),
                  torch.tensor(b, dtype=torch.float64, requires_grad=True))
    return out


def mlp(om, P, pos, dirs, params):
    g = om.n_parameters[0]
    pos_map, dir_map = fourier(pos, om.pos_freqs), fourier(dirs, om.dir_freqs)
    if g > 0:
        pos_map = torch.cat([pos_map, fourier(params[:, :g], om.param_freqs)], -1)
    if om.n_parameters[1] > 0:
        dir_map = torch.cat([dir_map, fourier(params[:, g:], om.param_freqs)], -1)
    return mlp_from_maps(om, P, pos_map, dir_map)


def mlp_from_maps(om, P, pos_map, dir_map):
    """The Dense stack of model.py:104-123 on already-encoded inputs."""
    h = pos_map
    for i in range(om.depth):
        W, b = P[f"trunk{i}"]
        h = torch.relu(h @ W + b)
        if i in om.skips:
            h = torch.cat([pos_map, h], -1)
    W, b = P["alpha"]
    alpha = h @ W + b
    W, b = P["feature"]
    h = torch.cat([dir_map, h @ W + b], -1)
    for i in range(om.color_depth):
        W, b = P[f"colorhid{i}"]
        h = torch.relu(h @ W + b)
    W, b = P["colorhalf"]
    h = torch.relu(h @ W + b)
    W, b = P["rgb"]
    return h @ W + b, alpha


def composite(color, sigma, z, rays_d, composite_bkgd=False, bkgd=(1., 1., 1.), map_exr=False, noise=None):
    dists = z[:, 1:] - z[:, :-1]
    dists = torch.cat([dists, dists[:, -1:]], -1) * rays_d.norm(dim=-1, keepdim=True)
    col = torch.nn.functional.elu(color) + 1 if map_exr else torch.sigmoid(color)
    s = sigma if noise is None else sigma + noise
    a = 1 - torch.exp(-torch.relu(s) * dists)
    T = torch.cumprod(torch.cat([torch.ones_like(a[:, :1]), 1 - a + 1e-10], -1), -1)[:, :-1]
    w = a * T
    C, A = (w[..., None] * col).sum(-2), w.sum(-1)
    if composite_bkgd:
        C = C + (1 - A)[:, None] * torch.tensor(bkgd, dtype=C.dtype)
    return C, A, w


def network_pass(om, P, rays_o, rays_d, z, parameters, cone_scale=None, blur_idx=None, **kw):
    R, S = z.shape
    pts = (rays_o[:, None, :] + rays_d[:, None, :] * z[:, :, None]).reshape(-1, 3)
    dn = rays_d / rays_d.norm(dim=-1, keepdim=True)
    dirs = dn.repeat_interleave(S, 0)
    prm = parameters.repeat_interleave(S, 0)
    if blur_idx is not None:
        scale = (cone_scale[:, None, :] * z[:, :, None]).reshape(-1)
        prm = prm.clone()
        prm[:, blur_idx] = prm[:, blur_idx] * scale
    color, alpha = mlp(om, P, pts, dirs, prm)
    return composite(color.reshape(R, S, 3), alpha.reshape(R, S), z, rays_d, **kw)


def smape(y_true, y_pred, eps=1e-2):
    return torch.mean(torch.abs(y_true - y_pred) / (y_true + y_pred + eps))


def mse(y_true, y_pred):
    return torch.mean((y_true - y_pred) ** 2)


def alpha_loss(color_true, alpha_true, color_pred, alpha_pred, color_pred_coarse=None, alpha_pred_coarse=None, gamma=1.0):
    mask = (alpha_true[..., None] > 0).to(color_true.dtype)
    ct = color_true * mask
    loss = smape(ct, color_pred * mask) + gamma * mse(alpha_true, alpha_pred)
    if color_pred_coarse is not None:
        loss = loss + smape(ct, color_pred_coarse * mask) + gamma * mse(alpha_true, alpha_pred_coarse)
    return loss


def flat_grads(om, P):
    """Gradients in the flat master order: per layer W then b."""
    return torch.cat([torch.cat([P[n][0].grad.reshape(-1), P[n][1].grad.reshape(-1)]) for n, _, _ in om.layers])

"""K3 parity: stratified depths, sample_pdf and the hierarchical merge are BIT-EXACT against the
oracle (NumPy, defined op order) given identical uniform draws; searchsorted indices identical."""

import numpy as np
import pytest
import torch

from oracle import nerftex_oracle as O

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def dev(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a)).to("cuda", dtype)


@pytest.mark.parametrize("R,S", [(1, 2), (37, 64), (4096, 64), (513, 256), (5, 1)])
@pytest.mark.parametrize("perturb", [False, True])
def test_stratified_bit_exact(R, S, perturb):
    from nerf_tex_b200 import ops
    rng = np.random.default_rng(R * 1000 + S)
    t = np.stack([rng.uniform(2, 3, R), rng.uniform(4, 6, R)], -1).astype(np.float32)
    zr = rng.uniform(size=(R, S)).astype(np.float32)
    want = O.stratified_z(t, S, perturb, zr)
    got = ops.stratified_z(dev(t), S, perturb, dev(zr)).cpu().numpy()
    assert np.array_equal(bits(got), bits(want))


@pytest.mark.parametrize("R,C,N", [(1, 2, 1), (3, 3, 7), (100, 63, 128), (4096, 63, 128), (33, 255, 64), (7, 1025, 40)])
@pytest.mark.parametrize("det", [True, False])
def test_sample_pdf_bit_exact(R, C, N, det):
    from nerf_tex_b200 import ops
    rng = np.random.default_rng(C * 7 + N + R)
    bins = np.sort(rng.uniform(2, 6, (R, C)).astype(np.float32), -1)
    w = (rng.uniform(size=(R, C - 1)) ** 8).astype(np.float32)      # peaky (SURVEY.md C1 stress input)
    w[0] = 0                                                          # all-zero weights -> the 1e-5 floor
    u = rng.uniform(size=(R, N)).astype(np.float32)
    u[:, 0] = 0.0
    u[:, -1] = np.float32(1.0)                                        # inds == len(cdf) edge
    want, winds = O.sample_pdf(bins, w, N, det, u, return_inds=True)
    got, ginds = ops.sample_pdf(dev(bins), dev(w), N, det, dev(u), return_inds=True)
    assert np.array_equal(ginds.cpu().numpy(), winds)
    assert np.array_equal(bits(got.cpu().numpy()), bits(want))


@pytest.mark.parametrize("R,S,N", [(4096, 64, 128), (17, 64, 128), (9, 3, 5), (65, 256, 256), (3, 512, 512)])
@pytest.mark.parametrize("det", [True, False])
def test_hierarchical_bit_exact(R, S, N, det):
    from nerf_tex_b200 import ops
    rng = np.random.default_rng(S + N + R)
    t = np.stack([rng.uniform(2, 3, R), rng.uniform(4, 6, R)], -1).astype(np.float32)
    z = O.stratified_z(t, S, True, rng.uniform(size=(R, S)).astype(np.float32))
    w = (rng.uniform(size=(R, S)) ** 8).astype(np.float32)
    u = rng.uniform(size=(R, N)).astype(np.float32)
    wz, ws, wi = O.hierarchical_z(z, w, N, det, u, return_inds=True)
    gz, gs, gi = ops.hierarchical_z(dev(z), dev(w), N, det, dev(u), return_aux=True)
    assert np.array_equal(gi.cpu().numpy(), wi)
    assert np.array_equal(bits(gs.cpu().numpy()), bits(ws))
    assert np.array_equal(bits(gz.cpu().numpy()), bits(wz))
    assert np.all(np.diff(gz.cpu().numpy(), axis=-1) >= 0)            # sortedness (size-independent property)


def test_empty_inputs():
    from nerf_tex_b200 import ops
    z = ops.stratified_z(torch.empty((0, 2), device="cuda"), 64, False)
    assert z.shape == (0, 64)
    s = ops.sample_pdf(torch.empty((0, 63), device="cuda"), torch.empty((0, 62), device="cuda"), 128, det=True)
    assert s.shape == (0, 128)


def test_errors_are_loud():
    from nerf_tex_b200 import ops, _lib
    with pytest.raises(_lib.NtxError):
        ops.hierarchical_z(torch.zeros((4, 600), device="cuda"), torch.zeros((4, 600), device="cuda"), 600, True)
    with pytest.raises(_lib.NtxError):
        _lib.ptr(torch.zeros((4, 2)))                        # CPU tensor at the ABI boundary: no CPU path

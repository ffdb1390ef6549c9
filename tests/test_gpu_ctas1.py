"""The single-CTA variants of the tensor-core kernels (NTX_MLP_CTAS=1: cta_group::1, plain bulk copies, no cluster) are
selected once per process, so they run in a child process here; the default CTA-pair path (cta_group::2, tensor-map
weight copies credited to the issuing CTA's barrier) runs in this one.  Same K order per row in both, so the forward
results are bit-equal; the weight gradient sums tiles in a different order and is compared to fp32 tolerance."""

import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys
import numpy as np
import torch
sys.path.insert(0, {root!r})
from nerf_tex_b200.model import FusedNerfTexMLP
from nerf_tex_b200 import autograd as A

def run():
    n = 3000
    m = FusedNerfTexMLP(n_geo=1, n_app=6, seed=3)
    g = torch.Generator(device="cuda").manual_seed(7)
    pos = torch.rand((n, 3), device="cuda", generator=g) * 2 - 1
    d = torch.nn.functional.normalize(torch.randn((n, 3), device="cuda", generator=g), dim=-1)
    prm = torch.rand((n, 7), device="cuda", generator=g)
    out = m.forward_rows(n_rows=n, pos=pos, dir=d, params=prm, rows_per_group=1)
    rgbs, saved = A.mlp_forward_train(m, n_rows=n, pos=pos, dir=d, params=prm, rows_per_group=1)
    d_rgbs = torch.randn((n, 4), device="cuda", generator=g)
    grad = A.mlp_backward_native(m, saved, d_rgbs)
    return out.cpu().numpy(), rgbs.cpu().numpy(), grad.cpu().numpy()

if __name__ == "__main__":
    o, r, g = run()
    np.savez(sys.argv[1], out=o, rgbs=r, grad=g)
"""


def test_single_cta_kernels_agree_with_the_pair_kernels(tmp_path):
    script = tmp_path / "ctas1.py"
    script.write_text(SCRIPT.format(root=ROOT))
    env = dict(os.environ, NTX_MLP_CTAS="1")
    res = subprocess.run([sys.executable, str(script), str(tmp_path / "c1.npz")], env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    one = np.load(tmp_path / "c1.npz")
    ns = {}
    exec(compile(SCRIPT.format(root=ROOT).split("if __name__")[0], "ctas_pair", "exec"), ns)
    out, rgbs, grad = ns["run"]()
    assert np.isfinite(out).all() and np.abs(out).max() > 0
    assert np.array_equal(out.view(np.uint32), one["out"].view(np.uint32))
    assert np.array_equal(rgbs.view(np.uint32), one["rgbs"].view(np.uint32))
    scale = np.abs(grad).max()
    assert np.abs(grad - one["grad"]).max() <= 2e-5 * scale, (np.abs(grad - one["grad"]).max(), scale)

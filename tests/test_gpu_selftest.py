"""tcgen05 / TMA self-tests: validates the UMMA shared-memory descriptors the fused MLP uses."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf16(x):
    return x.to(torch.bfloat16).to(torch.float32)


def test_bulk_copy_roundtrip():
    from nerf_tex_b200 import ops
    src = torch.randint(0, 255, (16384,), dtype=torch.uint8, device="cuda")
    dst = ops.selftest_bulk_copy(src)
    torch.cuda.synchronize()
    assert torch.equal(src, dst)


@pytest.mark.parametrize("layout", [0, 1, 2])
@pytest.mark.parametrize("N,K", [(256, 64), (256, 128), (128, 256)])
def test_umma_gemm_layouts(layout, N, K):
    from nerf_tex_b200 import ops
    g = torch.Generator(device="cpu").manual_seed(layout * 100 + N + K)
    A = torch.randn((128, K), generator=g)
    B = torch.randn((N, K), generator=g)
    D = ops.selftest_umma(A.cuda(), B.cuda(), layout)
    torch.cuda.synchronize()
    ref = _bf16(A).double() @ _bf16(B).double().T
    err = (D.cpu().double() - ref).abs().max().item()
    assert err < 1e-3, f"layout {layout} N={N} K={K}: max abs err {err}"

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


@pytest.mark.parametrize("N,K", [(256, 128), (256, 64), (128, 128)])
def test_umma_gemm_mn_major_operands(N, K):
    """Both operands MN-major (stored [K][M], [K][N]) on the forward kernel's SWIZZLE_64B row images:
    the operand form of the weight-gradient GEMM dW = X^T dY."""
    from nerf_tex_b200 import ops
    g = torch.Generator(device="cpu").manual_seed(N + K)
    At = torch.randn((K, 128), generator=g)
    Bt = torch.randn((K, N), generator=g)
    D = ops.selftest_umma_mn(At.cuda(), Bt.cuda())
    torch.cuda.synchronize()
    ref = _bf16(At).double().T @ _bf16(Bt).double()
    err = (D.cpu().double() - ref).abs().max().item()
    assert err < 1e-3, f"MN-major N={N} K={K}: max abs err {err}"


@pytest.mark.parametrize("N,K", [(128, 64), (256, 64), (256, 128)])
def test_umma_gemm_mn_major_fp8_operands(N, K):
    """kind::f8f6f4 with e4m3 operands stored MN-major (the form 8-bit saved activation / gradient images would have for
    the weight-gradient GEMM): exact against the product of the e4m3-rounded operands (fp32 accumulate)."""
    from nerf_tex_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(N + K)
    At = torch.randn((K, 128), device="cuda", generator=g) * 4
    Bt = torch.randn((K, N), device="cuda", generator=g) * 0.05
    D = ops.selftest_umma_mn8(At, Bt)
    q = lambda x: x.to(torch.float8_e4m3fn).to(torch.float64)
    want = (q(At).T @ q(Bt)).float()
    torch.testing.assert_close(D, want, rtol=1e-5, atol=1e-5)

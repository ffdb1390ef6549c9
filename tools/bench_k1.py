"""K1 alone: the fused encoding + MLP forward on one bench-sized launch (1.6 M rows, model [1,6]); TFLOP/s from CUDA events."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_tex_b200.model import FusedNerfTexMLP  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1607680
    npar = (1, 6)
    m = FusedNerfTexMLP(*npar, seed=1, device=dev)
    g = torch.Generator(device=dev).manual_seed(0)
    pos = torch.rand((n, 3), device=dev, generator=g) * 2 - 1
    d = torch.nn.functional.normalize(torch.randn((n, 3), device=dev, generator=g), dim=-1)
    prm = torch.rand((n, sum(npar)), device=dev, generator=g)
    out = torch.empty((n, 4), device=dev)
    for _ in range(5):
        m.forward_rows(n_rows=n, pos=pos, dir=d, params=prm, rows_per_group=1, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        m.forward_rows(n_rows=n, pos=pos, dir=d, params=prm, rows_per_group=1, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"lib {os.environ.get('NTX_LIB', 'default')} rows {n} ms {ms:.4f} TFLOP/s {n * 1361664 / ms / 1e9:.1f} checksum {float(out.double().abs().mean()):.6f}")

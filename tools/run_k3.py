"""One K3 (hierarchical_z, 64 + 128) launch on bench-sized inputs: the target of `ncu -k regex:sample_pdf`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_tex_b200 import ops  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    R = 1 << 21
    z0 = torch.sort(torch.rand((R, 64), device=dev, generator=g) * 4 + 2, -1).values
    w0 = torch.rand((R, 64), device=dev, generator=g) ** 8
    for _ in range(3):
        out = ops.hierarchical_z(z0, w0, 128, True)
    torch.cuda.synchronize()
    print("ok", float(out[0].sum() if isinstance(out, tuple) else out.sum()))

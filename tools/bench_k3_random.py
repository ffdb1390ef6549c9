"""K3 (hierarchical_z, 64 + 128) with random u (the perturb=False render path): rays/s."""
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
    u = torch.rand((R, 128), device=dev, generator=g)
    for det, uu in ((True, None), (False, u)):
        for _ in range(3):
            ops.hierarchical_z(z0, w0, 128, det, uu)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            ops.hierarchical_z(z0, w0, 128, det, uu)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print("det" if det else "random u", f"{ms:.3f} ms", f"{R / ms / 1e3:.1f} M rays/s")

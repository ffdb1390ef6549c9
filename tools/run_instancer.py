"""A few device-instancer calls on a dense chunk of the bench's carpet frame: the target of `ncu -k regex:instancer`."""
import math
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nerf_tex_b200.dataset import look_at  # noqa: E402
from nerf_tex_b200.instancer import Instancer  # noqa: E402
from nerf_tex_b200.proxy import AABB  # noqa: E402
from nerf_tex_b200.ray_sampler import Proxy  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
    inst = Instancer(b_0=[-1.4, -1.2, -.1], b_1=[1.2, 1.2, 1.8], textures=["", "", "", "", "light"],
                     transformations=bench.carpet_patches(1500), patch_scale=0.09, instance_sampling_method="nearest", device=dev)
    focal = 800 / math.tan(0.55 / 2) / 2
    z = 0.4
    pose = look_at([6 * math.sqrt(1 - z * z), 0.0, 6 * z])
    rays_o, rays_d, t, cone = Proxy(800, 800, focal, AABB([-1.5] * 3, [1.5] * 3))(None, pose)
    hit = torch.isfinite(t[:, 0])
    mid = int(hit.sum()) // 2
    ro, rd = rays_o[hit][mid - n // 2:mid + n // 2].contiguous(), rays_d[hit][mid - n // 2:mid + n // 2].contiguous()
    pf = torch.tensor([[1, 1, 1, .1, 0, 0, 1.]], device=dev).repeat(ro.shape[0], 1)
    for _ in range(int(os.environ.get('NTX_ITERS', '3'))):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        csr = inst.get_model_input_csr(ro, rd, pf, 1024, 0.002, seed=0)
        torch.cuda.synchronize()
        print("rays", ro.shape[0], "samples", csr["pts"].shape[0], "hit rays", csr["ray_index"].shape[0], "ms", (time.perf_counter() - t0) * 1e3)

"""Trace + march of the device instancer alone: 131 072 rays over a procedural carpet of n patches at the bench's areal
density, with the uniform grid and with the all-instances loop (use_grid=False)."""
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nerf_tex_b200.instancer import Instancer  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    R = 131072
    out = {}
    for n in (1500, 20000):
        extent = 1.45 * math.sqrt(n / 1500)
        T = bench.carpet_patches(n, extent=extent)
        g = torch.Generator(device=dev).manual_seed(0)
        # camera-like bundle: origins on a small patch 5 above the carpet, directions towards points on it
        tgt = (torch.rand((R, 3), device=dev, generator=g) * 2 - 1) * torch.tensor([1.45, 1.45, 0.0], device=dev)
        org = torch.tensor([3.0, 2.0, 4.0], device=dev).expand(R, 3).contiguous()
        d = torch.nn.functional.normalize(tgt - org, dim=-1)
        prm = torch.tensor([bench.DATASET_PARAMS], device=dev).repeat(R, 1)
        for use_grid in (True, False):
            inst = Instancer(b_0=bench.B0, b_1=bench.B1, textures=["", "", "", "", "light"], transformations=T, patch_scale=0.09,
                             instance_sampling_method="nearest", device=dev, use_grid=use_grid)
            for _ in range(2):
                csr = inst.get_model_input_csr(org, d, prm, 1024, 0.002, seed=0)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                inst.get_model_input_csr(org, d, prm, 1024, 0.002, seed=0)
            e1.record()
            torch.cuda.synchronize()
            out[f"n={n} grid={use_grid}"] = {"ms_trace_plus_march": e0.elapsed_time(e1) / 5, "samples": int(csr["pts"].shape[0]),
                                             "hit_rays": int(csr["ray_index"].shape[0])}
    print(json.dumps(out, indent=1))

"""A few C3 training steps (4096 rays x 256 samples, model [2,3]) for `ncu --metrics gpu__time_duration.sum`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_tex_b200.loss import AlphaLoss  # noqa: E402
from nerf_tex_b200.model import FusedNerfTexMLP  # noqa: E402
from nerf_tex_b200.renderer import Renderer  # noqa: E402
from nerf_tex_b200.train import Trainer  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    R = 4096
    ro = torch.tensor([[0., 0.5, 5.]], device=dev).repeat(R, 1)
    rd = torch.rand((R, 3), device=dev, generator=g) * 2 - 1 - ro
    rd = rd / rd.norm(dim=-1, keepdim=True)
    data = {"rays_o": ro[None], "rays_d": rd[None],
            "t": torch.stack([3 + torch.rand(R, device=dev, generator=g), 6 + torch.rand(R, device=dev, generator=g)], -1)[None],
            "parameters": torch.rand((1, 5), device=dev, generator=g), "cone_scale": torch.rand((1, R, 1), device=dev, generator=g) * .01,
            "color": torch.rand((1, R, 3), device=dev, generator=g), "alpha": torch.rand((1, R), device=dev, generator=g)}
    models = {"model": FusedNerfTexMLP(2, 3, seed=1)}
    if os.environ.get("NTX_FP8_SAVED") == "1":
        models["model"].enable_fp8_saved_operands()
    r = Renderer(n_samples=256, n_importance=0, perturb=True, blur_idx=0, raw_noise_std=0.1, seed=0, **models)
    tr = Trainer(models, r, AlphaLoss("network.loss.smape", "network.loss.mse"))
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(n):
        if i == 2:
            e0.record()
        if i == n - 1:
            torch.cuda.synchronize()
            torch.cuda.nvtx.range_push("last_step")
        loss = tr.train_step(data)
    e1.record()
    torch.cuda.synchronize()
    print("loss", float(loss), "ms/step", e0.elapsed_time(e1) / max(n - 2, 1))

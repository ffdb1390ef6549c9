"""One frame across several GPUs: rays are independent (every op of network/renderer.py:47-354 is per ray or per
sample), so each rank renders a contiguous block of the frame's rays and the only communication is the gather of the
finished per-ray outputs — equal, padded blocks in one all-gather per output, over NCCL / NVLink on GPUs."""

from __future__ import annotations

from typing import Dict

import torch
import torch.distributed as dist

from .train import shard_rays


def render_sharded(renderer, rays_o, rays_d, t, parameters, cone_scale, *, group=None, block: int = 16,
                   **kwargs) -> Dict[str, torch.Tensor]:
    """`renderer(rays_o[B,Nr,3], rays_d, t[B,Nr,2], parameters[B,p], cone_scale[B,Nr,1], **kwargs)` with the Nr axis
    split over the ranks of `group`; returns the same dict as the renderer, complete on every rank.  Per-ray keyword
    tensors (z_rand, u, noise ... shaped [B*Nr, .] like Renderer.__call__ takes them) are not split: pass draws through
    the renderer's seed instead."""
    if not (dist.is_available() and dist.is_initialized()):
        return renderer(rays_o, rays_d, t, parameters, cone_scale, **kwargs)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    B, Nr = rays_o.shape[0], rays_o.shape[1]
    sl = shard_rays(Nr, rank, world, block)
    per = shard_rays(Nr, 0, world, block).stop                                  # block length of every rank but the last
    out = renderer(rays_o[:, sl].contiguous(), rays_d[:, sl].contiguous(), t[:, sl].contiguous(), parameters,
                   cone_scale[:, sl].contiguous(), **kwargs)
    full: Dict[str, torch.Tensor] = {}
    n_mine = sl.stop - sl.start
    for key in sorted(out):
        v = out[key]                                                            # [B, n_mine, ...]
        tail = tuple(v.shape[2:])
        mine = torch.zeros((per, B) + tail, dtype=v.dtype, device=v.device)
        mine[:n_mine] = v.transpose(0, 1)                                       # ray axis first: blocks concatenate
        gathered = torch.empty((world * per, B) + tail, dtype=v.dtype, device=v.device)
        dist.all_gather_into_tensor(gathered, mine, group=group)
        full[key] = gathered[:Nr].transpose(0, 1).contiguous()                  # ranks' blocks tile [0, Nr) in order
    return full

"""One frame across several GPUs: rays are independent (every op of network/renderer.py:47-354 is per ray or per
sample), so each rank renders its share of the frame's rays and the only communication is the gather of the finished
per-ray outputs — ONE all-gather of equal, padded per-rank records (all output keys packed side by side) over NCCL /
NVLink on GPUs.

Two ways to split the ray axis:
  * contiguous blocks (`interleave=False`): the plain renderer, whose work per ray is constant;
  * round-robin stripes of `block` rays (`interleave=True`): the instanced renderer, whose work per ray follows the
    shell coverage of the image, so neighbouring stripes on different ranks even it out.

Every rank takes part in the same collective with the same shapes whatever it rendered: a rank whose rays all miss
the proxy (Renderer.__call__ then returns an empty dict, like renderer.py:72-90) or whose share is empty contributes
the values the single-GPU call would have scattered for missed rays (zeros; the background colour for colour keys
under composite_bkgd, renderer.py:85-87), so the gathered frame is identical to the single-GPU one.
"""

from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .train import shard_rays


def output_keys(renderer) -> List[str]:
    """Keys Renderer.__call__ produces (renderer.py:122,138): the coarse pair exists iff the renderer resamples.  The
    instanced renderers never do (renderer.py:236-245)."""
    keys = ["color_pred", "alpha_pred"]
    from .renderer import InstanceRenderer
    if getattr(renderer, "n_importance", 0) > 0 and not isinstance(renderer, InstanceRenderer):
        keys += ["color_pred_coarse", "alpha_pred_coarse"]
    return keys


def stripe_indices(n_rays: int, rank: int, world: int, block: int, device=None) -> torch.Tensor:
    """Ray indices of `rank` under round-robin stripes of `block` rays: stripe s = rays [s*block, (s+1)*block) goes
    to rank s % world."""
    r = torch.arange(n_rays, device=device)
    return r[(r // block) % world == rank]


def stripe_count(n_rays: int, rank: int, world: int, block: int) -> int:
    n_stripes = (n_rays + block - 1) // block
    mine = (n_stripes - rank + world - 1) // world if n_stripes > rank else 0
    n = mine * block
    last = n_stripes - 1
    if mine and last % world == rank:
        n -= n_stripes * block - n_rays                                        # the frame's last stripe may be short
    return n


def shard_index(n_rays: int, rank: int, world: int, block: int, interleave: bool, device=None) -> torch.Tensor:
    if interleave:
        return stripe_indices(n_rays, rank, world, block, device)
    sl = shard_rays(n_rays, rank, world, block)
    return torch.arange(sl.start, sl.stop, device=device)


_PLAN_CACHE: dict = {}


def shard_plan(n_rays: int, world: int, block: int, interleave: bool, device):
    """(per, [index tensor of every rank], restore) for one way of splitting a frame, built once per shape: `per` is the
    padded share length, `restore(gathered[world * per, ...])` puts the gathered rows back into the frame's ray order.
    Both splits are regular, so restoring is a reshape (blocks) or a permuted copy (stripes: the row of local stripe
    l of rank r is global stripe l * world + r), never an index gather."""
    key = (n_rays, world, block, bool(interleave), str(device))
    plan = _PLAN_CACHE.get(key)
    if plan is None:
        idx = [shard_index(n_rays, r, world, block, interleave, device) for r in range(world)]
        if interleave:
            n_stripes = (n_rays + block - 1) // block
            per = ((n_stripes + world - 1) // world) * block

            def restore(g, per=per):
                g = g.reshape(world, per // block, block, *g.shape[1:])
                return g.transpose(0, 1).reshape(world * per, *g.shape[3:])[:n_rays]
        else:
            per = shard_rays(n_rays, 0, world, block).stop

            def restore(g):
                return g[:n_rays]
        assert all(int(i.shape[0]) <= per for i in idx)
        if len(_PLAN_CACHE) > 16:
            _PLAN_CACHE.clear()
        plan = _PLAN_CACHE[key] = (per, idx, restore)
    return plan


def render_sharded(renderer, rays_o, rays_d, t, parameters, cone_scale, *, group=None, block: int = 16,
                   interleave: Optional[bool] = None, **kwargs) -> Dict[str, torch.Tensor]:
    """`renderer(rays_o[B,Nr,3], rays_d, t[B,Nr,2], parameters[B,p], cone_scale[B,Nr,1], **kwargs)` with the Nr axis
    split over the ranks of `group`; returns the same dict as the renderer, complete on every rank.  `interleave`
    defaults to round-robin stripes for instanced renderers and contiguous blocks otherwise.  Per-ray keyword tensors
    (z_rand, u, noise ... shaped [B*Nr, .] like Renderer.__call__ takes them) are not split: pass draws through the
    renderer's seed instead."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return renderer(rays_o, rays_d, t, parameters, cone_scale, **kwargs)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    B, Nr = rays_o.shape[0], rays_o.shape[1]
    dev = rays_o.device
    if interleave is None:
        from .renderer import InstanceRenderer
        interleave = isinstance(renderer, InstanceRenderer)
    per, all_idx, restore = shard_plan(Nr, world, block, interleave, dev)
    idx = all_idx[rank]
    n_mine = int(idx.shape[0])
    out: Dict[str, torch.Tensor] = {}
    from .renderer import InstanceRenderer
    if isinstance(renderer, InstanceRenderer) and "ray_ids" not in kwargs and n_mine > 0:
        # name the rays by their position in the whole frame: the instancer's per-ray draws then do not depend on the split
        kwargs = dict(kwargs, ray_ids=(torch.arange(B, device=dev)[:, None] * Nr + idx[None, :]).reshape(-1).to(torch.int32))
    if n_mine > 0:
        pick = (lambda a: a[:, idx[0]:idx[0] + n_mine]) if not interleave else (lambda a: a[:, idx])
        out = renderer(pick(rays_o).contiguous(), pick(rays_d).contiguous(), pick(t).contiguous(), parameters,
                       pick(cone_scale).contiguous(), **kwargs)

    keys = output_keys(renderer)
    widths = [3 if "color" in k else 1 for k in keys]
    C = sum(widths)
    # record = [per, B, C] outputs + one trailing flag "this rank's renderer returned something"
    rec = torch.zeros(per * B * C + 1, dtype=torch.float32, device=dev)
    body = rec[:-1].view(per, B, C)
    if out:
        col = 0
        for k, w in zip(keys, widths):
            body[:n_mine, :, col:col + w] = out[k].reshape(B, n_mine, w).transpose(0, 1)   # ray axis first
            col += w
        rec[-1] = 1.0
    elif kwargs.get("composite_bkgd", False):                                   # what the scatter at :85-87 leaves
        bk = torch.as_tensor([float(c) for c in kwargs.get("bkgd_color", (1., 1., 1.))], dtype=torch.float32, device=dev)
        col = 0
        for k, w in zip(keys, widths):
            if w == 3:
                body[:n_mine, :, col:col + 3] = bk
            col += w
    gathered = torch.empty((world, per * B * C + 1), dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(gathered.view(-1), rec, group=group)
    if not bool((gathered[:, -1] > 0).any()):                                   # no ray of the frame hit: renderer.py returns {}
        return {}
    full = restore(gathered[:, :-1].reshape(world * per, B, C))                 # ray order of the frame
    res: Dict[str, torch.Tensor] = {}
    col = 0
    for k, w in zip(keys, widths):
        v = full[:, :, col:col + w].transpose(0, 1).contiguous()               # [B, Nr, w]
        res[k] = v if w == 3 else v.reshape(B, Nr)
        col += w
    return res

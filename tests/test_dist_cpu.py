"""Host-side multi-process logic on CPU (gloo, world_size 2): gradient all-reduce averaging and
ray sharding.  The kernels are not involved; this covers the N > 1 plumbing of train.py."""

import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _FakeModel:
    def __init__(self, n, rank):
        self.master = torch.zeros(n, requires_grad=True)
        self.master.grad = torch.full((n,), float(rank + 1))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from nerf_tex_b200.train import Trainer, shard_rays
    tr = Trainer.__new__(Trainer)
    tr.models, tr.pg = {"model": _FakeModel(1000, rank), "model_fine": _FakeModel(10, rank)}, None
    tr.allreduce_gradients()
    g = tr.models["model"].master.grad
    sl = shard_rays(640000, rank, world, block=16)
    q.put((rank, float(g[0]), float(g[-1]), sl.start, sl.stop))
    dist.destroy_process_group()


def test_allreduce_mean_and_ray_shards_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(abs(r[1] - 1.5) < 1e-6 and abs(r[2] - 1.5) < 1e-6 for r in res)      # mean of 1 and 2
    assert res[0][3] == 0 and res[0][4] == res[1][3] and res[1][4] == 640000          # shards tile the frame


def test_shard_rays_covers_everything():
    from nerf_tex_b200.train import shard_rays
    for n, w in [(640000, 8), (4096, 3), (10, 4), (1, 8)]:
        covered = []
        for r in range(w):
            s = shard_rays(n, r, w, block=16)
            covered += list(range(s.start, s.stop))
        assert covered == list(range(n))


class _FakeRenderer:
    """Per-ray function of the inputs (CPU tensors), shaped like Renderer.__call__'s result, with its culling
    (network/renderer.py:58-88): rays with t == inf get zeros (the background colour under composite_bkgd) and a call
    in which no ray hits returns an empty dict."""

    n_importance = 0

    def __call__(self, rays_o, rays_d, t, parameters, cone_scale, composite_bkgd=False, bkgd_color=(1., 1., 1.), **kw):
        hit = t[..., 0] != float("inf")
        if not bool(hit.any()):
            return {}
        c = rays_o * 2 + rays_d + parameters[:, None, :3] + cone_scale
        c = torch.where(hit[..., None], c, torch.tensor(bkgd_color) if composite_bkgd else torch.zeros(3))
        a = torch.where(hit, t[..., 0] * t[..., 1], torch.zeros(()))
        out = {"color_pred": c, "alpha_pred": a}
        if self.n_importance > 0:
            out.update({"color_pred_coarse": c * 0.5, "alpha_pred_coarse": a * 0.5})
        return out


def _shard_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from nerf_tex_b200.sharded import render_sharded
    g = torch.Generator().manual_seed(0)
    B, Nr = 2, 1003                                       # not a multiple of the block size or the world size
    args = (torch.rand((B, Nr, 3), generator=g), torch.rand((B, Nr, 3), generator=g), torch.rand((B, Nr, 2), generator=g),
            torch.rand((B, 5), generator=g), torch.rand((B, Nr, 1), generator=g))
    got = render_sharded(_FakeRenderer(), *args, block=16)
    want = _FakeRenderer()(*args)
    ok = set(got) == set(want) and all(torch.equal(got[k], want[k]) for k in want)
    # round-robin stripes (the instanced renderer's split) and the coarse pair of a resampling renderer
    fr = _FakeRenderer()
    fr.n_importance = 128
    got2 = render_sharded(fr, *args, block=16, interleave=True)
    want2 = fr(*args)
    ok = ok and set(got2) == set(want2) and all(torch.equal(got2[k], want2[k]) for k in want2)
    # rank 1's whole contiguous block misses the proxy (t == inf): it must still join the gather
    t_miss = args[2].clone()
    t_miss[:, 512:] = float("inf")
    for bk in (False, True):
        a3 = (args[0], args[1], t_miss, args[3], args[4])
        got3 = render_sharded(_FakeRenderer(), *a3, block=16, composite_bkgd=bk, bkgd_color=(0.25, 0.5, 1.0))
        want3 = _FakeRenderer()(*a3, composite_bkgd=bk, bkgd_color=(0.25, 0.5, 1.0))
        ok = ok and set(got3) == set(want3) and all(torch.equal(got3[k], want3[k]) for k in want3)
    # nobody hits: the reference returns an empty dict
    t_none = torch.full_like(args[2], float("inf"))
    ok = ok and render_sharded(_FakeRenderer(), args[0], args[1], t_none, args[3], args[4], block=16) == {}
    # fewer rays than world * block: rank 1's share is empty
    small = tuple(a[:, :10] if a.dim() == 3 else a for a in args)
    got5 = render_sharded(_FakeRenderer(), *small, block=16)
    want5 = _FakeRenderer()(*small)
    ok = ok and all(torch.equal(got5[k], want5[k]) for k in want5)
    q.put((rank, ok, tuple(got["color_pred"].shape)))
    dist.destroy_process_group()


def test_render_sharded_gathers_the_whole_frame_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_shard_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(ok and shape == (2, 1003, 3) for _, ok, shape in res)


def test_stripe_indices_partition_the_frame():
    from nerf_tex_b200.sharded import stripe_count, stripe_indices
    for n, w, b in [(640000, 8, 256), (1003, 2, 16), (10, 4, 16), (33, 3, 16), (4096, 8, 800)]:
        seen = torch.cat([stripe_indices(n, r, w, b) for r in range(w)])
        assert sorted(seen.tolist()) == list(range(n))
        assert [stripe_count(n, r, w, b) for r in range(w)] == [int(stripe_indices(n, r, w, b).numel()) for r in range(w)]


def test_shard_plan_restore_is_the_inverse_of_the_split():
    """Every ray lands on exactly one rank and `restore` puts the gathered, padded shares back in frame order — for
    stripes and blocks, frames that are no multiple of the stripe, more ranks than stripes, a short last stripe."""
    import torch
    from nerf_tex_b200.sharded import shard_plan
    for n_rays, world, block, interleave in [(1000, 3, 16, True), (1000, 3, 16, False), (640000, 8, 256, True), (257, 8, 256, True),
                                             (256, 2, 256, True), (5, 4, 1, True), (5, 4, 2, False), (4096, 8, 16, True),
                                             (1023, 2, 64, True), (1023, 2, 64, False), (7, 8, 16, False)]:
        per, idx, restore = shard_plan(n_rays, world, block, interleave, torch.device("cpu"))
        seen = torch.cat(idx).sort().values
        assert torch.equal(seen, torch.arange(n_rays)), (n_rays, world, block, interleave)
        gathered = torch.full((world * per, 1, 2), -1.0)
        for r, i in enumerate(idx):
            assert i.shape[0] <= per
            gathered[r * per:r * per + i.shape[0], 0, 0] = i.float()
            gathered[r * per:r * per + i.shape[0], 0, 1] = 2.0 * i.float()
        full = restore(gathered)
        assert full.shape == (n_rays, 1, 2)
        assert torch.equal(full[:, 0, 0], torch.arange(n_rays).float()) and torch.equal(full[:, 0, 1], 2.0 * torch.arange(n_rays).float())

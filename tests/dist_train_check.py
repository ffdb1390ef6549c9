"""Data-parallel training check, run with torchrun on >= 2 GPUs:
per-rank ray batches (seed 1000 + rank), identical initial weights, one NCCL all-reduce of the flat
gradient buffer per model per step, Adam on every rank.
Checks: (1) the all-reduced gradient equals the mean of the per-rank gradients (gathered separately),
(2) after 5 steps the replicas' weights are bit-identical.
"""

import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from nerf_tex_b200.loss import AlphaLoss
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.train import Trainer

    models = {"model": FusedNerfTexMLP(1, 4, seed=1), "model_fine": FusedNerfTexMLP(1, 4, seed=2, name="model_fine")}
    r = Renderer(n_samples=64, n_importance=128, perturb=True, seed=1000 + rank, **models)
    tr = Trainer(models, r, AlphaLoss("network.loss.smape", "network.loss.mse"))
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    R = 1024
    ro = torch.tensor([[0., 0.5, 5.]], device=dev).repeat(R, 1)
    rd = torch.rand((R, 3), device=dev, generator=g) * 2 - 1 - ro
    rd = rd / rd.norm(dim=-1, keepdim=True)
    data = {"rays_o": ro[None], "rays_d": rd[None],
            "t": torch.stack([3 + torch.rand(R, device=dev, generator=g), 6 + torch.rand(R, device=dev, generator=g)], -1)[None],
            "parameters": torch.rand((1, 5), device=dev, generator=g), "cone_scale": torch.full((1, R, 1), 1e-3, device=dev),
            "color": torch.rand((1, R, 3), device=dev, generator=g), "alpha": torch.rand((1, R), device=dev, generator=g)}

    # (1) one step by hand: local gradients, then the trainer's all-reduce
    pred = r(**{k: data[k] for k in ("rays_o", "rays_d", "t", "parameters", "cone_scale")}, training=True)
    loss = tr.loss_fn(color_true=data["color"], alpha_true=data["alpha"], **pred)
    loss.backward()
    ok = True
    means = {}
    for name, m in models.items():
        local_g = m.master.grad.detach().clone()
        gathered = [torch.empty_like(local_g) for _ in range(world)]
        dist.all_gather(gathered, local_g)
        means[name] = (torch.stack(gathered).sum(0) / world, bool((gathered[0] != gathered[-1]).any()))
    tr.allreduce_gradients()
    for name, m in models.items():
        mean, differ = means[name]
        err = float((m.master.grad - mean).abs().max() / (mean.abs().max() + 1e-30))
        ok &= err < 1e-6 and differ
        if rank == 0:
            print(f"{name}: |allreduce - mean of per-rank grads| / max = {err:.2e}; per-rank grads differ: {differ}")
        m.master.grad = None
    # (2) five real steps, then compare replicas
    for _ in range(5):
        tr.train_step(data)
    for name, m in models.items():
        w = m.master.detach()
        ref = w.clone()
        dist.broadcast(ref, src=0)
        same = bool(torch.equal(w, ref))
        flags = [None] * world
        dist.all_gather_object(flags, same)
        ok &= all(flags)
        if rank == 0:
            print(f"{name}: replicas bit-identical after 5 steps: {all(flags)}")
    # (3) one frame sharded over the ranks (render_sharded): identical to the single-GPU render of the same rays
    from nerf_tex_b200.sharded import render_sharded
    g2 = torch.Generator(device=dev).manual_seed(7)                      # same rays on every rank
    Rf = 5003
    fo = torch.tensor([[0., 0.5, 5.]], device=dev).repeat(Rf, 1)
    fd = torch.rand((Rf, 3), device=dev, generator=g2) * 2 - 1 - fo
    fd = fd / fd.norm(dim=-1, keepdim=True)
    ft = torch.stack([3 + torch.rand(Rf, device=dev, generator=g2), 6 + torch.rand(Rf, device=dev, generator=g2)], -1)
    fp, fc = torch.rand((1, 5), device=dev, generator=g2), torch.full((1, Rf, 1), 1e-3, device=dev)
    rr = Renderer(n_samples=64, n_importance=128, perturb=False, **models)        # deterministic depths; u drawn per call
    rr.n_importance = 0                                                            # no random draws at all
    whole = rr(fo[None], fd[None], ft[None], fp, fc, composite_bkgd=True, training=False)
    parts = render_sharded(rr, fo[None], fd[None], ft[None], fp, fc, composite_bkgd=True, training=False)
    same = all(torch.equal(whole[k], parts[k]) for k in whole)
    flags = [None] * world
    dist.all_gather_object(flags, same)
    ok &= all(flags)
    if rank == 0:
        print(f"render_sharded == single-GPU render on every rank: {all(flags)}")
    # (4) ntx_nccl_allreduce_grads on a communicator of our own (raw NCCL through ctypes; the id travels over torch.distributed)
    ok &= c_abi_allreduce(rank, world, dev)
    dist.barrier()
    dist.destroy_process_group()
    if not ok:
        sys.exit(1)
    if rank == 0:
        print("DIST_TRAIN_OK")


def c_abi_allreduce(rank, world, dev) -> bool:
    import ctypes as C
    import glob
    from nerf_tex_b200 import _lib
    from nerf_tex_b200._lib import check, ptr, stream_ptr
    cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "nccl", "lib", "libnccl.so*")) + ["libnccl.so.2"]
    nccl = None
    for cand in cands:
        try:
            nccl = C.CDLL(cand, mode=C.RTLD_GLOBAL)
            break
        except OSError:
            continue
    if nccl is None:
        print("no libnccl found: skipping the C-ABI all-reduce check")
        return True
    uid = (C.c_byte * 128)()
    if rank == 0:
        assert nccl.ncclGetUniqueId(C.byref(uid)) == 0
    t = torch.tensor(list(uid), dtype=torch.int8, device=dev)
    dist.broadcast(t, 0)
    uid = (C.c_byte * 128)(*t.cpu().tolist())
    comm = C.c_void_p()

    class UniqueId(C.Structure):
        _fields_ = [("internal", C.c_byte * 128)]
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, UniqueId, C.c_int]
    assert nccl.ncclCommInitRank(C.byref(comm), world, UniqueId(uid), rank) == 0
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    grad = torch.randn(678916, device=dev, generator=g)
    gathered = [torch.empty_like(grad) for _ in range(world)]
    dist.all_gather(gathered, grad)
    want = torch.stack(gathered).sum(0) / world
    check(_lib.load().ntx_nccl_allreduce_grads(comm, ptr(grad), grad.numel(), world, stream_ptr()))
    torch.cuda.synchronize()
    err = float((grad - want).abs().max() / want.abs().max())
    nccl.ncclCommDestroy.argtypes = [C.c_void_p]
    nccl.ncclCommDestroy(comm)
    flags = [None] * world
    dist.all_gather_object(flags, bool(err < 1e-6))
    if rank == 0:
        print(f"ntx_nccl_allreduce_grads == mean of the per-rank buffers: {all(flags)} (max rel err {err:.1e})")
    return all(flags)


if __name__ == "__main__":
    main()

"""Training step around the hot path (network/train.py:44-67): renderer forward under autograd,
AlphaLoss / NerfLoss, Adam with the reference's exponential learning-rate decay, and — new, the
reference is single-device — data-parallel training: every rank renders its own ray batch and the
flat fp32 gradient buffers are summed with one NCCL all-reduce per model over NVLink before the
(identical) optimizer step on every rank.
"""

from __future__ import annotations

from typing import Any, Dict

import torch
import torch.distributed as dist

from . import util


class Trainer:
    """Owns the optimizer state of the models in ``models`` ({name: FusedNerfTexMLP})."""

    def __init__(self, models: Dict[str, Any], renderer, loss_fn, lrate: float = 5e-4, lrate_decay: float = 500,
                 process_group=None) -> None:
        self.models, self.renderer, self.loss_fn = models, renderer, loss_fn
        self.lrate, self.lrate_decay = float(lrate), float(lrate_decay)
        self.step = 0
        self.pg = process_group
        for m in models.values():
            m.requires_grad_(True)
        # tf.keras.optimizers.Adam defaults: beta (0.9, 0.999), epsilon 1e-7 (train.py:52); Keras adds
        # epsilon to sqrt(v) BEFORE the bias correction (lr_t form), unlike torch.optim.Adam, so the
        # update is written out here on the flat buffers.
        self.beta1, self.beta2, self.eps = 0.9, 0.999, 1e-7
        self.m = {k: torch.zeros_like(v.master) for k, v in models.items()}
        self.v = {k: torch.zeros_like(v.master) for k, v in models.items()}
        self.sync_replicas()

    def sync_replicas(self) -> None:
        """Data-parallel replicas must start identical (the all-reduce only averages gradients): rank 0's weights and
        Adam moments are broadcast to every rank.  Unseeded constructors (seed=None) would otherwise leave each rank
        with its own glorot draw."""
        if self.world_size() == 1:
            return
        with torch.no_grad():
            for k in sorted(self.models):
                model = self.models[k]
                for buf in (model.master, self.m[k], self.v[k]):
                    dist.broadcast(buf.detach(), 0, group=self.pg)
                model.mark_dirty()

    def current_lr(self) -> float:
        """ExponentialDecay(lrate, decay_steps=lrate_decay*1e3, decay_rate=0.1), non-staircase (train.py:49-50)."""
        if self.lrate_decay <= 0:
            return self.lrate
        return self.lrate * 0.1 ** (self.step / (self.lrate_decay * 1e3))

    def world_size(self) -> int:
        return dist.get_world_size(self.pg) if dist.is_available() and dist.is_initialized() else 1

    def train_step(self, data: Dict[str, torch.Tensor], composite_bkgd: bool = False, bkgd_color=(1., 1., 1.),
                   **draws) -> torch.Tensor:
        """One iteration of the loop at train.py:60-67.  ``data`` holds rays_o, rays_d, t, parameters,
        cone_scale, color, alpha (dataset.py:28-57)."""
        lr = self.current_lr()
        for m in self.models.values():
            m.master.grad = None
        pred = self.renderer(rays_o=data["rays_o"], rays_d=data["rays_d"], t=data["t"], parameters=data["parameters"],
                             cone_scale=data["cone_scale"], composite_bkgd=composite_bkgd, bkgd_color=bkgd_color,
                             training=True, **draws)
        loss = self.loss_fn(color_true=data["color"], alpha_true=data["alpha"], **pred)
        loss.backward()
        self.allreduce_gradients()
        self.step += 1
        self.adam_update(lr)
        return loss.detach()

    @torch.no_grad()
    def adam_update(self, lr: float) -> None:
        t = self.step
        lr_t = lr * (1.0 - self.beta2 ** t) ** 0.5 / (1.0 - self.beta1 ** t)
        for k, model in self.models.items():
            g = model.master.grad
            if g is None:
                continue
            m, v = self.m[k], self.v[k]
            from . import _lib
            from ._lib import check, ptr, stream_ptr
            with torch.cuda.device(model.master.device):  # one launch (ntx_adam_step); ptr() refuses CPU tensors
                check(_lib.load().ntx_adam_step(ptr(model.master.detach()), ptr(g.contiguous()), ptr(m), ptr(v), m.numel(),
                                                float(lr_t), self.beta1, self.beta2, self.eps, stream_ptr()))
            model.mark_dirty()                            # the packed weight images (forward and transposed) are stale

    def allreduce_gradients(self) -> None:
        """Mean of the per-rank gradients: one all-reduce per model on its flat buffer."""
        world = self.world_size()
        if world == 1:
            return
        for m in self.models.values():
            g = m.master.grad
            if g is None:
                g = m.master.grad = torch.zeros_like(m.master)
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=self.pg)
            g.div_(world)


def shard_rays(n_rays: int, rank: int, world: int, block: int = 1) -> slice:
    """Contiguous ray block of `rank` for ray-sharded rendering (rays are independent units;
    no collective is needed until the final image gather)."""
    per = (n_rays + world - 1) // world
    per = ((per + block - 1) // block) * block
    lo = min(rank * per, n_rays)
    return slice(lo, min(lo + per, n_rays))


def Train(target_path: str, train_dataset_config, val_dataset_config, model_config, loss_config, n_iters: int,
          lrate: float, lrate_decay: float, renderer_config, logger_config=None, train_dataset=None, **kwargs):
    """Signature of network/train.py:7-18.  Dataset loading, logging and checkpointing are outside
    the hot path: pass an iterable ``train_dataset`` of ray-batch dicts with
    ``n_parameters`` / ``composite_bkgd`` / ``bkgd_color`` attributes."""
    if train_dataset is None:
        raise NotImplementedError("the tf.data pipeline of network/dataset.py is out of scope; pass train_dataset=")
    model_config = util.EasyDict(model_config)
    model_config.setdefault("n_parameters", getattr(train_dataset, "n_parameters", None))   # train.py:30
    models = util.instantiate(model_config)
    renderer_config = util.EasyDict(renderer_config)
    renderer_config.update(models)                                                           # train.py:42
    renderer = util.instantiate(renderer_config)
    loss_fn = util.instantiate(loss_config)
    trainer = Trainer(models, renderer, loss_fn, lrate, lrate_decay)
    logger = None
    if logger_config is not None:                                                            # train.py:54-57
        logger_config = util.EasyDict(logger_config)
        val_dataset = util.instantiate(val_dataset_config) if isinstance(val_dataset_config, dict) else val_dataset_config
        logger_config.update({"target_path": target_path, "checkpoint_variables": dict(models, step=0, optimizer=trainer),
                              "dataset": val_dataset, "renderer": renderer, "n_iters": n_iters})
        logger = util.instantiate(logger_config)                                             # restores the latest checkpoint
    losses = []
    for i, data in enumerate(train_dataset):
        if trainer.step >= n_iters:                                                          # train.py:60
            break
        losses.append(trainer.train_step(data, getattr(train_dataset, "composite_bkgd", False),
                                         getattr(train_dataset, "bkgd_color", (1., 1., 1.))))
        if logger is not None:
            logger({"loss": losses[-1]})                                                     # train.py:69
    return trainer, losses

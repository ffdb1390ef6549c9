"""Logger with the reference's constructor (network/logger.py:10-56) for what sits directly after the renderer:
full-frame rendering of a dataset into images (render_images / render_image / write_image, :88-144) and
checkpoints of the flat weight buffers.  TensorBoard summaries are out of scope.

Checkpoints are plain ``.npz`` files (``checkpoints/ckpt-<step>.npz``: one flat fp32 array per model, Adam
moments if a trainer is attached, and the step).  If no such file exists but the directory holds a checkpoint
written by the reference (TensorFlow tensor bundle), its Dense kernels / biases and step are read through
``tf_checkpoint`` (see the status note there)."""

from __future__ import annotations

import glob
import os
import re
import time
from typing import Any, Dict, List, Optional

import numpy as np
import torch

from . import image, util


class Logger:
    def __init__(self, target_path: str, checkpoint_variables: dict, source_path: str = None, dataset: Any = None,
                 is_training: bool = True, renderer: Any = None, n_iters: int = 5e5, i_summary: int = 10, i_print: int = 100,
                 i_img: int = 5e3, i_checkpoint: int = 1e3, max_to_keep: int = 3, keep_every_n_hours: int = 12,
                 write_exr: bool = False, downsampling_factor: int = 1, **kwargs) -> None:
        if write_exr:
            raise NotImplementedError("Logger: OpenEXR output (pyexr) is not available here; PNG only")
        self.target_path = target_path
        self.source_path = source_path if source_path is not None else target_path      # :15
        self.dataset, self.is_training, self.renderer = dataset, is_training, renderer
        self.n_iters, self.i_print, self.i_img, self.i_checkpoint = int(n_iters), int(i_print), int(i_img), int(i_checkpoint)
        self.max_to_keep = int(max_to_keep)
        self.downsampling_factor = int(downsampling_factor)
        self.write_exr = False
        self.models = {k: v for k, v in checkpoint_variables.items() if hasattr(v, "master")}
        self.trainer = checkpoint_variables.get("optimizer", None)                      # a train.Trainer, if any
        self.step = int(checkpoint_variables.get("step", 0) or 0)
        self.time_print = time.perf_counter()
        self.checkpoint_path = os.path.join(self.source_path, "checkpoints")           # :31
        os.makedirs(self.checkpoint_path, exist_ok=True)
        latest = self.latest_checkpoint()
        if latest:                                                                     # :38-40
            self.restore(latest)
            print("Restored model & optimizer from {}.".format(latest))
        else:
            from . import tf_checkpoint
            tf_prefix = tf_checkpoint.latest_checkpoint(self.checkpoint_path)          # a checkpoint the reference wrote
            if tf_prefix and os.path.exists(tf_prefix + ".index"):
                for k, m in self.models.items():
                    tf_checkpoint.load_model_weights(tf_prefix, m, k)
                step = tf_checkpoint.read_step(tf_prefix)
                self.step = self.step if step is None else step
                if self.trainer is not None:
                    self.trainer.step = self.step                                      # Adam moments restart from zero
                print("Restored model weights from the TensorFlow checkpoint {}.".format(tf_prefix))
        self.imgs_path = os.path.join(self.target_path, "media/validation" if is_training else "media/test")   # :48,:52
        os.makedirs(self.imgs_path, exist_ok=True)
        if not is_training:
            self.render_images(self.imgs_path)                                         # :56

    # -- checkpoints --------------------------------------------------------------------------------------------
    def latest_checkpoint(self) -> Optional[str]:
        files = glob.glob(os.path.join(self.checkpoint_path, "ckpt-*.npz"))
        key = lambda f: int(re.search(r"ckpt-(\d+)\.npz$", f).group(1))
        return max(files, key=key) if files else None

    def save(self, checkpoint_number: int) -> str:
        data = {"step": np.int64(self.step)}
        for k, m in self.models.items():
            data[k] = m.master.detach().cpu().numpy()
            if self.trainer is not None and k in getattr(self.trainer, "m", {}):
                data[k + "/adam_m"] = self.trainer.m[k].cpu().numpy()
                data[k + "/adam_v"] = self.trainer.v[k].cpu().numpy()
        path = os.path.join(self.checkpoint_path, "ckpt-{}.npz".format(int(checkpoint_number)))
        np.savez(path, **data)
        files = sorted(glob.glob(os.path.join(self.checkpoint_path, "ckpt-*.npz")),
                       key=lambda f: int(re.search(r"ckpt-(\d+)\.npz$", f).group(1)))
        for f in files[:-self.max_to_keep]:                                            # :35 max_to_keep
            os.remove(f)
        return path

    def restore(self, path: str) -> None:
        with np.load(path) as data:
            self.step = int(data["step"])
            for k, m in self.models.items():
                if k not in data:
                    continue                                                           # expect_partial (:38)
                assert data[k].shape[0] == m.n_params, f"checkpoint {path}: {k} has {data[k].shape[0]} weights, model {m.n_params}"
                with torch.no_grad():
                    m.master.copy_(torch.from_numpy(data[k]))
                m.mark_dirty()
                if self.trainer is not None and k + "/adam_m" in data:
                    self.trainer.m[k].copy_(torch.from_numpy(data[k + "/adam_m"]))
                    self.trainer.v[k].copy_(torch.from_numpy(data[k + "/adam_v"]))
        if self.trainer is not None:
            self.trainer.step = self.step

    # -- training-loop hook (:57-86) ----------------------------------------------------------------------------
    def __call__(self, loss: Dict[str, Any]) -> None:
        self.step += 1
        if self.step % self.i_print == 0:                                              # :68-73
            msg = "Step {}".format(self.step)
            for key, value in loss.items():
                msg += " | {} {:.3g}".format(key, float(value))
            print(msg + " | Duration {:.3g}".format(time.perf_counter() - self.time_print))
            self.time_print = time.perf_counter()
        if self.step % self.i_img == 0 and self.dataset is not None:                   # :76-78
            print("Rendering validation images.")
            self.render_images(os.path.join(self.imgs_path, util.format_name("", self.step, self.n_iters, "")), True)
        if self.step % self.i_checkpoint == 0:                                         # :84-86
            print("Saved checkpoint to {}.".format(self.save(self.step)))

    # -- images (:88-144) ---------------------------------------------------------------------------------------
    def render_images(self, imgs_path: str, return_imgs: bool = False) -> Optional[List[torch.Tensor]]:
        os.makedirs(imgs_path, exist_ok=True)
        try:
            max_idx = len(self.dataset)
        except TypeError:
            max_idx = 256                                                              # :94-95
        imgs = []
        for i, data in enumerate(self.dataset):
            img_path = os.path.join(imgs_path, util.format_name("", i, max_idx, ".png"))   # :106
            img = self.render_image(data, img_path)
            if return_imgs:
                imgs.append(img)
        return imgs if return_imgs else None

    def render_image(self, data: dict, img_path: Optional[str] = None) -> torch.Tensor:
        with torch.no_grad():
            pred = self.renderer(**data, composite_bkgd=self.dataset.composite_bkgd, bkgd_color=self.dataset.bkgd_color,
                                 training=False)                                       # :123
        h, w = self.dataset.height, self.dataset.width
        img, u8 = image.finish_image(pred["color_pred"], pred["alpha_pred"], h, w, self.downsampling_factor,
                                     straight_alpha=True, want_u8=img_path is not None)   # :126-135
        if img_path is not None:
            with open(img_path, "wb") as f:                                            # :144
                f.write(image.encode_png(u8))
        return img

"""Model constructors with the reference's signatures (network/model.py).

Each constructor returns ``{name: model}`` like the reference (``model.py:45,56,125``) so that
``renderer_config.update(model)`` works unchanged.  A model is a ``FusedNerfTexMLP``: flat fp32
master weights on the GPU + the packed bf16 image the sm_100a kernel consumes; calling it runs
``ntx_mlp_forward`` (encoding + all Dense layers in one kernel).
"""

from __future__ import annotations

import ctypes as C
import math
from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch

from . import _lib, util
from ._lib import MlpConfig, MlpInputs, as_cuda, check, ptr, stream_ptr
from .layer import FourierFeatures, IntegratedPositionalEncoding  # noqa: F401  (configs name network.model.<layer>)


class FusedNerfTexMLP:
    """Callable ``((pos, dir, params), training=) -> (color[N,3], alpha[N,1])`` (raw outputs,
    network/model.py:111,123), exposing ``trainable_variables`` like a Keras model (train.py:34)."""

    def __init__(self, n_geo: int, n_app: int, depth: int = 8, width: int = 256, skips: Sequence[int] = (4,),
                 color_depth: int = 1, pos_freqs: int = 10, dir_freqs: int = 4, param_freqs: int = 4, n_pos: int = 3,
                 name: str = "model", seed: Optional[int] = None, device=None) -> None:
        self.name = name
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        skip_mask = 0
        for s in skips:
            skip_mask |= 1 << int(s)
        self.cfg = MlpConfig(n_pos, pos_freqs, dir_freqs, param_freqs, n_geo, n_app, depth, width, skip_mask, color_depth)
        self.n_parameters = [n_geo, n_app]
        lib = _lib.load()
        h = C.c_void_p()
        check(lib.ntx_mlp_create(C.byref(self.cfg), C.byref(h)))
        self._h = h
        self.n_params = lib.ntx_mlp_num_params(h)
        self.layers: List[Tuple[int, int, int, int]] = []          # (w_off, b_off, fan_in, fan_out)
        for i in range(lib.ntx_mlp_num_layers(h)):
            wo, bo, fi, fo = C.c_int64(), C.c_int64(), C.c_int32(), C.c_int32()
            check(lib.ntx_mlp_layer_info(h, i, C.byref(wo), C.byref(bo), C.byref(fi), C.byref(fo)))
            self.layers.append((wo.value, bo.value, fi.value, fo.value))
        names = [f"trunk{i}" for i in range(depth)] + ["alpha", "feature"] + [f"colorhid{i}" for i in range(color_depth)]
        self.layer_names = names + ["colorhalf", "rgb"]
        with torch.cuda.device(self.device):
            self.master = torch.zeros(self.n_params, dtype=torch.float32, device=self.device)
            self.packed = torch.empty(lib.ntx_mlp_packed_bytes(h), dtype=torch.uint8, device=self.device)
        self._packed_version = -1
        self.reset_parameters(seed)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.load().ntx_mlp_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # -- parameters -------------------------------------------------------------------------
    def reset_parameters(self, seed: Optional[int] = None) -> None:
        """Keras Dense defaults: glorot_uniform kernels, zero biases."""
        gen = torch.Generator(device="cpu")
        if seed is not None:
            gen.manual_seed(int(seed))
        flat = torch.zeros(self.n_params, dtype=torch.float32)
        for wo, bo, fi, fo in self.layers:
            lim = math.sqrt(6.0 / (fi + fo))
            flat[wo:wo + fi * fo] = (torch.rand(fi * fo, generator=gen) * 2 - 1) * lim
        with torch.no_grad():                                # master may already be a trainable leaf (Trainer)
            self.master.copy_(flat)
        self.mark_dirty()

    def load_named_weights(self, named: Dict[str, Tuple]) -> None:
        """``{layer_name: (W[in,out], b[out])}`` in the reference's Dense order / concat order."""
        flat = torch.zeros(self.n_params, dtype=torch.float32)
        for name, (wo, bo, fi, fo) in zip(self.layer_names, self.layers):
            W, b = named[name]
            W, b = torch.as_tensor(W, dtype=torch.float32), torch.as_tensor(b, dtype=torch.float32)
            assert tuple(W.shape) == (fi, fo) and tuple(b.shape) == (fo,), (name, W.shape, (fi, fo))
            flat[wo:wo + fi * fo] = W.reshape(-1)
            flat[bo:bo + fo] = b
        with torch.no_grad():                                # Logger restores checkpoints after Trainer made master a leaf
            self.master.copy_(flat)
        self.mark_dirty()

    @property
    def trainable_variables(self) -> List[torch.Tensor]:
        out = []
        for wo, bo, fi, fo in self.layers:
            out.append(self.master[wo:wo + fi * fo].view(fi, fo))
            out.append(self.master[bo:bo + fo])
        return out

    def schedule(self):
        """Host mirror of the kernel's layer schedule (used by the backward pass)."""
        if getattr(self, "_schedule", None) is None:
            from .autograd import MlpSchedule
            self._schedule = MlpSchedule(self)
        return self._schedule

    def requires_grad_(self, flag: bool = True):
        """Make the flat master buffer a trainable leaf (its .grad is filled by loss.backward())."""
        self.master.requires_grad_(flag)
        return self

    def enable_fp8_saved_operands(self, enable: bool = True):
        """Training: keep the saved tile images of the backward in fp8 e4m3 (half the bytes of the step's three big
        kernels); see autograd.enable_fp8_saved_operands."""
        from .autograd import enable_fp8_saved_operands
        enable_fp8_saved_operands(self, enable)
        return self

    def mark_dirty(self) -> None:
        """Call after changing ``master`` in place (optimizer step, checkpoint load)."""
        self._packed_version = -1
        self._packed_t_version = None                    # transposed image of the backward (autograd.py)

    def ensure_packed(self) -> torch.Tensor:
        if self._packed_version != self.master._version:
            check(_lib.load().ntx_mlp_pack_weights(self._h, ptr(self.master.detach()), ptr(self.packed), stream_ptr()))
            self._packed_version = self.master._version
        return self.packed

    # -- forward ----------------------------------------------------------------------------
    def forward_rows(self, *, n_rows: int, dir: torch.Tensor, params: Optional[torch.Tensor], rows_per_group: int,
                     pos: Optional[torch.Tensor] = None, rays_o=None, rays_d=None, z=None, row_scale=None,
                     group_scale=None, blur_idx: Optional[int] = None, out: Optional[torch.Tensor] = None,
                     pos_cov: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Raw rgbs[N,4] for N rows; see ``ntx_mlp_inputs`` in include/nerftex_b200.h."""
        packed = self.ensure_packed()
        if out is None:
            out = torch.empty((n_rows, 4), dtype=torch.float32, device=self.device)
        inp = MlpInputs(ptr(pos), ptr(rays_o), ptr(rays_d), ptr(z), ptr(dir), ptr(params), ptr(row_scale),
                        ptr(group_scale), -1 if blur_idx is None else int(blur_idx), int(rows_per_group), int(n_rows),
                        ptr(pos_cov))
        check(_lib.load().ntx_mlp_forward(self._h, ptr(packed), C.byref(inp), ptr(out), stream_ptr()))
        return out

    def __call__(self, inputs, training: bool = False):
        pos, dirs, params = inputs
        pos, dirs = as_cuda(pos, device=self.device), as_cuda(dirs, device=self.device)
        n = pos.shape[0]
        npar = sum(self.n_parameters)
        params = as_cuda(params, device=self.device) if npar > 0 else None
        if params is not None:
            assert params.shape == (n, npar), f"params must be [N,{npar}]"
        cov = None
        if self.cfg.n_pos == 6:                                  # [mean, diagonal covariance] (renderer.py:390)
            assert pos.shape[1] == 6, "an n_pos = 6 model takes pos[N,6] = [mean, cov_diag]"
            pos, cov = pos[:, :3].contiguous(), pos[:, 3:].contiguous()
        rgbs = self.forward_rows(n_rows=n, pos=pos, pos_cov=cov, dir=dirs, params=params, rows_per_group=1)
        return rgbs[:, :3], rgbs[:, 3:4]


def _freqs(cfg, default: int, kind: str = "FourierFeatures") -> int:
    """n_freq_bands of an embedding config (dict naming network.layer.* or an instance); checks its class."""
    if cfg is None:
        return default
    if isinstance(cfg, dict):
        name = util.resolve_path(cfg.get("module", "")).split(".")[-1]
        n = int(cfg["n_freq_bands"])
    else:
        name, n = type(cfg).__name__, int(cfg.n_freq_bands)
    if name != kind:
        raise NotImplementedError(f"embedding {name} is not implemented here on the sm_100a path (expected {kind})")
    return n


def ParamNerf(pos_embedding, dir_embedding, param_embedding, n_parameters: Union[int, list], n_pos: int = 3,
              param_depth: int = 0, param_width: int = 128, depth: int = 8, width: int = 256, skips: list = [4],
              color_depth: int = 1, embedding_config=None, include_param_dims: bool = False, name: str = "model",
              seed: Optional[int] = None, **kwargs) -> dict:
    """network/model.py:58-125."""
    if isinstance(n_parameters, int):
        n_parameters = [n_parameters, 0]                     # model.py:63-64
    if param_depth != 0:
        raise NotImplementedError("param_depth > 0 (extra parameter MLPs, model.py:90-91,98-99) is unused by every "
                                  "shipped config and not implemented in the fused kernel")
    if embedding_config is not None:
        raise NotImplementedError("embedding_config names an embedding class that does not exist in the reference")
    if n_pos not in (3, 6):
        raise NotImplementedError("n_pos must be 3 (FourierFeatures) or 6 (IntegratedPositionalEncoding of mean + variance)")
    pos_kind = "IntegratedPositionalEncoding" if n_pos == 6 else "FourierFeatures"     # layer.py:8-41
    m = FusedNerfTexMLP(n_geo=int(n_parameters[0]), n_app=int(n_parameters[1]), depth=depth, width=width, skips=skips,
                        color_depth=color_depth, pos_freqs=_freqs(pos_embedding, 10, pos_kind), dir_freqs=_freqs(dir_embedding, 4),
                        param_freqs=_freqs(param_embedding, 4), n_pos=n_pos, name=name, seed=seed)
    return {name: m}


def Nerf(pos_embedding, dir_embedding, depth: int = 8, width: int = 256, skips: list = [4], name: str = "model",
         seed: Optional[int] = None, **kwargs) -> dict:
    """network/model.py:9-45: ParamNerf without parameters and without the extra colour layer."""
    m = FusedNerfTexMLP(n_geo=0, n_app=0, depth=depth, width=width, skips=skips, color_depth=0,
                        pos_freqs=_freqs(pos_embedding, 10), dir_freqs=_freqs(dir_embedding, 4), name=name, seed=seed)
    return {name: m}


def CoarseFine(model_config, **kwargs) -> dict:
    """network/model.py:47-56: the same architecture twice, named ``model`` and ``model_fine``."""
    model_config = util.EasyDict(model_config)
    for key, value in kwargs.items():
        model_config.setdefault(key, value)
    seed = model_config.get("seed", None)
    model_coarse = util.instantiate(model_config)
    model_config["name"] = next(iter(model_coarse)) + "_fine"
    if seed is not None:
        model_config["seed"] = seed + 1
    model_fine = util.instantiate(model_config)
    return dict(model_coarse, **model_fine)

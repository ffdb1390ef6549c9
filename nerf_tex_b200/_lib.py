"""ctypes binding of libnerftex_b200.so — the only way Python reaches the CUDA kernels.

There is no fallback: if the shared library is missing, ``load()`` raises with the build command;
if a call fails, ``NtxError`` carries ``ntx_last_error()``.  Tensors cross the boundary as raw
device pointers taken from torch CUDA tensors (torch is plumbing: allocation, streams, DLPack
import of foreign arrays via ``as_cuda``).
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

import torch

_PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("NTX_LIB", _PKG / "libnerftex_b200.so"))

_lib: Optional[C.CDLL] = None


class NtxError(RuntimeError):
    pass


class MlpConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_pos", "pos_freqs", "dir_freqs", "param_freqs", "n_geo", "n_app", "depth",
                                          "width", "skip_mask", "color_depth")]


class RenderConfig(C.Structure):
    _fields_ = [("n_samples", C.c_int32), ("n_importance", C.c_int32), ("perturb", C.c_int32), ("map_exr", C.c_int32),
                ("composite_bkgd", C.c_int32), ("blur_idx", C.c_int32), ("bkgd_color", C.c_float * 3)]


class MlpInputs(C.Structure):
    _fields_ = [("pos", C.c_void_p), ("rays_o", C.c_void_p), ("rays_d", C.c_void_p), ("z", C.c_void_p),
                ("dir", C.c_void_p), ("params", C.c_void_p), ("row_scale", C.c_void_p), ("group_scale", C.c_void_p),
                ("blur_idx", C.c_int32), ("rows_per_group", C.c_int32), ("n_rows", C.c_int64), ("pos_cov", C.c_void_p)]


class SceneGrid(C.Structure):
    _fields_ = [("cell_start", C.c_void_p), ("items", C.c_void_p), ("lo", C.c_float * 3), ("cell", C.c_float * 3),
                ("res", C.c_int32 * 3)]


class Texture(C.Structure):
    _fields_ = [("offset", C.c_int64), ("rows", C.c_int32), ("cols", C.c_int32)]


MAX_PARAM_TEXTURES = MAX_AUX_MESHES = 8


class InstancerShading(C.Structure):
    _fields_ = [("pool", C.c_void_p), ("mesh_uv", C.c_void_p), ("mesh_n", C.c_void_p), ("n_base_faces", C.c_int32),
                ("n_param_tex", C.c_int32), ("param_idx", C.c_int32 * MAX_PARAM_TEXTURES),
                ("param_tex", Texture * MAX_PARAM_TEXTURES), ("radius", C.c_float), ("min_samples", C.c_int32),
                ("n_samples", C.c_int32), ("n_aux", C.c_int32), ("aux_face_begin", C.c_int32 * (MAX_AUX_MESHES + 1)),
                ("aux_channels", C.c_int32 * MAX_AUX_MESHES), ("aux_tex", (Texture * 3) * MAX_AUX_MESHES)]


class InstancerScene(C.Structure):
    _fields_ = [("minv", C.c_void_p), ("dirt", C.c_void_p), ("origins", C.c_void_p), ("spheres", C.c_void_p),
                ("n_inst", C.c_int32), ("n_faces", C.c_int32), ("mesh_v", C.c_void_p), ("mesh_f", C.c_void_p),
                ("inst_grid", SceneGrid), ("tri_grid", SceneGrid), ("shading", C.POINTER(InstancerShading))]


# name -> (restype, argtypes); must list every symbol include/nerftex_b200.h declares.
_P, _I, _F, _I64, _U32 = C.c_void_p, C.c_int, C.c_float, C.c_int64, C.c_uint32
SIGNATURES = {
    "ntx_version": (_I, []),
    "ntx_last_error": (C.c_char_p, []),
    "ntx_stratified_z": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "ntx_sample_pdf": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P]),
    "ntx_hierarchical_z": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P]),
    "ntx_composite_forward": (_I, [_P, _P, _I, _P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P]),
    "ntx_composite_backward": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P]),
    "ntx_composite_instanced_forward": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _F, _F, _I, _I, _P, _P, _P, _P]),
    "ntx_mlp_create": (_I, [C.POINTER(MlpConfig), C.POINTER(_P)]),
    "ntx_mlp_destroy": (_I, [_P]),
    "ntx_mlp_num_params": (_I64, [_P]),
    "ntx_mlp_num_layers": (_I, [_P]),
    "ntx_mlp_layer_info": (_I, [_P, _I, C.POINTER(_I64), C.POINTER(_I64), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "ntx_mlp_packed_bytes": (_I64, [_P]),
    "ntx_mlp_pack_weights": (_I, [_P, _P, _P, _P]),
    "ntx_mlp_forward": (_I, [_P, _P, C.POINTER(MlpInputs), _P, _P]),
    "ntx_mlp_train_tiles": (_I64, [_I64]),
    "ntx_mlp_forward_train": (_I, [_P, _P, C.POINTER(MlpInputs), _P, _P, _P, _P, _P, _P, _I64, _P]),
    "ntx_untile": (_I, [_P, _I64, _I, _I64, _P, _P]),
    "ntx_mlp_packed_t_bytes": (_I64, [_P]),
    "ntx_mlp_pack_weights_t": (_I, [_P, _P, _P, _P]),
    "ntx_mlp_backward_data": (_I, [_P, _P, _P, _P, _P, _I64, _I64, _P]),
    "ntx_mlp_backward_weights": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64, _P]),
    "ntx_mlp_forward_train_fp8": (_I, [_P, _P, C.POINTER(MlpInputs), _P, _P, _P, _P, _P, _P, _P, _I64, _P]),
    "ntx_mlp_backward_data_fp8": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I64, _I64, _P]),
    "ntx_mlp_backward_weights_fp8": (_I, [_P] * 11 + [_I64, _I64, _P]),
    "ntx_fp8_update_scales": (_I, [_P, _P, _P, _I, _F, _P]),
    "ntx_mlp_num_steps": (_I, [_P]),
    "ntx_mlp_step_info": (_I, [_P, _I] + [C.POINTER(C.c_int32)] * 6),
    "ntx_mlp_enc_columns": (_I, [_P, _I, C.POINTER(C.c_int32)]),
    "ntx_rays_from_camera": (_I, [_P, _I64, _I, _I, _F, _P, _I, _P, _P, _F, _F, _P, _P, _P, _P, _P]),
    "ntx_normalize_dirs": (_I, [_P, _I, _P, _P]),
    "ntx_alpha_loss": (_I, [_P] * 6 + [_I64, _I, _I, _F, _I, _I] + [_P] * 5 + [_P]),
    "ntx_adam_step": (_I, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _P]),
    "ntx_nccl_allreduce_grads": (_I, [_P, _P, _I64, _I, _P]),
    "ntx_mip_cone_gaussians": (_I, [_P, _P, _P, _P, _I, _I, _P, _P, _P]),
    "ntx_mip_cone_gaussians_instanced": (_I, [_P, _P, _P, _P, _I64, _P, _P]),
    "ntx_composite_edges_forward": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P]),
    "ntx_composite_edges_backward": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P]),
    "ntx_image_finish": (_I, [_P, _P, _I, _I, _I, _P, _I, _I, _P, _P, _P]),
    "ntx_instancer_workspace_bytes": (_I64, [_I, _I]),
    "ntx_instancer_trace": (_I, [_P, _P, _P, _P, _I, C.POINTER(InstancerScene), _P, _P, _F, _I, _U32, _I, _I, _P, _I, _I,
                                 _P, _P, _P, _P, _P, _P]),
    "ntx_instancer_march": (_I, [_P, _P, _P, _P, _I, _I, C.POINTER(InstancerScene), _F, _I, _U32, _I, _I, _I, _F, _I,
                                 _I, _I, _I, _P, _P, _P, _P, _I64] + [_P] * 7 + [_P]),
    "ntx_render_rays_workspace_bytes": (_I64, [_I, _I, _I]),
    "ntx_render_rays": (_I, [_P, _P, _P, _P, C.POINTER(RenderConfig)] + [_P] * 9 + [_I] + [_P] * 4 + [_P, _I64, _P]),
    "ntx_render_rays_static": (_I, [_P, _P, _P, _P, C.POINTER(RenderConfig)] + [_P] * 9 + [_I] + [_P] * 4 + [_P]),
    "ntx_selftest_umma": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "ntx_selftest_umma_mn": (_I, [_P, _P, _I, _I, _P, _P]),
    "ntx_selftest_umma_mn8": (_I, [_P, _P, _I, _I, _P, _P]),
    "ntx_selftest_bulk_copy": (_I, [_P, _P, _I, _P]),
}


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise NtxError(f"{LIB_PATH} not found: the CUDA library is required (no CPU fallback). "
                       f"Build it with `python -m nerf_tex_b200._build`.")
    lib = C.CDLL(str(LIB_PATH))
    missing = [name for name in SIGNATURES if not hasattr(lib, name)]
    if missing:                                           # a stale prebuilt library must fail here, not at first use
        raise NtxError(f"{LIB_PATH} does not export {', '.join(missing)}: rebuild it with "
                       f"`python -m nerf_tex_b200._build -f`")
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().ntx_last_error().decode("utf-8", "replace")
        raise NtxError(f"libnerftex_b200 error {rc}: {msg}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    """Device pointer of a contiguous CUDA tensor (None passes NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise NtxError("expected a CUDA tensor: libnerftex_b200 has no CPU path")
    if not t.is_contiguous():
        raise NtxError("expected a contiguous tensor (the C ABI never copies silently)")
    return t.data_ptr()


def as_cuda(x, dtype=torch.float32, device=None) -> torch.Tensor:
    """Accept a torch tensor or any DLPack exporter (CuPy, JAX, NumPy ...) and return a
    contiguous CUDA tensor of `dtype` on `device` (default: current CUDA device)."""
    if not isinstance(x, torch.Tensor):
        if hasattr(x, "__dlpack__"):
            x = torch.from_dlpack(x)
        else:
            x = torch.as_tensor(x)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if x.device != dev or x.dtype != dtype:
        x = x.to(device=dev, dtype=dtype)
    return x.contiguous()


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream

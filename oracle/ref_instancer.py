"""TEST INFRASTRUCTURE ONLY.  The reference's own instancer (instancer/src/instancer.cpp) compiled here against the
Embree stand-in of oracle/embree_shim/ and driven through ctypes — the counterpart of instancer/instancer.pyx.

    build()                        g++ on the reference's source where it lies under /root/reference -> oracle/_ref/
    RefInstancer(...)              constructor arguments of instancer.pyx:9
    .get_model_input(o, d, p, S, step)   the ten arrays of instancer.pyx:38-54 as NumPy (dense [R, S, ...] form)

Only tests/ and tests/golden/make_ref_instancer.py may use this; /root/reference does not exist on the GPU box, the
built library travels there like any other prebuilt .so (oracle/_ref is git-ignored, not gpurun-ignored)."""

from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REF = Path("/root/reference/instancer")
LIB = HERE / "_ref" / "libref_instancer.so"


def build(force: bool = False) -> Path:
    """Compile oracle/ref_instancer_capi.cpp (which includes the reference's instancer.cpp from /root/reference) with the
    include directories of the reference's instancer/setup.py, Embree's replaced by oracle/embree_shim."""
    src = HERE / "ref_instancer_capi.cpp"
    shim = HERE / "embree_shim" / "embree3" / "rtcore.h"
    if LIB.exists() and not force:
        if not REF.exists() or LIB.stat().st_mtime >= max(src.stat().st_mtime, shim.stat().st_mtime):
            return LIB
    if not REF.exists():
        raise FileNotFoundError("the reference tree is not here and oracle/_ref/libref_instancer.so was not prebuilt")
    LIB.parent.mkdir(exist_ok=True)
    inc = [HERE / "embree_shim", REF, REF / "submodules/eigen", REF / "submodules/libigl/include", REF / "submodules/stb",
           REF / "submodules/json/single_include/nlohmann"]
    cmd = ["g++", "-O2", "-std=c++17", "-w", "-fPIC", "-shared", "-ffp-contract=off"]
    for i in inc:
        cmd += ["-I", str(i)]
    cmd += ["-o", str(LIB), str(src)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("building the reference instancer failed:\n" + res.stderr[-4000:])
    return LIB


def available() -> bool:
    return LIB.exists() or REF.exists()


_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = C.CDLL(str(build()))
        P, F, I, U = C.c_void_p, C.c_float, C.c_int, C.c_uint
        lib.ref_instancer_new.restype = P
        lib.ref_instancer_new.argtypes = [P, P, I, P, I, I, I, I, I, F, I, I, U]
        lib.ref_instancer_delete.argtypes = [P]
        lib.ref_instancer_add_instance.argtypes = [P, P]
        lib.ref_instancer_distribute.restype = I
        lib.ref_instancer_distribute.argtypes = [P, C.c_char_p, F, C.c_char_p]
        lib.ref_instancer_add_mesh.restype = I
        lib.ref_instancer_add_mesh.argtypes = [P, C.c_char_p, C.c_char_p]
        lib.ref_instancer_commit.argtypes = [P]
        lib.ref_instancer_count.restype = U
        lib.ref_instancer_count.argtypes = [P]
        lib.ref_instancer_export.argtypes = [P, C.c_char_p]
        lib.ref_instancer_get_model_input.argtypes = [P] * 11 + [U, U, F, P]
        _lib = lib
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class RefInstancer:
    """instancer.pyx:6-36 (same arguments, same order of AddInstance / DistributeInstancesOnMesh / AddMesh / CommitScene)."""

    METHODS = {"random": 0, "nearest": 1, "nearest_blend": 2}

    def __init__(self, b_0, b_1, cast_shadow_rays=False, textures=(), transformations=(), mesh_path=None, patch_scale=1.0,
                 patch_origins_path="", min_shadow_samples=4, n_shadow_samples=512, min_texture_samples=4,
                 n_texture_samples=512, jitter_amount=0.0, instance_sampling_method="random", use_mean_distance=False,
                 auxiliary_meshes=(), transformation_export_path=None, seed=0):
        lib = _load()
        b0, b1 = np.asarray(b_0, np.float32), np.asarray(b_1, np.float32)
        tex = [t.encode() for t in textures]
        arr = (C.c_char_p * max(len(tex), 1))(*tex)
        self.n_parameters = None
        self._h = lib.ref_instancer_new(_ptr(b0), _ptr(b1), int(cast_shadow_rays), C.cast(arr, C.c_void_p), len(tex),
                                        min_shadow_samples, n_shadow_samples, min_texture_samples, n_texture_samples,
                                        float(jitter_amount), self.METHODS[instance_sampling_method], int(use_mean_distance), seed)
        for t in transformations:
            m = np.ascontiguousarray(np.asarray(t, np.float32).reshape(16))
            lib.ref_instancer_add_instance(self._h, _ptr(m))
        if mesh_path is not None:
            if lib.ref_instancer_distribute(self._h, str(mesh_path).encode(), float(patch_scale), str(patch_origins_path).encode()):
                raise RuntimeError("DistributeInstancesOnMesh threw")
            if transformation_export_path is not None:
                lib.ref_instancer_export(self._h, str(transformation_export_path).encode())
        for mp, tp in auxiliary_meshes:
            if lib.ref_instancer_add_mesh(self._h, str(mp).encode(), str(tp).encode()):
                raise RuntimeError("AddMesh threw")
        lib.ref_instancer_commit(self._h)

    def n_instances(self) -> int:
        return int(_load().ref_instancer_count(self._h))

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.ref_instancer_delete(self._h)
            self._h = None

    def get_model_input(self, rays_o, rays_d, parameters, n_samples, step_size):
        """instancer.pyx:38-54 without the tf.constant wrapping; `hit` is the boolean mask (the binding returns tf.where)."""
        rays_o = np.ascontiguousarray(rays_o, np.float32)
        n_rays = rays_o.shape[0]
        rays_d_np = np.ascontiguousarray(np.repeat(np.asarray(rays_d, np.float32)[:, None, :], n_samples, axis=1))
        t = np.zeros((n_rays, n_samples), np.float32)
        dist = np.zeros((n_rays, n_samples), np.float32)
        pts = np.zeros((n_rays, n_samples, 3), np.float32)
        color = np.zeros((n_rays, 1, 3), np.float32)
        density = np.zeros((n_rays, 1), np.float32)
        density_weight = np.ones((n_rays, n_samples), np.float32)
        instance_id = np.zeros((n_rays, n_samples), np.int32)
        hit = np.zeros(n_rays, np.bool_)
        params = np.ascontiguousarray(np.repeat(np.asarray(parameters, np.float32)[:, None, :], n_samples, axis=1))
        _load().ref_instancer_get_model_input(self._h, _ptr(rays_o), _ptr(rays_d_np), _ptr(t), _ptr(dist), _ptr(pts), _ptr(color),
                                              _ptr(density), _ptr(density_weight), _ptr(instance_id), _ptr(hit), n_rays, n_samples,
                                              float(step_size), _ptr(params))
        return dict(rays_d=rays_d_np, pts=pts, t=t, dist=dist, color=color, density=density, density_weight=density_weight,
                    instance_id=instance_id, hit=hit, parameters=params)

"""ctypes binding of oracle/libntx_oracle.so (the C restatement).  TEST INFRASTRUCTURE ONLY."""

from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "libntx_oracle.so"
_lib = None


def build(force: bool = False) -> Path:
    src = HERE / "ntx_oracle.c"
    if force or not LIB.exists() or LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE), "-B" if force else "-s", "libntx_oracle.so"], check=True,
                       capture_output=True)
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(LIB))
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def stratified_z(t, n_samples, perturb, z_rand=None):
    t, z_rand = _f(t), _f(z_rand)
    R = t.shape[0]
    z = np.empty((R, n_samples), np.float32)
    load().ntxo_stratified_z(_p(t), _p(z_rand), C.c_int(R), C.c_int(n_samples), C.c_int(int(perturb)), _p(z))
    return z


def composite(color, sigma, z, rays_d, noise=None, map_exr=False, composite_bkgd=False, bkgd=(1., 1., 1.)):
    color, sigma, z, rays_d, noise = _f(color), _f(sigma), _f(z), _f(rays_d), _f(noise)
    R, S = z.shape
    bk = np.asarray(bkgd, np.float32)
    cmap, amap = np.empty((R, 3), np.float32), np.empty((R,), np.float32)
    w, depth = np.empty((R, S), np.float32), np.empty((R,), np.float32)
    load().ntxo_composite(_p(color), _p(sigma), _p(z), _p(rays_d), _p(noise), C.c_int(R), C.c_int(S),
                          C.c_int(int(map_exr)), C.c_int(int(composite_bkgd)), _p(bk), _p(cmap), _p(amap), _p(w),
                          _p(depth))
    return cmap, amap, w, depth


def sample_pdf(bins, weights, n, det=False, u=None):
    bins, weights, u = _f(bins), _f(weights), _f(u)
    R, Cb = bins.shape
    s = np.empty((R, n), np.float32)
    inds = np.empty((R, n), np.int32)
    load().ntxo_sample_pdf(_p(bins), _p(weights), _p(u), C.c_int(R), C.c_int(Cb), C.c_int(n), C.c_int(int(det)),
                           _p(s), _p(inds))
    return s, inds


def hierarchical_z(z, weights, n, det, u=None):
    z, weights, u = _f(z), _f(weights), _f(u)
    R, S = z.shape
    zo = np.empty((R, S + n), np.float32)
    zs = np.empty((R, n), np.float32)
    inds = np.empty((R, n), np.int32)
    load().ntxo_hierarchical_z(_p(z), _p(weights), _p(u), C.c_int(R), C.c_int(S), C.c_int(n), C.c_int(int(det)),
                               _p(zo), _p(zs), _p(inds))
    return zo, zs, inds

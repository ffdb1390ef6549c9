"""Generates tests/golden/widened_small.npz from the oracles: instancer records, mip cone gaussians and edge
compositing, the integrated encoding, camera rays and the image post-processing.  ORACLE-generated regression
vectors (the reference cannot run here): they freeze the oracles' behaviour and give the GPU tests committed
inputs / outputs.  Re-run:  python tests/golden/make_golden_wide.py"""

import math
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import instancer_oracle as IO  # noqa: E402
from oracle import nerftex_oracle as O  # noqa: E402

B0, B1 = [-0.5, -0.5, -0.1], [0.5, 0.5, 0.1]


def patches(n, rng):
    Ts = []
    for i in range(n):
        for j in range(n):
            a = rng.uniform(0, 2 * math.pi)
            T = np.eye(4)
            T[:3, :3] = np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]]) * rng.uniform(0.8, 1.3)
            T[:3, 3] = [i * 0.8, j * 0.8, 0.1 * rng.uniform()]
            Ts.append(T)
    return np.array(Ts)


def main():
    rng = np.random.default_rng(20261107)
    out = {}
    # instancer: 3 x 3 patches, 120 rays, directional light at parameter 2
    Ts = patches(3, rng)
    R = 120
    o = np.stack([rng.uniform(-1, 2.5, R), rng.uniform(-1, 2.5, R), np.full(R, 3.0)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-.3, .3, R), rng.uniform(-.3, .3, R), np.full(R, -1.0)], 1).astype(np.float32)
    prm = rng.normal(size=(R, 5)).astype(np.float32)
    for method in ("random", "nearest", "nearest_blend"):
        rec = IO.get_model_input(o, d, prm, 24, 0.02, Ts, B0, B1, seed=9, light_dir_idx=2, method=method, patch_scale=1.0)   # explicit patches: the reference keeps its patch_scale member at 1 (instancer.cpp:53,245)
        for k, v in rec.items():
            out[f"inst_{method}_{k}"] = v
    out.update(inst_T=Ts, inst_o=o, inst_d=d, inst_prm=prm)
    # mip: cone gaussians + edge compositing + integrated encoding
    Rm, S = 40, 32
    z = np.sort(rng.uniform(2, 6, (Rm, S + 1)).astype(np.float32), -1)
    mo, md = rng.normal(size=(Rm, 3)).astype(np.float32), rng.normal(size=(Rm, 3)).astype(np.float32)
    radii = rng.uniform(1e-3, 3e-2, (Rm, 1)).astype(np.float32)
    mean, cov = O.mip_cone_gaussians(mo, md, z, radii)
    rgbs = rng.normal(size=(Rm, S, 4)).astype(np.float32)
    rgbs[..., 3] = rgbs[..., 3] * 3 + 1
    ec, ea, ew = O.composite_edges(rgbs[..., :3], rgbs[..., 3], z, md, True, (1., .5, .25))
    ipe = O.integrated_positional_encoding(np.concatenate([mean, cov], -1).reshape(-1, 6)[:64], 10)
    out.update(mip_z=z, mip_o=mo, mip_d=md, mip_radii=radii, mip_mean=mean, mip_cov=cov, mip_rgbs=rgbs, mip_cmap=ec, mip_amap=ea,
               mip_w=ew, mip_ipe=ipe)
    # camera rays against an AABB, look_at pose given explicitly
    H, W = 12, 16
    c2w = np.array([[0.6, -0.48, 0.64, 3.2], [0.8, 0.36, -0.48, -2.4], [0, 0.8, 0.6, 3.0], [0, 0, 0, 1]], np.float32)
    ii, jj = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    loc = np.stack([ii.ravel(), jj.ravel()], -1).astype(np.float32)
    ro, rd, t, cone = O.proxy_rays(loc, H, W, 20.0, c2w, [-1, -1, -1], [1, 1, 1])
    out.update(ray_c2w=c2w, ray_o=ro, ray_d=rd, ray_t=t, ray_cone=cone)
    # image post-processing, factor 3 on a 17 x 22 frame
    alpha = rng.uniform(0, 1, (17 * 22,)).astype(np.float32)
    color = (rng.uniform(0, 1, (17 * 22, 3)).astype(np.float32) * alpha[:, None]).astype(np.float32)
    out.update(img_color=color, img_alpha=alpha, img_out=O.render_image_post(color, alpha, 17, 22, 3))
    np.savez_compressed(Path(__file__).with_name("widened_small.npz"), **out)
    print("wrote", Path(__file__).with_name("widened_small.npz"))


if __name__ == "__main__":
    main()

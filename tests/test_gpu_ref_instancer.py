"""The DEVICE instancer (csrc/ntx_instancer.cu through the C ABI) against the REFERENCE'S OWN C++ instancer
(oracle/ref_instancer.py: instancer/src/instancer.cpp compiled against the Embree stand-in) on the same scenes and the
same per-ray offsets — the reference's std::mt19937 stream, reproduced and handed to the kernel as `ray_offsets`.
Scenes, tolerances and their reasons are those of tests/test_ref_instancer.py (which holds the oracle to the same
reference on the CPU); the library is prebuilt in the container where the reference tree is and travels to the GPU box."""

import numpy as np
import pytest
import torch

from oracle import ref_instancer as RI
import test_ref_instancer as T

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not RI.available(), reason="no prebuilt oracle/_ref library")]
F = np.float32


def device_records(inst, o, d, prm, S, step, u):
    cu = lambda a: torch.as_tensor(np.ascontiguousarray(a)).cuda()
    csr = inst.get_model_input_csr(cu(o), cu(d), cu(prm), S, step, seed=0, ray_offsets=cu(u))
    n_all = csr["n_samples_all"].cpu().numpy()
    hit = np.zeros(o.shape[0], bool)
    hit[csr["ray_index"].cpu().numpy()] = True
    alpha_last = np.zeros(o.shape[0], F)
    alpha_last[hit] = csr["alpha_last"].cpu().numpy()
    color_last = np.zeros((o.shape[0], 3), F)
    color_last[hit] = csr["color_last"].cpu().numpy()
    g = lambda k: csr[k].cpu().numpy()
    return {"n_samples": n_all, "hit": hit, "alpha_last": alpha_last, "color_last": color_last, "pts": g("pts"), "dirs": g("dirs"),
            "params": g("params"), "dists": g("dists"), "t": g("t"), "alpha_weight": g("alpha_weight"), "instance_id": g("instance_id")}


def reference_run(ref_kwargs, o, d, prm, S, step):
    inst = RI.RefInstancer(T.B0, T.B1, **ref_kwargs)
    ref = inst.get_model_input(o, d, prm, S, step)
    drew = ref["dist"][:, 0] > 0
    skip = 0
    if ref_kwargs.get("jitter_amount", 0) > 0 and ref_kwargs.get("mesh_path"):
        skip = inst.n_instances() - len(ref_kwargs.get("transformations", ()))
    u = np.zeros(o.shape[0], F)
    u[drew] = T.mt_stream(0, skip + int(drew.sum()))[skip:]
    return ref, u


@pytest.mark.parametrize("mean,light", [(False, "light"), (True, "point")])
def test_explicit_patches(mean, light):
    from nerf_tex_b200.instancer import Instancer
    Tm = T.patches(60, 1)
    R, S, step = 300, 96, 0.01
    o, d = T.camera_rays(R, 2)
    prm = np.tile(np.asarray([[0.5, 0.7, 0.3, 0.2, 1.0]] if light == "light" else [[0.5, 0.7, 40.0, 0.3, 0.2, 3.0]], F), (R, 1))
    ref, u = reference_run(dict(textures=["", "", light], transformations=Tm.tolist(), instance_sampling_method="nearest",
                                use_mean_distance=mean), o, d, prm, S, step)
    inst = Instancer(T.B0, T.B1, textures=["", "", light], transformations=Tm, instance_sampling_method="nearest", use_mean_distance=mean)
    got = device_records(inst, o, d, prm, S, step, u)
    assert T.compare(ref, got, S) > 5000


def test_mesh_scene_with_textures_shadows_and_auxiliary_meshes(tmp_path):
    """Everything at once, from files: base mesh + anchors (PLY), a PNG parameter texture, shadow rays, a textured table
    and a wall as auxiliary meshes.  The product reads the same files the reference reads."""
    from nerf_tex_b200 import ply
    from nerf_tex_b200.image import encode_png
    from nerf_tex_b200.instancer import Instancer
    mesh, _, ref_kw = T.mesh_scene(tmp_path, 11)
    table = T.plane_mesh(4, 0.5, 0.0, z=0.45)
    table["V"][:, 0] += 0.3
    wall = T.plane_mesh(3, 0.8, 0.0)
    wall["V"] = np.stack([np.full(16, -1.15), wall["V"][:, 0], wall["V"][:, 1] + 0.6], 1).astype(F)
    wall["N"] = np.tile(np.asarray([[1, 0, 0]], F), (16, 1))
    rgb = np.stack([T.checker_texture(16, 2, 3), T.checker_texture(16, 2, 4), T.checker_texture(16, 2, 5)], -1)
    ply.write_ply(str(tmp_path / "table.ply"), table["V"], table["F"], table["N"], table["UV"])
    ply.write_ply(str(tmp_path / "wall.ply"), wall["V"], wall["F"], wall["N"], wall["UV"])
    (tmp_path / "table.png").write_bytes(encode_png(rgb))
    (tmp_path / "a.png").write_bytes(encode_png(T.checker_texture(32, 4, 1)[:, :, None]))
    R, S, step = 300, 96, 0.004
    o, d = T.rays_from_above(R, 12)
    d[::4] = np.asarray([-1.0, 0.1, -0.2], F) + np.random.default_rng(13).uniform(-.1, .1, (len(d[::4]), 3)).astype(F)
    prm = np.tile(np.asarray([[0.9, 0.3, 0.2, 0.9]], F), (R, 1))
    kw = dict(textures=[str(tmp_path / "a.png"), "light"], cast_shadow_rays=True, min_shadow_samples=8, n_shadow_samples=4000,
              min_texture_samples=8, n_texture_samples=4000,
              auxiliary_meshes=[(str(tmp_path / "table.ply"), str(tmp_path / "table.png")), (str(tmp_path / "wall.ply"), "")])
    ref, u = reference_run(dict(kw, **ref_kw), o, d, prm, S, step)
    inst = Instancer(T.B0, T.B1, **dict(kw, **ref_kw))
    got = device_records(inst, o, d, prm, S, step, u)
    li = 1
    flips, total = T.reconcile_flips(ref, got, li)
    assert total > 1000
    T.compare(ref, got, S, tol_pts=5e-4, params_tol=2e-3, max_id_mismatch=0.01)
    close = np.abs(ref["color"][:, 0] - got["color_last"]).max(-1) < 2e-3
    assert close.mean() > 0.98 and (got["color_last"].max(-1) > 0.05).sum() > 20
    shadowed = (got["params"][:, li:li + 3] == np.asarray([0, 0, -1], F)).all(-1)
    assert 0.01 < shadowed.mean() < 0.99

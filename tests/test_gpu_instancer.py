"""Device shell instancer (ntx_instancer_trace / ntx_instancer_march) against oracle/instancer_oracle.py:
bit-exact records, then end to end through InstanceRenderer."""

import math

import numpy as np
import pytest
import torch

from oracle import instancer_oracle as IO

pytestmark = pytest.mark.gpu

B0, B1 = [-0.5, -0.5, -0.1], [0.5, 0.5, 0.1]


def patch_grid(n, seed, spacing=0.9, tilt=0.0):
    """n x n patches on the z = 0 plane, randomly rotated about z (and tilted), randomly scaled: they overlap."""
    rng = np.random.default_rng(seed)
    Ts = []
    for i in range(n):
        for j in range(n):
            a, b = rng.uniform(0, 2 * math.pi), rng.uniform(-tilt, tilt)
            rz = np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]])
            rx = np.array([[1, 0, 0], [0, math.cos(b), -math.sin(b)], [0, math.sin(b), math.cos(b)]])
            T = np.eye(4)
            T[:3, :3] = (rz @ rx) * rng.uniform(0.8, 1.3)
            T[:3, 3] = [i * spacing, j * spacing, 0.1 * rng.uniform()]
            Ts.append(T)
    return np.array(Ts)


def rays_over(n, R, seed, spacing=0.9, grazing=False):
    rng = np.random.default_rng(seed)
    ext = n * spacing
    o = np.stack([rng.uniform(-1, ext, R), rng.uniform(-1, ext, R), np.full(R, 3.0)], 1).astype(np.float32)
    s = 3.0 if grazing else 0.3
    d = np.stack([rng.uniform(-s, s, R), rng.uniform(-s, s, R), np.full(R, -1.0)], 1).astype(np.float32)
    return o, d


def check_records(got, want, R):
    ns = got["n_samples_all"].cpu().numpy()
    assert np.array_equal(ns, want["n_samples"])
    assert np.array_equal(got["ray_index"].cpu().numpy(), np.nonzero(want["hit"])[0])
    ptr_all = np.concatenate([[0], np.cumsum(want["n_samples"])])
    assert np.array_equal(got["ray_ptr"].cpu().numpy(), np.concatenate([ptr_all[:-1][want["hit"]], ptr_all[-1:]]))
    assert np.array_equal(got["instance_id"].cpu().numpy(), want["instance_id"])
    for k in ("pts", "dirs", "params", "dists", "t", "alpha_weight"):
        g = got[k].cpu().numpy()
        assert g.shape == want[k].shape, k
        assert np.array_equal(g.view(np.uint32), want[k].view(np.uint32)), (k, np.abs(g - want[k]).max())


@pytest.mark.parametrize("method,mean_dist,light", [("random", False, 1), ("nearest", False, -1), ("random", True, 4),
                                                    ("nearest", True, 0)])
def test_records_bit_exact(method, mean_dist, light):
    from nerf_tex_b200.instancer import Instancer
    Ts = patch_grid(5, seed=1, tilt=0.3)
    R = 700
    o, d = rays_over(5, R, seed=2)
    prm = np.random.default_rng(3).normal(size=(R, 7)).astype(np.float32)
    textures = ["light" if i == light else "" for i in range(5)] if light >= 0 else []
    inst = Instancer(B0, B1, textures=textures, transformations=Ts.tolist(), instance_sampling_method=method,
                     use_mean_distance=mean_dist)
    assert inst.n_instances() == 25 and inst.light_dir_idx == light
    want = IO.get_model_input(o, d, prm, 48, 0.02, Ts, B0, B1, seed=11, light_dir_idx=light, method=method,
                              use_mean_distance=mean_dist)
    got = inst.get_model_input_csr(torch.as_tensor(o), torch.as_tensor(d), torch.as_tensor(prm), 48, 0.02, seed=11)
    torch.cuda.synchronize()
    check_records(got, want, R)
    assert want["n_samples"].sum() > 3000 and (want["alpha_weight"] > 1).any() == (method == "random")
    assert 0 < want["hit"].sum() < R


def test_point_light_and_parameter_layout():
    """'point' occupies four parameters (strength, position): the direction becomes position - sample in the patch
    frame, the strength the received radiance (instancer.cpp:75-92, :572-590, :954-956)."""
    from nerf_tex_b200.instancer import Instancer
    Ts = patch_grid(4, seed=3, tilt=0.25)
    R = 400
    o, d = rays_over(4, R, seed=4)
    rng = np.random.default_rng(5)
    prm = rng.uniform(0.1, 1.0, (R, 6)).astype(np.float32)
    prm[:, 3:] = rng.uniform(-2, 4, (R, 3)).astype(np.float32) + np.array([0, 0, 3], np.float32)   # light position
    inst = Instancer(B0, B1, textures=["", "", "point"], transformations=Ts.tolist())
    assert (inst.light_strength_idx, inst.light_dir_idx, inst.n_parameters) == (2, 3, 6)
    assert Instancer(B0, B1, textures=["light", "", ""], transformations=Ts.tolist()).light_dir_idx == 0
    assert Instancer(B0, B1, textures=["", "point", "light"], transformations=Ts.tolist()).light_dir_idx == 5   # last one wins
    want = IO.get_model_input(o, d, prm, 40, 0.02, Ts, B0, B1, seed=2, light_dir_idx=3, light_strength_idx=2)
    got = inst.get_model_input_csr(torch.as_tensor(o), torch.as_tensor(d), torch.as_tensor(prm), 40, 0.02, seed=2)
    check_records(got, want, R)
    p = want["params"]
    assert np.allclose(np.linalg.norm(p[:, 3:], axis=1), 1, atol=1e-5) and (p[:, 2] > 0).all() and not np.allclose(p[:, 2], p[0, 2])


def test_nearest_blend():
    """instance_sampling_method='nearest_blend' (instancer.cpp:694-713): bit-exact records; the density weight is
    1 / probability, so it is 1 where one patch dominates and larger inside the transition band."""
    from nerf_tex_b200.instancer import Instancer
    Ts = patch_grid(5, seed=8, tilt=0.2)
    R = 500
    o, d = rays_over(5, R, seed=9)
    prm = np.zeros((R, 2), np.float32)
    # explicit patches: the reference's patch_scale member is only set by DistributeInstancesOnMesh (instancer.cpp:245) and
    # stays 1 whatever the keyword says (checked on the reference's own code, tests/test_ref_instancer.py)
    inst = Instancer(B0, B1, transformations=Ts.tolist(), instance_sampling_method="nearest_blend", patch_scale=1.5)
    want = IO.get_model_input(o, d, prm, 40, 0.02, Ts, B0, B1, seed=6, method="nearest_blend", patch_scale=1.0)
    got = inst.get_model_input_csr(torch.as_tensor(o), torch.as_tensor(d), torch.as_tensor(prm), 40, 0.02, seed=6)
    check_records(got, want, R)
    w = want["alpha_weight"]
    assert w.min() >= 1 and (w == 1).any() and (w > 1.05).any()
    near = IO.get_model_input(o, d, prm, 40, 0.02, Ts, B0, B1, seed=6, method="nearest")
    assert 0.5 < (near["instance_id"] == want["instance_id"]).mean() < 1.0      # mostly, not always, the nearest patch


def test_grazing_rays_clip_and_many_instances():
    """Rays nearly parallel to the patch plane cross many patches: the per-ray hit list saturates at 200 entries,
    n_pts clips the sample count, and more instances than one shared-memory tile (128) are walked."""
    from nerf_tex_b200.instancer import Instancer
    Ts = patch_grid(25, seed=5, spacing=0.2)                       # 625 instances, heavy overlap
    R = 160
    o, d = rays_over(25, R, seed=6, spacing=0.2, grazing=True)
    o[:, 2] = 0.3
    d[:, 2] = -0.05
    prm = np.zeros((R, 1), np.float32)
    inst = Instancer(B0, B1, transformations=Ts.tolist())
    for n_pts, step in ((16, 0.05), (256, 0.004)):
        want = IO.get_model_input(o, d, prm, n_pts, step, Ts, B0, B1, seed=5)
        got = inst.get_model_input_csr(torch.as_tensor(o), torch.as_tensor(d), torch.as_tensor(prm), n_pts, step, seed=5)
        check_records(got, want, R)
        assert want["n_samples"].max() == n_pts
    hits = [len(IO.ray_hits(o[i], d[i], IO.prepare_instances(Ts)[0], np.float32(B0), np.float32(B1))) for i in range(R)]
    assert max(hits) == IO.MAX_HITS


def test_edge_inputs():
    from nerf_tex_b200 import ops
    from nerf_tex_b200.instancer import Instancer
    Ts = patch_grid(2, seed=0)
    inst = Instancer(B0, B1, transformations=Ts.tolist())
    # no ray at all
    e = inst.get_model_input_csr(torch.zeros(0, 3), torch.zeros(0, 3), torch.zeros(0, 2), 8, 0.05)
    assert e["pts"].shape == (0, 3) and e["ray_ptr"].tolist() == [0] and e["ray_index"].numel() == 0
    # every ray misses; axis-parallel rays (division by zero inside the slab test) behave like the oracle
    o = np.array([[10, 10, 3], [0.2, 0.1, 3], [0.2, 0.1, 0.05], [-3, 0.1, 0.0]], np.float32)
    d = np.array([[0, 0, -1], [0, 0, -1], [1, 0, 0], [1, 0, 0]], np.float32)
    prm = np.ones((4, 2), np.float32)
    want = IO.get_model_input(o, d, prm, 8, 0.05, Ts, B0, B1, seed=1)
    got = inst.get_model_input_csr(torch.as_tensor(o), torch.as_tensor(d), torch.as_tensor(prm), 8, 0.05, seed=1)
    check_records(got, want, 4)
    assert not want["hit"][0] and want["hit"][1]
    # refusals: a mesh file that is not there, and what is not on the device yet
    with pytest.raises((OSError, ValueError)):
        Instancer(B0, B1, mesh_path="meshes/no_such_mesh.ply")
    with pytest.raises((OSError, ValueError)):
        Instancer(B0, B1, transformations=Ts.tolist(), auxiliary_meshes=[("meshes/no_such_table.ply", "")])
    # a texture file that cannot be read adds no parameter (stbi_load returns NULL, instancer.cpp:38-50)
    assert Instancer(B0, B1, transformations=Ts.tolist(), textures=["", "textures/no_such.png", "light"]).n_parameters == 4
    with pytest.raises(ValueError):
        Instancer(B0, B1)
    with pytest.raises(RuntimeError):
        ops.instancer_get_model_input(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 17), 8, 0.05, inst.scene, B0, B1)


def test_dense_tuple_and_instance_renderer():
    """instancer.pyx:54 dense tuple from the device records, and the whole InstanceRenderer on top of it against
    the oracle instancer + oracle network + oracle compositor."""
    from nerf_tex_b200.instancer import Instancer
    from nerf_tex_b200.renderer import InstanceRenderer
    from oracle import nerftex_oracle as O
    from test_gpu_renderer import build_models, psnr
    Ts = patch_grid(4, seed=9, tilt=0.2)
    models, oracles = build_models((1, 6), False, seed=7)
    kw = dict(b_0=B0, b_1=B1, textures=["", "", "", "", "light"], transformations=Ts.tolist())
    r = InstanceRenderer(instancer_config=Instancer(seed=4, **kw), model=models["model"], n_samples=32, render_chunk=16384,
                         step_size=0.01, density_scale=100.0)
    R = 500
    o, d = rays_over(4, R, seed=8)
    params = np.array([[1, 1, 1, .1, .3, -.2, 1]], np.float32)
    cone = np.random.default_rng(1).uniform(0.01, 0.05, (1, R, 1)).astype(np.float32)
    t = np.tile(np.array([[[1., 3.]]], np.float32), (1, R, 1))
    cu = lambda a: torch.as_tensor(a).cuda()
    got = r(cu(o[None]), cu(d[None]), cu(t), cu(params), cu(cone), composite_bkgd=True, training=False)
    pf = np.repeat(params, R, 0)
    want = IO.get_model_input(o, d, pf, 32, 0.01, Ts, B0, B1, seed=4, light_dir_idx=4)
    dense = Instancer(seed=4, **kw).get_model_input(cu(o), cu(d), cu(pf), 32, 0.01)
    dense = [x.cpu().numpy() for x in dense]
    # the dense tuple holds exactly the oracle records
    ns = want["n_samples"]
    assert np.array_equal(dense[8][:, 0], np.nonzero(want["hit"])[0])
    k = 0
    for ray in range(R):
        n = ns[ray]
        assert np.array_equal(dense[1][ray, :n], want["pts"][k:k + n]) and np.array_equal(dense[0][ray, :n], want["dirs"][k:k + n])
        assert np.array_equal(dense[3][ray, :n], want["dists"][k:k + n]) and not dense[3][ray, n:].any()
        assert np.array_equal(dense[9][ray, :n], want["params"][k:k + n])
        k += n
    wc, wa = O.instance_evaluate_model(oracles["model"], dense, R, cone_scale=cone[0], blur_idx=None, patch_scale=1.0,
                                       density_scale=100.0, composite_bkgd=True, instance_color=None)
    gc, ga = got["color_pred"].cpu().numpy()[0], got["alpha_pred"].cpu().numpy()[0]
    assert (wa > 0.05).mean() > 0.05 and wa.max() > 0.4
    # density_scale 100 x 0.2 path length amplifies the bf16 error of sigma 20-fold (tolerance section)
    print("instancer e2e: max |d alpha|", np.abs(ga - wa).max(), "max |d color|", np.abs(gc - wc).max(), "psnr", psnr(gc, wc))
    assert np.abs(ga - wa).max() < 1e-2 and np.abs(gc - wc).max() < 1e-2, (np.abs(ga - wa).max(), np.abs(gc - wc).max())
    assert psnr(gc, wc) > 45


# ------------------------------------------------------------------------------------------ acceleration grid, base mesh
def big_carpet(n, seed, density_extent=1.45, base_n=1500, patch_scale=0.09):
    """`n` patches at the areal density of the bench's 1 500-patch carpet (bench.carpet_patches), box of the carpet config."""
    import bench
    extent = density_extent * math.sqrt(n / base_n)
    return bench.carpet_patches(n, seed=seed, extent=extent, patch_scale=patch_scale), extent


def plane_mesh(n, extent, z=0.0, wobble=0.0, seed=0):
    """(n+1)^2 vertex grid over [-extent, extent]^2 with normals and UVs: a procedural stand-in for the LFS meshes."""
    rng = np.random.default_rng(seed)
    xs = np.linspace(-extent, extent, n + 1)
    X, Y = np.meshgrid(xs, xs, indexing="ij")
    Z = z + wobble * np.sin(2.1 * X) * np.cos(1.7 * Y)
    V = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    dzdx, dzdy = wobble * 2.1 * np.cos(2.1 * X) * np.cos(1.7 * Y), -wobble * 1.7 * np.sin(2.1 * X) * np.sin(1.7 * Y)
    N = np.stack([-dzdx, -dzdy, np.ones_like(X)], -1).reshape(-1, 3)
    N = (N / np.linalg.norm(N, axis=-1, keepdims=True)).astype(np.float32)
    UV = np.stack([(X + extent) / (2 * extent), (Y + extent) / (2 * extent)], -1).reshape(-1, 2).astype(np.float32)
    idx = lambda i, j: i * (n + 1) + j
    F = []
    for i in range(n):
        for j in range(n):
            F += [[idx(i, j), idx(i + 1, j), idx(i + 1, j + 1)], [idx(i, j), idx(i + 1, j + 1), idx(i, j + 1)]]
    return {"V": V, "F": np.asarray(F, np.int32), "N": N, "UV": UV}


CB0, CB1 = [-1.4, -1.2, -.1], [1.2, 1.2, 1.8]


def test_grid_equals_brute_force_and_oracle_at_20000_instances():
    """The uniform grid must not change a single record: against the all-instances loop on every ray, and against the
    oracle (which knows no grid) on a subset."""
    from nerf_tex_b200.instancer import Instancer
    Ts, extent = big_carpet(20000, seed=3)
    R = 6000
    rng = np.random.default_rng(4)
    o = np.stack([rng.uniform(-extent, extent, R), rng.uniform(-extent, extent, R), rng.uniform(1.0, 3.0, R)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-1.2, 1.2, R), rng.uniform(-1.2, 1.2, R), np.full(R, -1.0)], 1).astype(np.float32)
    d[::7, 2] = -0.05                                             # grazing rays cross many cells and many patches
    d[::11] *= -1                                                 # some look away from the carpet
    o[::13] = [0.0, 0.0, 0.05]                                    # origins inside the shell layer
    prm = np.tile(np.array([[1, 1, 1, .1, 0, 0, 1]], np.float32), (R, 1))
    kw = dict(b_0=CB0, b_1=CB1, textures=["", "", "", "", "light"], transformations=Ts, patch_scale=0.09,
              instance_sampling_method="nearest")
    grid, brute = Instancer(**kw), Instancer(use_grid=False, **kw)
    assert grid.scene.inst_grid.cell_start and not brute.scene.inst_grid.cell_start
    cu = lambda a: torch.as_tensor(a).cuda()
    a = grid.get_model_input_csr(cu(o), cu(d), cu(prm), 1024, 0.002, seed=2)
    b = brute.get_model_input_csr(cu(o), cu(d), cu(prm), 1024, 0.002, seed=2)
    assert a["pts"].shape[0] > 100 * R / 4
    for k in a:
        assert torch.equal(a[k], b[k]), k
    sub = np.concatenate([np.arange(0, 200), np.arange(0, R, 7)[:60], np.arange(0, R, 13)[:20]])
    want = IO.get_model_input(o[sub], d[sub], prm[sub], 1024, 0.002, Ts, CB0, CB1, seed=2, light_dir_idx=4, method="nearest",
                              patch_scale=0.09, ray_ids=sub)
    got = grid.get_model_input_csr(cu(o[sub]), cu(d[sub]), cu(prm[sub]), 1024, 0.002, seed=2, ray_ids=cu(sub.astype(np.int32)))
    check_records(got, want, len(sub))
    assert want["n_samples"].max() > 300


def test_ray_names_make_records_independent_of_chunking():
    from nerf_tex_b200.instancer import Instancer
    Ts = patch_grid(5, seed=1, tilt=0.3)
    R = 900
    o, d = rays_over(5, R, seed=3)
    prm = np.zeros((R, 1), np.float32)
    inst = Instancer(B0, B1, transformations=Ts.tolist(), instance_sampling_method="random")
    cu = lambda a: torch.as_tensor(a).cuda()
    ids = np.arange(R, dtype=np.int32)
    whole = inst.get_model_input_csr(cu(o), cu(d), cu(prm), 40, 0.02, seed=9, ray_ids=cu(ids))
    perm = np.random.default_rng(0).permutation(R)                 # any subset, any order
    for part in (perm[:300], perm[300:]):
        got = inst.get_model_input_csr(cu(o[part]), cu(d[part]), cu(prm[part]), 40, 0.02, seed=9, ray_ids=cu(ids[part]))
        want = IO.get_model_input(o[part], d[part], prm[part], 40, 0.02, Ts, B0, B1, seed=9, ray_ids=ids[part])
        check_records(got, want, len(part))
    # and the whole call is the concatenation of its rays' records
    want = IO.get_model_input(o, d, prm, 40, 0.02, Ts, B0, B1, seed=9)
    check_records(whole, want, R)


@pytest.mark.parametrize("use_grid", [True, False])
def test_base_mesh_ends_rays(use_grid):
    """Patches distributed on a wobbly plane (DistributeInstancesOnMesh), rays from above, from below and grazing:
    box hits behind the first mesh hit are dropped, the mesh closes the open segment, rays that end on the mesh carry
    alpha_last = 1 — all bit-exact against the oracle's brute-force watertight test."""
    from nerf_tex_b200.instancer import Instancer
    mesh = plane_mesh(24, 1.5, wobble=0.08)
    rng = np.random.default_rng(5)
    anchors = np.stack([rng.uniform(-1.3, 1.3, 300), rng.uniform(-1.3, 1.3, 300)], 1)
    anchors = np.concatenate([anchors, 0.08 * np.sin(2.1 * anchors[:, :1]) * np.cos(1.7 * anchors[:, 1:])], 1).astype(np.float32)
    inst = Instancer(CB0, CB1, textures=["", "light"], mesh=mesh, patch_origins=anchors, patch_scale=0.09, jitter_amount=1.0,
                     instance_sampling_method="nearest", use_grid=use_grid)
    assert inst.n_instances() == 300 and inst.scene.n_faces == 2 * 24 * 24
    Ts = inst.transformations
    R = 500
    o = np.stack([rng.uniform(-1.6, 1.6, R), rng.uniform(-1.6, 1.6, R), rng.uniform(0.5, 2.0, R)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-1, 1, R), rng.uniform(-1, 1, R), np.full(R, -1.0)], 1).astype(np.float32)
    o[::5, 2] *= -1                                               # from below: only the sliver of the boxes under the mesh
    d[::5, 2] = 1.0
    d[1::9, 2] = -0.03                                            # grazing
    o[2::17] = [0.1, 0.2, 0.05]                                   # start inside the shell layer
    prm = np.tile(np.array([[0.7, 0, 0, 1]], np.float32), (R, 1))
    want = IO.get_model_input(o, d, prm, 256, 0.004, Ts, CB0, CB1, seed=1, light_dir_idx=1, method="nearest", patch_scale=0.09,
                              mesh=(mesh["V"], mesh["F"]))
    cu = lambda a: torch.as_tensor(a).cuda()
    got = inst.get_model_input_csr(cu(o), cu(d), cu(prm), 256, 0.004, seed=1)
    check_records(got, want, R)
    hit = want["hit"]
    assert np.array_equal(got["alpha_last"].cpu().numpy(), want["alpha_last"][hit]) and not got["color_last"].any()
    below = np.arange(R) % 5 == 0
    assert want["alpha_last"][below & hit].mean() > 0.8 and np.median(want["n_samples"][below]) < 10  # 0.009 deep under the mesh
    assert (want["alpha_last"][~below] == 1).mean() > 0.4 and (want["alpha_last"] == 0)[hit].any()
    assert want["n_samples"][~below].max() > 30


def test_instancer_from_ply_files_matches_distribute_oracle(tmp_path):
    """mesh_path / patch_origins_path through the PLY reader (binary and ascii), placement against the loop oracle."""
    from nerf_tex_b200 import ply
    from nerf_tex_b200.instancer import Instancer
    mesh = plane_mesh(6, 1.0, wobble=0.1)
    rng = np.random.default_rng(2)
    anchors = np.stack([rng.uniform(-.9, .9, 40), rng.uniform(-.9, .9, 40)], 1)
    anchors = np.concatenate([anchors, 0.1 * np.sin(2.1 * anchors[:, :1]) * np.cos(1.7 * anchors[:, 1:]) + 0.01], 1).astype(np.float32)
    ply.write_ply(str(tmp_path / "mesh.ply"), mesh["V"], mesh["F"], mesh["N"], mesh["UV"], binary=True)
    ply.write_ply(str(tmp_path / "anchors.ply"), anchors, binary=False)
    back = ply.read_ply(str(tmp_path / "mesh.ply"))
    assert all(np.array_equal(back[k], mesh[k]) for k in ("V", "F", "N", "UV"))
    assert np.allclose(ply.read_ply(str(tmp_path / "anchors.ply"))["V"], anchors, atol=0)
    for jitter in (0.0, 1.0):
        inst = Instancer(CB0, CB1, mesh_path=str(tmp_path / "mesh.ply"), patch_origins_path=str(tmp_path / "anchors.ply"),
                         patch_scale=0.09, jitter_amount=jitter)
        want, scale = IO.distribute_instances_on_mesh(mesh["V"], mesh["F"], mesh["N"], mesh["UV"], anchors, 0.09, jitter, seed=0)
        assert scale == inst.patch_scale and inst.n_instances() == 40
        np.testing.assert_allclose(inst.transformations, want, atol=2e-6)
        R = want[:, :3, :3] / 0.09                                # orthonormal frames, third axis = surface normal
        np.testing.assert_allclose(np.einsum("iab,iac->ibc", R, R), np.tile(np.eye(3), (40, 1, 1)), atol=1e-5)
    # no anchor file: one patch per distinct vertex (:350-372); scale <= 0: the average edge length (:244)
    inst = Instancer(CB0, CB1, mesh_path=str(tmp_path / "mesh.ply"), patch_origins_path=str(tmp_path / "missing.ply"),
                     patch_scale=0.0, jitter_amount=0.5)
    want, scale = IO.distribute_instances_on_mesh(mesh["V"], mesh["F"], mesh["N"], mesh["UV"], None, 0.0, 0.5, seed=0)
    assert inst.n_instances() == 49 and abs(inst.patch_scale - scale) < 1e-7
    np.testing.assert_allclose(inst.transformations, want, atol=2e-6)


def shadow_scene(seed=7):
    """Patches on a hilly plane (the plane is the base mesh) plus a few patches floating above it: with a low sun the
    hills (front faces of the mesh), the floating boxes' bottom faces and tilted boxes' top faces cast shadows."""
    mesh = plane_mesh(16, 1.5, wobble=0.15)
    rng = np.random.default_rng(seed)
    xy = rng.uniform(-1.3, 1.3, (200, 2))
    anchors = np.concatenate([xy, 0.15 * np.sin(2.1 * xy[:, :1]) * np.cos(1.7 * xy[:, 1:])], 1).astype(np.float32)
    floating = np.tile(np.eye(4), (25, 1, 1))
    floating[:, :3, :3] *= 0.09
    floating[:, :3, 3] = np.concatenate([rng.uniform(-1.2, 1.2, (25, 2)), rng.uniform(0.3, 0.5, (25, 1))], 1)
    return mesh, anchors, floating


@pytest.mark.parametrize("light,use_grid,n_shadow", [("light", True, 64), ("point", True, 64), ("light", False, 4000)])
def test_shadow_rays(light, use_grid, n_shadow):
    """cast_shadow_rays (instancer.cpp:593-603, filter :539-551, sampling :863,:936-951,:992-1000): interpolated shadow
    samples (n_shadow < n_pts) and the per-sample path (n_shadow >= n_pts); directional and point light.  Bit-exact."""
    from nerf_tex_b200.instancer import Instancer
    mesh, anchors, floating = shadow_scene()
    inst = Instancer(CB0, CB1, textures=["", light], transformations=floating, mesh=mesh, patch_origins=anchors, patch_scale=0.09,
                     jitter_amount=1.0, instance_sampling_method="nearest", use_grid=use_grid, cast_shadow_rays=True,
                     min_shadow_samples=8, n_shadow_samples=n_shadow)
    rng = np.random.default_rng(8)
    R = 300
    o = np.stack([rng.uniform(-1.4, 1.4, R), rng.uniform(-1.4, 1.4, R), rng.uniform(0.7, 2.0, R)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-1, 1, R), rng.uniform(-1, 1, R), np.full(R, -1.0)], 1).astype(np.float32)
    if light == "light":
        prm = np.tile(np.array([[0.7, 0.95, 0.1, 0.08]], np.float32), (R, 1))          # low sun
        kw = dict(light_dir_idx=1)
    else:
        prm = np.tile(np.array([[0.7, 3.0, 1.2, 0.3, 0.1]], np.float32), (R, 1))       # strength, position
        kw = dict(light_dir_idx=2, light_strength_idx=1)
    n_pts = 128
    ref = dict(seed=3, method="nearest", patch_scale=0.09, mesh=(mesh["V"], mesh["F"]), **kw)
    want = IO.get_model_input(o, d, prm, n_pts, 0.004, inst.transformations, CB0, CB1, cast_shadow_rays=True,
                              min_shadow_samples=8, n_shadow_samples=n_shadow, **ref)
    cu = lambda a: torch.as_tensor(a).cuda()
    got = inst.get_model_input_csr(cu(o), cu(d), cu(prm), n_pts, 0.004, seed=3)
    check_records(got, want, R)
    li = kw["light_dir_idx"]
    dark = np.all(want["params"][:, li:li + 3] == np.array([0, 0, -1], np.float32), -1)
    assert 0.05 < dark.mean() < 0.95, dark.mean()                                      # both outcomes occur
    unshadowed = IO.get_model_input(o, d, prm, n_pts, 0.004, inst.transformations, CB0, CB1, **ref)
    assert np.array_equal(unshadowed["params"][~dark], want["params"][~dark])          # lit samples are untouched


def checker_texture(n=32, cells=4, seed=0):
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    img = (((xx * cells // n) + (yy * cells // n)) % 2) * 200 + rng.integers(0, 50, (n, n))
    return img.astype(np.uint8)


@pytest.mark.parametrize("use_grid,n_tex", [(True, 64), (False, 64), (True, 4000)])
def test_parameter_textures(use_grid, n_tex, tmp_path):
    """Image entries of `textures` (getParameters, instancer.cpp:640-667): nearest base-mesh point -> UV -> bilinear
    lookup -> the parameter is multiplied; sampled along the ray's segments and linearly interpolated (n_tex < n_pts) or
    looked up per sample.  One texture comes from a PNG file, one from an array.  Bit-exact."""
    from nerf_tex_b200.image import encode_png
    from nerf_tex_b200.instancer import Instancer
    mesh, anchors, floating = shadow_scene(seed=11)
    img0, img1 = checker_texture(32, 4, 1), checker_texture(17, 3, 2)
    with open(tmp_path / "checker.png", "wb") as f:
        f.write(encode_png(img0[:, :, None]))
    textures = [str(tmp_path / "checker.png"), "", img1, "light"]
    inst = Instancer(CB0, CB1, textures=textures, mesh=mesh, patch_origins=anchors, patch_scale=0.09, jitter_amount=1.0,
                     instance_sampling_method="nearest", use_grid=use_grid, min_texture_samples=8, n_texture_samples=n_tex)
    assert inst.n_parameters == 6 and inst.light_dir_idx == 3 and inst.tex_param_idxs == [0, 2]
    rng = np.random.default_rng(12)
    R = 250
    o = np.stack([rng.uniform(-1.4, 1.4, R), rng.uniform(-1.4, 1.4, R), rng.uniform(0.7, 2.0, R)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-1, 1, R), rng.uniform(-1, 1, R), np.full(R, -1.0)], 1).astype(np.float32)
    prm = np.tile(np.array([[0.9, 0.4, 0.7, 0.2, 0.1, 1.0]], np.float32), (R, 1))
    tex = dict(UV=mesh["UV"], textures=IO.texture_channels(img0) + IO.texture_channels(img1), idxs=[0, 2],
               radius=inst.patch_max_extent, min_samples=8, n_samples=n_tex)
    n_pts = 128
    want = IO.get_model_input(o, d, prm, n_pts, 0.004, inst.transformations, CB0, CB1, seed=3, light_dir_idx=3, method="nearest",
                              patch_scale=0.09, mesh=(mesh["V"], mesh["F"]), texture=tex)
    cu = lambda a: torch.as_tensor(a).cuda()
    got = inst.get_model_input_csr(cu(o), cu(d), cu(prm), n_pts, 0.004, seed=3)
    check_records(got, want, R)
    p = want["params"]
    assert np.allclose(p[:, 1], 0.4, atol=1e-6) and p[:, 0].min() < 0.2 < 0.6 < p[:, 0].max() <= np.float32(0.9)   # both checker colours
    assert len(np.unique(p[:, 2])) > 50 and p[:, 2].max() <= np.float32(0.7)


def test_auxiliary_meshes_end_rays_and_are_shaded(tmp_path):
    """AddMesh / shadeMesh (instancer.cpp:393-418, 716-744): a textured table top and an untextured wall beside the
    carpet, directional light, shadows on; rays that end on them carry the shaded colour, on the base mesh black."""
    from nerf_tex_b200 import ply
    from nerf_tex_b200.instancer import Instancer
    mesh, anchors, floating = shadow_scene(seed=13)
    table = plane_mesh(4, 0.5, z=0.45)                                       # above part of the carpet, facing +z
    table["V"][:, 0] += 0.3
    wall = plane_mesh(3, 0.8, z=0.0)
    wall["V"] = np.stack([np.full(16, -1.45), wall["V"][:, 0], wall["V"][:, 1] + 0.6], 1).astype(np.float32)   # x = -1.45 plane
    wall["N"] = np.tile(np.array([[1, 0, 0]], np.float32), (16, 1))
    ply.write_ply(str(tmp_path / "wall.ply"), wall["V"], wall["F"], wall["N"], wall["UV"])
    rgb = np.stack([checker_texture(16, 2, 3), checker_texture(16, 2, 4), checker_texture(16, 2, 5)], -1)
    aux = [(table, rgb), (str(tmp_path / "wall.ply"), "")]
    inst = Instancer(CB0, CB1, textures=["", "light"], mesh=mesh, patch_origins=anchors, patch_scale=0.09, jitter_amount=1.0,
                     instance_sampling_method="nearest", cast_shadow_rays=True, min_shadow_samples=8, n_shadow_samples=64,
                     auxiliary_meshes=aux)
    rng = np.random.default_rng(14)
    R = 400
    o = np.stack([rng.uniform(-1.4, 1.4, R), rng.uniform(-1.4, 1.4, R), rng.uniform(0.9, 2.0, R)], 1).astype(np.float32)
    d = np.stack([rng.uniform(-1, 1, R), rng.uniform(-1, 1, R), np.full(R, -1.0)], 1).astype(np.float32)
    d[::4] = np.array([-1.0, 0.1, -0.2], np.float32) + rng.uniform(-.1, .1, (len(d[::4]), 3)).astype(np.float32)   # towards the wall
    prm = np.tile(np.array([[0.7, 0.3, 0.2, 0.9]], np.float32), (R, 1))
    aux_o = [dict(V=table["V"], F=table["F"], N=table["N"], UV=table["UV"], textures=IO.texture_channels(rgb)),
             dict(V=wall["V"], F=wall["F"], N=wall["N"], UV=wall["UV"], textures=[])]
    want = IO.get_model_input(o, d, prm, 128, 0.004, inst.transformations, CB0, CB1, seed=3, light_dir_idx=1, method="nearest",
                              patch_scale=0.09, mesh=(mesh["V"], mesh["F"]), cast_shadow_rays=True, min_shadow_samples=8,
                              n_shadow_samples=64, aux_meshes=aux_o)
    cu = lambda a: torch.as_tensor(a).cuda()
    got = inst.get_model_input_csr(cu(o), cu(d), cu(prm), 128, 0.004, seed=3)
    check_records(got, want, R)
    hit = want["hit"]
    gc = got["color_last"].cpu().numpy()
    assert np.array_equal(gc.view(np.uint32), want["color_last"][hit].view(np.uint32))
    assert np.array_equal(got["alpha_last"].cpu().numpy(), want["alpha_last"][hit])
    coloured = want["color_last"].any(-1)
    assert coloured.sum() > 40 and (want["alpha_last"][~coloured] == 1).sum() > 40        # auxiliary hits and base-mesh hits
    grey = coloured & (want["color_last"][:, 0] == want["color_last"][:, 1]) & (want["color_last"][:, 1] == want["color_last"][:, 2])
    assert grey.any() and (coloured & ~grey).any()                                         # the wall (0.8 grey) and the textured table

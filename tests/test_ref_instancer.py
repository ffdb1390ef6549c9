"""The oracle of the shell instancer (oracle/instancer_oracle.py) against the REFERENCE'S OWN C++ (instancer/src/
instancer.cpp), compiled from where it lies under /root/reference by oracle/ref_instancer.py.  Embree — the one dependency
that is absent — is replaced by a brute-force stand-in (oracle/embree_shim: our code, no BVH, Moeller-Trumbore in double),
so what is pinned here is everything the reference does with the hits: sorting and duplicate removal, patch toggling,
in-shell length, step count, the std::mt19937 offset, the walk over concatenated segments, 'nearest' patch choice, the
mapping into the patch frame, light mapping, mesh termination, shadow and texture sampling along segments, shadeMesh,
DistributeInstancesOnMesh (libigl PLY reader, tangent frames, closest points, jitter stream) — not the ray / box and
ray / triangle arithmetic itself (the oracle uses fp32 slab and watertight tests, the stand-in double precision: hit
distances agree to a few ulps, which is why the bars below are tolerances and not bit equality).

CPU only.  The library is built on demand when /root/reference exists; the prebuilt oracle/_ref/libref_instancer.so
is used otherwise (it travels to the GPU box); with neither the tests skip."""

import json

import numpy as np
import pytest

from oracle import instancer_oracle as IO
from oracle import ref_instancer as RI

F = np.float32
pytestmark = pytest.mark.skipif(not RI.available(), reason="neither /root/reference nor a prebuilt oracle/_ref library")

B0, B1 = [-1.0, -1.0, -0.1], [1.0, 1.0, 1.8]


def mt_stream(seed, n):
    """What std::uniform_real_distribution<float>(0, 1) draws from std::mt19937(seed) (instancer.cpp:767, :849, :858)."""
    g = IO.MT19937(seed)
    return [IO.uniform_real_float(g, 0.0, 1.0) for _ in range(n)]


def rotations(n, rng):
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    return np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                     np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                     np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], 1)


def patches(n, seed, scale=0.2, spread=1.0):
    rng = np.random.default_rng(seed)
    T = np.tile(np.eye(4), (n, 1, 1))
    T[:, :3, :3] = rotations(n, rng) * scale
    T[:, :3, 3] = rng.uniform(-spread, spread, (n, 3))
    return T.astype(F)


def camera_rays(R, seed, eye=(0.0, 0.3, 5.0), spread=1.0):
    rng = np.random.default_rng(seed)
    o = np.tile(np.asarray([eye], F), (R, 1))
    d = rng.uniform(-spread, spread, (R, 3)).astype(F) - o
    return o, (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(F)


def run_both(ref_kwargs, ora_kwargs, o, d, prm, S, step, T):
    inst = RI.RefInstancer(B0, B1, **ref_kwargs)
    ref = inst.get_model_input(o, d, prm, S, step)
    drew = ref["dist"][:, 0] > 0                                   # rays with in-shell length draw one offset, in ray order
    u = np.zeros(o.shape[0], F)
    # the one generator of the instance has already served the placement: one jitter angle per patch (:329, :354)
    skip = 0
    if ref_kwargs.get("jitter_amount", 0) > 0 and ref_kwargs.get("mesh_path"):
        skip = inst.n_instances() - len(ref_kwargs.get("transformations", ()))
    u[drew] = mt_stream(0, skip + int(drew.sum()))[skip:]
    ora = IO.get_model_input(o, d, prm, S, step, T, B0, B1, ray_u=u, **ora_kwargs)
    return inst, ref, ora


def compare(ref, ora, S, *, tol_t=2e-5, tol_pts=2e-4, max_count_mismatch=0.01, max_id_mismatch=0.005, params_tol=1e-5):
    R = ref["hit"].shape[0]
    assert np.array_equal(ref["hit"], ora["hit"])
    n_ref = np.minimum((ref["dist"] > 0).sum(1), (ref["t"] != 0).sum(1))   # samples the reference really emitted
    n_ora = ora["n_samples"]
    same = n_ref == n_ora
    assert (~same).mean() <= max_count_mismatch, (int((~same).sum()), R)
    ptr = np.concatenate([[0], np.cumsum(n_ora)])
    n_cmp = n_id_bad = 0
    for i in np.nonzero(same & (n_ora > 0))[0]:
        a, b = ptr[i], ptr[i + 1]
        n = b - a
        assert np.abs(ref["t"][i, :n] - ora["t"][a:b]).max() <= tol_t
        assert np.abs(ref["dist"][i, :n] - ora["dists"][a:b]).max() <= tol_t
        ok = ref["instance_id"][i, :n] == ora["instance_id"][a:b]
        n_cmp, n_id_bad = n_cmp + n, n_id_bad + int((~ok).sum())
        if ok.any():
            assert np.abs(ref["pts"][i, :n] - ora["pts"][a:b])[ok].max() <= tol_pts
            assert np.abs(ref["rays_d"][i, :n] - ora["dirs"][a:b])[ok].max() <= 1e-5
            assert np.abs(ref["density_weight"][i, :n] - ora["alpha_weight"][a:b])[ok].max() <= 1e-6
            assert np.abs(ref["parameters"][i, :n] - ora["params"][a:b])[ok].max() <= params_tol, i
        assert not ref["t"][i, n:].any()                            # nothing behind the last sample
    assert n_cmp > 0 and n_id_bad <= max_id_mismatch * n_cmp, (n_id_bad, n_cmp)
    assert np.array_equal(ref["density"][:, 0], ora["alpha_last"])
    return n_cmp


def reconcile_flips(ref, rec, li, max_frac=0.01):
    """A shadow edge crossed within the hit-distance rounding flips a sample between lit and ambient light: such samples
    are counted (<= max_frac of all) and then taken out of the value comparison.  Returns (flips, samples looked at)."""
    amb = np.asarray([0, 0, -1], F)
    ptr = np.concatenate([[0], np.cumsum(rec["n_samples"])])
    flips = total = 0
    for i in np.nonzero(rec["n_samples"] > 0)[0]:
        a, b = ptr[i], ptr[i + 1]
        if (ref["dist"][i] > 0).sum() != b - a:
            continue
        amb_r = (ref["parameters"][i, :b - a, li:li + 3] == amb).all(-1)
        amb_o = (rec["params"][a:b, li:li + 3] == amb).all(-1)
        flips, total = flips + int((amb_r != amb_o).sum()), total + (b - a)
        rec["params"][a:b][amb_r != amb_o] = ref["parameters"][i, :b - a][amb_r != amb_o]
    assert total > 0 and flips <= max_frac * total, (flips, total)
    return flips, total


@pytest.mark.parametrize("mean", [False, True])
@pytest.mark.parametrize("light", ["light", "point"])
def test_explicit_patches_nearest(mean, light):
    """Rotated patches given as matrices (instancer.pyx:17-18), 'nearest' choice, directional and point light."""
    T = patches(60, 1)
    R, S, step = 300, 96, 0.01
    o, d = camera_rays(R, 2)
    if light == "light":
        prm = np.tile(np.asarray([[0.5, 0.7, 0.3, 0.2, 1.0]], F), (R, 1))
        kw = dict(light_dir_idx=2)
    else:
        prm = np.tile(np.asarray([[0.5, 0.7, 40.0, 0.3, 0.2, 3.0]], F), (R, 1))
        kw = dict(light_dir_idx=3, light_strength_idx=2)
    _, ref, ora = run_both(dict(textures=["", "", light], transformations=T.tolist(), instance_sampling_method="nearest",
                                use_mean_distance=mean),
                           dict(method="nearest", use_mean_distance=mean, **kw), o, d, prm, S, step, T)
    n = compare(ref, ora, S)
    assert n > 5000 and ora["hit"].sum() > 100 and not ora["alpha_last"].any()
    assert (ora["n_samples"] == S).any() and (ora["n_samples"][ora["hit"]] < S).any()      # clipped and unclipped rays


def test_short_crossings_take_the_single_sample_branch():
    """Steps longer than most crossings: n_steps == 0 -> one sample with the whole length as its distance (:848-851)."""
    T = patches(40, 3, scale=0.05)
    R, S = 400, 16
    o, d = camera_rays(R, 4)
    prm = np.tile(np.asarray([[1.0, 0.0, 0.0, 1.0]], F), (R, 1))
    _, ref, ora = run_both(dict(textures=["", "light"], transformations=T.tolist(), instance_sampling_method="nearest"),
                           dict(method="nearest", light_dir_idx=1), o, d, prm, S, 0.25, T)
    compare(ref, ora, S)
    assert (ora["n_samples"][ora["hit"]] == 1).mean() > 0.5


def plane_mesh(n, extent, wobble, z=0.0):
    xs = np.linspace(-extent, extent, n + 1)
    X, Y = np.meshgrid(xs, xs, indexing="ij")
    Z = z + wobble * np.sin(2.1 * X) * np.cos(1.7 * Y)
    V = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(F)
    dzdx, dzdy = wobble * 2.1 * np.cos(2.1 * X) * np.cos(1.7 * Y), -wobble * 1.7 * np.sin(2.1 * X) * np.sin(1.7 * Y)
    N = np.stack([-dzdx, -dzdy, np.ones_like(X)], -1).reshape(-1, 3)
    N = (N / np.linalg.norm(N, axis=-1, keepdims=True)).astype(F)
    UV = np.stack([(X + extent) / (2 * extent), (Y + extent) / (2 * extent)], -1).reshape(-1, 2).astype(F)
    idx = lambda i, j: i * (n + 1) + j
    Fc = [[idx(i, j), idx(i + 1, j), idx(i + 1, j + 1)] for i in range(n) for j in range(n)]
    Fc += [[idx(i, j), idx(i + 1, j + 1), idx(i, j + 1)] for i in range(n) for j in range(n)]
    return {"V": V, "F": np.asarray(Fc, np.int32), "N": N, "UV": UV}


def anchors_on(mesh_fn, n, seed, lim, lift=0.0):
    rng = np.random.default_rng(seed)
    xy = rng.uniform(-lim, lim, (n, 2))
    return np.concatenate([xy, mesh_fn(xy[:, :1], xy[:, 1:]) + lift], 1).astype(F)


@pytest.mark.parametrize("jitter,with_anchors", [(0.0, True), (1.0, True), (0.5, False)])
def test_distribute_instances_on_mesh(jitter, with_anchors, tmp_path):
    """DistributeInstancesOnMesh (:233-391) run by the reference on PLY files our writer produced (read back by libigl's
    reader), exported with ExportTransformations (:1040-1061): placement, frames and the mt19937 jitter stream of
    oracle.distribute_instances_on_mesh and of the product's nerf_tex_b200.distribute agree with it."""
    from nerf_tex_b200 import distribute, ply
    mesh = plane_mesh(6, 1.0, 0.1)
    anchors = anchors_on(lambda x, y: 0.1 * np.sin(2.1 * x) * np.cos(1.7 * y), 40, 2, 0.9, lift=0.01)
    ply.write_ply(str(tmp_path / "mesh.ply"), mesh["V"], mesh["F"], mesh["N"], mesh["UV"], binary=True)
    ply.write_ply(str(tmp_path / "anchors.ply"), anchors, binary=False)
    origins = str(tmp_path / "anchors.ply") if with_anchors else ""
    scale = 0.09 if with_anchors else 0.0                            # <= 0: the average edge length (:244)
    inst = RI.RefInstancer(B0, B1, mesh_path=str(tmp_path / "mesh.ply"), patch_origins_path=origins, patch_scale=scale,
                           jitter_amount=jitter, transformation_export_path=str(tmp_path / "t.json"))
    got = np.asarray(json.loads((tmp_path / "t.json").read_text()), np.float64)
    want, used_scale = IO.distribute_instances_on_mesh(mesh["V"], mesh["F"], mesh["N"], mesh["UV"], anchors if with_anchors else None,
                                                       scale, jitter, seed=0)
    assert inst.n_instances() == want.shape[0] == (40 if with_anchors else 49)
    np.testing.assert_allclose(got, want, atol=3e-6)                 # (the export inverts the stored inverse: two roundings)
    prod, prod_scale = distribute.distribute_instances_on_mesh(mesh["V"], mesh["F"], mesh["N"], mesh["UV"],
                                                               anchors if with_anchors else None, scale, jitter, seed=0)
    np.testing.assert_allclose(got, prod, atol=3e-6)
    assert abs(prod_scale - used_scale) < 1e-7


def checker_texture(n=32, cells=4, seed=0):
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    return ((((xx * cells // n) + (yy * cells // n)) % 2) * 200 + rng.integers(0, 50, (n, n))).astype(np.uint8)


def mesh_scene(tmp_path, seed, n_anchor=150, jitter=1.0, scale=0.09):
    """A hilly plane as base mesh + anchors on it, written as PLY files for the reference; the oracle gets the arrays and
    the patch matrices of its own DistributeInstancesOnMesh (pinned to the reference's in the test above)."""
    from nerf_tex_b200 import ply
    hill = lambda x, y: 0.15 * np.sin(2.1 * x) * np.cos(1.7 * y)
    mesh = plane_mesh(12, 1.2, 0.15)
    anchors = anchors_on(hill, n_anchor, seed, 1.0)
    ply.write_ply(str(tmp_path / "mesh.ply"), mesh["V"], mesh["F"], mesh["N"], mesh["UV"], binary=True)
    ply.write_ply(str(tmp_path / "anchors.ply"), anchors, binary=True)
    T, _ = IO.distribute_instances_on_mesh(mesh["V"], mesh["F"], mesh["N"], mesh["UV"], anchors, scale, jitter, seed=0)
    ref_kw = dict(mesh_path=str(tmp_path / "mesh.ply"), patch_origins_path=str(tmp_path / "anchors.ply"), patch_scale=scale,
                  jitter_amount=jitter, instance_sampling_method="nearest")
    return mesh, T.astype(F), ref_kw


def rays_from_above(R, seed, lim=1.3):
    rng = np.random.default_rng(seed)
    o = np.stack([rng.uniform(-lim, lim, R), rng.uniform(-lim, lim, R), rng.uniform(0.7, 2.0, R)], 1).astype(F)
    d = np.stack([rng.uniform(-1, 1, R), rng.uniform(-1, 1, R), np.full(R, -1.0)], 1).astype(F)
    return o, d


def test_base_mesh_ends_rays(tmp_path):
    """Patches distributed by the reference on a PLY mesh; rays from above, from below and grazing: the first mesh hit
    closes the open segment and ends the ray with density 1 and a black last sample (:804-814, :1019-1029)."""
    mesh, T, ref_kw = mesh_scene(tmp_path, 5)
    R, S, step = 400, 128, 0.004
    o, d = rays_from_above(R, 6)
    o[::5, 2] *= -1                                                  # from below
    d[::5, 2] = 1.0
    d[1::9, 2] = -0.05                                               # grazing
    prm = np.tile(np.asarray([[0.7, 0.0, 0.0, 1.0]], F), (R, 1))
    _, ref, ora = run_both(dict(textures=["", "light"], **ref_kw),
                           dict(method="nearest", light_dir_idx=1, patch_scale=0.09, mesh=(mesh["V"], mesh["F"])), o, d, prm, S, step, T)
    compare(ref, ora, S, tol_pts=5e-4)
    assert 0.3 < ora["alpha_last"].mean() < 1.0 and (ora["alpha_last"][ora["hit"]] == 0).any()
    assert not ref["color"].any() and not ora["color_last"].any()


@pytest.mark.parametrize("n_tex", [64, 4000])
def test_parameter_textures(n_tex, tmp_path):
    """getParameters (:640-667): a PNG decoded by the reference's stb_image against the product's decoder + the oracle's
    orientation / bilinear lookup; looked up per sample (n_tex >= n_pts) or along the segments and interpolated."""
    from nerf_tex_b200.image import encode_png
    mesh, T, ref_kw = mesh_scene(tmp_path, 7)
    img0, img1 = checker_texture(32, 4, 1), checker_texture(17, 3, 2)
    for name, img in (("a.png", img0), ("b.png", img1)):
        (tmp_path / name).write_bytes(encode_png(img[:, :, None]))
    R, S, step = 200, 96, 0.004
    o, d = rays_from_above(R, 8)
    prm = np.tile(np.asarray([[0.9, 0.4, 0.7, 0.2, 0.1, 1.0]], F), (R, 1))
    radius = F(np.linalg.norm(np.maximum(np.asarray(B0, F), np.asarray(B1, F))) * F(0.09))
    tex = dict(UV=mesh["UV"], textures=IO.texture_channels(img0) + IO.texture_channels(img1), idxs=[0, 2], radius=radius,
               min_samples=8, n_samples=n_tex)
    _, ref, ora = run_both(dict(textures=[str(tmp_path / "a.png"), "", str(tmp_path / "b.png"), "light"], min_texture_samples=8,
                                n_texture_samples=n_tex, **ref_kw),
                           dict(method="nearest", light_dir_idx=3, patch_scale=0.09, mesh=(mesh["V"], mesh["F"]), texture=tex),
                           o, d, prm, S, step, T)
    # a texel edge crossed within the hit-distance rounding moves a looked-up value by a texel difference: rare, bounded
    compare(ref, ora, S, tol_pts=5e-4, params_tol=2e-3)
    p = ora["params"]
    assert p[:, 0].min() < 0.2 < 0.6 < p[:, 0].max() and len(np.unique(p[:, 2])) > 50


@pytest.mark.parametrize("light,n_shadow", [("light", 64), ("light", 4000), ("point", 64)])
def test_shadow_rays(light, n_shadow, tmp_path):
    """isShadowed / getShadowedLightDir (:572-603) with the occlusion filter of :544-551, sampled along the segments
    (:992-1000, the nearer neighbour counts) or per sample; a low sun so that hills and boxes cast shadows."""
    mesh, T, ref_kw = mesh_scene(tmp_path, 9)
    # a few patches floating above the carpet, given as matrices: they come first (instancer.pyx:17-21); their bottom
    # faces and the hills' slopes are what blocks the light
    floating = np.tile(np.eye(4), (25, 1, 1))
    floating[:, :3, :3] *= 0.09
    floating[:, :3, 3] = np.concatenate([np.random.default_rng(3).uniform(-1.0, 1.0, (25, 2)),
                                         np.random.default_rng(4).uniform(0.3, 0.5, (25, 1))], 1)
    T = np.concatenate([floating.astype(F), T], 0)
    ref_kw = dict(ref_kw, transformations=floating.astype(F).tolist())
    R, S, step = 250, 96, 0.004
    o, d = rays_from_above(R, 10)
    if light == "light":
        prm = np.tile(np.asarray([[0.7, 1.0, 0.2, 0.3]], F), (R, 1))
        kw = dict(light_dir_idx=1)
        tex = ["", "light"]
    else:
        prm = np.tile(np.asarray([[0.7, 30.0, 2.5, 0.3, 0.7]], F), (R, 1))
        kw = dict(light_dir_idx=2, light_strength_idx=1)
        tex = ["", "point"]
    _, ref, ora = run_both(dict(textures=tex, cast_shadow_rays=True, min_shadow_samples=8, n_shadow_samples=n_shadow, **ref_kw),
                           dict(method="nearest", patch_scale=0.09, mesh=(mesh["V"], mesh["F"]), cast_shadow_rays=True,
                                min_shadow_samples=8, n_shadow_samples=n_shadow, **kw), o, d, prm, S, step, T)
    li = kw["light_dir_idx"]
    reconcile_flips(ref, ora, li)
    compare(ref, ora, S, tol_pts=5e-4, params_tol=1e-4)
    shadowed = (ora["params"][:, li:li + 3] == np.asarray([0, 0, -1], F)).all(-1)
    assert 0.02 < shadowed.mean() < 0.98, float(shadowed.mean())


def test_auxiliary_meshes_are_shaded(tmp_path):
    """AddMesh / shadeMesh (:393-418, :716-744): a textured table top above the carpet and an untextured wall beside it;
    rays that end on them carry the shaded colour (texture x diffuse / ambient term, shadowed or not)."""
    from nerf_tex_b200 import ply
    from nerf_tex_b200.image import encode_png
    mesh, T, ref_kw = mesh_scene(tmp_path, 11)
    table = plane_mesh(4, 0.5, 0.0, z=0.45)
    table["V"][:, 0] += 0.3
    wall = plane_mesh(3, 0.8, 0.0)
    wall["V"] = np.stack([np.full(16, -1.15), wall["V"][:, 0], wall["V"][:, 1] + 0.6], 1).astype(F)
    wall["N"] = np.tile(np.asarray([[1, 0, 0]], F), (16, 1))
    rgb = np.stack([checker_texture(16, 2, 3), checker_texture(16, 2, 4), checker_texture(16, 2, 5)], -1)
    ply.write_ply(str(tmp_path / "table.ply"), table["V"], table["F"], table["N"], table["UV"])
    ply.write_ply(str(tmp_path / "wall.ply"), wall["V"], wall["F"], wall["N"], wall["UV"])
    (tmp_path / "table.png").write_bytes(encode_png(rgb))
    R, S, step = 300, 96, 0.004
    o, d = rays_from_above(R, 12)
    d[::4] = np.asarray([-1.0, 0.1, -0.2], F) + np.random.default_rng(13).uniform(-.1, .1, (len(d[::4]), 3)).astype(F)
    prm = np.tile(np.asarray([[0.7, 0.3, 0.2, 0.9]], F), (R, 1))
    aux_o = [dict(V=table["V"], F=table["F"], N=table["N"], UV=table["UV"], textures=IO.texture_channels(rgb)),
             dict(V=wall["V"], F=wall["F"], N=wall["N"], UV=wall["UV"], textures=[])]
    _, ref, ora = run_both(dict(textures=["", "light"], cast_shadow_rays=True, min_shadow_samples=8, n_shadow_samples=4000,
                                auxiliary_meshes=[(str(tmp_path / "table.ply"), str(tmp_path / "table.png")),
                                                  (str(tmp_path / "wall.ply"), "")], **ref_kw),
                           dict(method="nearest", light_dir_idx=1, patch_scale=0.09, mesh=(mesh["V"], mesh["F"]), cast_shadow_rays=True,
                                min_shadow_samples=8, n_shadow_samples=4000, aux_meshes=aux_o), o, d, prm, S, step, T)
    # (the table's and the wall's second triangles block light from either side: the `primID == 1` clause of :550)
    flips, _ = reconcile_flips(ref, ora, 1, max_frac=0.002)
    compare(ref, ora, S, tol_pts=5e-4, params_tol=1e-4, max_id_mismatch=0.01)
    rc, oc = ref["color"][:, 0], ora["color_last"]
    close = np.abs(rc - oc).max(-1) < 2e-3
    assert close.mean() > 0.98, float(close.mean())                  # texel / shadow edges within the rounding aside
    assert (oc.max(-1) > 0.05).sum() > 20


@pytest.mark.parametrize("method", ["random", "nearest_blend"])
def test_drawn_choices_carry_the_weights_the_restatement_gives(method):
    """'random' and 'nearest_blend' draw per sample from the reference's std::mt19937 through std::uniform_int_distribution /
    std::discrete_distribution; those streams are not reproduced.  What is checked instead, sample by sample on the
    reference's own output: the chosen patch contains the sample, and the density weight it carries is what the
    restated rule gives for THAT choice — the number of patches containing the point ('random', :670-675), or
    sum(w) / w_chosen with w_i = max(0.2 * patch_scale + min_dist - dist_i, 0) ('nearest_blend', :694-713) — with the
    containing set found independently of any hit logic (point-in-box in every patch frame)."""
    Tm = patches(80, 21, scale=0.3, spread=0.8)                     # heavy overlap
    R, S, step = 150, 64, 0.01
    o, d = camera_rays(R, 22, spread=0.8)
    prm = np.tile(np.asarray([[0.5, 0.0, 0.0, 1.0]], F), (R, 1))
    inst = RI.RefInstancer(B0, B1, textures=["", "light"], transformations=Tm.tolist(), instance_sampling_method=method)
    ref = inst.get_model_input(o, d, prm, S, step)
    minv, _, origins = IO.prepare_instances(Tm)
    b0, b1 = np.asarray(B0, np.float64), np.asarray(B1, np.float64)
    checked = multi = 0
    for i in range(R):
        n = int((ref["dist"][i] > 0).sum())
        for k in range(n):
            p = o[i].astype(np.float64) + float(ref["t"][i, k]) * d[i].astype(np.float64)
            loc = np.einsum("iab,b->ia", minv[:, :, :3].astype(np.float64), p) + minv[:, :, 3]
            margin = np.minimum(loc - b0, b1 - loc).min(-1)
            if (np.abs(margin) < 1e-4).any():                        # too close to a face to call
                continue
            inside = np.nonzero(margin > 0)[0]
            sel = int(ref["instance_id"][i, k])
            assert sel in inside, (i, k, sel, inside)
            w_ref = float(ref["density_weight"][i, k])
            if method == "random":
                assert w_ref == len(inside)
            else:
                dist = np.linalg.norm(p[None, :] - origins[inside].astype(np.float64), axis=-1)
                w = np.maximum(0.2 * 1.0 + dist.min() - dist, 0.0)   # patch_scale stays 1 without DistributeInstancesOnMesh
                assert w[list(inside).index(sel)] > 0
                assert abs(w_ref - w.sum() / w[list(inside).index(sel)]) <= 2e-3 * w_ref, (i, k)
            checked += 1
            multi += len(inside) > 1
            # the sample is mapped into the chosen patch's frame
            assert np.abs(ref["pts"][i, k] - loc[sel]).max() < 1e-4
    assert checked > 2000 and multi > 0.3 * checked

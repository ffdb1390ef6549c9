"""Host side of the instancer scene (CPU only): PLY reader / writer, DistributeInstancesOnMesh against the loop oracle,
the uniform-grid builder, and closed-form cases that pin the oracle's mesh hit and mesh termination."""

import numpy as np

from oracle import instancer_oracle as IO
from nerf_tex_b200 import distribute as D
from nerf_tex_b200 import ply, scene_grid


def plane(n=6, extent=1.0, wobble=0.1):
    xs = np.linspace(-extent, extent, n + 1)
    X, Y = np.meshgrid(xs, xs, indexing="ij")
    Z = wobble * np.sin(2.1 * X) * np.cos(1.7 * Y)
    V = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    N = np.stack([-wobble * 2.1 * np.cos(2.1 * X) * np.cos(1.7 * Y), wobble * 1.7 * np.sin(2.1 * X) * np.sin(1.7 * Y),
                  np.ones_like(X)], -1).reshape(-1, 3)
    N = (N / np.linalg.norm(N, axis=-1, keepdims=True)).astype(np.float32)
    UV = np.stack([(X + extent) / (2 * extent), (Y + extent) / (2 * extent)], -1).reshape(-1, 2).astype(np.float32)
    idx = lambda i, j: i * (n + 1) + j
    F = [t for i in range(n) for j in range(n) for t in ([idx(i, j), idx(i + 1, j), idx(i + 1, j + 1)],
                                                          [idx(i, j), idx(i + 1, j + 1), idx(i, j + 1)])]
    return V, np.asarray(F, np.int32), N, UV


def test_mt19937_and_libstdcxx_uniform_float():
    g = IO.MT19937(5489)                                            # the C++ standard's check value for mt19937
    x = [g() for _ in range(10000)]
    assert x[0] == 3499211612 and x[-1] == 4123659995
    g = IO.MT19937(0)
    want = np.array([IO.uniform_real_float(g, 0, np.pi) for _ in range(100)])
    assert np.array_equal(D.mt19937_uniform_float(0, 100, 0, np.pi), want) and want.min() >= 0 and want.max() < np.float32(np.pi)


def test_ply_round_trip(tmp_path):
    V, F, N, UV = plane()
    for binary in (True, False):
        ply.write_ply(str(tmp_path / "m.ply"), V, F, N, UV, binary=binary)
        back = ply.read_ply(str(tmp_path / "m.ply"))
        assert np.array_equal(back["V"], V) and np.array_equal(back["F"], F) and np.array_equal(back["N"], N) and np.array_equal(back["UV"], UV)
    (tmp_path / "quad.ply").write_text("ply\nformat ascii 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n"
                                       "element face 1\nproperty list uchar int vertex_indices\nend_header\n"
                                       "0 0 0\n1 0 0\n1 1 0\n0 1 0\n4 0 1 2 3\n")
    q = ply.read_ply(str(tmp_path / "quad.ply"))
    assert q["F"].tolist() == [[0, 1, 2], [0, 2, 3]] and q["N"] is None and q["UV"] is None      # polygons are fanned


def test_distribute_matches_loop_oracle_and_flat_plane_is_closed_form():
    V, F, N, UV = plane()
    rng = np.random.default_rng(2)
    xy = rng.uniform(-.9, .9, (40, 2))
    anchors = np.concatenate([xy, 0.1 * np.sin(2.1 * xy[:, :1]) * np.cos(1.7 * xy[:, 1:]) + 0.01], 1).astype(np.float32)
    for jitter in (0.0, 1.0):
        a, _ = D.distribute_instances_on_mesh(V, F, N, UV, anchors, 0.09, jitter)
        b, _ = IO.distribute_instances_on_mesh(V, F, N, UV, anchors, 0.09, jitter)
        np.testing.assert_allclose(a, b, atol=1e-6)
    a, s = D.distribute_instances_on_mesh(V, F, N, UV, None, 0.0, 0.5)
    b, s2 = IO.distribute_instances_on_mesh(V, F, N, UV, None, 0.0, 0.5)
    np.testing.assert_allclose(a, b, atol=1e-6)
    assert abs(s - s2) < 1e-7 and abs(s - D.avg_edge_length(V, F)) < 1e-12
    # flat plane, UV = xy: tangent = +x, bitangent = +y, normal = +z, so a patch is scale * I at its anchor
    Vf, Ff, Nf, UVf = plane(wobble=0.0)
    flat, _ = D.distribute_instances_on_mesh(Vf, Ff, Nf, UVf, np.array([[0.3, -0.2, 0.0]], np.float32), 0.25, 0.0)
    want = np.eye(4, dtype=np.float32)
    want[:3, :3] *= 0.25
    want[:3, 3] = [0.3, -0.2, 0.0]
    np.testing.assert_allclose(flat[0], want, atol=1e-6)


def test_grid_lists_every_item_in_every_cell_it_overlaps():
    rng = np.random.default_rng(0)
    lo = rng.uniform(-3, 3, (500, 3))
    hi = lo + rng.uniform(0.05, 0.6, (500, 3))
    g = scene_grid.build_grid(lo, hi)
    res, cell, glo = g["res"].astype(np.int64), g["cell"].astype(np.float64), g["lo"].astype(np.float64)
    assert g["cell_start"][-1] == len(g["items"]) and np.all(np.diff(g["cell_start"]) >= 0)
    for cid in rng.integers(0, res.prod(), 300):
        z, y, x = cid // (res[0] * res[1]), (cid // res[0]) % res[1], cid % res[0]
        c0 = glo + np.array([x, y, z]) * cell
        overlap = np.all((hi >= c0) & (lo <= c0 + cell), -1)
        listed = set(g["items"][g["cell_start"][cid]:g["cell_start"][cid + 1]].tolist())
        assert set(np.nonzero(overlap)[0].tolist()) <= listed


def test_oracle_mesh_hit_closed_forms():
    V = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], np.float32)
    F = np.array([[0, 1, 2], [0, 2, 3]])
    t, f, u, v = IO.mesh_first_hit([0.2, -0.5, 3], [0, 0, -1], V, F)
    assert t == 3 and f == 0 and abs(u - 0.35) < 1e-6 and abs(v - 0.25) < 1e-6      # P = v0 + u (v1 - v0) + v (v2 - v0)
    assert IO.mesh_first_hit([0.0, 0.0, 3], [0, 0, -1], V, F)[1] == 0                  # on the shared diagonal: smaller face
    assert IO.mesh_first_hit([2, 0, 3], [0, 0, -1], V, F) is None
    assert IO.mesh_first_hit([0, 0, 3], [0, 0, 1], V, F) is None                       # behind the origin
    assert IO.mesh_first_hit([0, 0, 3], [0, 0, -2], V, F)[0] == 1.5                    # t is in units of |d|
    assert IO.mesh_first_hit([0, 0, 300], [0, 0, -1], V, F) is None                    # beyond tfar = 100
    # watertight along the shared edge and at vertices: a fan of rays through the diagonal never slips through
    for s in np.linspace(-0.999, 0.999, 401):
        assert IO.mesh_first_hit([s, s, 1], [0, 0, -1], V, F) is not None
        assert IO.mesh_first_hit([0.3, 0.7, 1], [s - 0.3, s - 0.7, -1], V, F) is not None


def test_oracle_mesh_termination_closed_form():
    """One identity patch box [-.5,.5]^2 x [-.1,.1]; a floor at z = 0 cuts the crossing [2.9, 3.1] in half."""
    V = np.array([[-2, -2, 0], [2, -2, 0], [2, 2, 0], [-2, 2, 0]], np.float32)
    F = np.array([[0, 1, 2], [0, 2, 3]])
    o, d = np.array([[0.1, -0.2, 3.0]], np.float32), np.array([[0, 0, -1.0]], np.float32)
    prm = np.zeros((1, 1), np.float32)
    B0, B1 = [-0.5, -0.5, -0.1], [0.5, 0.5, 0.1]
    free = IO.get_model_input(o, d, prm, 64, 0.02, [np.eye(4)], B0, B1, seed=1)
    cut = IO.get_model_input(o, d, prm, 64, 0.02, [np.eye(4)], B0, B1, seed=1, mesh=(V, F))
    assert free["alpha_last"][0] == 0 and cut["alpha_last"][0] == 1 and cut["hit"][0]
    np.testing.assert_allclose(cut["dists"].sum(), 0.1, atol=2e-6)                    # in-shell length up to the floor
    assert abs(int(cut["n_samples"][0]) - 5) <= 1 and cut["t"].max() < 3.0 and np.all(cut["pts"][:, 2] > 0)
    np.testing.assert_allclose(free["dists"].sum(), 0.2, atol=2e-6)
    # from below: the half of the box under the floor, then the floor
    up = IO.get_model_input(np.array([[0.1, -0.2, -3.0]], np.float32), -d, prm, 64, 0.02, [np.eye(4)], B0, B1, seed=1, mesh=(V, F))
    assert up["hit"][0] and up["alpha_last"][0] == 1 and abs(int(up["n_samples"][0]) - 5) <= 1 and np.all(up["pts"][:, 2] < 0)
    # a floor in front of the box: a hit ray with an opaque last sample and no shell sample
    V2 = V + np.array([0, 0, 1.0], np.float32)
    front = IO.get_model_input(o, d, prm, 64, 0.02, [np.eye(4)], B0, B1, seed=1, mesh=(V2, F))
    assert front["hit"][0] and front["alpha_last"][0] == 1 and front["n_samples"][0] == 0
    # a ray that misses every box but hits the mesh is a hit ray too (hits.size() > 0, instancer.cpp:783-785)
    far = IO.get_model_input(np.array([[1.5, 1.5, 3.0]], np.float32), d, prm, 64, 0.02, [np.eye(4)], B0, B1, seed=1, mesh=(V, F))
    assert far["hit"][0] and far["alpha_last"][0] == 1 and far["n_samples"][0] == 0


def test_oracle_shadow_rays_closed_form():
    """shadow_ray_filter_function (instancer.cpp:539-551): own top face left from inside does not block; a bottom face
    does (either side); another patch's top face entered from above does; side faces do not."""
    B0, B1 = [-0.5, -0.5, -0.1], [0.5, 0.5, 0.1]
    M1 = IO.prepare_instances([np.eye(4)])[0]
    b0, b1 = np.float32(B0), np.float32(B1)
    inside = [0.1, 0.0, 0.0]
    assert not IO.is_shadowed(inside, [0, 0, 1.0], M1, b0, b1)             # up through the own top face, from inside
    assert IO.is_shadowed(inside, [0, 0, -1.0], M1, b0, b1)                # down through the own bottom face
    assert not IO.is_shadowed(inside, [1.0, 0, 0], M1, b0, b1)             # out through a side face
    assert IO.is_shadowed([0.1, 0, 1.0], [0, 0, -1.0], M1, b0, b1)         # from above into the top face
    assert not IO.is_shadowed([0.1, 0, 1.0], [0, 0, 1.0], M1, b0, b1)      # pointing away
    assert not IO.is_shadowed([3.0, 0, 1.0], [0, 0, -1.0], M1, b0, b1)     # beside the patch
    up = np.eye(4)
    up[2, 3] = 1.0
    M2 = IO.prepare_instances([np.eye(4), up])[0]
    assert IO.is_shadowed(inside, [0, 0, 1.0], M2, b0, b1)                 # the second patch's bottom face is in the way
    V = np.array([[-2, -2, 0.5], [2, -2, 0.5], [2, 2, 0.5], [-2, 2, 0.5]], np.float32)
    F = np.array([[0, 1, 2], [0, 2, 3]])                                   # counter-clockwise seen from +z: front = +z
    assert not IO.is_shadowed(inside, [0, 0, 1.0], M1, b0, b1, (V, F))     # reaches the sheet from behind
    assert IO.is_shadowed([0.1, 0, 2.0], [0.3, 0, -1.0], M1 * 0 + M1, b0, b1, (V, F))      # from its front side
    # through the whole pipeline: light from straight above a single patch lights every sample, from below none
    o, d = np.array([[0.1, -0.2, 3.0]], np.float32), np.array([[0.2, 0, -1.0]], np.float32)
    kw = dict(seed=1, light_dir_idx=1, cast_shadow_rays=True, min_shadow_samples=4, n_shadow_samples=64)
    lit = IO.get_model_input(o, d, np.array([[0.5, 0, 0, 2.0]], np.float32), 64, 0.02, [np.eye(4)], B0, B1, **kw)
    dark = IO.get_model_input(o, d, np.array([[0.5, 0, 0, -2.0]], np.float32), 64, 0.02, [np.eye(4)], B0, B1, **kw)
    plain = IO.get_model_input(o, d, np.array([[0.5, 0, 0, 2.0]], np.float32), 64, 0.02, [np.eye(4)], B0, B1, seed=1, light_dir_idx=1)
    assert np.array_equal(lit["params"], plain["params"]) and np.array_equal(lit["t"], dark["t"])
    assert np.all(dark["params"][:, 1:] == [0, 0, -1]) and np.all(lit["params"][:, 3] == 1)
    # per-sample path (n_shadow >= n_pts) agrees here, since the occlusion is constant along the ray
    dark2 = IO.get_model_input(o, d, np.array([[0.5, 0, 0, -2.0]], np.float32), 8, 0.02, [np.eye(4)], B0, B1,
                               **dict(kw, min_shadow_samples=100))
    assert np.all(dark2["params"][:, 1:] == [0, 0, -1])


def test_oracle_textures_and_aux_mesh_closed_form():
    """interpolate2d / loadTexture orientation, getParameters through the base mesh's UVs, shadeMesh."""
    img = (np.arange(12, dtype=np.uint8).reshape(3, 4) * 20)                # img[y, x] = (4 y + x) * 20, y from the top
    tex = IO.texture_channels(img)[0]
    assert tex.shape == (4, 3)                                               # [width, height]
    assert IO.interpolate2d([0, 0], tex) == np.float32(160 / 255)            # u = v = 0: bottom-left pixel
    assert IO.interpolate2d([1, 1], tex) == np.float32(60 / 255)             # top-right
    np.testing.assert_allclose(IO.interpolate2d([0.5, 0.5], tex), 110 / 255, rtol=1e-6)    # centre: mean of the four middle pixels
    np.testing.assert_allclose(IO.interpolate2d([1 / 3, 0.0], tex), 180 / 255, rtol=1e-6)  # on a pixel centre: that pixel
    V = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], np.float32)
    F = np.array([[0, 1, 2], [0, 2, 3]])
    UV = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    # a point above (x, y): nearest mesh point is (x, y, 0), UV = ((x + 1) / 2, (y + 1) / 2)
    out = IO.get_parameters([-1.0, -1.0, 0.3], np.array([2.0, 5.0], np.float32), V, F, UV, [tex], [1], radius=1.0)
    assert out[0] == 2 and out[1] == np.float32(5.0) * np.float32(160 / 255)
    far = IO.get_parameters([0.0, 0.0, 3.0], np.array([2.0, 5.0], np.float32), V, F, UV, [tex], [1], radius=1.0)
    assert np.array_equal(far, [2.0, 5.0])                                   # the mesh is out of reach: unchanged
    f, w = IO.closest_face([0.2, -0.5, 0.3], V, F, 5.0)
    assert f == 0 and np.allclose(V[F[0]].T @ w, [0.2, -0.5, 0.0], atol=1e-6)
    assert IO.closest_face([3.0, 0.0, 0.0], V, F, 5.0)[0] == 0 and IO.closest_face([3.0, 0.0, 0.0], V, F, 1.0) is None
    # shadeMesh: a floor facing +z lit from (0, 0, 1): diffuse 1 -> 0.8 * min(1 + 0.2, 1); lit at 60 degrees: 0.8 * 0.7
    am = {"V": V, "F": F, "N": np.tile(np.array([[0, 0, 1]], np.float32), (4, 1)), "UV": UV, "textures": []}
    M = IO.prepare_instances([np.eye(4)])[0][:0]                             # no patch boxes around
    b0, b1 = np.float32([-.5, -.5, -.1]), np.float32([.5, .5, .1])
    c = IO.shade_mesh([0.2, 0.1, 0.0], am, 0, [0.3, 0.3, 0.4], np.float32([0, 0, 2.0]), M, b0, b1, (V, F))
    np.testing.assert_allclose(c, [0.8, 0.8, 0.8], rtol=1e-6)
    c = IO.shade_mesh([0.2, 0.1, 0.0], am, 0, [0.3, 0.3, 0.4], np.float32([np.sqrt(3.0), 0, 1.0]), M, b0, b1, (V, F))
    np.testing.assert_allclose(c, [0.56, 0.56, 0.56], rtol=1e-5)
    c = IO.shade_mesh([0.2, 0.1, 0.0], am, 0, [0.3, 0.3, 0.4], np.float32([0, 0, -1.0]), M, b0, b1, (V, F))
    np.testing.assert_allclose(c, [0.16, 0.16, 0.16], rtol=1e-6)            # light from behind: ambient only
    am_t = dict(am, textures=[tex])                                          # one channel colours all three
    c = IO.shade_mesh([-1.0, -1.0, 0.0], am_t, 0, [1.0, 0.0, 0.0], np.float32([0, 0, 1.0]), M, b0, b1, (V, F))
    np.testing.assert_allclose(c, [160 / 255] * 3, rtol=1e-6)
    # through get_model_input: a ray that ends on an auxiliary floor carries its shaded colour, on the base mesh black
    o, d = np.array([[0.1, -0.2, 3.0]], np.float32), np.array([[0, 0, -1.0]], np.float32)
    prm = np.array([[0.5, 0, 0, 1.0]], np.float32)
    Vl = V - np.array([0, 0, 0.5], np.float32)                                # a floor below the patch box
    am = dict(am, V=Vl)
    aux = IO.get_model_input(o, d, prm, 64, 0.02, [np.eye(4)], [-.5, -.5, -.1], [.5, .5, .1], seed=1, light_dir_idx=1, aux_meshes=[am])
    base = IO.get_model_input(o, d, prm, 64, 0.02, [np.eye(4)], [-.5, -.5, -.1], [.5, .5, .1], seed=1, light_dir_idx=1, mesh=(Vl, F))
    assert aux["alpha_last"][0] == base["alpha_last"][0] == 1 and np.array_equal(aux["t"], base["t"])
    assert not base["color_last"].any()
    np.testing.assert_allclose(aux["color_last"][0], [0.16] * 3, rtol=1e-6)  # under the patch: its bottom face shadows the floor
    side = IO.get_model_input(np.array([[0.9, 0.9, 3.0]], np.float32), d, prm, 64, 0.02, [np.eye(4)], [-.5, -.5, -.1], [.5, .5, .1],
                              seed=1, light_dir_idx=1, aux_meshes=[am])
    np.testing.assert_allclose(side["color_last"][0], [0.8] * 3, rtol=1e-6)  # beside the patch: fully lit

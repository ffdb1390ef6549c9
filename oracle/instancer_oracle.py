"""CPU oracle for the shell instancer (instancer/src/instancer.cpp:751-1037, GetModelInput).
TEST INFRASTRUCTURE ONLY.  PINNING: the reference's own instancer.cpp is compiled here against a brute-force
stand-in for Embree's API and run on generated assets (oracle/ref_instancer.py, oracle/embree_shim); this restatement
agrees with it (tests/test_ref_instancer.py: explicit patches, patches distributed on a mesh, mesh termination,
parameter textures, shadow rays, auxiliary meshes; the device kernels are held to the same runs in
tests/test_gpu_ref_instancer.py).  Not pinned: Embree's own intersection arithmetic and traversal order, and the
`random` / `nearest_blend` draw streams.

What is restated (file:line of instancer.cpp):
  * all hits of a ray with every instanced patch box, sorted by (t, instance)            :780-799
    (Embree's quad hits are restated as a slab test in the patch frame, t in [0, 100]);
  * union of the in-shell intervals via the toggling `active_instances` set              :801-827
  * number of marching steps, per-sample `dists`, random start offset                     :843-861
  * the march itself: samples are laid out along the concatenated segments              :871-1003
  * instance choice when patches overlap: 'random' (weight n), 'nearest' (weight 1),       :671-713,:896-915
    'nearest_blend' (draw from max(range + min_dist - dist, 0), weight 1 / probability)
  * world -> patch mapping of points, view directions and the light direction; point lights  :556-590,:929-956
  * `use_mean_distance` (:746-748), no mesh hit: colour 0 / density 0                      :1020-1027
  * a base mesh in the scene (:374-391): its first hit along the ray ends the in-shell length (:804-814) and the
    march (:983), and gives the ray's last sample colour 0 / density 1 (:1019-1023).  Embree's triangle test is not
    in the reference tree; it is restated as the watertight test of Woop, Benthin and Wald (JCGT 2013) in fp32
    with the double-precision fallback on zero edge functions, ties between triangles to the smaller face index
    (the reference's Hit order, :444-455)
  * DistributeInstancesOnMesh (:233-391): per-vertex tangent frames from the UV parametrisation, closest point on
    the mesh for every anchor, interpolated frame, optional rotation about the normal by jitter_amount * U(0, pi)
    drawn from std::mt19937(seed = 0) exactly as libstdc++ draws it, patch matrix [T B N | anchor] * scale

The reference draws offsets / instance choices from a member `std::mt19937`; here they come from a counter-based
hash of (seed, ray[, step]) so that a GPU kernel can reproduce them — or, for the per-ray offsets, from `ray_u`, which
lets a test feed the reference's stream (reproduced exactly by `MT19937` / `uniform_real_float` below).  All arithmetic is fp32 in a fixed order (the kernel mirrors it with round-to-nearest
intrinsics), so the kernel is expected to match this oracle bit for bit.
"""

from __future__ import annotations

import numpy as np

F = np.float32
MAX_HITS = 200                       # MAX_TOTAL_HITS, instancer.cpp:22
T_FAR = F(100.0)


def hash_u32(x: int) -> int:
    """PCG-style integer hash (32 bit)."""
    x &= 0xFFFFFFFF
    x = (x * 747796405 + 2891336453) & 0xFFFFFFFF
    w = ((x >> ((x >> 28) + 4)) ^ x) & 0xFFFFFFFF
    w = (w * 277803737) & 0xFFFFFFFF
    return ((w >> 22) ^ w) & 0xFFFFFFFF


def ray_uniform(seed: int, ray: int) -> np.float32:
    """Uniform [0,1) with 24 random bits: the per-ray start offset draw (instancer.cpp:849,861)."""
    return F((hash_u32(hash_u32(seed) ^ (ray * 2 + 1)) >> 8) * (1.0 / 16777216.0))


def pick(seed: int, ray: int, step: int, n: int) -> int:
    """Index of the instance chosen among n overlapping ones (instancer.cpp:671-677)."""
    return hash_u32(hash_u32(seed ^ 0x9E3779B9) ^ ((ray * 1315423911 + step * 2 + 2) & 0xFFFFFFFF)) % n


def step_uniform(seed: int, ray: int, step: int) -> np.float32:
    """Uniform [0,1) per (ray, step): the discrete-distribution draw of nearest_blend (instancer.cpp:709-710)."""
    return F((hash_u32(hash_u32(seed ^ 0x85EBCA6B) ^ ((ray * 1315423911 + step * 2 + 3) & 0xFFFFFFFF)) >> 8) * (1.0 / 16777216.0))


def prepare_instances(transformations):
    """AddInstance (instancer.cpp:125-141): inverse transform, normalised transposed rotation, origin."""
    T = np.asarray(transformations, dtype=np.float64).reshape(-1, 4, 4)
    minv = np.linalg.inv(T)[:, :3, :].astype(F)                       # [I,3,4]
    rt = np.transpose(T[:, :3, :3], (0, 2, 1))
    dirt = (rt / np.linalg.norm(rt, axis=2, keepdims=True)).astype(F)  # rowwise normalised transpose
    origins = T[:, :3, 3].astype(F)
    return np.ascontiguousarray(minv), np.ascontiguousarray(dirt), np.ascontiguousarray(origins)


def _dot3(m, v, add=None):
    r = F(F(m[0] * v[0]) + F(m[1] * v[1]))
    r = F(r + F(m[2] * v[2]))
    return r if add is None else F(r + add)


def _norm3(v):
    return F(np.sqrt(F(F(F(v[0] * v[0]) + F(v[1] * v[1])) + F(v[2] * v[2]))))


def _vdot3(M, v, add=None):
    """Rows of M[.., 3] dotted with v in the fixed fp32 order of _dot3 (vectorised over the leading axes)."""
    r = (M[..., 0] * v[0]).astype(F) + (M[..., 1] * v[1]).astype(F)
    r = r.astype(F) + (M[..., 2] * v[2]).astype(F)
    r = r.astype(F)
    return r if add is None else (r + add).astype(F)


def ray_hits(o, d, minv, b0, b1, t_max=None):
    """Sorted, bounded list of (t, instance) box hits of one ray (slab test in every patch frame, all instances at
    once; every elementwise numpy op rounds to fp32 exactly like the scalar form).  `t_max`: keep only hits with
    t <= t_max (the first mesh hit; boxes sort before the mesh at equal t, instancer.cpp:444-455)."""
    o, d = np.asarray(o, F), np.asarray(d, F)
    ol = _vdot3(minv[:, :, :3], o, minv[:, :, 3])                     # [I,3]
    dl = _vdot3(minv[:, :, :3], d)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        inv = (F(1.0) / dl).astype(F)
        t0 = ((b0[None, :] - ol).astype(F) * inv).astype(F)
        t1 = ((b1[None, :] - ol).astype(F) * inv).astype(F)
        lt = t0 < t1
        lo, hi = np.where(lt, t0, t1), np.where(lt, t1, t0)
        tin, tout = lo[:, 0], hi[:, 0]
        for a in (1, 2):
            tin = np.where(lo[:, a] > tin, lo[:, a], tin)
            tout = np.where(hi[:, a] < tout, hi[:, a], tout)
        ok = (tin < tout) & (tout >= 0) & (tin <= T_FAR)
    hits = []
    for inst in np.nonzero(ok)[0]:
        if tin[inst] >= 0:
            hits.append((float(tin[inst] + F(0.0)), int(inst)))
        if tout[inst] <= T_FAR:
            hits.append((float(tout[inst] + F(0.0)), int(inst)))
    if t_max is not None:
        hits = [h for h in hits if h[0] <= float(t_max)]
    hits.sort()
    return hits[:MAX_HITS]


def ray_frame(d):
    """Per-ray constants of the watertight triangle test: dominant axis kz, the other two in winding-preserving
    order, and the shear (Woop et al., section 3)."""
    d = np.asarray(d, F)
    ad = np.abs(d)
    kz = 0 if (ad[0] >= ad[1] and ad[0] >= ad[2]) else (1 if ad[1] >= ad[2] else 2)
    kx = (kz + 1) % 3
    ky = (kx + 1) % 3
    if d[kz] < 0:
        kx, ky = ky, kx
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        Sx, Sy, Sz = F(d[kx] / d[kz]), F(d[ky] / d[kz]), F(F(1.0) / d[kz])
    return kx, ky, kz, Sx, Sy, Sz


def mesh_hits(o, d, V, Fc, want_tuv=False):
    """Mask of the triangles the ray hits with t in [0, 100] (watertight test)."""
    return _mesh_test(np.asarray(o, F), np.asarray(d, F), V, Fc)[0]


def mesh_first_hit(o, d, V, Fc):
    """Nearest hit of the ray with the triangles (V[n,3] fp32, Fc[m,3] int) for t in [0, 100]:
    (t, face, u, v) with u, v the barycentric weights of the face's 2nd and 3rd vertex, or None."""
    o, d = np.asarray(o, F), np.asarray(d, F)
    if not np.any(d != 0):
        return None
    hit, t, u, v = _mesh_test(o, d, V, Fc)
    cand = np.nonzero(hit)[0]
    if cand.shape[0] == 0:
        return None
    best = cand[np.lexsort((cand, t[cand]))[0]]                        # smallest t, then smallest face index
    return F(t[best]), int(best), F(u[best]), F(v[best])


def _mesh_test(o, d, V, Fc):
    kx, ky, kz, Sx, Sy, Sz = ray_frame(d)
    P = [(V[Fc[:, k]] - o[None, :]).astype(F) for k in range(3)]       # vertices relative to the origin
    X = [(p[:, kx] - (Sx * p[:, kz]).astype(F)).astype(F) for p in P]
    Y = [(p[:, ky] - (Sy * p[:, kz]).astype(F)).astype(F) for p in P]
    (Ax, Bx, Cx), (Ay, By, Cy) = X, Y
    U = ((Cx * By).astype(F) - (Cy * Bx).astype(F)).astype(F)
    Vv = ((Ax * Cy).astype(F) - (Ay * Cx).astype(F)).astype(F)
    W = ((Bx * Ay).astype(F) - (By * Ax).astype(F)).astype(F)
    z = (U == 0) | (Vv == 0) | (W == 0)
    if z.any():                                                        # exact products in double, rounded once
        D = np.float64
        U = np.where(z, (Cx.astype(D) * By.astype(D) - Cy.astype(D) * Bx.astype(D)).astype(F), U)
        Vv = np.where(z, (Ax.astype(D) * Cy.astype(D) - Ay.astype(D) * Cx.astype(D)).astype(F), Vv)
        W = np.where(z, (Bx.astype(D) * Ay.astype(D) - By.astype(D) * Ax.astype(D)).astype(F), W)
    miss = ((U < 0) | (Vv < 0) | (W < 0)) & ((U > 0) | (Vv > 0) | (W > 0))
    det = ((U + Vv).astype(F) + W).astype(F)
    miss |= det == 0
    Z = [(Sz * p[:, kz]).astype(F) for p in P]
    T = (((U * Z[0]).astype(F) + (Vv * Z[1]).astype(F)).astype(F) + (W * Z[2]).astype(F)).astype(F)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        t = (T / det).astype(F)
        u = (Vv / det).astype(F)
        v = (W / det).astype(F)
    miss |= ~((t >= 0) & (t <= T_FAR))
    return ~miss, t, u, v


def is_shadowed(pt, direction, minv, b0, b1, mesh=None, prim1=()):
    """isShadowed (instancer.cpp:593-603) with shadow_ray_filter_function (:539-551): a ray from `pt` along `direction`
    (not normalised), t in [0, 100], is blocked by
      * the TOP face of a patch box (quad 4 of createAABB, :117: z = b_1[2]) entered from outside, i.e. travelling
        towards -z in the patch frame,
      * the BOTTOM face of a patch box (quad 1: z = b_0[2]) from either side,
      * a mesh triangle hit on its front side (direction . Ng < 0, Ng = (v1 - v0) x (v2 - v0)),
      * and — because the filter's last clause `hit->primID == 1` (:550) does not look at the geometry — the SECOND
        triangle of the base mesh and of every auxiliary mesh from either side (`prim1`: their indices in the merged
        face list).  Found by running the reference's own code (tests/test_ref_instancer.py).
    Side faces and top faces left from inside let the ray pass."""
    pt, direction = np.asarray(pt, F), np.asarray(direction, F)
    ol = _vdot3(minv[:, :, :3], pt, minv[:, :, 3])
    dl = _vdot3(minv[:, :, :3], direction)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        inv = (F(1.0) / dl[:, 2]).astype(F)
        for plane, need_down in ((b1[2], True), (b0[2], False)):
            t = ((plane - ol[:, 2]).astype(F) * inv).astype(F)
            x = (ol[:, 0] + (t * dl[:, 0]).astype(F)).astype(F)
            y = (ol[:, 1] + (t * dl[:, 1]).astype(F)).astype(F)
            ok = (t >= 0) & (t <= T_FAR) & (x >= b0[0]) & (x <= b1[0]) & (y >= b0[1]) & (y <= b1[1])
            ok &= (dl[:, 2] < 0) if need_down else (dl[:, 2] != 0)
            if ok.any():
                return True
    if mesh is not None and np.any(direction != 0):
        V, Fc = mesh
        hit = mesh_hits(pt, direction, V, Fc)
        if hit.any():
            P = V[Fc[hit]]
            e1, e2 = (P[:, 1] - P[:, 0]).astype(F), (P[:, 2] - P[:, 0]).astype(F)
            ng = np.stack([((e1[:, 1] * e2[:, 2]).astype(F) - (e1[:, 2] * e2[:, 1]).astype(F)).astype(F),
                           ((e1[:, 2] * e2[:, 0]).astype(F) - (e1[:, 0] * e2[:, 2]).astype(F)).astype(F),
                           ((e1[:, 0] * e2[:, 1]).astype(F) - (e1[:, 1] * e2[:, 0]).astype(F)).astype(F)], -1)
            if (_vdot3(ng, direction) < 0).any():
                return True
            if any(hit[f] for f in prim1 if f < hit.shape[0]):
                return True
    return False


def texture_channels(img_u8):
    """loadTexture (instancer.cpp:34-52): one [width, height] fp32 matrix per channel, value / 255, rows indexed by the
    pixel column x and columns by the image row counted from the BOTTOM (the rowwise().reverse() of :46)."""
    img = np.asarray(img_u8)
    if img.ndim == 2:
        img = img[:, :, None]
    return [np.ascontiguousarray((img[::-1, :, c].T.astype(F) / F(255.0)).astype(F)) for c in range(img.shape[2])]


def interpolate2d(uv, tex):
    """instancer.cpp:606-626: bilinear lookup of tex[rows, cols] at uv in [0, 1]^2 (indices clamped at the far edge,
    where the reference would read one past the end with weight 0)."""
    rows, cols = tex.shape
    x0, x1 = F(F(uv[0]) * F(rows - 1)), F(F(uv[1]) * F(cols - 1))
    i0, j0 = int(x0), int(x1)
    w0, w1 = F(x0 - F(np.floor(x0))), F(x1 - F(np.floor(x1)))
    i0, j0 = min(max(i0, 0), rows - 1), min(max(j0, 0), cols - 1)
    i1, j1 = min(i0 + 1, rows - 1), min(j0 + 1, cols - 1)
    m0, m1 = F(F(1) - w0), F(F(1) - w1)
    y = F(F(tex[i0, j0] * m0) * m1)
    y = F(y + F(F(tex[i0, j1] * m0) * w1))
    y = F(y + F(F(tex[i1, j0] * w0) * m1))
    return F(y + F(F(tex[i1, j1] * w0) * w1))


def closest_face(pt, V, Fc, radius):
    """The rtcPointQuery of instancer.cpp:628-647 with closest_point_query_function (:201-231): the face nearest to
    `pt` within `radius` (strictly), ties to the smaller face index, and the barycentric weights of the closest point;
    fp32 in the fixed order below.  None if no face is in range."""
    pt = np.asarray(pt, F)
    a, b, c = V[Fc[:, 0]], V[Fc[:, 1]], V[Fc[:, 2]]
    dot = lambda x, y: _vdot3(x, [y[:, 0], y[:, 1], y[:, 2]]) if False else (
        ((x[:, 0] * y[:, 0]).astype(F) + (x[:, 1] * y[:, 1]).astype(F)).astype(F) + (x[:, 2] * y[:, 2]).astype(F)).astype(F)
    sub = lambda x, y: (x - y).astype(F)
    ab, ac, ap = sub(b, a), sub(c, a), sub(pt[None, :], a)
    d1, d2 = dot(ab, ap), dot(ac, ap)
    bp = sub(pt[None, :], b)
    d3, d4 = dot(ab, bp), dot(ac, bp)
    cp = sub(pt[None, :], c)
    d5, d6 = dot(ab, cp), dot(ac, cp)
    mul = lambda x, y: (x * y).astype(F)
    vc = sub(mul(d1, d4), mul(d3, d2))
    vb = sub(mul(d5, d2), mul(d1, d6))
    va = sub(mul(d3, d6), mul(d5, d4))
    n = Fc.shape[0]
    uvw = np.zeros((n, 3), F)
    done = np.zeros(n, bool)

    def take(mask, w0, w1, w2):
        m = mask & ~done
        uvw[m, 0], uvw[m, 1], uvw[m, 2] = w0[m], w1[m], w2[m]
        done[m] = True
    one, zero = np.ones(n, F), np.zeros(n, F)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        take((d1 <= 0) & (d2 <= 0), one, zero, zero)
        take((d3 >= 0) & (d4 <= d3), zero, one, zero)
        take((d6 >= 0) & (d5 <= d6), zero, zero, one)
        v = (d1 / sub(d1, d3)).astype(F)
        take((vc <= 0) & (d1 >= 0) & (d3 <= 0), sub(one, v), v, zero)
        v = (d2 / sub(d2, d6)).astype(F)
        take((vb <= 0) & (d2 >= 0) & (d6 <= 0), sub(one, v), zero, v)
        e43, e56 = sub(d4, d3), sub(d5, d6)
        v = (e43 / (e43 + e56).astype(F)).astype(F)
        take((va <= 0) & (e43 >= 0) & (e56 >= 0), zero, sub(one, v), v)
        denom = (F(1) / ((va + vb).astype(F) + vc).astype(F)).astype(F)
        v, w = mul(vb, denom), mul(vc, denom)
        take(np.ones(n, bool), sub(sub(one, v), w), v, w)
    # closest point = a * u + b * v + c * w (the same form for every region), distance to it
    q = ((a * uvw[:, 0:1]).astype(F) + (b * uvw[:, 1:2]).astype(F)).astype(F)
    q = (q + (c * uvw[:, 2:3]).astype(F)).astype(F)
    df = sub(pt[None, :], q)
    d = np.sqrt(dot(df, df)).astype(F)
    ok = d < F(radius)
    if not ok.any():
        return None
    cand = np.nonzero(ok)[0]
    best = cand[np.lexsort((cand, d[cand]))[0]]
    return int(best), uvw[best].copy()


def get_parameters(pt, parameters, V, Fc, UV, textures, texture_idxs, radius):
    """getParameters (instancer.cpp:628-667): the parameters of a point, each textured one multiplied by its texture at
    the UV of the nearest point of the base mesh (unchanged if the mesh is farther than `radius`).  Literal to :661:
    the i-th texture FILE multiplies parameter texture_idxs[i] by channel i of the flat channel list."""
    out = np.asarray(parameters, F).copy()
    hit = closest_face(pt, V, Fc, radius)
    if hit is None:
        return out
    f, w = Fc[hit[0]], hit[1]
    uv = [F(F(F(UV[f[0], k] * w[0]) + F(UV[f[1], k] * w[1])) + F(UV[f[2], k] * w[2])) for k in range(2)]
    for i, idx in enumerate(texture_idxs):
        out[idx] = F(out[idx] * interpolate2d(uv, textures[i]))
    return out


def shade_mesh(pt, am, prim, uvw, light_raw, minv, b0, b1, mesh, diffuse=F(1.0), ambient=F(0.2), prim1=()):
    """shadeMesh (instancer.cpp:716-744): albedo (texture at the hit's UV, or 0.8) times min(diffuse + ambient, 1);
    diffuse = max(n . l, 0) unless the point, lifted 1e-6 along the normal, is shadowed.  Without a light parameter the
    reference reads an uninitialised vector; here that case is ambient only."""
    f = np.asarray(am["F"], np.int64)[prim]
    Nm = np.asarray(am["N"], F)
    n = [F(F(F(Nm[f[0], k] * uvw[0]) + F(Nm[f[1], k] * uvw[1])) + F(Nm[f[2], k] * uvw[2])) for k in range(3)]
    nn = _norm3(n)
    n = [F(n[k] / nn) for k in range(3)]
    tex = am.get("textures") or []
    if not tex:
        albedo = [F(0.8)] * 3
    else:
        UVm = np.asarray(am["UV"], F)
        uv = [F(F(F(UVm[f[0], k] * uvw[0]) + F(UVm[f[1], k] * uvw[1])) + F(UVm[f[2], k] * uvw[2])) for k in range(2)]
        a = [interpolate2d(uv, ch) for ch in tex]
        albedo = a if len(a) == 3 else [a[0]] * 3
    if light_raw is None:
        diffuse = F(0)
    else:
        lifted = [F(F(pt[k]) + F(n[k] * F(1e-6))) for k in range(3)]
        if not is_shadowed(lifted, light_raw, minv, b0, b1, mesh, prim1):
            ln = _norm3(light_raw)
            ld = [F(light_raw[k] / ln) for k in range(3)]
            diffuse = F(diffuse * max(_dot3(n, ld), F(0)))
        else:
            diffuse = F(0)
    shade = min(F(diffuse + ambient), F(1))
    return [F(albedo[k] * shade) for k in range(3)]


def mean_distance(mu, hw):
    """instancer.cpp:746-748 (std::pow promotes to double)."""
    mu, hw = float(mu), float(hw)
    return F(mu + 2 * mu * hw ** 2 / (3 * mu ** 2 + hw ** 2))


def get_model_input(rays_o, rays_d, parameters, n_pts, step_size, transformations, b_0, b_1, *, seed=0,
                    light_dir_idx=-1, method="random", use_mean_distance=False, light_strength_idx=-1, patch_scale=1.0,
                    mesh=None, ray_ids=None, cast_shadow_rays=False, min_shadow_samples=4, n_shadow_samples=512,
                    texture=None, aux_meshes=None, ray_u=None):
    """Returns the CSR form over ALL rays: n_samples[R], hit[R], alpha_last[R], color_last[R,3] and per-sample
    records.  `mesh` = (V[n,3], F[m,3]): the base mesh the patches sit on (a ray ends at its first hit).  `ray_ids`:
    the identity of every ray for the random draws (default: its index in this call), so that a frame cut into chunks
    or spread over GPUs draws the same numbers.  `texture` = dict(UV=[n,2] of the base mesh, textures=[channel
    matrices], idxs=[parameter index per texture file], radius, min_samples, n_samples): parameter textures
    (:640-667, :916-926, :985-991).  `aux_meshes` = [dict(V, F, N, UV or None, textures=[channel matrices])]: meshes
    added with AddMesh (:393-418) — they end rays like the base mesh and colour the last sample with shadeMesh (:716-744).
    `ray_u` = [R] per-ray offsets in [0, 1) to use instead of the counter hash (tests/test_ref_instancer.py feeds the
    std::mt19937 stream the reference draws, :849/:858)."""
    rays_o, rays_d, parameters = np.asarray(rays_o, F), np.asarray(rays_d, F), np.asarray(parameters, F)
    base = None
    if mesh is not None:
        base = (np.ascontiguousarray(mesh[0], F), np.ascontiguousarray(mesh[1], np.int64))
    n_base_faces = 0 if base is None else base[1].shape[0]
    aux_ranges = []
    if aux_meshes:                                                   # one face list in (geomID, primID) order (:444-455)
        Vs, Fs = ([base[0]], [base[1]]) if base is not None else ([], [])
        n_v, n_f = sum(v.shape[0] for v in Vs), n_base_faces
        for am in aux_meshes:
            Vs.append(np.ascontiguousarray(am["V"], F))
            Fs.append(np.ascontiguousarray(am["F"], np.int64) + n_v)
            aux_ranges.append((n_f, n_f + Fs[-1].shape[0]))
            n_v, n_f = n_v + Vs[-1].shape[0], n_f + Fs[-1].shape[0]
        mesh = (np.concatenate(Vs), np.concatenate(Fs))
    else:
        mesh = base
    # the second triangle of every mesh, in the merged face list (the shadow filter's `primID == 1` clause, :550)
    prim1 = ([1] if n_base_faces >= 2 else []) + [f0 + 1 for f0, f1 in aux_ranges if f1 - f0 >= 2]
    minv, dirt, origins = prepare_instances(transformations)
    b0, b1 = np.asarray(b_0, F), np.asarray(b_1, F)
    step_size = F(step_size)
    R = rays_o.shape[0]
    n_samples = np.zeros(R, np.int32)
    hit = np.zeros(R, bool)
    alpha_last = np.zeros(R, F)
    color_last = np.zeros((R, 3), F)
    recs = dict(pts=[], dirs=[], params=[], dists=[], t=[], weight=[], inst=[])
    for i in range(R):
        o, d = rays_o[i], rays_d[i]
        rid = i if ray_ids is None else int(ray_ids[i])
        mh = mesh_first_hit(o, d, *mesh) if mesh is not None else None
        hits = ray_hits(o, d, minv, b0, b1, None if mh is None else mh[0])
        if mh is not None and len(hits) >= MAX_HITS:                  # the mesh hit would be entry 201 of the sorted list
            mh = None
        if not hits and mh is None:
            continue
        hit[i] = True
        if mh is not None:                                          # :1019-1023 (instancer mesh: black, opaque)
            alpha_last[i] = F(1)
            for am, (f0, f1) in zip(aux_meshes or [], aux_ranges):   # an auxiliary mesh: shaded (:1022)
                if f0 <= mh[1] < f1:
                    raw = parameters[i, light_dir_idx:light_dir_idx + 3] if light_dir_idx >= 0 else None
                    color_last[i] = shade_mesh([F(o[a] + F(mh[0] * d[a])) for a in range(3)], am, mh[1] - f0,
                                               [F(F(F(1) - mh[2]) - mh[3]), mh[2], mh[3]], raw, minv, b0, b1, mesh, prim1=prim1)
        # :801-827 total in-shell length
        active, total, t_entry = set(), F(0), F(0)
        segment_lengths = []
        for tt, inst in hits:
            tt = F(tt)
            if inst in active:
                active.discard(inst)
                if not active:
                    total = F(total + F(tt - t_entry))
                    segment_lengths.append(F(tt - t_entry))
            else:
                if not active:
                    t_entry = tt
                active.add(inst)
        if mh is not None and active:                               # :805-810 the mesh closes the open segment
            total = F(total + F(mh[0] - t_entry))
            segment_lengths.append(F(mh[0] - t_entry))
        if not (total > 0):
            continue
        necessary = int(F(total / step_size))                       # :845
        n_steps = min(necessary, n_pts)
        u = ray_uniform(seed, rid) if ray_u is None else F(ray_u[i])
        dists = []
        if n_steps == 0:                                            # :848-851
            dists = [total]
            t_offset = F(u * total)
            n_steps = 1
        else:                                                       # :852-861
            dists = [step_size] * (n_steps - 1) + [F(F(step_size + total) - F(F(n_steps) * step_size))]
            t_offset = F(u * step_size)
        dn_norm = _norm3(d)
        dn = [F(d[a] / dn_norm) for a in range(3)]
        light = None
        if light_dir_idx >= 0 and light_strength_idx < 0:                              # directional light
            lv = parameters[i, light_dir_idx:light_dir_idx + 3]
            ln = _norm3(lv)
            light = [F(lv[a] / ln) for a in range(3)]
        # :871-1003 march
        active = set()
        segment_offset, cleared, t_entry, step = F(0), F(0), F(0), 0
        emitted = 0
        # shadow rays (:863, :936-952, :992-1000): with fewer shadow samples than marching samples the occlusion is
        # looked up at points spaced evenly along each segment and the nearer of the two neighbours is used
        total_len = total
        shadow_on = cast_shadow_rays and light_dir_idx >= 0
        n_shadow = max(int(min_shadow_samples), int(F(F(n_shadow_samples) * total_len))) if shadow_on else 0
        shadow_interp = shadow_on and n_shadow < n_pts
        raw_light = parameters[i, light_dir_idx:light_dir_idx + 3].copy() if light_dir_idx >= 0 else None
        seg_l, sh_step, sh_entry = 0, F(0), F(0)
        # parameter textures (:864, :916-926, :985-991): like the shadows, but linearly interpolated
        tex_on = texture is not None and len(texture["idxs"]) > 0 and base is not None
        n_tex = max(int(texture["min_samples"]), int(F(F(texture["n_samples"]) * total_len))) if tex_on else 0
        tex_interp = tex_on and n_tex < n_pts
        tx_step, tx_entry = F(0), F(0)
        textured = lambda tq: get_parameters([F(o[a] + F(F(tq) * d[a])) for a in range(3)], parameters[i], base[0], base[1],
                                             texture["UV"], texture["textures"], texture["idxs"], texture["radius"])
        occluded = lambda tq: is_shadowed([F(o[a] + F(F(tq) * d[a])) for a in range(3)], raw_light, minv, b0, b1, mesh, prim1)
        for tt, inst in hits + ([(float(mh[0]), -1)] if mh is not None else []):
            tt = F(tt)
            if step >= n_steps:
                break

            def t_of(s):
                t_mu = F(F(F(F(s) * step_size) + t_offset) + segment_offset)
                return t_mu, (mean_distance(t_mu, step_size) if use_mean_distance else t_mu)

            t_mu, t_pt = t_of(step)
            while active and t_pt < tt and step < n_steps:
                pt = [F(o[a] + F(t_pt * d[a])) for a in range(3)]
                act = sorted(active)
                if len(act) == 1:
                    sel, w = act[0], F(1)
                elif method == "random":
                    sel, w = act[pick(seed, rid, step, len(act))], F(len(act))
                elif method == "nearest_blend":                     # :694-713, transition_range 0.2 (instancer.hpp:84)
                    ds = [_norm3([F(pt[c] - origins[a_][c]) for c in range(3)]) for a_ in act]
                    lim = F(F(F(0.2) * F(patch_scale)) + min(ds))
                    ws = [max(F(lim - d_), F(0)) for d_ in ds]
                    total = F(0)
                    for w_ in ws:
                        total = F(total + w_)
                    target = F(step_uniform(seed, rid, step) * total)
                    run, sel, w_sel = F(0), act[0], F(1)
                    for a_, w_ in zip(act, ws):
                        run = F(run + w_)
                        if w_ > 0:
                            sel, w_sel = a_, w_
                            if target < run:
                                break
                    w = F(total / w_sel)
                else:                                               # nearest, :680-692
                    best, sel = F(np.inf), act[0]
                    for a_ in act:
                        dist = _norm3([F(pt[c] - origins[a_][c]) for c in range(3)])
                        if dist < best:
                            best, sel = dist, a_
                    w = F(1)
                M = minv[sel]
                recs["pts"].append([_dot3(M[a, :3], pt, M[a, 3]) for a in range(3)])        # getPt :556-558
                recs["dirs"].append([_dot3(dirt[sel][a], dn) for a in range(3)])           # getDir :561-563
                prm = parameters[i].copy()
                if tex_interp:                                      # :916-923
                    k = 1
                    while t_pt > F(tx_entry + F(F(k) * tx_step)):
                        k += 1
                    t0 = tx_entry if k == 1 else F(tx_entry + F(F(k - 1) * tx_step))
                    t1 = F(tx_entry + F(F(k) * tx_step))
                    wt = F(F(t_pt - t0) / tx_step)
                    s0, s1 = textured(t0), textured(t1)
                    prm = ((s0 * F(F(1) - wt)).astype(F) + (s1 * wt).astype(F)).astype(F)
                elif tex_on:                                        # :924-926
                    prm = get_parameters(pt, parameters[i], base[0], base[1], texture["UV"], texture["textures"],
                                         texture["idxs"], texture["radius"])
                shadowed = False
                if shadow_interp:                                   # :937-948 nearest of the two shadow samples around t_pt
                    k = 1
                    while t_pt > F(sh_entry + F(F(k) * sh_step)):
                        k += 1
                    t0 = sh_entry if k == 1 else F(sh_entry + F(F(k - 1) * sh_step))
                    t1 = F(sh_entry + F(F(k) * sh_step))
                    shadowed = occluded(t1 if F(F(t_pt - t0) / sh_step) >= F(0.5) else t0)
                elif shadow_on:                                     # :949-951 at the sample itself
                    shadowed = is_shadowed(pt, raw_light, minv, b0, b1, mesh, prim1)
                if shadowed:                                        # :572-575 ambient only
                    prm[light_dir_idx:light_dir_idx + 3] = [F(0), F(0), F(-1)]
                    if light_strength_idx >= 0:                     # :954-956 the strength is mapped regardless
                        lv = [F(parameters[i, light_dir_idx + a] - pt[a]) for a in range(3)]
                        dsq = F(F(F(lv[0] * lv[0]) + F(lv[1] * lv[1])) + F(lv[2] * lv[2]))
                        prm[light_strength_idx] = F(parameters[i, light_strength_idx] / F(F(F(12.566370614359172) * dsq) + F(1e-6)))
                elif light is not None:
                    prm[light_dir_idx:light_dir_idx + 3] = [_dot3(dirt[sel][a], light) for a in range(3)]
                elif light_dir_idx >= 0:                            # point light: :572-590, :954-956
                    lv = [F(parameters[i, light_dir_idx + a] - pt[a]) for a in range(3)]
                    dsq = F(F(F(lv[0] * lv[0]) + F(lv[1] * lv[1])) + F(lv[2] * lv[2]))
                    prm[light_strength_idx] = F(parameters[i, light_strength_idx] / F(F(F(12.566370614359172) * dsq) + F(1e-6)))
                    ln = _norm3(lv)
                    lvn = [F(lv[a] / ln) for a in range(3)]
                    prm[light_dir_idx:light_dir_idx + 3] = [_dot3(dirt[sel][a], lvn) for a in range(3)]
                recs["params"].append(prm)
                recs["dists"].append(dists[step])
                recs["t"].append(t_mu)
                recs["weight"].append(w)
                recs["inst"].append(sel)
                emitted += 1
                step += 1
                t_mu, t_pt = t_of(step)
            if inst < 0:                                            # :983 the mesh culls whatever lies behind it
                break
            if inst in active:
                active.discard(inst)
                if not active:
                    cleared = F(cleared + F(tt - t_entry))
            else:
                if not active:
                    segment_offset = F(tt - cleared)
                    t_entry = tt
                    if shadow_interp:                               # :992-1000
                        seg_len = segment_lengths[seg_l]
                        n_seg = max(int(min_shadow_samples), int(F(F(F(n_shadow) * seg_len) / total_len)))
                        sh_step, sh_entry = F(seg_len / F(n_seg - 1)), t_entry
                    if tex_interp:                                  # :985-991
                        seg_len = segment_lengths[seg_l]
                        n_seg = max(int(texture["min_samples"]), int(F(F(F(n_tex) * seg_len) / total_len)))
                        tx_step, tx_entry = F(seg_len / F(n_seg - 1)), t_entry
                    seg_l += 1
                active.add(inst)
        n_samples[i] = emitted
    p = parameters.shape[1]
    out = {
        "n_samples": n_samples, "hit": hit, "alpha_last": alpha_last, "color_last": color_last,
        "pts": np.asarray(recs["pts"], F).reshape(-1, 3), "dirs": np.asarray(recs["dirs"], F).reshape(-1, 3),
        "params": np.asarray(recs["params"], F).reshape(len(recs["params"]), p), "dists": np.asarray(recs["dists"], F),
        "t": np.asarray(recs["t"], F), "alpha_weight": np.asarray(recs["weight"], F),
        "instance_id": np.asarray(recs["inst"], np.int32),
    }
    return out


# ----------------------------------------------------------------------------------------------------------------------
# DistributeInstancesOnMesh (instancer.cpp:233-391), restated with plain loops (small meshes only)
# ----------------------------------------------------------------------------------------------------------------------

class MT19937:
    """std::mt19937 (Matsumoto & Nishimura's reference recurrence; init_genrand seeding)."""

    def __init__(self, seed=5489):
        self.mt = [0] * 624
        self.mt[0] = seed & 0xFFFFFFFF
        for i in range(1, 624):
            self.mt[i] = (1812433253 * (self.mt[i - 1] ^ (self.mt[i - 1] >> 30)) + i) & 0xFFFFFFFF
        self.idx = 624

    def __call__(self):
        if self.idx >= 624:
            mt = self.mt
            for k in range(624):
                y = (mt[k] & 0x80000000) | (mt[(k + 1) % 624] & 0x7FFFFFFF)
                mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0)
            self.idx = 0
        y = self.mt[self.idx]
        self.idx += 1
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        y ^= y >> 18
        return y & 0xFFFFFFFF


def uniform_real_float(gen, a, b):
    """libstdc++ std::uniform_real_distribution<float>(a, b)(gen): generate_canonical<float, 24> takes one 32-bit
    output, float(x) / 2^32, and replaces a result of 1 by the float below it."""
    c = F(F(gen()) / F(4294967296.0))
    if c >= F(1):
        c = np.nextafter(F(1), F(0))
    return F(F(c * F(F(b) - F(a))) + F(a))


def closest_point_on_triangle(p, a, b, c):
    """instancer.cpp:151-199 (Ericson, Real-Time Collision Detection 5.1.5): (point, barycentric weights)."""
    ab, ac, ap = b - a, c - a, p - a
    d1, d2 = ab @ ap, ac @ ap
    if d1 <= 0 and d2 <= 0:
        return a, np.array([1, 0, 0], F)
    bp = p - b
    d3, d4 = ab @ bp, ac @ bp
    if d3 >= 0 and d4 <= d3:
        return b, np.array([0, 1, 0], F)
    cp = p - c
    d5, d6 = ab @ cp, ac @ cp
    if d6 >= 0 and d5 <= d6:
        return c, np.array([0, 0, 1], F)
    vc = d1 * d4 - d3 * d2
    if vc <= 0 and d1 >= 0 and d3 <= 0:
        v = d1 / (d1 - d3)
        return a + v * ab, np.array([1 - v, v, 0], F)
    vb = d5 * d2 - d1 * d6
    if vb <= 0 and d2 >= 0 and d6 <= 0:
        v = d2 / (d2 - d6)
        return a + v * ac, np.array([1 - v, 0, v], F)
    va = d3 * d6 - d5 * d4
    if va <= 0 and (d4 - d3) >= 0 and (d5 - d6) >= 0:
        v = (d4 - d3) / ((d4 - d3) + (d5 - d6))
        return b + v * (c - b), np.array([0, 1 - v, v], F)
    denom = 1 / (va + vb + vc)
    v, w = vb * denom, vc * denom
    return a + v * ab + w * ac, np.array([1 - v - w, v, w], F)


def distribute_instances_on_mesh(V, Fc, N, UV, patch_origins=None, scale=1.0, jitter_amount=0.0, seed=0):
    """instancer.cpp:233-372: list of 4x4 patch matrices (what AddInstance receives) and the patch scale."""
    V, N, UV, Fc = np.asarray(V, F), np.asarray(N, F).copy(), np.asarray(UV, F), np.asarray(Fc, np.int64)
    nv, nf = V.shape[0], Fc.shape[0]
    ael, edges = 0.0, set()                                            # igl::avg_edge_length: every undirected edge ONCE
    for f in Fc:                                                       # (igl::edges), lengths in fp32, sum in double
        for k in range(3):
            e = (min(int(f[k]), int(f[(k + 1) % 3])), max(int(f[k]), int(f[(k + 1) % 3])))
            if e in edges:
                continue
            edges.add(e)
            dv = [F(V[e[0]][c] - V[e[1]][c]) for c in range(3)]
            ael += float(np.sqrt(F(F(F(dv[0] * dv[0]) + F(dv[1] * dv[1])) + F(dv[2] * dv[2]))))
    ael /= len(edges)
    if scale <= 0:
        scale = ael
    T = np.zeros((nv, 3), F)
    B = np.zeros((nv, 3), F)
    for f in Fc:                                                       # :250-265
        e0, e1 = V[f[1]] - V[f[0]], V[f[2]] - V[f[0]]
        uv0, uv1 = UV[f[1]] - UV[f[0]], UV[f[2]] - UV[f[0]]
        r = F(1) / (uv0[0] * uv1[1] - uv0[1] * uv1[0])
        t = (e0 * uv1[1] - e1 * uv0[1]) * r
        for k in range(3):
            T[f[k]] += t
    for i in range(nv):                                                # :266-277
        N[i] = N[i] / np.linalg.norm(N[i])
        n, t = N[i], T[i]
        t = t - n * (n @ t)
        t = t / np.linalg.norm(t)
        T[i], B[i] = t, np.cross(n, t)
    gen = MT19937(seed)
    out = []
    if patch_origins is not None and len(patch_origins) > 0:
        for pt in np.asarray(patch_origins, F):                        # :299-345
            radius, best, uvw = F(ael), -1, None
            for j, f in enumerate(Fc):                                 # point query: strictly closer wins
                cp, w = closest_point_on_triangle(pt, V[f[0]], V[f[1]], V[f[2]])
                d = F(np.linalg.norm(pt - cp))
                if d < radius:
                    radius, best, uvw = d, j, w
            assert best >= 0, "patch origin farther than the average edge length from the mesh"
            f = Fc[best]
            iN = N[f[0]] * uvw[0] + N[f[1]] * uvw[1] + N[f[2]] * uvw[2]
            iT = T[f[0]] * uvw[0] + T[f[1]] * uvw[1] + T[f[2]] * uvw[2]
            iN, iT = iN / np.linalg.norm(iN), iT / np.linalg.norm(iT)
            iB = np.cross(iN, iT)
            if jitter_amount > 0:
                ang = F(F(jitter_amount) * uniform_real_float(gen, 0.0, np.pi))
                iB = iB * np.cos(ang) + np.cross(iN, iB) * np.sin(ang) + iN * (iN @ iB) * (1 - np.cos(ang))
            iT = np.cross(iB, iN)
            M = np.eye(4, dtype=F)
            M[:3, 0], M[:3, 1], M[:3, 2], M[:3, 3] = iT * F(scale), iB * F(scale), iN * F(scale), pt
            out.append(M)
    else:
        seen = []
        for i in range(nv):                                            # :350-372
            if any(np.array_equal(V[i], s) for s in seen):
                continue
            tT, tB, tN = T[i], B[i], N[i]
            if jitter_amount > 0:
                ang = F(F(jitter_amount) * uniform_real_float(gen, 0.0, np.pi))
                tB = tB * np.cos(ang) + np.cross(tN, tB) * np.sin(ang) + tN * (tN @ tB) * (1 - np.cos(ang))
                tc = np.cross(tN, tB)
                tT = (F(-1) if tT @ tc < 0 else F(1)) * tc
            M = np.eye(4, dtype=F)
            M[:3, 0], M[:3, 1], M[:3, 2], M[:3, 3] = tT * F(scale), tB * F(scale), tN * F(scale), V[i]
            out.append(M)
            seen.append(V[i])
    return np.asarray(out, F), float(scale)

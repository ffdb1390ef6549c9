"""DistributeInstancesOnMesh (instancer/src/instancer.cpp:233-391) on the host, once per scene: tangent frames of
the base mesh from its UV parametrisation, the closest mesh point of every anchor, the interpolated frame, the optional
rotation about the normal, and the patch matrix [T B N | anchor] * scale that AddInstance (:125-141) receives.
Vectorised NumPy in fp32 (the reference computes in Eigen floats); the jitter angles are drawn from std::mt19937 the way
libstdc++'s uniform_real_distribution<float> draws them, so a scene built here places and turns its patches like the
reference does for the same files and seed."""

from __future__ import annotations

import numpy as np

from .scene_grid import build_grid

F32 = np.float32


def mt19937_uniform_float(seed: int, n: int, a: float, b: float) -> np.ndarray:
    """n draws of std::uniform_real_distribution<float>(a, b) from std::mt19937(seed) (libstdc++: one 32-bit output per
    float, canonical = float(x) / 2^32, a result of 1 is replaced by the float below 1)."""
    bg = np.random.MT19937()
    bg._legacy_seeding(int(seed))                        # init_genrand, the std::mt19937(seed) constructor
    x = bg.random_raw(n).astype(np.uint32)
    c = (x.astype(F32) / F32(4294967296.0)).astype(F32)
    c = np.where(c >= F32(1), np.nextafter(F32(1), F32(0)), c).astype(F32)
    return (c * F32(F32(b) - F32(a)) + F32(a)).astype(F32)


def avg_edge_length(V: np.ndarray, F: np.ndarray) -> float:
    """igl::avg_edge_length: mean over the mesh's UNIQUE undirected edges (igl::edges), every length taken in the
    vertices' own precision (fp32 in the reference, instancer.cpp:244) and summed in double.  (Checked against the
    reference's own code in tests/test_ref_instancer.py.)"""
    F = np.asarray(F, np.int64)
    E = np.concatenate([F[:, [0, 1]], F[:, [1, 2]], F[:, [2, 0]]], 0)
    E = np.unique(np.sort(E, axis=1), axis=0)
    d = (np.asarray(V, F32)[E[:, 0]] - np.asarray(V, F32)[E[:, 1]]).astype(F32)
    sq = ((d[:, 0] * d[:, 0]).astype(F32) + (d[:, 1] * d[:, 1]).astype(F32)).astype(F32)
    sq = (sq + (d[:, 2] * d[:, 2]).astype(F32)).astype(F32)
    return float(np.sqrt(sq).astype(F32).astype(np.float64).sum() / E.shape[0])


def _unit(v):
    return (v / np.sqrt((v * v).sum(-1, keepdims=True, dtype=F32)).astype(F32)).astype(F32)


def vertex_frames(V, F, N, UV):
    """Per-vertex normal, tangent, bitangent (:248-277)."""
    V, N, UV, F = np.asarray(V, F32), np.asarray(N, F32), np.asarray(UV, F32), np.asarray(F, np.int64)
    e0, e1 = V[F[:, 1]] - V[F[:, 0]], V[F[:, 2]] - V[F[:, 0]]
    uv0, uv1 = UV[F[:, 1]] - UV[F[:, 0]], UV[F[:, 2]] - UV[F[:, 0]]
    with np.errstate(divide="ignore", invalid="ignore"):
        r = (F32(1) / (uv0[:, 0] * uv1[:, 1] - uv0[:, 1] * uv1[:, 0])).astype(F32)
    t = ((e0 * uv1[:, 1:2] - e1 * uv0[:, 1:2]) * r[:, None]).astype(F32)
    T = np.zeros_like(V)
    for k in range(3):                                   # faces in order, like the reference's accumulation
        np.add.at(T, F[:, k], t)
    n = _unit(N)
    T = T - n * (n * T).sum(-1, keepdims=True, dtype=F32)
    T = _unit(T.astype(F32))
    return n, T, np.cross(n, T).astype(F32)


def closest_point_triangle(p, a, b, c):
    """Ericson's closest point on a triangle (:151-199), vectorised over rows; returns (point, uvw)."""
    ab, ac, ap = b - a, c - a, p - a
    d1, d2 = (ab * ap).sum(-1), (ac * ap).sum(-1)
    bp = p - b
    d3, d4 = (ab * bp).sum(-1), (ac * bp).sum(-1)
    cp = p - c
    d5, d6 = (ab * cp).sum(-1), (ac * cp).sum(-1)
    vc, vb, va = d1 * d4 - d3 * d2, d5 * d2 - d1 * d6, d3 * d6 - d5 * d4
    n = p.shape[0]
    uvw = np.zeros((n, 3), p.dtype)
    done = np.zeros(n, bool)

    def take(mask, w):
        m = mask & ~done
        uvw[m] = w[m] if w.ndim == 2 else w
        done[m] = True

    one = np.ones(n, p.dtype)
    zero = np.zeros(n, p.dtype)
    with np.errstate(divide="ignore", invalid="ignore"):
        take((d1 <= 0) & (d2 <= 0), np.stack([one, zero, zero], -1))
        take((d3 >= 0) & (d4 <= d3), np.stack([zero, one, zero], -1))
        take((d6 >= 0) & (d5 <= d6), np.stack([zero, zero, one], -1))
        v = d1 / (d1 - d3)
        take((vc <= 0) & (d1 >= 0) & (d3 <= 0), np.stack([1 - v, v, zero], -1))
        v = d2 / (d2 - d6)
        take((vb <= 0) & (d2 >= 0) & (d6 <= 0), np.stack([1 - v, zero, v], -1))
        v = (d4 - d3) / ((d4 - d3) + (d5 - d6))
        take((va <= 0) & ((d4 - d3) >= 0) & ((d5 - d6) >= 0), np.stack([zero, 1 - v, v], -1))
        denom = 1 / (va + vb + vc)
        v, w = vb * denom, vc * denom
        take(np.ones(n, bool), np.stack([1 - v - w, v, w], -1))
    pt = a * uvw[:, 0:1] + b * uvw[:, 1:2] + c * uvw[:, 2:3]
    return pt, uvw


def closest_faces(points, V, F, radius: float):
    """For every point the mesh face nearest to it within `radius` (ties: smaller face index) and the barycentric
    weights of the closest point (the rtcPointQuery of :300-312).  Raises if a point has no face in range."""
    V, F = np.asarray(V, F32), np.asarray(F, np.int64)
    P = V[F]
    grid = build_grid(P.min(1) - radius, P.max(1) + radius)
    pts = np.asarray(points, F32)
    c = np.floor((pts.astype(np.float64) - grid["lo"]) / grid["cell"]).astype(np.int64)
    inside = np.all((c >= 0) & (c < grid["res"]), -1)
    c = np.clip(c, 0, grid["res"] - 1)
    cell = (c[:, 2] * grid["res"][1] + c[:, 1]) * grid["res"][0] + c[:, 0]
    start, stop = grid["cell_start"][cell].astype(np.int64), grid["cell_start"][cell + 1].astype(np.int64)
    cnt = np.where(inside, stop - start, 0)
    owner = np.repeat(np.arange(pts.shape[0]), cnt)
    k = np.arange(int(cnt.sum())) - np.repeat(np.cumsum(cnt) - cnt, cnt)
    face = grid["items"][np.repeat(start, cnt) + k].astype(np.int64)
    cp, uvw = closest_point_triangle(pts[owner], P[face, 0], P[face, 1], P[face, 2])
    d = np.sqrt(((pts[owner] - cp) ** 2).sum(-1, dtype=F32)).astype(F32)
    ok = d < F32(radius)
    owner, face, uvw, d = owner[ok], face[ok], uvw[ok], d[ok]
    order = np.lexsort((face, d, owner))
    first = np.ones(order.shape[0], bool)
    first[1:] = owner[order][1:] != owner[order][:-1]
    sel = order[first]
    if sel.shape[0] != pts.shape[0]:
        missing = np.setdiff1d(np.arange(pts.shape[0]), owner[sel])
        raise ValueError(f"{missing.shape[0]} patch origins (first: #{int(missing[0])}) have no mesh face within the "
                         f"average edge length {radius:g}: the reference's point query fails for them too")
    return face[sel], uvw[sel].astype(F32)


def rotate_about(n, b, angle):
    """Rodrigues rotation of b about the unit axis n (:326-328)."""
    c, s = np.cos(angle).astype(F32)[:, None], np.sin(angle).astype(F32)[:, None]
    return (b * c + np.cross(n, b) * s + n * (n * b).sum(-1, keepdims=True, dtype=F32) * (1 - c)).astype(F32)


def distribute_instances_on_mesh(V, F, N, UV, patch_origins=None, scale: float = 1.0, jitter_amount: float = 0.0,
                                 seed: int = 0):
    """Returns (transformations [I,4,4] fp32, patch_scale).  `patch_origins` [A,3] or None (then one patch per distinct
    mesh vertex, :350-372)."""
    V, F = np.asarray(V, F32), np.asarray(F, np.int64)
    ael = avg_edge_length(V, F)
    if scale <= 0:                                       # :244
        scale = ael
    n, T, B = vertex_frames(V, F, N, UV)
    if patch_origins is not None and len(patch_origins) > 0:
        pts = np.asarray(patch_origins, F32)
        face, uvw = closest_faces(pts, V, F, ael)
        f = F[face]
        mix = lambda X: (X[f[:, 0]] * uvw[:, 0:1] + X[f[:, 1]] * uvw[:, 1:2] + X[f[:, 2]] * uvw[:, 2:3]).astype(F32)
        iN, iT = _unit(mix(n)), _unit(mix(T))
        iB = np.cross(iN, iT).astype(F32)
        if jitter_amount > 0:
            ang = (F32(jitter_amount) * mt19937_uniform_float(seed, pts.shape[0], 0.0, np.pi)).astype(F32)
            iB = rotate_about(iN, iB, ang)
        iT = np.cross(iB, iN).astype(F32)
        origins = pts
    else:
        _, first = np.unique(V, axis=0, return_index=True)
        keep = np.sort(first)                            # first occurrence of every distinct position, in vertex order
        iN, iT, iB, origins = n[keep], T[keep], B[keep], V[keep]
        if jitter_amount > 0:
            ang = (F32(jitter_amount) * mt19937_uniform_float(seed, keep.shape[0], 0.0, np.pi)).astype(F32)
            iB = rotate_about(iN, iB, ang)
            tc = np.cross(iN, iB).astype(F32)
            sgn = np.where((iT * tc).sum(-1, keepdims=True) < 0, F32(-1), F32(1))
            iT = (sgn * tc).astype(F32)
    M = np.tile(np.eye(4, dtype=F32), (origins.shape[0], 1, 1))
    M[:, :3, 0], M[:, :3, 1], M[:, :3, 2], M[:, :3, 3] = iT * F32(scale), iB * F32(scale), iN * F32(scale), origins
    return M, float(scale)

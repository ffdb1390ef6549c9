"""CPU oracle for the shell instancer (instancer/src/instancer.cpp:751-1037, GetModelInput).
TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED: the reference instancer needs Embree 3 (absent) and its
meshes/textures are git-LFS stubs, so it can neither be built nor fed here; this is a restatement of
its algorithm for the subset that needs no mesh: explicit patch instances (`transformations`), no
parameter textures, no shadow rays, no auxiliary meshes.

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

Not reproducible by construction: the reference draws offsets / instance choices from a member
`std::mt19937`; here they come from a counter-based hash of (seed, ray[, step]) so that a GPU kernel can
reproduce them.  All arithmetic is fp32 in a fixed order (the kernel mirrors it with round-to-nearest
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


def ray_hits(o, d, minv, b0, b1):
    """Sorted, bounded list of (t, instance) box hits of one ray."""
    hits = []
    for inst in range(minv.shape[0]):
        M = minv[inst]
        ol = [_dot3(M[a, :3], o, M[a, 3]) for a in range(3)]
        dl = [_dot3(M[a, :3], d) for a in range(3)]
        tin, tout = F(-np.inf), F(np.inf)
        with np.errstate(divide="ignore", invalid="ignore"):
            for a in range(3):
                inv = F(F(1.0) / dl[a])
                t0 = F(F(b0[a] - ol[a]) * inv)
                t1 = F(F(b1[a] - ol[a]) * inv)
                lo, hi = (t0, t1) if t0 < t1 else (t1, t0)
                tin = lo if (a == 0 or lo > tin) else tin
                tout = hi if (a == 0 or hi < tout) else tout
        if not (tin < tout) or not (tout >= 0) or not (tin <= T_FAR):
            continue
        if tin >= 0:
            hits.append((float(tin + F(0.0)), inst))
        if tout <= T_FAR:
            hits.append((float(tout + F(0.0)), inst))
    hits.sort()
    return hits[:MAX_HITS]


def mean_distance(mu, hw):
    """instancer.cpp:746-748 (std::pow promotes to double)."""
    mu, hw = float(mu), float(hw)
    return F(mu + 2 * mu * hw ** 2 / (3 * mu ** 2 + hw ** 2))


def get_model_input(rays_o, rays_d, parameters, n_pts, step_size, transformations, b_0, b_1, *, seed=0,
                    light_dir_idx=-1, method="random", use_mean_distance=False, light_strength_idx=-1, patch_scale=1.0):
    """Returns the CSR form over ALL rays: n_samples[R], hit[R] and per-sample records."""
    rays_o, rays_d, parameters = np.asarray(rays_o, F), np.asarray(rays_d, F), np.asarray(parameters, F)
    minv, dirt, origins = prepare_instances(transformations)
    b0, b1 = np.asarray(b_0, F), np.asarray(b_1, F)
    step_size = F(step_size)
    R = rays_o.shape[0]
    n_samples = np.zeros(R, np.int32)
    hit = np.zeros(R, bool)
    recs = dict(pts=[], dirs=[], params=[], dists=[], t=[], weight=[], inst=[])
    for i in range(R):
        o, d = rays_o[i], rays_d[i]
        hits = ray_hits(o, d, minv, b0, b1)
        if not hits:
            continue
        hit[i] = True
        # :801-827 total in-shell length
        active, total, t_entry = set(), F(0), F(0)
        for tt, inst in hits:
            tt = F(tt)
            if inst in active:
                active.discard(inst)
                if not active:
                    total = F(total + F(tt - t_entry))
            else:
                if not active:
                    t_entry = tt
                active.add(inst)
        if not (total > 0):
            continue
        necessary = int(F(total / step_size))                       # :845
        n_steps = min(necessary, n_pts)
        u = ray_uniform(seed, i)
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
        for tt, inst in hits:
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
                    sel, w = act[pick(seed, i, step, len(act))], F(len(act))
                elif method == "nearest_blend":                     # :694-713, transition_range 0.2 (instancer.hpp:84)
                    ds = [_norm3([F(pt[c] - origins[a_][c]) for c in range(3)]) for a_ in act]
                    lim = F(F(F(0.2) * F(patch_scale)) + min(ds))
                    ws = [max(F(lim - d_), F(0)) for d_ in ds]
                    total = F(0)
                    for w_ in ws:
                        total = F(total + w_)
                    target = F(step_uniform(seed, i, step) * total)
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
                if light is not None:
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
            if inst in active:
                active.discard(inst)
                if not active:
                    cleared = F(cleared + F(tt - t_entry))
            else:
                if not active:
                    segment_offset = F(tt - cleared)
                    t_entry = tt
                active.add(inst)
        n_samples[i] = emitted
    p = parameters.shape[1]
    out = {
        "n_samples": n_samples, "hit": hit,
        "pts": np.asarray(recs["pts"], F).reshape(-1, 3), "dirs": np.asarray(recs["dirs"], F).reshape(-1, 3),
        "params": np.asarray(recs["params"], F).reshape(len(recs["params"]), p), "dists": np.asarray(recs["dists"], F),
        "t": np.asarray(recs["t"], F), "alpha_weight": np.asarray(recs["weight"], F),
        "instance_id": np.asarray(recs["inst"], np.int32),
    }
    return out

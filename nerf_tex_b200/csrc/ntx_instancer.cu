// Shell instancer on the device (row f1 of the scope table): what C_Instancer::GetModelInput
// (instancer/src/instancer.cpp:751-1037) does on the host with Embree, for the subset that needs no mesh —
// explicit patch instances, no parameter textures, no shadow rays:
//   pass 1 (ntx_instancer_trace): one thread per ray tests every instanced patch box — a conservative bounding-sphere
//          test on tiles staged through shared memory, then for the survivors the slab test in the patch frame —
//          keeps the sorted hit list, walks the in-shell intervals and counts the marching samples;
//   pass 2 (ntx_instancer_march): after a prefix sum over the counts, one thread per SAMPLE (so every store is
//          coalesced) finds its ray and hit interval by binary search, replays the hits before it to get the patches it
//          is inside, and emits the record CSR-packed: patch-local point / view direction / parameters (light
//          direction mapped into the patch frame), step length, ray parameter, overlap weight, instance id.
// Arithmetic is fp32 with explicit round-to-nearest ops in the order of oracle/instancer_oracle.py (bit-exact);
// random draws come from a counter hash of (seed, ray[, step]) instead of the reference's std::mt19937.
#include <math_constants.h>

#include "ntx_common.cuh"

namespace {

constexpr int kMaxHits = 200;        // MAX_TOTAL_HITS, instancer.cpp:22
constexpr int kFirstOpenStride = 208; // kMaxHits bytes per ray, rounded up to 16
constexpr int kInstTile = 128;
#ifndef NTX_TRACE_THREADS
#define NTX_TRACE_THREADS 128
#endif
constexpr int kTraceThreads = NTX_TRACE_THREADS;   // rays per block of the trace pass
constexpr int kSmemHits = 56;        // hit-list entries per ray kept in shared memory; the rest lives in the ray's workspace row
constexpr int kMailbox = 16;         // per-ray direct-mapped cache of instances already tested (grid walk)
constexpr float kTFar = 100.0f;

__host__ __device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
    x = x * 747796405u + 2891336453u;
    uint32_t w = ((x >> ((x >> 28) + 4u)) ^ x) * 277803737u;
    return (w >> 22) ^ w;
}
__device__ __forceinline__ float ray_uniform(uint32_t seed, uint32_t ray) {
    return (float)(hash_u32(hash_u32(seed) ^ (ray * 2u + 1u)) >> 8) * (1.0f / 16777216.0f);
}
__device__ __forceinline__ float step_uniform(uint32_t seed, uint32_t ray, uint32_t step) {
    return (float)(hash_u32(hash_u32(seed ^ 0x85EBCA6Bu) ^ (ray * 1315423911u + step * 2u + 3u)) >> 8) * (1.0f / 16777216.0f);
}
__device__ __forceinline__ uint32_t pick(uint32_t seed, uint32_t ray, uint32_t step, uint32_t n) {
    return hash_u32(hash_u32(seed ^ 0x9E3779B9u) ^ (ray * 1315423911u + step * 2u + 2u)) % n;
}
__device__ __forceinline__ float dot3(const float* m, const float* v) {
    return __fadd_rn(__fadd_rn(__fmul_rn(m[0], v[0]), __fmul_rn(m[1], v[1])), __fmul_rn(m[2], v[2]));
}
__device__ __forceinline__ float norm3(const float* v) {
    return sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(v[0], v[0]), __fmul_rn(v[1], v[1])), __fmul_rn(v[2], v[2])));
}
__device__ __forceinline__ float mean_distance(float mu_f, float hw_f) {        // instancer.cpp:746-748
    const double mu = mu_f, hw = hw_f;
    const double num = __dmul_rn(__dmul_rn(2.0, mu), __dmul_rn(hw, hw));
    const double den = __dadd_rn(__dmul_rn(3.0, __dmul_rn(mu, mu)), __dmul_rn(hw, hw));
    return (float)__dadd_rn(mu, __ddiv_rn(num, den));
}

struct MarchOut {
    float* pts; float* dirs; float* params; float* dists; float* t; float* weight; int32_t* inst;
};

// What pass 1 leaves per ray for pass 2 (besides the sorted hit list): for every hit j the number of samples emitted
// before the walk consumed it and the segment offset in force after it, plus the per-ray constants.
struct RayInfo {
    float t_offset, last_dist;
    int32_t n_steps;         // negative: the whole crossing is shorter than one step (one sample, :848-851)
    float total;             // total in-shell length of the ray (shadow / texture sample counts, :863-864)
};

struct Workspace {
    uint64_t* hits;      // [R][kMaxHits]
    int32_t* step_at;    // [R][kMaxHits]
    float* segoff;       // [R][kMaxHits]
    RayInfo* info;       // [R]
    int32_t* n_hits;     // [R]
    uint8_t* first_open; // [R][kFirstOpenStride]: oldest hit still open behind hit j (j + 1 if none): where pass 2's scans start
    float2* seg;         // [R][kMaxHits] (only with shadow rays): (entry t, length) of the segment open after hit j
};
__host__ __device__ inline Workspace carve_workspace(void* base, int64_t R) {
    uint8_t* p = (uint8_t*)base;
    Workspace w;
    w.hits = (uint64_t*)p;   p += R * kMaxHits * 8;
    w.step_at = (int32_t*)p; p += R * kMaxHits * 4;
    w.segoff = (float*)p;    p += R * kMaxHits * 4;
    w.info = (RayInfo*)p;    p += R * 16;
    w.n_hits = (int32_t*)p;  p += ((R * 4 + 15) / 16) * 16;
    w.first_open = p;        p += R * kFirstOpenStride;
    w.seg = (float2*)p;
    return w;
}

// A hit is a 64-bit key: fp32 bits of t in the high word (t >= 0, so integer order is t order), then the instance id
// (24 bits) — and, once the list is final, in the top byte of the low word the index of the instance's OTHER hit on this
// ray (kNoPartner if it has none): a box is convex, so an instance contributes at most an entry and an exit.
constexpr uint32_t kInstMask = 0x00FFFFFFu;
constexpr uint32_t kNoPartner = 0xFFu;
__device__ __forceinline__ uint32_t hit_inst(uint64_t h) { return (uint32_t)h & kInstMask; }
__device__ __forceinline__ uint32_t hit_partner(uint64_t h) { return ((uint32_t)h) >> 24; }
__device__ __forceinline__ float hit_t(uint64_t h) { return __uint_as_float((uint32_t)(h >> 32)); }

// The instances a ray is currently inside (std::set<unsigned> active_instances, instancer.cpp:801), each with the
// index of the hit that put it there.  Pass 1 only needs membership and the size, so the list is unordered
// (append / swap-remove); the ascending-id order the reference's set has matters for the choice among overlapping
// patches only, which pass 2 makes.
struct OpenList {
    uint32_t id[kMaxHits];
    uint8_t from[kMaxHits];
    int n = 0;
    // returns the index of the hit that had inserted `inst` (now removed), or -1 if it was not a member (now inserted)
    __device__ int toggle(uint32_t inst, int hit_index) {
        for (int k = 0; k < n; ++k) {
            if (id[k] == inst) {
                const int f = from[k];
                --n;
                id[k] = id[n];
                from[k] = from[n];
                return f;
            }
        }
        id[n] = inst;
        from[n++] = (uint8_t)hit_index;
        return -1;
    }
};

// Pass 1, per ray: total in-shell length, step count, then the walk along the concatenated segments
// (instancer.cpp:801-1003) — counting only; what each sample needs later is recorded per hit.
// `hits(j)` is a reference to the j-th key of the ray in ascending order; the first loop also pairs entries with exits.
// `t_mesh` < +inf: the ray ends on the base mesh there (the list holds no box hit behind it): the mesh closes an open
// segment (:805-810) and samples go on behind the last box hit up to the mesh (:983).
template <class Hits>
__device__ int walk_ray(Hits& hits, int n_hits, int n_pts, float step_size, uint32_t seed, uint32_t ray, const float* u_given, int use_mean,
                        float t_mesh, int32_t* __restrict__ step_at, float* __restrict__ segoff,
                        RayInfo* __restrict__ info, float2* __restrict__ seg, uint8_t* __restrict__ first_open) {
    OpenList act;
    // :801-827 total in-shell length
    float total = 0.0f, t_entry = 0.0f;
    int seg_begin = 0;                                   // first hit of the segment being walked
    auto close_segment = [&](int end, float len) {       // hits [seg_begin, end) see this segment behind them
        if (seg)
            for (int q = seg_begin; q < end; ++q) seg[q] = make_float2(t_entry, len);
    };
    for (int j = 0; j < n_hits; ++j) {
        const uint64_t hj = hits(j);
        const float tt = hit_t(hj);
        const bool was_empty = act.n == 0;
        const int f = act.toggle(hit_inst(hj), j);
        if (f >= 0) {
            hits(j) = hj | ((uint64_t)f << 24);
            hits(f) = (hits(f) & ~((uint64_t)kNoPartner << 24)) | ((uint64_t)j << 24);
            if (act.n == 0) {
                const float len = __fsub_rn(tt, t_entry);
                total = __fadd_rn(total, len);
                close_segment(j, len);
            }
        } else {
            hits(j) = hj | ((uint64_t)kNoPartner << 24);
            if (was_empty) { t_entry = tt; seg_begin = j; }
        }
        int fo = j + 1;                                   // a sample behind hit j is inside no patch entered before hit fo
        for (int k = 0; k < act.n; ++k) fo = min(fo, (int)act.from[k]);
        first_open[j] = (uint8_t)fo;
    }
    if (t_mesh < CUDART_INF_F && act.n > 0) {
        const float len = __fsub_rn(t_mesh, t_entry);
        total = __fadd_rn(total, len);
        close_segment(n_hits, len);
    }
    RayInfo ri = {0.f, 0.f, 0, 0.f};
    ri.total = total;
    if (!(total > 0.0f)) { *info = ri; return 0; }
    const uint32_t necessary = (uint32_t)__fdiv_rn(total, step_size);             // :845
    int n_steps = (int)min(necessary, (uint32_t)n_pts);
    const float u = u_given ? *u_given : ray_uniform(seed, ray);
    if (n_steps == 0) {                                                           // :848-851
        ri.last_dist = total; ri.t_offset = __fmul_rn(u, total); n_steps = 1;
        ri.n_steps = -1;
    } else {                                                                      // :852-861
        ri.last_dist = __fsub_rn(__fadd_rn(step_size, total), __fmul_rn((float)n_steps, step_size));
        ri.t_offset = __fmul_rn(u, step_size);
    }
    if (ri.n_steps == 0) ri.n_steps = n_steps;
    *info = ri;
    // One loop whose iteration either emits (here: counts) the next sample or consumes the next hit — the
    // reference nests a sample loop inside the hit loop (:871-1003); same order of events.
    float segment_offset = 0.0f, cleared = 0.0f;
    t_entry = 0.0f;
    int step = 0, j = 0, n_act = 0;
    uint64_t hj = n_hits > 0 ? hits(0) : 0;
    while (j < n_hits && step < n_steps) {
        const float tt = hit_t(hj);
        const float t_mu = __fadd_rn(__fadd_rn(__fmul_rn((float)step, step_size), ri.t_offset), segment_offset);
        const float t_pt = use_mean ? mean_distance(t_mu, step_size) : t_mu;
        if (n_act > 0 && t_pt < tt) {
            ++step;
        } else {
            // a hit whose partner came earlier closes that instance, any other hit opens one (the pairing above)
            const uint32_t partner = hit_partner(hj);
            if (partner != kNoPartner && (int)partner < j) {
                if (--n_act == 0) cleared = __fadd_rn(cleared, __fsub_rn(tt, t_entry));
            } else if (n_act++ == 0) {
                segment_offset = __fsub_rn(tt, cleared);
                t_entry = tt;
            }
            step_at[j] = step;
            segoff[j] = segment_offset;
            if (++j < n_hits) hj = hits(j);
        }
    }
    for (; j < n_hits; ++j) step_at[j] = 0x7fffffff;     // never consumed: no sample lies behind these hits
    if (t_mesh < CUDART_INF_F && n_act > 0) {            // behind the last box hit, in front of the mesh
        while (step < n_steps) {
            const float t_mu = __fadd_rn(__fadd_rn(__fmul_rn((float)step, step_size), ri.t_offset), segment_offset);
            const float t_pt = use_mean ? mean_distance(t_mu, step_size) : t_mu;
            if (!(t_pt < t_mesh)) break;
            ++step;
        }
    }
    return step;
}

// ---- uniform grid walk (Amanatides & Woo): calls visit(cell, t_exit) for every cell the ray crosses within
// [tmin, tmax], front to back; visit returns false to stop.  The grid's item boxes were inflated when they were binned,
// so the rounding of the walk (which may clip a corner cell by a few ulps) cannot lose an item the ray really hits.
struct Grid {
    const int32_t* __restrict__ start;
    const int32_t* __restrict__ items;
    float lo[3], cell[3];
    int res[3];
};
__host__ inline Grid make_grid(const ntx_scene_grid& g) {
    Grid r;
    r.start = g.cell_start; r.items = g.items;
    for (int a = 0; a < 3; ++a) { r.lo[a] = g.lo[a]; r.cell[a] = g.cell[a]; r.res[a] = g.res[a]; }
    return r;
}
template <class Visit>
__device__ __forceinline__ void grid_walk(const Grid& g, const float o[3], const float d[3], float tmin, float tmax,
                                          Visit&& visit) {
    float t0 = tmin, t1 = tmax;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float hi = g.lo[a] + (float)g.res[a] * g.cell[a];
        if (d[a] != 0.0f) {
            const float inv = 1.0f / d[a];
            float ta = (g.lo[a] - o[a]) * inv, tb = (hi - o[a]) * inv;
            if (ta > tb) { const float s = ta; ta = tb; tb = s; }
            t0 = fmaxf(t0, ta);
            t1 = fminf(t1, tb);
        } else if (o[a] < g.lo[a] || o[a] > hi) {
            return;
        }
    }
    if (!(t0 <= t1)) return;
    int c[3], stp[3];
    float tnext[3], tdelta[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float p = o[a] + d[a] * t0;
        int ci = (int)floorf((p - g.lo[a]) / g.cell[a]);
        ci = max(0, min(g.res[a] - 1, ci));
        c[a] = ci;
        if (d[a] > 0.0f) {
            stp[a] = 1; tnext[a] = (g.lo[a] + (float)(ci + 1) * g.cell[a] - o[a]) / d[a]; tdelta[a] = g.cell[a] / d[a];
        } else if (d[a] < 0.0f) {
            stp[a] = -1; tnext[a] = (g.lo[a] + (float)ci * g.cell[a] - o[a]) / d[a]; tdelta[a] = -g.cell[a] / d[a];
        } else {
            stp[a] = 0; tnext[a] = CUDART_INF_F; tdelta[a] = CUDART_INF_F;
        }
    }
    for (;;) {
        const int a = (tnext[0] <= tnext[1]) ? (tnext[0] <= tnext[2] ? 0 : 2) : (tnext[1] <= tnext[2] ? 1 : 2);
        const float tn = a == 0 ? tnext[0] : (a == 1 ? tnext[1] : tnext[2]);
        if (!visit((c[2] * g.res[1] + c[1]) * g.res[0] + c[0], fminf(tn, t1))) return;
        if (!(tn <= t1)) return;
        if (a == 0) { c[0] += stp[0]; if ((unsigned)c[0] >= (unsigned)g.res[0]) return; tnext[0] += tdelta[0]; }
        else if (a == 1) { c[1] += stp[1]; if ((unsigned)c[1] >= (unsigned)g.res[1]) return; tnext[1] += tdelta[1]; }
        else { c[2] += stp[2]; if ((unsigned)c[2] >= (unsigned)g.res[2]) return; tnext[2] += tdelta[2]; }
    }
}

// ---- watertight ray / triangle test (Woop, Benthin, Wald, JCGT 2013) in the operation order of
// oracle/instancer_oracle.py::mesh_first_hit.  RayFrame holds the per-ray constants.
struct RayFrame { int kx, ky, kz; float Sx, Sy, Sz; };
__device__ __forceinline__ float sel3(const float v[3], int k) { return k == 0 ? v[0] : (k == 1 ? v[1] : v[2]); }
__device__ __forceinline__ RayFrame ray_frame(const float d[3]) {
    RayFrame f;
    const float ax = fabsf(d[0]), ay = fabsf(d[1]), az = fabsf(d[2]);
    f.kz = (ax >= ay && ax >= az) ? 0 : (ay >= az ? 1 : 2);
    f.kx = (f.kz + 1) % 3;
    f.ky = (f.kx + 1) % 3;
    const float dz = sel3(d, f.kz);
    if (dz < 0.0f) { const int s = f.kx; f.kx = f.ky; f.ky = s; }
    f.Sx = __fdiv_rn(sel3(d, f.kx), dz);
    f.Sy = __fdiv_rn(sel3(d, f.ky), dz);
    f.Sz = __fdiv_rn(1.0f, dz);
    return f;
}
// true + (t, u, v) if the ray hits the triangle with t in [0, 100]
__device__ __forceinline__ bool tri_hit(const RayFrame& f, const float o[3], const float* __restrict__ V,
                                        const int32_t* __restrict__ face, float& t, float& u, float& v) {
    float P[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float* vp = V + 3 * (size_t)__ldg(face + k);
#pragma unroll
        for (int a = 0; a < 3; ++a) P[k][a] = __fsub_rn(__ldg(vp + a), o[a]);
    }
    float X[3], Y[3], Zk[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        Zk[k] = sel3(P[k], f.kz);
        X[k] = __fsub_rn(sel3(P[k], f.kx), __fmul_rn(f.Sx, Zk[k]));
        Y[k] = __fsub_rn(sel3(P[k], f.ky), __fmul_rn(f.Sy, Zk[k]));
    }
    float U = __fsub_rn(__fmul_rn(X[2], Y[1]), __fmul_rn(Y[2], X[1]));
    float Vv = __fsub_rn(__fmul_rn(X[0], Y[2]), __fmul_rn(Y[0], X[2]));
    float W = __fsub_rn(__fmul_rn(X[1], Y[0]), __fmul_rn(Y[1], X[0]));
    if (U == 0.0f || Vv == 0.0f || W == 0.0f) {
        U = __double2float_rn(__dsub_rn(__dmul_rn((double)X[2], (double)Y[1]), __dmul_rn((double)Y[2], (double)X[1])));
        Vv = __double2float_rn(__dsub_rn(__dmul_rn((double)X[0], (double)Y[2]), __dmul_rn((double)Y[0], (double)X[2])));
        W = __double2float_rn(__dsub_rn(__dmul_rn((double)X[1], (double)Y[0]), __dmul_rn((double)Y[1], (double)X[0])));
    }
    if ((U < 0.0f || Vv < 0.0f || W < 0.0f) && (U > 0.0f || Vv > 0.0f || W > 0.0f)) return false;
    const float det = __fadd_rn(__fadd_rn(U, Vv), W);
    if (det == 0.0f) return false;
    const float T = __fadd_rn(__fadd_rn(__fmul_rn(U, __fmul_rn(f.Sz, Zk[0])), __fmul_rn(Vv, __fmul_rn(f.Sz, Zk[1]))),
                              __fmul_rn(W, __fmul_rn(f.Sz, Zk[2])));
    t = __fdiv_rn(T, det);
    if (!(t >= 0.0f && t <= kTFar)) return false;
    u = __fdiv_rn(Vv, det);
    v = __fdiv_rn(W, det);
    return true;
}

struct SceneDev {
    const float* __restrict__ minv;
    const float* __restrict__ dirt;
    const float* __restrict__ origins;
    const float4* __restrict__ spheres;
    int n_inst, n_faces;
    const float* __restrict__ mesh_v;
    const int32_t* __restrict__ mesh_f;
    Grid ig, tg;
    ntx_instancer_shading sh;          // pool == NULL and n_aux == 0: nothing to shade
};

// ---- parameter textures (instancer.cpp:606-667) in the operation order of oracle/instancer_oracle.py
__device__ __forceinline__ float tex_lookup(const float* __restrict__ pool, const ntx_texture& tx, float u, float v) {
    const int rows = tx.rows, cols = tx.cols;
    const float x0 = __fmul_rn(u, (float)(rows - 1)), x1 = __fmul_rn(v, (float)(cols - 1));
    const float w0 = __fsub_rn(x0, floorf(x0)), w1 = __fsub_rn(x1, floorf(x1));
    const int i0 = min(max((int)x0, 0), rows - 1), j0 = min(max((int)x1, 0), cols - 1);
    const int i1 = min(i0 + 1, rows - 1), j1 = min(j0 + 1, cols - 1);
    const float m0 = __fsub_rn(1.0f, w0), m1 = __fsub_rn(1.0f, w1);
    const float* t = pool + tx.offset;
    float y = __fmul_rn(__fmul_rn(__ldg(t + (size_t)i0 * cols + j0), m0), m1);
    y = __fadd_rn(y, __fmul_rn(__fmul_rn(__ldg(t + (size_t)i0 * cols + j1), m0), w1));
    y = __fadd_rn(y, __fmul_rn(__fmul_rn(__ldg(t + (size_t)i1 * cols + j0), w0), m1));
    return __fadd_rn(y, __fmul_rn(__fmul_rn(__ldg(t + (size_t)i1 * cols + j1), w0), w1));
}

// closest point of triangle `face` to p (Ericson; closest_point_triangle, instancer.cpp:151-199): barycentric weights
// and the distance, every op rounded as in oracle closest_face
__device__ __forceinline__ float closest_on_face(const float* __restrict__ V, const int32_t* __restrict__ fp, const float p[3],
                                                 float uvw[3]) {
    float a[3], b[3], c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        a[k] = __ldg(V + 3 * (size_t)__ldg(fp) + k);
        b[k] = __ldg(V + 3 * (size_t)__ldg(fp + 1) + k);
        c[k] = __ldg(V + 3 * (size_t)__ldg(fp + 2) + k);
    }
    float ab[3], ac[3], ap[3], bp[3], cp[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        ab[k] = __fsub_rn(b[k], a[k]); ac[k] = __fsub_rn(c[k], a[k]);
        ap[k] = __fsub_rn(p[k], a[k]); bp[k] = __fsub_rn(p[k], b[k]); cp[k] = __fsub_rn(p[k], c[k]);
    }
    const float d1 = dot3(ab, ap), d2 = dot3(ac, ap), d3 = dot3(ab, bp), d4 = dot3(ac, bp), d5 = dot3(ab, cp), d6 = dot3(ac, cp);
    const float vc = __fsub_rn(__fmul_rn(d1, d4), __fmul_rn(d3, d2));
    const float vb = __fsub_rn(__fmul_rn(d5, d2), __fmul_rn(d1, d6));
    const float va = __fsub_rn(__fmul_rn(d3, d6), __fmul_rn(d5, d4));
    if (d1 <= 0.0f && d2 <= 0.0f) { uvw[0] = 1.0f; uvw[1] = 0.0f; uvw[2] = 0.0f; }
    else if (d3 >= 0.0f && d4 <= d3) { uvw[0] = 0.0f; uvw[1] = 1.0f; uvw[2] = 0.0f; }
    else if (d6 >= 0.0f && d5 <= d6) { uvw[0] = 0.0f; uvw[1] = 0.0f; uvw[2] = 1.0f; }
    else if (vc <= 0.0f && d1 >= 0.0f && d3 <= 0.0f) {
        const float v = __fdiv_rn(d1, __fsub_rn(d1, d3));
        uvw[0] = __fsub_rn(1.0f, v); uvw[1] = v; uvw[2] = 0.0f;
    } else if (vb <= 0.0f && d2 >= 0.0f && d6 <= 0.0f) {
        const float v = __fdiv_rn(d2, __fsub_rn(d2, d6));
        uvw[0] = __fsub_rn(1.0f, v); uvw[1] = 0.0f; uvw[2] = v;
    } else {
        const float e43 = __fsub_rn(d4, d3), e56 = __fsub_rn(d5, d6);
        if (va <= 0.0f && e43 >= 0.0f && e56 >= 0.0f) {
            const float v = __fdiv_rn(e43, __fadd_rn(e43, e56));
            uvw[0] = 0.0f; uvw[1] = __fsub_rn(1.0f, v); uvw[2] = v;
        } else {
            const float denom = __fdiv_rn(1.0f, __fadd_rn(__fadd_rn(va, vb), vc));
            const float v = __fmul_rn(vb, denom), w = __fmul_rn(vc, denom);
            uvw[0] = __fsub_rn(__fsub_rn(1.0f, v), w); uvw[1] = v; uvw[2] = w;
        }
    }
    float df[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float q = __fadd_rn(__fadd_rn(__fmul_rn(a[k], uvw[0]), __fmul_rn(b[k], uvw[1])), __fmul_rn(c[k], uvw[2]));
        df[k] = __fsub_rn(p[k], q);
    }
    return sqrtf(dot3(df, df));
}

// the base-mesh face nearest to p within `radius` (strictly), ties to the smaller face index; -1 if none
__device__ int closest_face(const SceneDev& sc, const float p[3], float radius, float uvw[3]) {
    int best = -1;
    float best_d = radius;
    const int n_base = sc.sh.n_base_faces;
    auto test = [&](int face) {
        if (face >= n_base) return;
        float w[3];
        const float d = closest_on_face(sc.mesh_v, sc.mesh_f + 3 * (size_t)face, p, w);
        if (d < best_d || (d == best_d && best >= 0 && face < best)) {
            best_d = d; best = face;
            uvw[0] = w[0]; uvw[1] = w[1]; uvw[2] = w[2];
        }
    };
    if (!sc.tg.start) {
        for (int face = 0; face < n_base; ++face) test(face);
        return best;
    }
    const Grid& g = sc.tg;
    int c0[3], c1[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        c0[a] = max(0, min(g.res[a] - 1, (int)floorf((p[a] - radius - g.lo[a]) / g.cell[a])));
        c1[a] = max(0, min(g.res[a] - 1, (int)floorf((p[a] + radius - g.lo[a]) / g.cell[a])));
    }
    // the cell of p first: its faces usually shrink the search radius to a fraction of the patch extent
    int pc[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) pc[a] = max(c0[a], min(c1[a], (int)floorf((p[a] - g.lo[a]) / g.cell[a])));
    auto visit = [&](int x, int y, int z) {
        // distance from p to the cell's box: a face listed here has a point in the cell, so it cannot be nearer than that
        float dd = 0.0f;
        const int cc[3] = {x, y, z};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float lo = g.lo[a] + (float)cc[a] * g.cell[a], hi = lo + g.cell[a];
            const float e = fmaxf(fmaxf(lo - p[a], p[a] - hi), 0.0f);
            dd += e * e;
        }
        if (dd > best_d * best_d * 1.0001f + 1e-12f) return;
        const int cell = (z * g.res[1] + y) * g.res[0] + x;
        const int e = __ldg(g.start + cell + 1);
        for (int q = __ldg(g.start + cell); q < e; ++q) test(__ldg(g.items + q));
    };
    visit(pc[0], pc[1], pc[2]);
    for (int z = c0[2]; z <= c1[2]; ++z)
        for (int y = c0[1]; y <= c1[1]; ++y)
            for (int x = c0[0]; x <= c1[0]; ++x)
                if (x != pc[0] || y != pc[1] || z != pc[2]) visit(x, y, z);
    return best;
}

// getParameters (:640-667): out[] = the ray's parameters, textured ones multiplied by their texture at the UV of the
// nearest base-mesh point
__device__ void get_parameters(const SceneDev& sc, const float p[3], const float* __restrict__ prm, int n_param, float* out) {
    for (int q = 0; q < n_param; ++q) out[q] = __ldg(prm + q);
    float w[3];
    const int face = closest_face(sc, p, sc.sh.radius, w);
    if (face < 0) return;
    const int32_t* fp = sc.mesh_f + 3 * (size_t)face;
    float uv[2];
#pragma unroll
    for (int k = 0; k < 2; ++k)
        uv[k] = __fadd_rn(__fadd_rn(__fmul_rn(__ldg(sc.sh.mesh_uv + 2 * (size_t)__ldg(fp) + k), w[0]),
                                    __fmul_rn(__ldg(sc.sh.mesh_uv + 2 * (size_t)__ldg(fp + 1) + k), w[1])),
                          __fmul_rn(__ldg(sc.sh.mesh_uv + 2 * (size_t)__ldg(fp + 2) + k), w[2]));
    for (int i = 0; i < sc.sh.n_param_tex; ++i) {
        const int idx = sc.sh.param_idx[i];
        out[idx] = __fmul_rn(out[idx], tex_lookup(sc.sh.pool, sc.sh.param_tex[i], uv[0], uv[1]));
    }
}

// isShadowed (instancer.cpp:593-603) with shadow_ray_filter_function (:539-551), in the operation order of
// oracle/instancer_oracle.py::is_shadowed: a ray from `pt` along `dir` (not normalised), t in [0, 100], is blocked by the
// top face of a patch box entered from above (patch frame), by the bottom face of a patch box from either side, or by a
// mesh triangle hit on its front side (or, like the reference, by the second triangle of any mesh from either side).
// Any-hit: the walk stops at the first blocker.
struct ShadowBox { float b0[3], b1[3]; };
__device__ bool shadow_blocked(const SceneDev& sc, const ShadowBox& bx, const float pt[3], const float dir[3]) {
    bool blocked = false;
    const float inv_dn = rsqrtf(fmaxf(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2], 1e-30f));
    const float dh[3] = {dir[0] * inv_dn, dir[1] * inv_dn, dir[2] * inv_dn};
    auto test_instance = [&](int inst) {
        if (sc.spheres) {
            const float4 sp = __ldg(sc.spheres + inst);
            const float cx = sp.x - pt[0], cy = sp.y - pt[1], cz = sp.z - pt[2];
            const float cc = cx * cx + cy * cy + cz * cz;
            const float along = cx * dh[0] + cy * dh[1] + cz * dh[2];
            if (fmaf(-along, along, cc) > fmaf(1e-5f, cc, sp.w)) return;
        }
        const float4* Mp = reinterpret_cast<const float4*>(sc.minv) + 3 * (size_t)inst;
        float ol[3], dl[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float4 m4 = __ldg(Mp + a);
            const float M[3] = {m4.x, m4.y, m4.z};
            ol[a] = __fadd_rn(dot3(M, pt), m4.w);
            dl[a] = dot3(M, dir);
        }
        const float inv = __fdiv_rn(1.0f, dl[2]);
#pragma unroll
        for (int face = 0; face < 2; ++face) {                                 // top (from above only), bottom
            if (face == 0 ? !(dl[2] < 0.0f) : !(dl[2] != 0.0f)) continue;
            const float t = __fmul_rn(__fsub_rn(face == 0 ? bx.b1[2] : bx.b0[2], ol[2]), inv);
            const float x = __fadd_rn(ol[0], __fmul_rn(t, dl[0])), y = __fadd_rn(ol[1], __fmul_rn(t, dl[1]));
            if (t >= 0.0f && t <= kTFar && x >= bx.b0[0] && x <= bx.b1[0] && y >= bx.b0[1] && y <= bx.b1[1]) blocked = true;
        }
    };
    if (sc.ig.start) {
        grid_walk(sc.ig, pt, dir, 0.0f, kTFar, [&](int cell, float) {
            const int e = __ldg(sc.ig.start + cell + 1);
            for (int q = __ldg(sc.ig.start + cell); q < e && !blocked; ++q) test_instance(__ldg(sc.ig.items + q));
            return !blocked;
        });
    } else {
        for (int inst = 0; inst < sc.n_inst && !blocked; ++inst) test_instance(inst);
    }
    if (blocked || sc.n_faces == 0 || (dir[0] == 0.0f && dir[1] == 0.0f && dir[2] == 0.0f)) return blocked;
    const RayFrame rf = ray_frame(dir);
    auto test_face = [&](int face) {
        float t, u, v;
        const int32_t* fp = sc.mesh_f + 3 * (size_t)face;
        if (!tri_hit(rf, pt, sc.mesh_v, fp, t, u, v)) return;
        float P[3][3];
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int a = 0; a < 3; ++a) P[k][a] = __ldg(sc.mesh_v + 3 * (size_t)__ldg(fp + k) + a);
        const float e1[3] = {__fsub_rn(P[1][0], P[0][0]), __fsub_rn(P[1][1], P[0][1]), __fsub_rn(P[1][2], P[0][2])};
        const float e2[3] = {__fsub_rn(P[2][0], P[0][0]), __fsub_rn(P[2][1], P[0][1]), __fsub_rn(P[2][2], P[0][2])};
        const float ng[3] = {__fsub_rn(__fmul_rn(e1[1], e2[2]), __fmul_rn(e1[2], e2[1])),
                             __fsub_rn(__fmul_rn(e1[2], e2[0]), __fmul_rn(e1[0], e2[2])),
                             __fsub_rn(__fmul_rn(e1[0], e2[1]), __fmul_rn(e1[1], e2[0]))};
        if (dot3(ng, dir) < 0.0f) blocked = true;
        // the reference's filter ends in `|| hit->primID == 1` without looking at the geometry (:550): the SECOND triangle of
        // the base mesh and of every auxiliary mesh blocks from either side (seen by running the reference's own code)
        if (face == 1 && sc.sh.n_base_faces >= 2) blocked = true;
        for (int m = 0; m < sc.sh.n_aux; ++m)
            if (face == sc.sh.aux_face_begin[m] + 1 && face < sc.sh.aux_face_begin[m + 1]) blocked = true;
    };
    if (sc.tg.start) {
        grid_walk(sc.tg, pt, dir, 0.0f, kTFar, [&](int cell, float) {
            const int e = __ldg(sc.tg.start + cell + 1);
            for (int q = __ldg(sc.tg.start + cell); q < e && !blocked; ++q) test_face(__ldg(sc.tg.items + q));
            return !blocked;
        });
    } else {
        for (int face = 0; face < sc.n_faces && !blocked; ++face) test_face(face);
    }
    return blocked;
}

struct ShadowArgs {
    int enabled, min_samples, n_samples, n_pts;
    ShadowBox box;
};

__global__ void __launch_bounds__(kTraceThreads)
instancer_trace_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d, const int32_t* __restrict__ ray_ids,
                       const float* __restrict__ ray_offsets, int R, SceneDev sc, float b00, float b01, float b02, float b10, float b11, float b12, float step_size,
                       int n_pts, uint32_t seed, int use_mean, Workspace ws, int32_t* __restrict__ n_samples,
                       int32_t* __restrict__ hit_flag, float* __restrict__ alpha_last, const float* __restrict__ params,
                       int n_param, int light_idx, float* __restrict__ color_last) {
    const int ray = blockIdx.x * blockDim.x + threadIdx.x;
    if (ray >= R) return;
    float o[3], d[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) { o[a] = rays_o[3 * ray + a]; d[a] = rays_d[3 * ray + a]; }
    const float b0[3] = {b00, b01, b02}, b1[3] = {b10, b11, b12};
    const float inv_dn = rsqrtf(fmaxf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2], 1e-30f));
    const float dh[3] = {d[0] * inv_dn, d[1] * inv_dn, d[2] * inv_dn};

    // ---- the base mesh first: its nearest hit bounds everything else (instancer.cpp:804-814)
    float t_mesh = CUDART_INF_F, u_mesh = 0.0f, v_mesh = 0.0f;
    int best = -1;
    if (sc.n_faces > 0 && (d[0] != 0.0f || d[1] != 0.0f || d[2] != 0.0f)) {
        const RayFrame rf = ray_frame(d);
        auto test = [&](int face) {
            float t, u, v;
            if (tri_hit(rf, o, sc.mesh_v, sc.mesh_f + 3 * (size_t)face, t, u, v) && (t < t_mesh || (t == t_mesh && face < best))) {
                t_mesh = t; u_mesh = u; v_mesh = v;
                best = face;
            }
        };
        if (sc.tg.start) {
            grid_walk(sc.tg, o, d, 0.0f, kTFar, [&](int cell, float t_exit) {
                const int e = __ldg(sc.tg.start + cell + 1);
                for (int q = __ldg(sc.tg.start + cell); q < e; ++q) test(__ldg(sc.tg.items + q));
                return !(t_mesh <= t_exit);            // a hit inside the cells walked so far cannot be beaten behind them
            });
        } else {
            for (int face = 0; face < sc.n_faces; ++face) test(face);
        }
    }
    const float t_limit = fminf(t_mesh, kTFar);        // box hits at t <= t_mesh sort before the mesh hit (:444-455)

    // the ray's sorted hit list: the first kSmemHits entries in shared memory (entry k of thread t at
    // s_hits[k * kTraceThreads + t], conflict-free), the tail in the ray's row of the global workspace.  (A per-thread
    // array would be local memory: 1.6 KB per ray falls out of L1, and holding all 200 entries in shared memory leaves
    // four warps per SM.)
    extern __shared__ __align__(16) uint64_t s_hits[];
    uint64_t* row = ws.hits + (size_t)ray * kMaxHits;
    auto H = [&](int k) -> uint64_t& { return k < kSmemHits ? s_hits[k * kTraceThreads + threadIdx.x] : row[k]; };
    int n_hits = 0;
    // keep the kMaxHits smallest (t, inst), sorted, each once (an instance listed in several grid cells is tested once
    // per cell the ray crosses; the same ray and box give the same t, so a repeat shows up as an equal key)
    auto insert = [&](float t, uint32_t inst) {
        const uint64_t key = ((uint64_t)__float_as_uint(t + 0.0f) << 32) | inst;
        if (n_hits < kSmemHits) {
            // the whole list is in shared memory (most rays): find the slot and open it in one backward pass (an exit
            // lands ~30 places before the end of the list).  Longer lists take the two-pass form below: fusing it as well
            // was measured slower (the shared / global split of every access costs more than the second pass).
            uint64_t* __restrict__ mine = s_hits + threadIdx.x;
            int q = n_hits;
            bool dup = false;
            while (q > 0) {
                const uint64_t h = mine[(q - 1) * kTraceThreads];
                if (h <= key) { dup = h == key; break; }
                mine[q * kTraceThreads] = h;
                --q;
            }
            if (dup) {                                   // met before in another cell (rare): close the gap again
                for (int r = q; r < n_hits; ++r) mine[r * kTraceThreads] = mine[(r + 1) * kTraceThreads];
                return;
            }
            mine[q * kTraceThreads] = key;
            ++n_hits;
            return;
        }
        int k = n_hits;
        while (k > 0 && H(k - 1) > key) --k;            // slot of the new key (hits arrive nearly in order: short scan)
        if (k > 0 && H(k - 1) == key) return;
        if (n_hits == kMaxHits) {
            if (k == kMaxHits) return;
        } else {
            ++n_hits;
        }
        for (int q = n_hits - 1; q > k; --q) H(q) = H(q - 1);
        H(k) = key;
    };
    // An instance is listed in every cell its box overlaps, so a ray meets it again in the next cells: a small
    // direct-mapped cache of ids already tested filters most repeats (the equal-key test above catches the rest).
    uint32_t* mail = reinterpret_cast<uint32_t*>(s_hits + (size_t)kSmemHits * kTraceThreads) + threadIdx.x;
    if (sc.ig.start)
        for (int k = 0; k < kMailbox; ++k) mail[k * kTraceThreads] = 0xFFFFFFFFu;
    auto test_instance = [&](int inst) {
        if (sc.ig.start) {
            uint32_t& slot = mail[((uint32_t)inst % kMailbox) * kTraceThreads];
            if (slot == (uint32_t)inst) return;
            slot = (uint32_t)inst;
        }
        if (sc.spheres) {
            // squared distance of the sphere centre from the ray's line against the (inflated) radius: rejects
            // only instances whose box the ray certainly misses, so the hit list is unchanged
            const float4 sp = __ldg(sc.spheres + inst);
            const float cx = sp.x - o[0], cy = sp.y - o[1], cz = sp.z - o[2];
            const float cc = cx * cx + cy * cy + cz * cz;
            const float along = cx * dh[0] + cy * dh[1] + cz * dh[2];          // dh: unit ray direction
            if (fmaf(-along, along, cc) > fmaf(1e-5f, cc, sp.w)) return;       // margin: rounding of cc - along^2
        }
        const float4* Mp = reinterpret_cast<const float4*>(sc.minv) + 3 * (size_t)inst;
        float tin = -CUDART_INF_F, tout = CUDART_INF_F;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float4 m4 = __ldg(Mp + a);
            const float M[3] = {m4.x, m4.y, m4.z};
            const float ol = __fadd_rn(dot3(M, o), m4.w);
            const float dl = dot3(M, d);
            const float inv = __fdiv_rn(1.0f, dl);
            const float t0 = __fmul_rn(__fsub_rn(b0[a], ol), inv), t1 = __fmul_rn(__fsub_rn(b1[a], ol), inv);
            const float lo = (t0 < t1) ? t0 : t1, hi = (t0 < t1) ? t1 : t0;
            tin = (a == 0 || lo > tin) ? lo : tin;
            tout = (a == 0 || hi < tout) ? hi : tout;
        }
        if (!(tin < tout) || !(tout >= 0.0f) || !(tin <= kTFar)) return;
        if (tin >= 0.0f && tin <= t_limit) insert(tin, (uint32_t)inst);
        if (tout <= t_limit) insert(tout, (uint32_t)inst);
    };
    if (sc.ig.start) {
        grid_walk(sc.ig, o, d, 0.0f, t_limit, [&](int cell, float) {
            const int e = __ldg(sc.ig.start + cell + 1);
            for (int q = __ldg(sc.ig.start + cell); q < e; ++q) test_instance(__ldg(sc.ig.items + q));
            return true;
        });
    } else {
        for (int inst = 0; inst < sc.n_inst; ++inst) test_instance(inst);
    }
    if (t_mesh < CUDART_INF_F && n_hits >= kMaxHits) t_mesh = CUDART_INF_F;   // the mesh hit would be entry 201 of the list
    ws.n_hits[ray] = n_hits;
    hit_flag[ray] = n_hits > 0 || t_mesh < CUDART_INF_F;
    alpha_last[ray] = t_mesh < CUDART_INF_F ? 1.0f : 0.0f;                    // instancer.cpp:1019-1027
    if (color_last) {
        float col[3] = {0.0f, 0.0f, 0.0f};                                    // base mesh: black (:1020)
        if (t_mesh < CUDART_INF_F && best >= sc.sh.n_base_faces && sc.sh.n_aux > 0) {     // auxiliary mesh: shadeMesh (:716-744)
            int m = 0;
            while (m + 1 < sc.sh.n_aux && best >= sc.sh.aux_face_begin[m + 1]) ++m;
            const int32_t* fp = sc.mesh_f + 3 * (size_t)best;
            const float w[3] = {__fsub_rn(__fsub_rn(1.0f, u_mesh), v_mesh), u_mesh, v_mesh};
            float n[3];
#pragma unroll
            for (int k = 0; k < 3; ++k)
                n[k] = __fadd_rn(__fadd_rn(__fmul_rn(__ldg(sc.sh.mesh_n + 3 * (size_t)__ldg(fp) + k), w[0]),
                                           __fmul_rn(__ldg(sc.sh.mesh_n + 3 * (size_t)__ldg(fp + 1) + k), w[1])),
                                 __fmul_rn(__ldg(sc.sh.mesh_n + 3 * (size_t)__ldg(fp + 2) + k), w[2]));
            const float nn = norm3(n);
#pragma unroll
            for (int k = 0; k < 3; ++k) n[k] = __fdiv_rn(n[k], nn);
            float albedo[3] = {0.8f, 0.8f, 0.8f};
            const int nch = sc.sh.aux_channels[m];
            if (nch > 0) {
                float uv[2];
#pragma unroll
                for (int k = 0; k < 2; ++k)
                    uv[k] = __fadd_rn(__fadd_rn(__fmul_rn(__ldg(sc.sh.mesh_uv + 2 * (size_t)__ldg(fp) + k), w[0]),
                                                __fmul_rn(__ldg(sc.sh.mesh_uv + 2 * (size_t)__ldg(fp + 1) + k), w[1])),
                                      __fmul_rn(__ldg(sc.sh.mesh_uv + 2 * (size_t)__ldg(fp + 2) + k), w[2]));
                for (int k = 0; k < 3; ++k) albedo[k] = tex_lookup(sc.sh.pool, sc.sh.aux_tex[m][nch == 3 ? k : 0], uv[0], uv[1]);
            }
            float diffuse = 0.0f;
            if (light_idx >= 0 && params) {
                const float* prm = params + (size_t)ray * n_param + light_idx;
                const float raw[3] = {__ldg(prm), __ldg(prm + 1), __ldg(prm + 2)};
                float lifted[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) lifted[k] = __fadd_rn(__fadd_rn(o[k], __fmul_rn(t_mesh, d[k])), __fmul_rn(n[k], 1e-6f));
                ShadowBox bx;
                for (int k = 0; k < 3; ++k) { bx.b0[k] = b0[k]; bx.b1[k] = b1[k]; }
                if (!shadow_blocked(sc, bx, lifted, raw)) {
                    const float ln = norm3(raw);
                    const float ld[3] = {__fdiv_rn(raw[0], ln), __fdiv_rn(raw[1], ln), __fdiv_rn(raw[2], ln)};
                    diffuse = __fmul_rn(1.0f, fmaxf(dot3(n, ld), 0.0f));
                }
            }
            const float shade = fminf(__fadd_rn(diffuse, 0.2f), 1.0f);
#pragma unroll
            for (int k = 0; k < 3; ++k) col[k] = __fmul_rn(albedo[k], shade);
        }
        color_last[3 * (size_t)ray] = col[0]; color_last[3 * (size_t)ray + 1] = col[1]; color_last[3 * (size_t)ray + 2] = col[2];
    }
    const uint32_t rid = ray_ids ? (uint32_t)ray_ids[ray] : (uint32_t)ray;
    n_samples[ray] = walk_ray(H, n_hits, n_pts, step_size, seed, rid, ray_offsets ? ray_offsets + ray : nullptr, use_mean, t_mesh,
                              ws.step_at + (size_t)ray * kMaxHits, ws.segoff + (size_t)ray * kMaxHits, ws.info + ray,
                              ws.seg ? ws.seg + (size_t)ray * kMaxHits : nullptr, ws.first_open + (size_t)ray * kFirstOpenStride);
    for (int k = 0; k < min(n_hits, kSmemHits); ++k) row[k] = H(k);     // with the partner indices walk_ray added
}

// Pass 2, one thread per sample: every store of a warp is contiguous.  The sample finds its ray (binary search in
// ray_ptr) and the hit interval it was emitted in (binary search in the ray's step_at row), replays the hits before it
// to get the set of patches it is inside, and writes its record.
// EXTRAS: shadow rays and parameter textures compiled in (their any-hit walks and parameter vectors cost registers:
// the plain variant runs at twice the occupancy)
template <bool EXTRAS>
__global__ void __launch_bounds__(256, EXTRAS ? 2 : 4)
instancer_emit_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d, const int32_t* __restrict__ ray_ids,
                      const float* __restrict__ params,
                      int n_param, int R, const SceneDev sc, float step_size, uint32_t seed, int light_idx, int strength_idx,
                      int method, float blend_range, int use_mean, Workspace ws, const int64_t* __restrict__ ray_ptr, int64_t T,
                      MarchOut out, const ShadowArgs sh) {
    const float* __restrict__ minv = sc.minv;
    const float* __restrict__ dirt = sc.dirt;
    const float* __restrict__ origins = sc.origins;
    const int64_t k0 = (int64_t)blockIdx.x * blockDim.x;
    const int64_t k = k0 + threadIdx.x;
    auto ray_of = [&](int64_t kk, int lo, int hi) {       // last ray in [lo, hi) with ray_ptr[ray] <= kk
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(ray_ptr + mid) <= kk) lo = mid; else hi = mid;
        }
        return lo;
    };
    // the rays of the block's first and last sample bound every thread's search
    __shared__ int s_win[2];
    if (threadIdx.x == 0) s_win[0] = ray_of(k0, 0, R);
    if (threadIdx.x == 32) s_win[1] = ray_of(min(k0 + (int64_t)blockDim.x, T) - 1, 0, R);
    __syncthreads();
    if (k >= T) return;
    const int ray = ray_of(k, s_win[0], s_win[1] + 1);
    const uint32_t rid = ray_ids ? (uint32_t)__ldg(ray_ids + ray) : (uint32_t)ray;     // the ray's name for the draws
    const int step = (int)(k - __ldg(ray_ptr + ray));
    const RayInfo ri = ws.info[ray];
    const int n_hits = ws.n_hits[ray];
    const uint64_t* hits = ws.hits + (size_t)ray * kMaxHits;
    const int32_t* step_at = ws.step_at + (size_t)ray * kMaxHits;
    int j = 0, jh = n_hits;                               // first hit consumed after this sample: step_at[j] > step
    while (j < jh) {
        const int mid = (j + jh) >> 1;
        if (__ldg(step_at + mid) <= step) j = mid + 1; else jh = mid;
    }
    const float segment_offset = j > 0 ? ws.segoff[(size_t)ray * kMaxHits + j - 1] : 0.0f;
    const int q0 = j > 0 ? (int)ws.first_open[(size_t)ray * kFirstOpenStride + j - 1] : 0;   // no open patch before hit q0
    float o[3], d[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) { o[a] = __ldg(rays_o + 3 * ray + a); d[a] = __ldg(rays_d + 3 * ray + a); }
    const float t_mu = __fadd_rn(__fadd_rn(__fmul_rn((float)step, step_size), ri.t_offset), segment_offset);
    const float t_pt = use_mean ? mean_distance(t_mu, step_size) : t_mu;
    float pt[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) pt[a] = __fadd_rn(o[a], __fmul_rn(t_pt, d[a]));
    // The patches this sample is inside: hit q < j whose partner (the instance's other hit) is not before j as well.
    auto inside = [&](uint64_t h) { const uint32_t p = hit_partner(h); return p == kNoPartner || (int)p >= j; };
    uint32_t sel = 0;
    float w = 1.0f;
    if (method == 0) {                                                        // random, :671-677
        int n_act = 0;
        for (int q = q0; q < j; ++q) {
            const uint64_t h = __ldg(hits + q);
            if (inside(h)) { ++n_act; sel = hit_inst(h); }
        }
        if (n_act > 1) {                                                      // the idx-th smallest id of the set
            const int idx = (int)pick(seed, rid, (uint32_t)step, (uint32_t)n_act);
            w = (float)n_act;
            for (int q = q0; q < j; ++q) {
                const uint64_t h = __ldg(hits + q);
                if (!inside(h)) continue;
                const uint32_t id = hit_inst(h);
                int rank = 0;
                for (int r = q0; r < j; ++r) {
                    const uint64_t g = __ldg(hits + r);
                    rank += inside(g) && hit_inst(g) < id;
                }
                if (rank == idx) { sel = id; break; }
            }
        }
    } else if (method == 2) {                                                 // nearest_blend, :694-713
        // weights max(range + min_dist - dist, 0) over the set in ascending id order, one draw from that discrete
        // distribution, density weight 1 / probability
        auto dist_to = [&](uint32_t id) {
            const float* og = origins + 3 * id;
            const float df[3] = {__fsub_rn(pt[0], __ldg(og)), __fsub_rn(pt[1], __ldg(og + 1)), __fsub_rn(pt[2], __ldg(og + 2))};
            return norm3(df);
        };
        auto next_id = [&](int64_t cur) {                                     // smallest member id above cur, or -1
            int64_t best = -1;
            for (int q = q0; q < j; ++q) {
                const uint64_t h = __ldg(hits + q);
                const int64_t id = hit_inst(h);
                if (inside(h) && id > cur && (best < 0 || id < best)) best = id;
            }
            return best;
        };
        float min_dist = CUDART_INF_F;
        for (int q = q0; q < j; ++q) {
            const uint64_t h = __ldg(hits + q);
            if (inside(h)) min_dist = fminf(min_dist, dist_to(hit_inst(h)));
        }
        const float lim = __fadd_rn(blend_range, min_dist);
        float total = 0.0f;
        for (int64_t id = next_id(-1); id >= 0; id = next_id(id))
            total = __fadd_rn(total, fmaxf(__fsub_rn(lim, dist_to((uint32_t)id)), 0.0f));
        const float target = __fmul_rn(step_uniform(seed, rid, (uint32_t)step), total);
        float run = 0.0f, w_sel = 1.0f;
        for (int64_t id = next_id(-1); id >= 0; id = next_id(id)) {
            const float wi = fmaxf(__fsub_rn(lim, dist_to((uint32_t)id)), 0.0f);
            run = __fadd_rn(run, wi);
            if (wi > 0.0f) {
                sel = (uint32_t)id;
                w_sel = wi;
                if (target < run) break;
            }
        }
        w = __fdiv_rn(total, w_sel);
    } else {                                                                  // nearest, :680-692
        // smallest rounded distance, ties to the smaller id (the reference walks the set in ascending id order and
        // keeps the first strict minimum); sqrt is monotone, so it is only taken for candidates that can win
        float best = CUDART_INF_F, best_sq = CUDART_INF_F;
        int n_act = 0;
        for (int q = q0; q < j; ++q) {
            const uint64_t h = __ldg(hits + q);
            if (!inside(h)) continue;
            const uint32_t id = hit_inst(h);
            if (n_act++ == 0) sel = id;
            const float* og = origins + 3 * id;
            const float d0 = __fsub_rn(pt[0], __ldg(og)), d1 = __fsub_rn(pt[1], __ldg(og + 1)), d2 = __fsub_rn(pt[2], __ldg(og + 2));
            const float sq = __fadd_rn(__fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1)), __fmul_rn(d2, d2));
            if (sq <= best_sq) {
                const float dist = sqrtf(sq);
                if (dist < best || (dist == best && id < sel)) { best = dist; best_sq = sq; sel = id; }
            }
        }
    }
    float M[12], Dt[9];
#pragma unroll
    for (int a = 0; a < 12; ++a) M[a] = __ldg(minv + 12 * sel + a);
#pragma unroll
    for (int a = 0; a < 9; ++a) Dt[a] = __ldg(dirt + 9 * sel + a);
    const float nn = norm3(d);
    const float dn[3] = {__fdiv_rn(d[0], nn), __fdiv_rn(d[1], nn), __fdiv_rn(d[2], nn)};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        out.pts[3 * k + a] = __fadd_rn(dot3(M + 4 * a, pt), M[4 * a + 3]);      // getPt :556-558
        out.dirs[3 * k + a] = dot3(Dt + 3 * a, dn);                           // getDir :561-563
    }
    const float* prm = params + (size_t)ray * n_param;
    if (EXTRAS && sc.sh.n_param_tex > 0 && sc.sh.n_base_faces > 0) {          // :864, :916-926, :985-991
        const uint32_t n_tex = max((uint32_t)sc.sh.min_samples, (uint32_t)__fmul_rn((float)sc.sh.n_samples, ri.total));
        float pv[16];
        if (n_tex < (uint32_t)sh.n_pts) {
            // textures are looked up at points spaced evenly along the segment and interpolated linearly
            const float2 sg = ws.seg[(size_t)ray * kMaxHits + (j > 0 ? j - 1 : 0)];
            const uint32_t n_seg = max((uint32_t)sc.sh.min_samples, (uint32_t)__fdiv_rn(__fmul_rn((float)n_tex, sg.y), ri.total));
            const float sstep = __fdiv_rn(sg.y, (float)(n_seg - 1u));
            auto t_at = [&](int kk) { return __fadd_rn(sg.x, __fmul_rn((float)kk, sstep)); };
            int kk = max(1, (int)ceilf((t_pt - sg.x) / sstep));
            if (!(kk < (1 << 24))) kk = 1;
            while (kk > 1 && !(t_pt > t_at(kk - 1))) --kk;
            for (int guard = 0; t_pt > t_at(kk) && guard < (1 << 20); ++guard) ++kk;
            const float t0 = kk == 1 ? sg.x : t_at(kk - 1), t1 = t_at(kk);
            const float wt = __fdiv_rn(__fsub_rn(t_pt, t0), sstep), mt = __fsub_rn(1.0f, wt);
            float q0[3], q1[3], p1[16];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                q0[a] = __fadd_rn(o[a], __fmul_rn(t0, d[a]));
                q1[a] = __fadd_rn(o[a], __fmul_rn(t1, d[a]));
            }
            get_parameters(sc, q0, prm, n_param, pv);
            get_parameters(sc, q1, prm, n_param, p1);
            for (int q = 0; q < n_param; ++q) pv[q] = __fadd_rn(__fmul_rn(pv[q], mt), __fmul_rn(p1[q], wt));
        } else {
            get_parameters(sc, pt, prm, n_param, pv);
        }
        for (int q = 0; q < n_param; ++q) out.params[k * n_param + q] = pv[q];
    } else {
        for (int q = 0; q < n_param; ++q) out.params[k * n_param + q] = __ldg(prm + q);
    }
    if (light_idx >= 0) {                                                     // :929-956
        // directional light: the parameter is the direction; point light ('point' texture entry): it is the position,
        // the direction becomes (position - sample) and the strength parameter the received radiance (:572-590)
        const float raw[3] = {__ldg(prm + light_idx), __ldg(prm + light_idx + 1), __ldg(prm + light_idx + 2)};
        bool shadowed = false;
        if (EXTRAS && sh.enabled) {                                           // :863, :936-951
            const uint32_t n_shadow = max((uint32_t)sh.min_samples, (uint32_t)__fmul_rn((float)sh.n_samples, ri.total));
            float q[3] = {pt[0], pt[1], pt[2]};
            if (n_shadow < (uint32_t)sh.n_pts) {
                // occlusion is sampled at points spaced evenly along the segment (:992-1000); the nearer neighbour counts
                const float2 sg = ws.seg[(size_t)ray * kMaxHits + (j > 0 ? j - 1 : 0)];
                const uint32_t n_seg = max((uint32_t)sh.min_samples,
                                           (uint32_t)__fdiv_rn(__fmul_rn((float)n_shadow, sg.y), ri.total));
                const float sstep = __fdiv_rn(sg.y, (float)(n_seg - 1u));
                auto t_at = [&](int kk) { return __fadd_rn(sg.x, __fmul_rn((float)kk, sstep)); };
                int kk = max(1, (int)ceilf((t_pt - sg.x) / sstep));
                if (!(kk < (1 << 24))) kk = 1;                                // degenerate step: walk from the start
                while (kk > 1 && !(t_pt > t_at(kk - 1))) --kk;
                for (int guard = 0; t_pt > t_at(kk) && guard < (1 << 20); ++guard) ++kk;
                const float t0 = kk == 1 ? sg.x : t_at(kk - 1);
                const float tq = (__fdiv_rn(__fsub_rn(t_pt, t0), sstep) >= 0.5f) ? t_at(kk) : t0;
#pragma unroll
                for (int a = 0; a < 3; ++a) q[a] = __fadd_rn(o[a], __fmul_rn(tq, d[a]));
            }
            shadowed = shadow_blocked(sc, sh.box, q, raw);
        }
        float lv[3] = {raw[0], raw[1], raw[2]};
        if (strength_idx >= 0) {
#pragma unroll
            for (int a = 0; a < 3; ++a) lv[a] = __fsub_rn(lv[a], pt[a]);
            const float dsq = __fadd_rn(__fadd_rn(__fmul_rn(lv[0], lv[0]), __fmul_rn(lv[1], lv[1])), __fmul_rn(lv[2], lv[2]));
            out.params[k * n_param + strength_idx] =
                __fdiv_rn(__ldg(prm + strength_idx), __fadd_rn(__fmul_rn(12.566370614359172f, dsq), 1e-6f));
        }
        if (shadowed) {                                                       // ambient only (:572-575)
            out.params[k * n_param + light_idx] = 0.0f;
            out.params[k * n_param + light_idx + 1] = 0.0f;
            out.params[k * n_param + light_idx + 2] = -1.0f;
        } else {
            const float ln = norm3(lv);
#pragma unroll
            for (int a = 0; a < 3; ++a) lv[a] = __fdiv_rn(lv[a], ln);
#pragma unroll
            for (int a = 0; a < 3; ++a) out.params[k * n_param + light_idx + a] = dot3(Dt + 3 * a, lv);
        }
    }
    out.dists[k] = (ri.n_steps < 0 || step == ri.n_steps - 1) ? ri.last_dist : step_size;
    out.t[k] = t_mu;
    out.weight[k] = w;
    out.inst[k] = (int32_t)sel;
}

}  // namespace

NTX_API int64_t ntx_instancer_workspace_bytes(int R, int cast_shadow_rays) {
    if (R < 0) return -1;
    return (int64_t)R * (kMaxHits * 16 + 16 + 4 + kFirstOpenStride) + 16 + (cast_shadow_rays ? (int64_t)R * kMaxHits * 8 : 0);
}

static int check_scene(const ntx_instancer_scene* s, SceneDev* out) {
    NTX_REQUIRE(s, "instancer: null scene");
    NTX_REQUIRE(s->n_inst >= 0 && s->n_inst <= (int)kInstMask && s->n_faces >= 0, "instancer: bad scene sizes");
    NTX_REQUIRE(s->n_inst == 0 || (s->minv && s->dirt && s->origins), "instancer: null instance arrays");
    NTX_REQUIRE(((uintptr_t)s->minv & 15) == 0, "instancer: minv must be 16-byte aligned");
    NTX_REQUIRE(!s->spheres || ((uintptr_t)s->spheres & 15) == 0, "instancer: spheres must be 16-byte aligned");
    NTX_REQUIRE(s->n_faces == 0 || (s->mesh_v && s->mesh_f), "instancer: a mesh needs vertices and faces");
    for (const ntx_scene_grid* g : {&s->inst_grid, &s->tri_grid}) {
        if (!g->cell_start) continue;
        NTX_REQUIRE(g->items, "instancer: grid without item list");
        for (int a = 0; a < 3; ++a)
            NTX_REQUIRE(g->res[a] >= 1 && g->res[a] <= 1024 && g->cell[a] > 0.0f, "instancer: bad grid axis %d (res %d, cell %g)",
                        a, g->res[a], (double)g->cell[a]);
    }
    out->minv = s->minv; out->dirt = s->dirt; out->origins = s->origins;
    out->spheres = reinterpret_cast<const float4*>(s->spheres);
    out->n_inst = s->n_inst; out->n_faces = s->n_faces; out->mesh_v = s->mesh_v; out->mesh_f = s->mesh_f;
    out->ig = make_grid(s->inst_grid);
    out->tg = make_grid(s->tri_grid);
    memset(&out->sh, 0, sizeof(out->sh));
    out->sh.n_base_faces = s->n_faces;                   // without a shading block every face is the base mesh's
    if (s->shading) {
        const ntx_instancer_shading& h = *s->shading;
        NTX_REQUIRE(h.n_base_faces >= 0 && h.n_base_faces <= s->n_faces && h.n_param_tex >= 0 &&
                    h.n_param_tex <= NTX_MAX_PARAM_TEXTURES && h.n_aux >= 0 && h.n_aux <= NTX_MAX_AUX_MESHES,
                    "instancer: bad shading block");
        NTX_REQUIRE(h.n_param_tex == 0 || (h.pool && h.mesh_uv && h.radius > 0.0f && h.min_samples >= 0 && h.n_samples >= 0),
                    "instancer: parameter textures need the pool, the mesh UVs and a search radius");
        NTX_REQUIRE(h.n_aux == 0 || h.mesh_n, "instancer: auxiliary meshes need vertex normals");
        for (int i = 0; i < h.n_aux; ++i)
            NTX_REQUIRE(h.aux_channels[i] == 0 || (h.pool && h.mesh_uv && (h.aux_channels[i] == 1 || h.aux_channels[i] == 3)),
                        "instancer: an auxiliary mesh texture has 1 or 3 channels and needs the pool and UVs");
        out->sh = h;
    }
    return NTX_OK;
}

NTX_API int ntx_instancer_trace(const float* rays_o, const float* rays_d, const int32_t* ray_ids, const float* ray_offsets, int R,
                                const ntx_instancer_scene* scene, const float* b_0, const float* b_1, float step_size,
                                int n_pts, uint32_t seed, int use_mean_distance, int cast_shadow_rays, const float* parameters,
                                int n_param, int light_dir_idx, void* workspace, int32_t* n_samples, int32_t* hit,
                                float* alpha_last, float* color_last, void* stream) {
    NTX_REQUIRE(R >= 0 && n_pts >= 1 && step_size > 0.0f, "instancer_trace: bad arguments");
    SceneDev sc;
    int rc = check_scene(scene, &sc);
    if (rc) return rc;
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(rays_o && rays_d && b_0 && b_1 && workspace && n_samples && hit && alpha_last, "instancer_trace: null pointer");
    NTX_REQUIRE(((uintptr_t)workspace & 15) == 0, "instancer_trace: the workspace must be 16-byte aligned");
    const size_t smem = (size_t)kSmemHits * kTraceThreads * sizeof(uint64_t) + (size_t)kMailbox * kTraceThreads * sizeof(uint32_t);
    NTX_CUDA(cudaFuncSetAttribute(instancer_trace_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NTX_REQUIRE(sc.sh.n_aux == 0 || color_last, "instancer_trace: auxiliary meshes need the color_last output");
    NTX_REQUIRE(light_dir_idx < 0 || (parameters && light_dir_idx + 3 <= n_param), "instancer_trace: light index out of range");
    Workspace wsp = carve_workspace(workspace, R);
    if (!cast_shadow_rays) wsp.seg = nullptr;            // that part of the workspace is not there
    instancer_trace_kernel<<<(R + kTraceThreads - 1) / kTraceThreads, kTraceThreads, smem, (cudaStream_t)stream>>>(
        rays_o, rays_d, ray_ids, ray_offsets, R, sc, b_0[0], b_0[1], b_0[2], b_1[0], b_1[1], b_1[2], step_size, n_pts, seed,
        use_mean_distance, wsp, n_samples, hit, alpha_last, parameters, n_param, light_dir_idx, color_last);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_instancer_march(const float* rays_o, const float* rays_d, const int32_t* ray_ids, const float* parameters,
                                int n_param, int R, const ntx_instancer_scene* scene, float step_size, int n_pts,
                                uint32_t seed, int light_dir_idx, int light_strength_idx, int method, float blend_range,
                                int use_mean_distance, int cast_shadow_rays, int min_shadow_samples, int n_shadow_samples,
                                const float* b_0, const float* b_1, const void* workspace, const int64_t* ray_ptr, int64_t T,
                                float* pts, float* dirs, float* params_out, float* dists, float* t, float* alpha_weight,
                                int32_t* instance_id, void* stream) {
    NTX_REQUIRE(R >= 0 && T >= 0 && n_param >= 0 && n_param <= 16, "instancer_march: bad arguments (at most 16 parameters)");
    NTX_REQUIRE(light_dir_idx < 0 || light_dir_idx + 3 <= n_param, "instancer_march: light direction index out of range");
    NTX_REQUIRE(light_strength_idx < 0 || (light_strength_idx < n_param && light_dir_idx >= 0),
                "instancer_march: a light strength parameter needs a light position (light_dir_idx) too");
    NTX_REQUIRE(method >= 0 && method <= 2, "instancer_march: instance sampling method must be 0 (random), 1 (nearest) or 2 (nearest_blend)");
    SceneDev sc;
    int rc = check_scene(scene, &sc);
    if (rc) return rc;
    if (R == 0 || T == 0) return NTX_OK;
    NTX_REQUIRE(rays_o && rays_d && (n_param == 0 || parameters) && workspace && ray_ptr, "instancer_march: null pointer");
    NTX_REQUIRE(pts && dirs && (n_param == 0 || params_out) && dists && t && alpha_weight && instance_id,
                "instancer_march: null output pointer");
    ShadowArgs sh;
    memset(&sh, 0, sizeof(sh));
    sh.n_pts = n_pts;
    if (cast_shadow_rays && light_dir_idx >= 0) {        // :936 shadows only matter where a light is mapped
        NTX_REQUIRE(b_0 && b_1 && min_shadow_samples >= 0 && n_shadow_samples >= 0, "instancer_march: shadow rays need b_0 / b_1 and sample counts");
        sh.enabled = 1; sh.min_samples = min_shadow_samples; sh.n_samples = n_shadow_samples; sh.n_pts = n_pts;
        for (int a = 0; a < 3; ++a) { sh.box.b0[a] = b_0[a]; sh.box.b1[a] = b_1[a]; }
    }
    MarchOut out = {pts, dirs, params_out, dists, t, alpha_weight, instance_id};
    const bool extras = sh.enabled || (sc.sh.n_param_tex > 0 && sc.sh.n_base_faces > 0);
    if (extras)
        instancer_emit_kernel<true><<<(unsigned)((T + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
            rays_o, rays_d, ray_ids, parameters, n_param, R, sc, step_size, seed, light_dir_idx,
            light_strength_idx, method, blend_range, use_mean_distance, carve_workspace(const_cast<void*>(workspace), R), ray_ptr,
            T, out, sh);
    else
        instancer_emit_kernel<false><<<(unsigned)((T + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
            rays_o, rays_d, ray_ids, parameters, n_param, R, sc, step_size, seed, light_dir_idx,
            light_strength_idx, method, blend_range, use_mean_distance, carve_workspace(const_cast<void*>(workspace), R), ray_ptr,
            T, out, sh);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

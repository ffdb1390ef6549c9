// K2 — alpha compositing (HBM-bound, one warp per ray).
//
// Per round a warp loads 32 consecutive samples of its ray (one 128-bit rgb-sigma load and one
// depth per lane, fully coalesced), maps sigma to alpha, forms the transmittance with a
// warp-level exclusive product scan (5 shuffles) chained across rounds by a carry, and keeps
// per-lane partial sums of w*rgb, w*z and w that are reduced once per ray.
//
// Algorithmic HBM bytes: plain 24*S + 28 per ray (rgb 12 + sigma 4 + z 4 read,
// weights 4 written per sample; rays_d 12 read, rgb 12 + acc 4 written per ray).
#include "ntx_common.cuh"

namespace {

__device__ __forceinline__ float color_act(float c, int map_exr) {
    // renderer.py:182-187
    // ex2.approx-based exp and an approximate divide: ~1e-6 relative, inside the 1e-5 parity bar
    if (map_exr) return c > 0.0f ? c + 1.0f : __expf(c);
    return __fdividef(1.0f, 1.0f + __expf(-c));
}

__device__ __forceinline__ float color_act_grad(float c, float y, int map_exr) {
    // derivative of color_act given its output y
    if (map_exr) return c > 0.0f ? 1.0f : y;
    return y * (1.0f - y);
}

// inclusive product scan across the warp
__device__ __forceinline__ float warp_scan_mul(float v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        float n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v *= n;
    }
    return v;
}

__device__ __forceinline__ float warp_scan_add(float v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        float n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}

__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

struct Sample {
    float r, g, b, s;
};

template <bool PACKED>
__device__ __forceinline__ Sample load_sample(const float* color, const float* sigma, int64_t k) {
    Sample o;
    if (PACKED) {
        float4 v = ldg_stream_f4(reinterpret_cast<const float4*>(color) + k);
        o.r = v.x; o.g = v.y; o.b = v.z; o.s = v.w;
    } else {
        o.r = __ldg(color + 3 * k); o.g = __ldg(color + 3 * k + 1); o.b = __ldg(color + 3 * k + 2);
        o.s = __ldg(sigma + k);
    }
    return o;
}

// distance of sample i to the next one; the last distance is duplicated (renderer.py:174-177)
__device__ __forceinline__ float sample_dist(const float* zrow, int i, int S, float zi, int edges) {
    if (edges) return __ldg(zrow + i + 1) - zi;          // S + 1 segment edges (mip renderer, renderer.py:440)
    if (S == 1) return 0.0f;
    if (i < S - 1) return __ldg(zrow + i + 1) - zi;
    return zi - __ldg(zrow + i - 1);
}

// one round of inputs of a lane: raw sample, depth, optional noise
struct RoundIn {
    Sample sm;
    float z, nz;
};

template <bool PACKED>
__device__ __forceinline__ RoundIn load_round(const float* color, const float* sigma, const float* zrow,
                                              const float* noise, int64_t row0, int i, int S, int edges) {
    RoundIn r;
    r.sm = {0.f, 0.f, 0.f, 0.f};
    r.z = 0.0f; r.nz = 0.0f;
    if (i < S) {
        r.sm = load_sample<PACKED>(color, sigma, row0 + i);
        if (noise) r.nz = __ldg(noise + row0 + i);
    }
    if (i < S + edges) r.z = __ldg(zrow + i);            // edges: one more depth than samples
    return r;
}

template <bool PACKED>
__global__ void __launch_bounds__(256)
composite_fwd_kernel(const float* __restrict__ color, const float* __restrict__ sigma, const float* __restrict__ z,
                     const float* __restrict__ rays_d, const float* __restrict__ noise, int R, int S, int map_exr,
                     int composite_bkgd, float bk0, float bk1, float bk2, float* __restrict__ color_map,
                     float* __restrict__ alpha_map, float* __restrict__ weights, float* __restrict__ depth, int edges) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < R; ray += nwarps) {
        const int64_t row0 = ray * S;
        const float* zrow = z + ray * (S + edges);
        // software pipeline: the loads of round r+1 are in flight while round r is scanned
        RoundIn cur = load_round<PACKED>(color, sigma, zrow, noise, row0, lane, S, edges);
        const float d0 = __ldg(rays_d + 3 * ray), d1 = __ldg(rays_d + 3 * ray + 1), d2 = __ldg(rays_d + 3 * ray + 2);
        const float norm = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);                 // renderer.py:180
        float carry = 1.0f, c0 = 0.0f, c1 = 0.0f, c2 = 0.0f, acc = 0.0f, dep = 0.0f, prev_z31 = 0.0f;
        for (int base = 0; base < S; base += 32) {
            const int i = base + lane;
            const bool valid = i < S;
            const RoundIn nxt = load_round<PACKED>(color, sigma, zrow, noise, row0, i + 32, S, edges);
            // neighbours' depths by shuffle: z[i+1] from lane+1 (lane 31: next round's lane 0),
            // z[i-1] from lane-1 (lane 0: previous round's lane 31)
            float z_up = __shfl_down_sync(0xffffffffu, cur.z, 1);
            const float z_next0 = __shfl_sync(0xffffffffu, nxt.z, 0);
            if (lane == 31) z_up = z_next0;
            float z_dn = __shfl_up_sync(0xffffffffu, cur.z, 1);
            if (lane == 0) z_dn = prev_z31;
            prev_z31 = __shfl_sync(0xffffffffu, cur.z, 31);
            float a = 0.0f;
            if (valid) {
                // last distance duplicated (renderer.py:174-177)
                float dist = edges ? z_up - cur.z : (S == 1) ? 0.0f : (i < S - 1 ? z_up - cur.z : cur.z - z_dn);
                dist *= norm;
                const float s = fmaxf(cur.sm.s + cur.nz, 0.0f);                 // renderer.py:190-195
                a = 1.0f - __expf(-s * dist);
            }
            const float om = (1.0f - a) + 1e-10f;                               // renderer.py:198
            const float incl = warp_scan_mul(om, lane);
            float excl = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) excl = 1.0f;
            const float w = a * (carry * excl);
            carry *= __shfl_sync(0xffffffffu, incl, 31);
            if (valid) {
                if (weights) weights[row0 + i] = w;
                c0 += w * color_act(cur.sm.r, map_exr);                         // renderer.py:201
                c1 += w * color_act(cur.sm.g, map_exr);
                c2 += w * color_act(cur.sm.b, map_exr);
                dep += w * cur.z;                                               // renderer.py:204
                acc += w;                                                       // renderer.py:207
            }
            cur = nxt;
        }
        c0 = ntx_warp_sum(c0); c1 = ntx_warp_sum(c1); c2 = ntx_warp_sum(c2);
        acc = ntx_warp_sum(acc);
        dep = ntx_warp_sum(dep);
        if (lane == 0) {
            if (composite_bkgd) {                                               // renderer.py:210-211
                const float om = 1.0f - acc;
                c0 += om * bk0; c1 += om * bk1; c2 += om * bk2;
            }
            color_map[3 * ray] = c0; color_map[3 * ray + 1] = c1; color_map[3 * ray + 2] = c2;
            alpha_map[ray] = acc;
            if (depth) depth[ray] = dep;
        }
    }
}

// ---- packed fast path: two samples per lane -------------------------------------------------------------------
// The one-sample-per-lane kernel above is issue-bound (ncu: ~170 warp instructions per 32 samples).  Here a lane
// owns two consecutive samples, so one warp round covers 64 samples: the product scan, the depth shuffles, the
// pointer arithmetic and the loop overhead are paid once per 64 samples, exp / reciprocal are single MUFU ops, and the
// per-ray reductions of (r, g, b, acc) share six shuffles.  Needs the packed layout and an even S.
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
template <bool MAP_EXR>
__device__ __forceinline__ float color_act2(float c) {                         // renderer.py:182-187
    constexpr float kLog2e = 1.4426950408889634f;
    if (MAP_EXR) return c > 0.0f ? c + 1.0f : ex2_approx(c * kLog2e);
    return rcp_approx(1.0f + ex2_approx(-c * kLog2e));
}
// v *= (value of lane - o), for lanes >= o (the shuffle's own range predicate guards the multiply)
__device__ __forceinline__ void scan_mul_step(float& v, int o) {
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t;\n\tshfl.sync.up.b32 t|p, %0, %1, 0, 0xffffffff;\n\t@p mul.f32 %0, %0, t;\n\t}"
        : "+f"(v) : "r"(o));
}

struct PairIn {
    float4 a, b;      // raw rgb-sigma of the lane's two samples
    float2 z, n;      // depths, sigma noise
};

__device__ __forceinline__ PairIn load_pair(const float4* prow, const float2* zrow, const float2* nrow, int pair, int npairs) {
    PairIn r;
    r.a = make_float4(0.f, 0.f, 0.f, 0.f);
    r.b = r.a;
    r.z = make_float2(0.f, 0.f);
    r.n = r.z;
    if (pair < npairs) {
        r.a = ldg_stream_f4(prow + 2 * pair);
        r.b = ldg_stream_f4(prow + 2 * pair + 1);
        r.z = __ldg(zrow + pair);
        if (nrow) r.n = __ldg(nrow + pair);
    }
    return r;
}

template <bool MAP_EXR>
__global__ void __launch_bounds__(256)
composite_fwd2_kernel(const float4* __restrict__ rgbs, const float* __restrict__ z, const float* __restrict__ rays_d,
                      const float* __restrict__ noise, int R, int S, int composite_bkgd, float bk0, float bk1, float bk2,
                      float* __restrict__ color_map, float* __restrict__ alpha_map, float* __restrict__ weights,
                      float* __restrict__ depth) {
    constexpr float kLog2e = 1.4426950408889634f;
    const int lane = threadIdx.x & 31;
    const int npairs = S >> 1;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < R; ray += nwarps) {
        const int64_t row0 = ray * S;
        const float4* prow = rgbs + row0;
        const float2* zrow = reinterpret_cast<const float2*>(z + row0);
        const float2* nrow = noise ? reinterpret_cast<const float2*>(noise + row0) : nullptr;
        float2* wrow = weights ? reinterpret_cast<float2*>(weights + row0) : nullptr;
        PairIn cur = load_pair(prow, zrow, nrow, lane, npairs);
        const float d0 = __ldg(rays_d + 3 * ray), d1 = __ldg(rays_d + 3 * ray + 1), d2 = __ldg(rays_d + 3 * ray + 2);
        const float scale = -kLog2e * sqrtf(d0 * d0 + d1 * d1 + d2 * d2);      // renderer.py:180, folded into the exp2
        float carry = 1.0f, c0 = 0.0f, c1 = 0.0f, c2 = 0.0f, acc = 0.0f, dep = 0.0f;
        for (int pbase = 0; pbase < npairs; pbase += 32) {
            const int pair = pbase + lane;
            const PairIn nxt = load_pair(prow, zrow, nrow, pair + 32, npairs);
            // z of the sample after the lane's second one: first depth of lane + 1 (lane 31: of the next round)
            float z_up = __shfl_down_sync(0xffffffffu, cur.z.x, 1);
            const float z_next0 = __shfl_sync(0xffffffffu, nxt.z.x, 0);
            if (lane == 31) z_up = z_next0;
            const float dist0 = cur.z.y - cur.z.x;
            const float dist1 = (pair == npairs - 1) ? dist0 : z_up - cur.z.y;   // last distance duplicated (:174-177)
            // lanes past the end hold zeros: sigma 0 -> alpha 0 -> factor 1 (1 + 1e-10 rounds to 1)
            const float s0 = fmaxf(cur.a.w + cur.n.x, 0.0f), s1 = fmaxf(cur.b.w + cur.n.y, 0.0f);   // renderer.py:190-195
            const float a0 = 1.0f - ex2_approx(s0 * dist0 * scale);
            const float a1 = 1.0f - ex2_approx(s1 * dist1 * scale);
            const float om0 = (1.0f - a0) + 1e-10f, om1 = (1.0f - a1) + 1e-10f;                    // renderer.py:198
            float incl = om0 * om1;
            scan_mul_step(incl, 1);
            scan_mul_step(incl, 2);
            scan_mul_step(incl, 4);
            scan_mul_step(incl, 8);
            scan_mul_step(incl, 16);
            float excl = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) excl = 1.0f;
            const float t0 = carry * excl;
            const float w0 = a0 * t0;
            const float w1 = a1 * (t0 * om0);
            carry *= __shfl_sync(0xffffffffu, incl, 31);
            if (wrow && pair < npairs) wrow[pair] = make_float2(w0, w1);
            c0 += w0 * color_act2<MAP_EXR>(cur.a.x) + w1 * color_act2<MAP_EXR>(cur.b.x);           // renderer.py:201
            c1 += w0 * color_act2<MAP_EXR>(cur.a.y) + w1 * color_act2<MAP_EXR>(cur.b.y);
            c2 += w0 * color_act2<MAP_EXR>(cur.a.z) + w1 * color_act2<MAP_EXR>(cur.b.z);
            dep += w0 * cur.z.x + w1 * cur.z.y;                                                     // renderer.py:204
            acc += w0 + w1;                                                                         // renderer.py:207
            cur = nxt;
        }
        // four sums in six shuffles: halves of the warp trade what the other half keeps
        const bool h16 = lane & 16, h8 = lane & 8;
        float x = (h16 ? c1 : c0) + __shfl_xor_sync(0xffffffffu, h16 ? c0 : c1, 16);
        float y = (h16 ? acc : c2) + __shfl_xor_sync(0xffffffffu, h16 ? c2 : acc, 16);
        float u = (h8 ? y : x) + __shfl_xor_sync(0xffffffffu, h8 ? x : y, 8);
        u += __shfl_xor_sync(0xffffffffu, u, 4);
        u += __shfl_xor_sync(0xffffffffu, u, 2);
        u += __shfl_xor_sync(0xffffffffu, u, 1);
        // lane 0: r, lane 16: g, lane 8: b, lane 24: acc
        const float acc_all = __shfl_sync(0xffffffffu, u, 24);
        if (depth) {                                                             // warp-uniform
            dep = ntx_warp_sum(dep);
            if (lane == 0) depth[ray] = dep;
        }
        if ((lane & 7) == 0) {
            const int comp = (lane >> 4) | ((lane >> 2) & 2);
            if (comp == 3) {
                alpha_map[ray] = u;
            } else {
                const float bk = comp == 0 ? bk0 : comp == 1 ? bk1 : bk2;
                color_map[3 * ray + comp] = composite_bkgd ? u + (1.0f - acc_all) * bk : u;       // renderer.py:210-211
            }
        }
    }
}

// Backward w.r.t. the packed raw outputs.  Pass 1 recomputes the forward and accumulates
// G = sum_k gw_k * w_k; pass 2 recomputes again and emits
//   d a_i   = gw_i * T_i - (G - prefix_incl_i) / om_i
//   d sigma = d a_i * dist_i * (1 - a_i) * [sigma_i + noise_i > 0]
//   d rgb   = w_i * dC * act'(raw)
__global__ void __launch_bounds__(256)
composite_bwd_kernel(const float* __restrict__ rgbs, const float* __restrict__ z, const float* __restrict__ rays_d,
                     const float* __restrict__ noise, int R, int S, int map_exr, int composite_bkgd, float bk0,
                     float bk1, float bk2, const float* __restrict__ d_color_map,
                     const float* __restrict__ d_alpha_map, float* __restrict__ d_rgbs, int edges) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; ray < R; ray += nwarps) {
        const float d0 = __ldg(rays_d + 3 * ray), d1 = __ldg(rays_d + 3 * ray + 1), d2 = __ldg(rays_d + 3 * ray + 2);
        const float norm = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
        const float g0 = __ldg(d_color_map + 3 * ray), g1 = __ldg(d_color_map + 3 * ray + 1),
                    g2 = __ldg(d_color_map + 3 * ray + 2);
        float gA = __ldg(d_alpha_map + ray);
        if (composite_bkgd) gA -= g0 * bk0 + g1 * bk1 + g2 * bk2;
        const float* zrow = z + ray * (S + edges);
        float G = 0.0f;
        for (int pass = 0; pass < 2; ++pass) {
            float carry = 1.0f, gcarry = 0.0f, gsum = 0.0f;
            for (int base = 0; base < S; base += 32) {
                const int i = base + lane;
                const bool valid = i < S;
                float a = 0.0f, dist = 0.0f, sraw = 0.0f;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (valid) {
                    const int64_t k = ray * S + i;
                    v = __ldg(reinterpret_cast<const float4*>(rgbs) + k);
                    const float zi = __ldg(zrow + i);
                    dist = sample_dist(zrow, i, S, zi, edges) * norm;
                    sraw = v.w;
                    if (noise) sraw += __ldg(noise + k);
                    a = 1.0f - expf(-fmaxf(sraw, 0.0f) * dist);
                }
                const float om = (1.0f - a) + 1e-10f;
                const float incl = warp_scan_mul(om, lane);
                float excl = __shfl_up_sync(0xffffffffu, incl, 1);
                if (lane == 0) excl = 1.0f;
                const float T = carry * excl;
                const float w = a * T;
                carry *= __shfl_sync(0xffffffffu, incl, 31);
                const float y0 = color_act(v.x, map_exr), y1 = color_act(v.y, map_exr), y2 = color_act(v.z, map_exr);
                const float gw = valid ? (g0 * y0 + g1 * y1 + g2 * y2 + gA) : 0.0f;
                const float gww = gw * w;
                if (pass == 0) {
                    gsum += gww;
                } else {
                    const float gscan = warp_scan_add(gww, lane);
                    const float pre = gcarry + gscan;
                    gcarry += __shfl_sync(0xffffffffu, gscan, 31);
                    if (valid) {
                        const float da = gw * T - (G - pre) / om;
                        const float ds = (sraw > 0.0f) ? da * dist * (1.0f - a) : 0.0f;
                        float4 o;
                        o.x = w * g0 * color_act_grad(v.x, y0, map_exr);
                        o.y = w * g1 * color_act_grad(v.y, y1, map_exr);
                        o.z = w * g2 * color_act_grad(v.z, y2, map_exr);
                        o.w = ds;
                        reinterpret_cast<float4*>(d_rgbs)[ray * S + i] = o;
                    }
                }
            }
            if (pass == 0) G = ntx_warp_sum(gsum);
        }
    }
}

// Instanced variant over CSR-packed samples with the mesh hit appended (renderer.py:318-354).  A ray's samples
// start anywhere in the packed arrays, so it stays one sample per lane; the rest follows the packed fast path: the
// next round's loads are issued before the current one is scanned, exp / sigmoid are single MUFU ops with the
// constant factors folded into one scale, the scan multiplies under the shuffle's own predicate, and the four per-ray
// sums share six shuffles.
struct InstIn {
    float4 v;
    float dist, aw;
};

__device__ __forceinline__ InstIn load_inst(const float4* __restrict__ rgbs, const float* __restrict__ dists,
                                            const float* __restrict__ alpha_weight, int64_t k, int64_t end) {
    InstIn r;
    r.v = make_float4(0.f, 0.f, 0.f, 0.f);
    r.dist = 0.0f;
    r.aw = 1.0f;
    if (k < end) {
        r.v = ldg_stream_f4(rgbs + k);
        r.dist = __ldg(dists + k);
        if (alpha_weight) r.aw = __ldg(alpha_weight + k);
    }
    return r;
}

template <bool MAP_EXR>
__global__ void __launch_bounds__(256)
composite_inst_kernel(const float* __restrict__ rgbs_, const float* __restrict__ dists,
                      const float* __restrict__ alpha_weight, const int64_t* __restrict__ ray_ptr,
                      const float* __restrict__ color_last, const float* __restrict__ alpha_last,
                      const int32_t* __restrict__ ray_index, const float* __restrict__ instance_color, int H,
                      float patch_scale, float density_scale, int composite_bkgd, float bk0, float bk1,
                      float bk2, float* __restrict__ color_map, float* __restrict__ alpha_map) {
    constexpr float kLog2e = 1.4426950408889634f;
    const float4* rgbs = reinterpret_cast<const float4*>(rgbs_);
    const int lane = threadIdx.x & 31;
    const float scale = -kLog2e / patch_scale;                                   // renderer.py:339, folded into the exp2
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t h = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; h < H; h += nwarps) {
        const int64_t beg = __ldg(ray_ptr + h), end = __ldg(ray_ptr + h + 1);
        float carry = 1.0f, c0 = 0.0f, c1 = 0.0f, c2 = 0.0f, acc = 0.0f;
        InstIn cur = load_inst(rgbs, dists, alpha_weight, beg + lane, end);
        for (int64_t base = beg; base < end; base += 32) {
            const int64_t k = base + lane;
            const InstIn nxt = load_inst(rgbs, dists, alpha_weight, k + 32, end);
            // lanes past the end hold sigma 0 -> alpha 0 -> factor 1 (1 + 1e-10 rounds to 1)
            const float s = fmaxf(cur.v.w * (cur.aw * density_scale), 0.0f);     // renderer.py:300
            const float a = 1.0f - ex2_approx((s * cur.dist) * scale);
            float incl = (1.0f - a) + 1e-10f;
            scan_mul_step(incl, 1);
            scan_mul_step(incl, 2);
            scan_mul_step(incl, 4);
            scan_mul_step(incl, 8);
            scan_mul_step(incl, 16);
            float excl = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) excl = 1.0f;
            const float w = a * (carry * excl);
            carry *= __shfl_sync(0xffffffffu, incl, 31);
            float y0, y1, y2;
            if (instance_color) {                                               // renderer.py:306-307,322-323
                y0 = y1 = y2 = 0.0f;
                if (k < end) {
                    y0 = __ldg(instance_color + 3 * k); y1 = __ldg(instance_color + 3 * k + 1);
                    y2 = __ldg(instance_color + 3 * k + 2);
                }
            } else {
                y0 = color_act2<MAP_EXR>(cur.v.x); y1 = color_act2<MAP_EXR>(cur.v.y); y2 = color_act2<MAP_EXR>(cur.v.z);
            }
            c0 += w * y0; c1 += w * y1; c2 += w * y2;
            acc += w;
            cur = nxt;
        }
        // four sums in six shuffles: halves of the warp trade what the other half keeps
        const bool h16 = lane & 16, h8 = lane & 8;
        float x = (h16 ? c1 : c0) + __shfl_xor_sync(0xffffffffu, h16 ? c0 : c1, 16);
        float y = (h16 ? acc : c2) + __shfl_xor_sync(0xffffffffu, h16 ? c2 : acc, 16);
        float u = (h8 ? y : x) + __shfl_xor_sync(0xffffffffu, h8 ? x : y, 8);
        u += __shfl_xor_sync(0xffffffffu, u, 4);
        u += __shfl_xor_sync(0xffffffffu, u, 2);
        u += __shfl_xor_sync(0xffffffffu, u, 1);
        // lane 0: r, lane 16: g, lane 8: b, lane 24: acc.  Appended sample S+1: the opaque mesh hit (renderer.py:331,339)
        const float wl = __ldg(alpha_last + h) * carry;
        const float acc_all = __shfl_sync(0xffffffffu, u, 24) + wl;
        if ((lane & 7) == 0) {
            const int comp = (lane >> 4) | ((lane >> 2) & 2);
            const int64_t dst = ray_index ? (int64_t)__ldg(ray_index + h) : h;
            if (comp == 3) {
                alpha_map[dst] = acc_all;
            } else {
                const float bk = comp == 0 ? bk0 : comp == 1 ? bk1 : bk2;
                float c = u + wl * __ldg(color_last + 3 * h + comp);
                if (composite_bkgd) c += (1.0f - acc_all) * bk;
                color_map[3 * dst + comp] = c;
            }
        }
    }
}

int grid_for_warps(int64_t n_warp_items, int threads) {
    const int warps_per_block = threads / 32;
    int64_t blocks = (n_warp_items + warps_per_block - 1) / warps_per_block;
    const int64_t cap = (int64_t)ntx_sm_count() * 8 * 4;   // a few waves of 8 resident blocks per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

static int composite_forward_impl(const float* color, const float* sigma, int packed, const float* z,
                                  const float* rays_d, const float* noise, int R, int S, int map_exr,
                                  int composite_bkgd, const float* bkgd_color, float* color_map, float* alpha_map,
                                  float* weights, float* depth, int edges, void* stream) {
    NTX_REQUIRE(R >= 0 && S >= 1, "composite_forward: bad shape R=%d S=%d", R, S);
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(color && z && rays_d && color_map && alpha_map, "composite_forward: null pointer");
    NTX_REQUIRE(packed || sigma, "composite_forward: split layout requires sigma");
    NTX_REQUIRE(!composite_bkgd || bkgd_color, "composite_forward: composite_bkgd requires bkgd_color");
    NTX_REQUIRE(!packed || ((uintptr_t)color & 15) == 0, "composite_forward: packed rgbs must be 16-byte aligned");
    const float b0 = composite_bkgd ? bkgd_color[0] : 0.f, b1 = composite_bkgd ? bkgd_color[1] : 0.f,
                b2 = composite_bkgd ? bkgd_color[2] : 0.f;
    const int threads = 256;
    const int grid = grid_for_warps(R, threads);
    cudaStream_t st = (cudaStream_t)stream;
    const bool pairs = !edges && packed && (S & 1) == 0 &&
                       ((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(noise) | reinterpret_cast<uintptr_t>(weights)) & 7) == 0;
    if (pairs && map_exr)
        composite_fwd2_kernel<true><<<grid, threads, 0, st>>>(reinterpret_cast<const float4*>(color), z, rays_d, noise, R, S,
                                                              composite_bkgd, b0, b1, b2, color_map, alpha_map, weights, depth);
    else if (pairs)
        composite_fwd2_kernel<false><<<grid, threads, 0, st>>>(reinterpret_cast<const float4*>(color), z, rays_d, noise, R, S,
                                                               composite_bkgd, b0, b1, b2, color_map, alpha_map, weights, depth);
    else if (packed)
        composite_fwd_kernel<true><<<grid, threads, 0, st>>>(color, sigma, z, rays_d, noise, R, S, map_exr,
                                                             composite_bkgd, b0, b1, b2, color_map, alpha_map,
                                                             weights, depth, edges);
    else
        composite_fwd_kernel<false><<<grid, threads, 0, st>>>(color, sigma, z, rays_d, noise, R, S, map_exr,
                                                              composite_bkgd, b0, b1, b2, color_map, alpha_map,
                                                              weights, depth, edges);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_composite_forward(const float* color, const float* sigma, int packed, const float* z,
                                  const float* rays_d, const float* noise, int R, int S, int map_exr,
                                  int composite_bkgd, const float* bkgd_color, float* color_map, float* alpha_map,
                                  float* weights, float* depth, void* stream) {
    return composite_forward_impl(color, sigma, packed, z, rays_d, noise, R, S, map_exr, composite_bkgd, bkgd_color,
                                  color_map, alpha_map, weights, depth, 0, stream);
}

NTX_API int ntx_composite_edges_forward(const float* rgbs, const float* z_edges, const float* rays_d, const float* noise,
                                        int R, int S, int map_exr, int composite_bkgd, const float* bkgd_color,
                                        float* color_map, float* alpha_map, float* weights, void* stream) {
    return composite_forward_impl(rgbs, nullptr, 1, z_edges, rays_d, noise, R, S, map_exr, composite_bkgd, bkgd_color,
                                  color_map, alpha_map, weights, nullptr, 1, stream);
}

static int composite_backward_impl(const float* rgbs, const float* z, const float* rays_d, const float* noise, int R,
                                   int S, int map_exr, int composite_bkgd, const float* bkgd_color,
                                   const float* d_color_map, const float* d_alpha_map, float* d_rgbs, int edges, void* stream) {
    NTX_REQUIRE(R >= 0 && S >= 1, "composite_backward: bad shape R=%d S=%d", R, S);
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(rgbs && z && rays_d && d_color_map && d_alpha_map && d_rgbs, "composite_backward: null pointer");
    NTX_REQUIRE(!composite_bkgd || bkgd_color, "composite_backward: composite_bkgd requires bkgd_color");
    NTX_REQUIRE((((uintptr_t)rgbs | (uintptr_t)d_rgbs) & 15) == 0, "composite_backward: rgbs must be 16-byte aligned");
    const float b0 = composite_bkgd ? bkgd_color[0] : 0.f, b1 = composite_bkgd ? bkgd_color[1] : 0.f,
                b2 = composite_bkgd ? bkgd_color[2] : 0.f;
    const int threads = 256;
    composite_bwd_kernel<<<grid_for_warps(R, threads), threads, 0, (cudaStream_t)stream>>>(
        rgbs, z, rays_d, noise, R, S, map_exr, composite_bkgd, b0, b1, b2, d_color_map, d_alpha_map, d_rgbs, edges);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_composite_backward(const float* rgbs, const float* z, const float* rays_d, const float* noise, int R,
                                   int S, int map_exr, int composite_bkgd, const float* bkgd_color,
                                   const float* d_color_map, const float* d_alpha_map, float* d_rgbs, void* stream) {
    return composite_backward_impl(rgbs, z, rays_d, noise, R, S, map_exr, composite_bkgd, bkgd_color, d_color_map,
                                   d_alpha_map, d_rgbs, 0, stream);
}

NTX_API int ntx_composite_edges_backward(const float* rgbs, const float* z_edges, const float* rays_d, const float* noise,
                                         int R, int S, int map_exr, int composite_bkgd, const float* bkgd_color,
                                         const float* d_color_map, const float* d_alpha_map, float* d_rgbs, void* stream) {
    return composite_backward_impl(rgbs, z_edges, rays_d, noise, R, S, map_exr, composite_bkgd, bkgd_color, d_color_map,
                                   d_alpha_map, d_rgbs, 1, stream);
}

NTX_API int ntx_composite_instanced_forward(const float* rgbs, const float* dists, const float* alpha_weight,
                                            const int64_t* ray_ptr, const float* color_last,
                                            const float* alpha_last, const int32_t* ray_index,
                                            const float* instance_color, int H, float patch_scale,
                                            float density_scale, int map_exr, int composite_bkgd,
                                            const float* bkgd_color, float* color_map, float* alpha_map,
                                            void* stream) {
    NTX_REQUIRE(H >= 0, "composite_instanced_forward: bad H=%d", H);
    if (H == 0) return NTX_OK;
    NTX_REQUIRE(ray_ptr && color_last && alpha_last && color_map && alpha_map,
                "composite_instanced_forward: null pointer");
    NTX_REQUIRE(patch_scale > 0.0f, "composite_instanced_forward: patch_scale must be positive");
    NTX_REQUIRE(!composite_bkgd || bkgd_color, "composite_instanced_forward: composite_bkgd requires bkgd_color");
    NTX_REQUIRE(((uintptr_t)rgbs & 15) == 0, "composite_instanced_forward: rgbs must be 16-byte aligned");
    const float b0 = composite_bkgd ? bkgd_color[0] : 0.f, b1 = composite_bkgd ? bkgd_color[1] : 0.f,
                b2 = composite_bkgd ? bkgd_color[2] : 0.f;
    const int threads = 256;
    if (map_exr)
        composite_inst_kernel<true><<<grid_for_warps(H, threads), threads, 0, (cudaStream_t)stream>>>(
            rgbs, dists, alpha_weight, ray_ptr, color_last, alpha_last, ray_index, instance_color, H, patch_scale,
            density_scale, composite_bkgd, b0, b1, b2, color_map, alpha_map);
    else
        composite_inst_kernel<false><<<grid_for_warps(H, threads), threads, 0, (cudaStream_t)stream>>>(
            rgbs, dists, alpha_weight, ray_ptr, color_last, alpha_last, ray_index, instance_color, H, patch_scale,
            density_scale, composite_bkgd, b0, b1, b2, color_map, alpha_map);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

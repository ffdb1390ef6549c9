// Shell instancer on the device (row f1 of the scope table): what C_Instancer::GetModelInput
// (instancer/src/instancer.cpp:751-1037) does on the host with Embree, for the subset that needs no mesh —
// explicit patch instances, no parameter textures, no shadow rays:
//   pass 1 (ntx_instancer_trace): one thread per ray tests every instanced patch box (slab test in the patch
//          frame, instance matrices staged through shared memory), keeps the sorted hit list, walks the
//          in-shell intervals and counts the marching samples;
//   pass 2 (ntx_instancer_march): after a prefix sum over the counts, the same walk emits the samples
//          CSR-packed: patch-local point / view direction / parameters (light direction mapped into the patch
//          frame), step length, ray parameter, overlap weight, instance id.
// Arithmetic is fp32 with explicit round-to-nearest ops in the order of oracle/instancer_oracle.py (bit-exact);
// random draws come from a counter hash of (seed, ray[, step]) instead of the reference's std::mt19937.
#include <math_constants.h>

#include "ntx_common.cuh"

namespace {

constexpr int kMaxHits = 64;
constexpr int kInstTile = 128;
constexpr float kTFar = 100.0f;

__host__ __device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
    x = x * 747796405u + 2891336453u;
    uint32_t w = ((x >> ((x >> 28) + 4u)) ^ x) * 277803737u;
    return (w >> 22) ^ w;
}
__device__ __forceinline__ float ray_uniform(uint32_t seed, uint32_t ray) {
    return (float)(hash_u32(hash_u32(seed) ^ (ray * 2u + 1u)) >> 8) * (1.0f / 16777216.0f);
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

// The interval walk + march shared by both passes.  EMIT = false only counts.
template <bool EMIT>
__device__ int march_ray(const uint64_t* hits, int n_hits, const float* o, const float* d, const float* prm, int n_param,
                         const float* __restrict__ minv, const float* __restrict__ dirt, const float* __restrict__ origins,
                         int n_pts, float step_size, uint32_t seed, uint32_t ray, int light_idx, int method, int use_mean,
                         int64_t out0, const MarchOut& out) {
    uint32_t active[kMaxHits];
    int n_act = 0;
    auto toggle = [&](uint32_t inst) -> bool {          // returns true if inst was active (now removed)
        for (int k = 0; k < n_act; ++k) {
            if (active[k] == inst) {
                for (int q = k; q + 1 < n_act; ++q) active[q] = active[q + 1];
                --n_act;
                return true;
            }
        }
        int k = n_act++;                                // sorted insert (ascending instance id, like std::set)
        while (k > 0 && active[k - 1] > inst) { active[k] = active[k - 1]; --k; }
        active[k] = inst;
        return false;
    };
    // :801-827 total in-shell length
    float total = 0.0f, t_entry = 0.0f;
    for (int j = 0; j < n_hits; ++j) {
        const float tt = __uint_as_float((uint32_t)(hits[j] >> 32));
        const bool was_empty = n_act == 0;
        if (toggle((uint32_t)hits[j])) {
            if (n_act == 0) total = __fadd_rn(total, __fsub_rn(tt, t_entry));
        } else if (was_empty) {
            t_entry = tt;
        }
    }
    n_act = 0;
    if (!(total > 0.0f)) return 0;
    const uint32_t necessary = (uint32_t)__fdiv_rn(total, step_size);             // :845
    int n_steps = (int)min(necessary, (uint32_t)n_pts);
    const float u = ray_uniform(seed, ray);
    float t_offset, last_dist;
    bool single = false;
    if (n_steps == 0) {                                                           // :848-851
        single = true; last_dist = total; t_offset = __fmul_rn(u, total); n_steps = 1;
    } else {                                                                      // :852-861
        last_dist = __fsub_rn(__fadd_rn(step_size, total), __fmul_rn((float)n_steps, step_size));
        t_offset = __fmul_rn(u, step_size);
    }
    float dn[3], light[3] = {0.f, 0.f, 0.f};
    if (EMIT) {
        const float nn = norm3(d);
        for (int a = 0; a < 3; ++a) dn[a] = __fdiv_rn(d[a], nn);
        if (light_idx >= 0) {
            const float ln = norm3(prm + light_idx);
            for (int a = 0; a < 3; ++a) light[a] = __fdiv_rn(prm[light_idx + a], ln);
        }
    }
    // :871-1003 march along the concatenated segments
    float segment_offset = 0.0f, cleared = 0.0f;
    t_entry = 0.0f;
    int step = 0;
    for (int j = 0; j < n_hits && step < n_steps; ++j) {
        const float tt = __uint_as_float((uint32_t)(hits[j] >> 32));
        float t_mu = __fadd_rn(__fadd_rn(__fmul_rn((float)step, step_size), t_offset), segment_offset);
        float t_pt = use_mean ? mean_distance(t_mu, step_size) : t_mu;
        while (n_act > 0 && t_pt < tt && step < n_steps) {
            if (EMIT) {
                float pt[3];
                for (int a = 0; a < 3; ++a) pt[a] = __fadd_rn(o[a], __fmul_rn(t_pt, d[a]));
                uint32_t sel = active[0];
                float w = 1.0f;
                if (n_act > 1) {
                    if (method == 0) {                                            // random, :671-677
                        sel = active[pick(seed, ray, (uint32_t)step, (uint32_t)n_act)];
                        w = (float)n_act;
                    } else {                                                      // nearest, :680-692
                        float best = CUDART_INF_F;
                        for (int k = 0; k < n_act; ++k) {
                            const float* og = origins + 3 * active[k];
                            const float df[3] = {__fsub_rn(pt[0], og[0]), __fsub_rn(pt[1], og[1]), __fsub_rn(pt[2], og[2])};
                            const float dist = norm3(df);
                            if (dist < best) { best = dist; sel = active[k]; }
                        }
                    }
                }
                const int64_t k = out0 + step;
                const float* M = minv + 12 * sel;
                const float* Dt = dirt + 9 * sel;
                for (int a = 0; a < 3; ++a) {
                    out.pts[3 * k + a] = __fadd_rn(dot3(M + 4 * a, pt), M[4 * a + 3]);      // getPt :556-558
                    out.dirs[3 * k + a] = dot3(Dt + 3 * a, dn);                           // getDir :561-563
                }
                for (int q = 0; q < n_param; ++q) out.params[k * n_param + q] = prm[q];
                if (light_idx >= 0)
                    for (int a = 0; a < 3; ++a) out.params[k * n_param + light_idx + a] = dot3(Dt + 3 * a, light);
                out.dists[k] = (step == n_steps - 1) ? last_dist : step_size;
                if (single) out.dists[k] = last_dist;
                out.t[k] = t_mu;
                out.weight[k] = w;
                out.inst[k] = (int32_t)sel;
            }
            ++step;
            t_mu = __fadd_rn(__fadd_rn(__fmul_rn((float)step, step_size), t_offset), segment_offset);
            t_pt = use_mean ? mean_distance(t_mu, step_size) : t_mu;
        }
        const bool was_empty = n_act == 0;
        if (toggle((uint32_t)hits[j])) {
            if (n_act == 0) cleared = __fadd_rn(cleared, __fsub_rn(tt, t_entry));
        } else if (was_empty) {
            segment_offset = __fsub_rn(tt, cleared);
            t_entry = tt;
        }
    }
    return step;
}

__global__ void __launch_bounds__(128)
instancer_trace_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d, int R,
                       const float* __restrict__ minv, const float* __restrict__ dirt, const float* __restrict__ origins,
                       int n_inst, float b00, float b01, float b02, float b10, float b11, float b12, float step_size,
                       int n_pts, uint32_t seed, int use_mean, uint64_t* __restrict__ ws_hits, int32_t* __restrict__ ws_nhits,
                       int32_t* __restrict__ n_samples, int32_t* __restrict__ hit_flag) {
    __shared__ float s_minv[kInstTile * 12];
    const int ray = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = ray < R;
    float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f};
    if (live) for (int a = 0; a < 3; ++a) { o[a] = rays_o[3 * ray + a]; d[a] = rays_d[3 * ray + a]; }
    const float b0[3] = {b00, b01, b02}, b1[3] = {b10, b11, b12};
    uint64_t hits[kMaxHits];
    int n_hits = 0;
    auto insert = [&](float t, uint32_t inst) {         // keep the kMaxHits smallest (t, inst), sorted
        const uint64_t key = ((uint64_t)__float_as_uint(t + 0.0f) << 32) | inst;
        int k;
        if (n_hits < kMaxHits) k = n_hits++;
        else if (key < hits[kMaxHits - 1]) k = kMaxHits - 1;
        else return;
        while (k > 0 && hits[k - 1] > key) { hits[k] = hits[k - 1]; --k; }
        hits[k] = key;
    };
    for (int base = 0; base < n_inst; base += kInstTile) {
        const int cnt = min(kInstTile, n_inst - base);
        __syncthreads();
        for (int i = threadIdx.x; i < cnt * 12; i += blockDim.x) s_minv[i] = minv[(size_t)base * 12 + i];
        __syncthreads();
        if (!live) continue;
        for (int j = 0; j < cnt; ++j) {
            const float* M = s_minv + 12 * j;
            float tin = -CUDART_INF_F, tout = CUDART_INF_F;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const float ol = __fadd_rn(dot3(M + 4 * a, o), M[4 * a + 3]);
                const float dl = dot3(M + 4 * a, d);
                const float inv = __fdiv_rn(1.0f, dl);
                const float t0 = __fmul_rn(__fsub_rn(b0[a], ol), inv), t1 = __fmul_rn(__fsub_rn(b1[a], ol), inv);
                const float lo = (t0 < t1) ? t0 : t1, hi = (t0 < t1) ? t1 : t0;
                tin = (a == 0 || lo > tin) ? lo : tin;
                tout = (a == 0 || hi < tout) ? hi : tout;
            }
            if (!(tin < tout) || !(tout >= 0.0f) || !(tin <= kTFar)) continue;
            if (tin >= 0.0f) insert(tin, (uint32_t)(base + j));
            if (tout <= kTFar) insert(tout, (uint32_t)(base + j));
        }
    }
    if (!live) return;
    for (int k = 0; k < n_hits; ++k) ws_hits[(size_t)ray * kMaxHits + k] = hits[k];
    ws_nhits[ray] = n_hits;
    hit_flag[ray] = n_hits > 0;
    MarchOut none = {};
    n_samples[ray] = march_ray<false>(hits, n_hits, o, d, nullptr, 0, minv, dirt, origins, n_pts, step_size, seed,
                                      (uint32_t)ray, -1, 0, use_mean, 0, none);
}

__global__ void __launch_bounds__(128)
instancer_march_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ params,
                       int n_param, int R, const float* __restrict__ minv, const float* __restrict__ dirt,
                       const float* __restrict__ origins, float step_size, int n_pts, uint32_t seed, int light_idx, int method,
                       int use_mean, const uint64_t* __restrict__ ws_hits, const int32_t* __restrict__ ws_nhits,
                       const int64_t* __restrict__ ray_ptr, MarchOut out) {
    const int ray = blockIdx.x * blockDim.x + threadIdx.x;
    if (ray >= R) return;
    const int n_hits = ws_nhits[ray];
    if (n_hits == 0) return;
    uint64_t hits[kMaxHits];
    for (int k = 0; k < n_hits; ++k) hits[k] = ws_hits[(size_t)ray * kMaxHits + k];
    float o[3], d[3], prm[16];
    for (int a = 0; a < 3; ++a) { o[a] = rays_o[3 * ray + a]; d[a] = rays_d[3 * ray + a]; }
    for (int q = 0; q < n_param; ++q) prm[q] = params[(size_t)ray * n_param + q];
    march_ray<true>(hits, n_hits, o, d, prm, n_param, minv, dirt, origins, n_pts, step_size, seed, (uint32_t)ray, light_idx,
                    method, use_mean, ray_ptr[ray], out);
}

}  // namespace

NTX_API int64_t ntx_instancer_workspace_bytes(int R) {
    if (R < 0) return -1;
    return (int64_t)R * kMaxHits * 8 + (int64_t)R * 4;
}

NTX_API int ntx_instancer_trace(const float* rays_o, const float* rays_d, int R, const float* minv, const float* dirt,
                                const float* origins, int n_inst, const float* b_0, const float* b_1, float step_size,
                                int n_pts, uint32_t seed, int use_mean_distance, void* workspace, int32_t* n_samples,
                                int32_t* hit, void* stream) {
    NTX_REQUIRE(R >= 0 && n_inst >= 0 && n_pts >= 1 && step_size > 0.0f, "instancer_trace: bad arguments");
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(rays_o && rays_d && minv && dirt && origins && b_0 && b_1 && workspace && n_samples && hit,
                "instancer_trace: null pointer");
    uint64_t* ws_hits = (uint64_t*)workspace;
    int32_t* ws_nhits = (int32_t*)((uint8_t*)workspace + (int64_t)R * kMaxHits * 8);
    instancer_trace_kernel<<<(R + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        rays_o, rays_d, R, minv, dirt, origins, n_inst, b_0[0], b_0[1], b_0[2], b_1[0], b_1[1], b_1[2], step_size, n_pts, seed,
        use_mean_distance, ws_hits, ws_nhits, n_samples, hit);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_instancer_march(const float* rays_o, const float* rays_d, const float* parameters, int n_param, int R,
                                const float* minv, const float* dirt, const float* origins, float step_size, int n_pts,
                                uint32_t seed, int light_dir_idx, int method, int use_mean_distance, const void* workspace,
                                const int64_t* ray_ptr, float* pts, float* dirs, float* params_out, float* dists, float* t,
                                float* alpha_weight, int32_t* instance_id, void* stream) {
    NTX_REQUIRE(R >= 0 && n_param >= 0 && n_param <= 16, "instancer_march: bad arguments (at most 16 parameters)");
    NTX_REQUIRE(light_dir_idx < 0 || light_dir_idx + 3 <= n_param, "instancer_march: light direction index out of range");
    NTX_REQUIRE(method == 0 || method == 1, "instancer_march: instance sampling method must be 0 (random) or 1 (nearest)");
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(rays_o && rays_d && (n_param == 0 || parameters) && minv && dirt && origins && workspace && ray_ptr,
                "instancer_march: null pointer");
    const uint64_t* ws_hits = (const uint64_t*)workspace;
    const int32_t* ws_nhits = (const int32_t*)((const uint8_t*)workspace + (int64_t)R * kMaxHits * 8);
    MarchOut out = {pts, dirs, params_out, dists, t, alpha_weight, instance_id};
    instancer_march_kernel<<<(R + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        rays_o, rays_d, parameters, n_param, R, minv, dirt, origins, step_size, n_pts, seed, light_dir_idx, method,
        use_mean_distance, ws_hits, ws_nhits, ray_ptr, out);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

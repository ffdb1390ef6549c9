// Cone-segment gaussians of the mip renderers (network/renderer.py:411-434 and :570-587): the mean and diagonal
// covariance that the integrated positional encoding of the fused MLP kernel (n_pos = 6) consumes.  Elementwise,
// HBM-bound: one thread per sample.
#include "ntx_common.cuh"

namespace {

struct SegVar {
    float t_mean, t_var, r_var;
};

// mu: segment centre, hw: half width, radius: cone radius at unit distance   (renderer.py:418-422 / :574-578)
__device__ __forceinline__ SegVar segment_moments(float mu, float hw, float radius) {
    const float mu2 = mu * mu, hw2 = hw * hw, hw4 = hw2 * hw2;
    const float den = 3.0f * mu2 + hw2;
    SegVar v;
    v.t_mean = mu + (2.0f * mu * hw2) / den;
    v.t_var = hw2 / 3.0f - (4.0f / 15.0f) * ((hw4 * (12.0f * mu2 - hw2)) / (den * den));
    v.r_var = radius * radius * (mu2 / 4.0f + (5.0f / 12.0f) * hw2 - (4.0f / 15.0f) * hw4 / den);
    return v;
}

// diagonal of t_var * d d^T + r_var * (I - d d^T / |d|^2)   (renderer.py:427-432)
__device__ __forceinline__ void lift_cov(const float d[3], float t_var, float r_var, float out[3]) {
    const float mag = fmaxf(1e-10f, d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float dd = d[a] * d[a];
        out[a] = t_var * dd + r_var * (1.0f - dd / mag);
    }
}

__global__ void cone_gaussians_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                      const float* __restrict__ z_edges, const float* __restrict__ radii, int64_t total, int S,
                                      float* __restrict__ mean, float* __restrict__ cov) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int64_t r = idx / S;
    const int i = (int)(idx - r * S);
    const float t0 = __ldg(z_edges + r * (S + 1) + i), t1 = __ldg(z_edges + r * (S + 1) + i + 1);
    const SegVar v = segment_moments((t0 + t1) / 2.0f, (t1 - t0) / 2.0f, __ldg(radii + r));
    float d[3], c[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) d[a] = __ldg(rays_d + 3 * r + a);
    lift_cov(d, v.t_var, v.r_var, c);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        mean[3 * idx + a] = __ldg(rays_o + 3 * r + a) + d[a] * v.t_mean;      // renderer.py:425
        cov[3 * idx + a] = c[a];
    }
}

__global__ void cone_gaussians_inst_kernel(const float* __restrict__ dirs, const float* __restrict__ t,
                                           const float* __restrict__ radii, const float* __restrict__ dists, int64_t total,
                                           float* __restrict__ cov) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const SegVar v = segment_moments(__ldg(t + idx), __ldg(dists + idx), __ldg(radii + idx));   // mu = t, hw = dists (:574-575)
    float d[3], c[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) d[a] = __ldg(dirs + 3 * idx + a);
    lift_cov(d, v.t_var, v.r_var, c);
#pragma unroll
    for (int a = 0; a < 3; ++a) cov[3 * idx + a] = c[a];
}

}  // namespace

NTX_API int ntx_mip_cone_gaussians(const float* rays_o, const float* rays_d, const float* z_edges, const float* radii, int R,
                                   int S, float* mean, float* cov, void* stream) {
    NTX_REQUIRE(R >= 0 && S >= 1, "mip_cone_gaussians: bad shape R=%d S=%d", R, S);
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(rays_o && rays_d && z_edges && radii && mean && cov, "mip_cone_gaussians: null pointer");
    const int64_t total = (int64_t)R * S;
    cone_gaussians_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rays_o, rays_d, z_edges, radii,
                                                                                            total, S, mean, cov);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_mip_cone_gaussians_instanced(const float* dirs, const float* t, const float* radii, const float* dists,
                                             int64_t T, float* cov, void* stream) {
    NTX_REQUIRE(T >= 0, "mip_cone_gaussians_instanced: bad size");
    if (T == 0) return NTX_OK;
    NTX_REQUIRE(dirs && t && radii && dists && cov, "mip_cone_gaussians_instanced: null pointer");
    cone_gaussians_inst_kernel<<<(unsigned)((T + 255) / 256), 256, 0, (cudaStream_t)stream>>>(dirs, t, radii, dists, T, cov);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

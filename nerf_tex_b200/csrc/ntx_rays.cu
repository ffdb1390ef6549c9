// Camera-ray generation on the device (row f3 of the scope table): pixel -> (origin, direction, [near, far],
// cone scale) for a whole frame or an explicit pixel list, with the optional AABB proxy, in one launch.
// Reference: network/ray_sampler.py:6-48 (Frustum, Proxy, rays_from_camera), network/proxy.py:13-35 (AABB).
// Sums are formed left to right with explicit round-to-nearest ops, like oracle/nerftex_oracle.py.
#include <math_constants.h>

#include "ntx_common.cuh"

namespace {

struct RayGenArgs {
    float m[9];          // c2w[:3,:3] row-major
    float o[3];          // c2w[:3,3]
    float b0[3], b1[3];
    float focal, near_, far_;
    int height, width, normalize, use_aabb;
};

__global__ void rays_from_camera_kernel(const float* __restrict__ loc, int64_t n, const __grid_constant__ RayGenArgs a,
                                        float* __restrict__ rays_o, float* __restrict__ rays_d, float* __restrict__ t,
                                        float* __restrict__ cone_scale) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    float li, lj;                                     // image_plane_loc[:,0] (row), [:,1] (column)
    if (loc) { li = loc[2 * idx]; lj = loc[2 * idx + 1]; }
    else { li = (float)(idx / a.width); lj = (float)(idx % a.width); }
    // ray_sampler.py:41
    const float dx = __fdiv_rn(__fsub_rn(__fadd_rn(lj, 0.5f), __fmul_rn(0.5f, (float)a.width)), a.focal);
    const float dy = -__fdiv_rn(__fsub_rn(__fadd_rn(li, 0.5f), __fmul_rn(0.5f, (float)a.height)), a.focal);
    const float dz = -1.0f;
    float d[3];
#pragma unroll
    for (int r = 0; r < 3; ++r)                       // :42  sum_k dirs[k] * c2w[r, k]
        d[r] = __fadd_rn(__fadd_rn(__fadd_rn(0.0f, __fmul_rn(dx, a.m[3 * r])), __fmul_rn(dy, a.m[3 * r + 1])),
                         __fmul_rn(dz, a.m[3 * r + 2]));
    // :46  cos(atan(|dirs_xy|)) / |dirs| / focal
    const float nxy = sqrtf(__fadd_rn(__fadd_rn(0.0f, __fmul_rn(dx, dx)), __fmul_rn(dy, dy)));
    const float nall = sqrtf(__fadd_rn(__fadd_rn(__fadd_rn(0.0f, __fmul_rn(dx, dx)), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    cone_scale[idx] = __fdiv_rn(__fdiv_rn(cosf(atanf(nxy)), nall), a.focal);
    if (a.normalize) {                                // ray_sampler.py:34
        const float nn = sqrtf(__fadd_rn(__fadd_rn(__fadd_rn(0.0f, __fmul_rn(d[0], d[0])), __fmul_rn(d[1], d[1])), __fmul_rn(d[2], d[2])));
#pragma unroll
        for (int r = 0; r < 3; ++r) d[r] = __fdiv_rn(d[r], nn);
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) { rays_o[3 * idx + r] = a.o[r]; rays_d[3 * idx + r] = d[r]; }
    if (!a.use_aabb) {                                // Frustum: constant near / far (ray_sampler.py:19)
        t[2 * idx] = a.near_; t[2 * idx + 1] = a.far_;
        return;
    }
    // proxy.py:13-35: slab test; misses give (inf, inf)
    float tn = -CUDART_INF_F, tf = CUDART_INF_F;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float inv = __fdiv_rn(1.0f, d[r]);
        const float t0 = __fmul_rn(__fsub_rn(a.b0[r], a.o[r]), inv), t1 = __fmul_rn(__fsub_rn(a.b1[r], a.o[r]), inv);
        const float lo = (t0 < t1) ? t0 : t1;         // tf.where(t_0 < t_1, t_0, t_1)
        const float hi = (t0 > t1) ? t0 : t1;         // tf.where(t_0 > t_1, t_0, t_1)
        tn = (r == 0) ? lo : fmaxf(tn, lo);
        tf = (r == 0) ? hi : fminf(tf, hi);
    }
    const bool hit = tn < tf;
    t[2 * idx] = hit ? tn : CUDART_INF_F;
    t[2 * idx + 1] = hit ? tf : CUDART_INF_F;
}

}  // namespace

NTX_API int ntx_rays_from_camera(const float* image_plane_loc, int64_t n, int height, int width, float focal,
                                 const float* c2w, int normalize, const float* b0, const float* b1, float near_,
                                 float far_, float* rays_o, float* rays_d, float* t, float* cone_scale, void* stream) {
    NTX_REQUIRE(n >= 0 && height > 0 && width > 0 && focal > 0.0f, "rays_from_camera: bad frame");
    if (n == 0) return NTX_OK;
    NTX_REQUIRE(c2w && rays_o && rays_d && t && cone_scale, "rays_from_camera: null pointer");
    NTX_REQUIRE((b0 == nullptr) == (b1 == nullptr), "rays_from_camera: give both box corners or neither");
    NTX_REQUIRE(image_plane_loc || n <= (int64_t)height * width, "rays_from_camera: full-frame mode with n > H*W");
    RayGenArgs a;
    for (int r = 0; r < 3; ++r) {
        for (int k = 0; k < 3; ++k) a.m[3 * r + k] = c2w[4 * r + k];
        a.o[r] = c2w[4 * r + 3];
        a.b0[r] = b0 ? b0[r] : 0.0f;
        a.b1[r] = b1 ? b1[r] : 0.0f;
    }
    a.focal = focal; a.near_ = near_; a.far_ = far_;
    a.height = height; a.width = width; a.normalize = normalize; a.use_aabb = b0 != nullptr;
    const int threads = 256;
    rays_from_camera_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        image_plane_loc, n, a, rays_o, rays_d, t, cone_scale);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

// One C call for a whole ray chunk of Renderer.render_rays (network/renderer.py:92-143):
// normalise view directions -> stratified depths (K3) -> fused MLP (K1) -> compositing (K2)
// [-> hierarchical resampling (K3) -> fine MLP (K1) -> compositing (K2)].
// Intermediates (depths, raw network outputs, weights) live in a caller-provided scratch buffer.
#include <mutex>

#include "ntx_common.cuh"

// Scratch of ntx_render_rays_static: lives in the library image, so every GPU that loads the
// library carries NTX_STATIC_SCRATCH_BYTES of it until the process exits.  In host code this name
// is only a handle for cudaGetSymbolAddress — never a pointer a kernel may be given.
static __device__ __align__(256) unsigned char g_render_scratch[NTX_STATIC_SCRATCH_BYTES];

namespace {

__global__ void normalize_dirs_kernel(const float* __restrict__ d, int R, float* __restrict__ out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const float x = d[3 * r], y = d[3 * r + 1], z = d[3 * r + 2];
    const float n = sqrtf(x * x + y * y + z * z);          // renderer.py:98
    out[3 * r] = x / n; out[3 * r + 1] = y / n; out[3 * r + 2] = z / n;
}

inline int64_t align256(int64_t v) { return (v + 255) & ~int64_t(255); }

struct Scratch {
    int64_t dir_n, z, rgbs, weights, z_fine, rgbs_fine, total;
};

Scratch plan(int64_t R, int64_t S, int64_t N) {
    Scratch s;
    int64_t o = 0;
    s.dir_n = o; o = align256(o + R * 3 * 4);
    s.z = o; o = align256(o + R * S * 4);
    s.rgbs = o; o = align256(o + R * S * 16);
    s.weights = o; o = align256(o + R * S * 4);
    s.z_fine = o; o = align256(o + (N > 0 ? R * (S + N) * 4 : 0));
    s.rgbs_fine = o; o = align256(o + (N > 0 ? R * (S + N) * 16 : 0));
    s.total = o;
    return s;
}

}  // namespace

NTX_API int ntx_normalize_dirs(const float* rays_d, int R, float* out, void* stream) {
    NTX_REQUIRE(R >= 0, "normalize_dirs: bad R=%d", R);
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(rays_d && out, "normalize_dirs: null pointer");
    normalize_dirs_kernel<<<(R + 255) / 256, 256, 0, (cudaStream_t)stream>>>(rays_d, R, out);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int64_t ntx_render_rays_workspace_bytes(int R, int n_samples, int n_importance) {
    if (R < 0 || n_samples < 1 || n_importance < 0) return -1;
    return plan(R, n_samples, n_importance).total;
}

// shared by the public entry points of this file
int ntx_render_rays_impl(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_t* mlp_fine, const void* packed_fine,
                         const ntx_render_config* cfg, const float* rays_o, const float* rays_d, const float* t,
                         const float* parameters, const float* cone_scale, const float* z_rand, const float* u,
                         const float* noise, const float* noise_fine, int R, float* color_pred, float* alpha_pred,
                         float* color_pred_coarse, float* alpha_pred_coarse, uint8_t* ws, void* stream) {
    const int S = cfg->n_samples, N = cfg->n_importance;
    const Scratch sc = plan(R, S, N);
    float* dir_n = (float*)(ws + sc.dir_n);
    float* z = (float*)(ws + sc.z);
    float* rgbs = (float*)(ws + sc.rgbs);
    float* weights = (float*)(ws + sc.weights);

    int rc = ntx_normalize_dirs(rays_d, R, dir_n, stream);                                      // :98
    if (rc) return rc;
    rc = ntx_stratified_z(t, z_rand, R, S, cfg->perturb, z, stream);                       // :101-111
    if (rc) return rc;

    ntx_mlp_inputs in;
    memset(&in, 0, sizeof(in));
    in.rays_o = rays_o; in.rays_d = rays_d; in.z = z; in.dir = dir_n; in.params = parameters;
    in.blur_idx = cfg->blur_idx;
    if (cfg->blur_idx >= 0) { in.row_scale = z; in.group_scale = cone_scale; }                  // :155-158
    in.rows_per_group = S;
    in.n_rows = (int64_t)R * S;
    rc = ntx_mlp_forward(mlp, packed, &in, rgbs, stream);                                       // :117
    if (rc) return rc;

    const bool fine = N > 0;
    float* c_out = fine ? color_pred_coarse : color_pred;
    float* a_out = fine ? alpha_pred_coarse : alpha_pred;
    rc = ntx_composite_forward(rgbs, nullptr, 1, z, rays_d, noise, R, S, cfg->map_exr, cfg->composite_bkgd,
                               cfg->bkgd_color, c_out, a_out, fine ? weights : nullptr, nullptr, stream);   // :120
    if (rc || !fine) return rc;

    float* z_fine = (float*)(ws + sc.z_fine);
    float* rgbs_fine = (float*)(ws + sc.rgbs_fine);
    rc = ntx_hierarchical_z(z, weights, u, R, S, N, cfg->perturb /* det = self.perturb, :128 */, z_fine, nullptr,
                            nullptr, stream);
    if (rc) return rc;
    in.z = z_fine;
    if (cfg->blur_idx >= 0) in.row_scale = z_fine;
    in.rows_per_group = S + N;
    in.n_rows = (int64_t)R * (S + N);
    const ntx_mlp_t* m2 = mlp_fine ? mlp_fine : mlp;                                            // :134
    const void* p2 = mlp_fine ? packed_fine : packed;
    rc = ntx_mlp_forward(m2, p2, &in, rgbs_fine, stream);
    if (rc) return rc;
    return ntx_composite_forward(rgbs_fine, nullptr, 1, z_fine, rays_d, noise_fine, R, S + N, cfg->map_exr,
                                 cfg->composite_bkgd, cfg->bkgd_color, color_pred, alpha_pred, nullptr, nullptr, stream);
}

int ntx_render_rays_check(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_t* mlp_fine, const void* packed_fine,
                          const ntx_render_config* cfg, const float* rays_o, const float* rays_d, const float* t,
                          const float* cone_scale, const float* z_rand, const float* u, int R, float* color_pred,
                          float* alpha_pred, float* color_pred_coarse, float* alpha_pred_coarse) {
    NTX_REQUIRE(mlp && packed && cfg, "render_rays: null model / config");
    NTX_REQUIRE(R >= 0 && cfg->n_samples >= 2 && cfg->n_importance >= 0, "render_rays: bad shape R=%d S=%d N=%d", R,
                cfg->n_samples, cfg->n_importance);
    NTX_REQUIRE(!mlp_fine || packed_fine, "render_rays: fine model without packed weights");
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(rays_o && rays_d && t && color_pred && alpha_pred, "render_rays: null pointer");
    NTX_REQUIRE(!cfg->perturb || z_rand, "render_rays: perturb requires z_rand");
    NTX_REQUIRE(cfg->blur_idx < 0 || cone_scale, "render_rays: blur_idx = %d scales that parameter by cone_scale * z "
                "(renderer.py:155-158): pass cone_scale[R]", cfg->blur_idx);
    if (cfg->n_importance > 0) {
        NTX_REQUIRE(color_pred_coarse && alpha_pred_coarse, "render_rays: n_importance > 0 needs the coarse outputs");
        NTX_REQUIRE(cfg->perturb || u, "render_rays: perturb == 0 draws random u (det = perturb): pass u");
        NTX_REQUIRE(cfg->n_samples >= 3 && cfg->n_samples + cfg->n_importance <= 1024, "render_rays: S + N <= 1024");
    }
    return NTX_OK;
}

NTX_API int ntx_render_rays(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_t* mlp_fine, const void* packed_fine,
                            const ntx_render_config* cfg, const float* rays_o, const float* rays_d, const float* t,
                            const float* parameters, const float* cone_scale, const float* z_rand, const float* u,
                            const float* noise, const float* noise_fine, int R, float* color_pred, float* alpha_pred,
                            float* color_pred_coarse, float* alpha_pred_coarse, void* workspace,
                            int64_t workspace_bytes, void* stream) {
    int rc = ntx_render_rays_check(mlp, packed, mlp_fine, packed_fine, cfg, rays_o, rays_d, t, cone_scale, z_rand, u, R,
                                   color_pred, alpha_pred, color_pred_coarse, alpha_pred_coarse);
    if (rc || R == 0) return rc;
    const int64_t need = ntx_render_rays_workspace_bytes(R, cfg->n_samples, cfg->n_importance);
    NTX_REQUIRE(workspace && workspace_bytes >= need, "render_rays: workspace of %lld B needed, %lld B given",
                (long long)need, (long long)workspace_bytes);
    NTX_REQUIRE(((uintptr_t)workspace & 255) == 0, "render_rays: workspace must be 256-byte aligned");
    return ntx_render_rays_impl(mlp, packed, mlp_fine, packed_fine, cfg, rays_o, rays_d, t, parameters, cone_scale,
                                z_rand, u, noise, noise_fine, R, color_pred, alpha_pred, color_pred_coarse,
                                alpha_pred_coarse, (uint8_t*)workspace, stream);
}

// Device address of g_render_scratch on the current device, looked up once per device.
static int static_scratch_address(uint8_t** out) {
    static std::mutex mu;
    static uint8_t* per_device[64] = {};
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    NTX_REQUIRE(dev >= 0 && dev < 64, "render_rays_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!per_device[dev]) {
        void* p = nullptr;
        NTX_CUDA(cudaGetSymbolAddress(&p, g_render_scratch));
        NTX_REQUIRE(p != nullptr, "render_rays_static: the scratch symbol has no address on device %d", dev);
        per_device[dev] = (uint8_t*)p;
    }
    *out = per_device[dev];
    return NTX_OK;
}

// ntx_render_rays for hosts that must not (or cannot) allocate device memory; see nerftex_b200.h.
NTX_API int ntx_render_rays_static(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_t* mlp_fine,
                                   const void* packed_fine, const ntx_render_config* cfg, const float* rays_o,
                                   const float* rays_d, const float* t, const float* parameters,
                                   const float* cone_scale, const float* z_rand, const float* u, const float* noise,
                                   const float* noise_fine, int R, float* color_pred, float* alpha_pred,
                                   float* color_pred_coarse, float* alpha_pred_coarse, void* stream) {
    int rc = ntx_render_rays_check(mlp, packed, mlp_fine, packed_fine, cfg, rays_o, rays_d, t, cone_scale, z_rand, u, R,
                                   color_pred, alpha_pred, color_pred_coarse, alpha_pred_coarse);
    if (rc || R == 0) return rc;
    const int64_t need = ntx_render_rays_workspace_bytes(R, cfg->n_samples, cfg->n_importance);
    NTX_REQUIRE(need <= (int64_t)NTX_STATIC_SCRATCH_BYTES,
                "render_rays_static: %d rays x (%d + %d) samples need %lld B of scratch but the built-in buffer holds "
                "%lld B; split the chunk or use ntx_render_rays with your own workspace",
                R, cfg->n_samples, cfg->n_importance, (long long)need, (long long)NTX_STATIC_SCRATCH_BYTES);
    uint8_t* ws = nullptr;
    rc = static_scratch_address(&ws);
    if (rc) return rc;
    // Nothing in the scratch has to start at zero: every region (view dirs, depths, raw outputs,
    // weights, merged depths) is written in full by an earlier kernel of this call before a later
    // one reads it, so whatever the previous call left behind is never observed.
    return ntx_render_rays_impl(mlp, packed, mlp_fine, packed_fine, cfg, rays_o, rays_d, t, parameters, cone_scale,
                                z_rand, u, noise, noise_fine, R, color_pred, alpha_pred, color_pred_coarse,
                                alpha_pred_coarse, ws, stream);
}

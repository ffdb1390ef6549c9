/* nerftex_b200.h — C ABI of libnerftex_b200.so (sm_100a).
 *
 * Drop-in boundary for the NeRF-Tex volume-rendering hot path.  The reference has no native
 * FFI for this path (its arithmetic is TensorFlow eager ops); each entry point below therefore
 * cites the reference *Python* call site it replaces (paths relative to hbaatz/nerf-tex).
 * The Python host mirror in nerf_tex_b200/ binds these with ctypes; INTEGRATION.md shows the
 * stub a reference maintainer would add.
 *
 * Conventions
 *   - every entry returns int: 0 = ok, <0 = error (NTX_E_*); ntx_last_error() gives the
 *     message of the calling thread's last failure.
 *   - all pointers are DEVICE pointers to contiguous row-major fp32 / int32 / int64 data unless
 *     the comment says "host".  The caller owns every buffer (inputs, outputs, workspace).
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it and never
 *     synchronise.  No entry allocates or frees device memory.
 *   - there is no CPU fallback: on a machine without a CUDA device every compute entry fails
 *     with NTX_E_CUDA.
 */
#ifndef NERFTEX_B200_H_
#define NERFTEX_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NTX_VERSION 100 /* 0.1.0 */

enum {
    NTX_OK = 0,
    NTX_E_INVALID = -1,     /* bad argument (null pointer, size out of range, unsupported config) */
    NTX_E_CUDA = -2,        /* CUDA runtime error (message in ntx_last_error) */
    NTX_E_UNSUPPORTED = -3, /* configuration outside what the sm_100a kernels implement */
    NTX_E_NCCL = -4
};

int ntx_version(void);
const char* ntx_last_error(void);

/* ------------------------------------------------------------------------------------------
 * K3 — sampling
 * ---------------------------------------------------------------------------------------- */

/* Stratified depths.  Replaces network/renderer.py:101-111 (Renderer.render_rays).
 *   t[R,2] near/far, z_rand[R,S] uniform draws (required iff perturb), z[R,S] out.
 * Bit-exact against the oracle's defined op order. */
int ntx_stratified_z(const float* t, const float* z_rand, int R, int S, int perturb, float* z, void* stream);

/* Inverse-CDF sampling.  Replaces network/renderer.py:589-618 (sample_pdf).
 *   bins[R,C], weights[R,C-1]; u[R,N] uniform draws, or NULL when det (u = linspace(0,1,N));
 *   samples[R,N] out; inds[R,N] out (searchsorted indices, may be NULL).  C <= 1025. */
int ntx_sample_pdf(const float* bins, const float* weights, const float* u, int R, int C, int N, int det,
                   float* samples, int32_t* inds, void* stream);

/* Hierarchical resampling.  Replaces network/renderer.py:127-130: z_mid, sample_pdf(z_mid,
 * weights[...,1:-1], N, det), sort(concat(z, z_samples)).
 *   z[R,S], weights[R,S], u[R,N] or NULL; z_out[R,S+N]; z_samples[R,N], inds[R,N] optional.
 *   S + N <= 1024. */
int ntx_hierarchical_z(const float* z, const float* weights, const float* u, int R, int S, int N, int det,
                       float* z_out, float* z_samples, int32_t* inds, void* stream);

/* ------------------------------------------------------------------------------------------
 * K2 — alpha compositing
 * ---------------------------------------------------------------------------------------- */

/* Replaces network/renderer.py:170-213 (Renderer.map_model_output).
 *   packed != 0: `color` is rgbs[R,S,4] (raw r,g,b,sigma per sample; the layout ntx_mlp_forward
 *                writes) and `sigma` is ignored.
 *   packed == 0: color[R,S,3], sigma[R,S] (the reference's two model outputs).
 *   z[R,S]; rays_d[R,3] (un-normalised; its norm scales dists); noise[R,S] already multiplied
 *   by raw_noise_std, or NULL; bkgd_color: HOST pointer to 3 floats (used iff composite_bkgd).
 *   Outputs: color_map[R,3], alpha_map[R], weights[R,S] (NULL to skip), depth[R] (NULL to
 *   skip; the reference computes and discards it, :204). */
int ntx_composite_forward(const float* color, const float* sigma, int packed, const float* z, const float* rays_d,
                          const float* noise, int R, int S, int map_exr, int composite_bkgd,
                          const float* bkgd_color, float* color_map, float* alpha_map, float* weights,
                          float* depth, void* stream);

/* Backward of ntx_composite_forward w.r.t. the raw model outputs (the reference obtains it from
 * tf.GradientTape, network/train.py:62-66).  d_color_map[R,3], d_alpha_map[R] in;
 * d_rgbs[R,S,4] out (packed raw r,g,b,sigma gradients).  Recomputes the forward per ray. */
int ntx_composite_backward(const float* rgbs, const float* z, const float* rays_d, const float* noise, int R,
                           int S, int map_exr, int composite_bkgd, const float* bkgd_color,
                           const float* d_color_map, const float* d_alpha_map, float* d_rgbs, void* stream);

/* Instanced compositing over CSR-packed in-shell samples.  Replaces network/renderer.py:300
 * (density reweighting) and :318-354 (InstanceRenderer.map_model_output): the mesh hit
 * (color_last, alpha_last) is appended as the final sample of every ray.
 *   rgbs[T,4] raw model outputs of the T packed samples; dists[T]; alpha_weight[T] or NULL;
 *   ray_ptr[H+1] int64 CSR offsets; color_last[H,3]; alpha_last[H];
 *   ray_index[H] int32 destination row of each hit ray in the outputs, or NULL (identity);
 *   instance_color[T,3] or NULL (false_color rendering, :306-307, :322-323);
 *   color_map[n_out,3], alpha_map[n_out] are written only at hit rows (caller pre-zeroes, :313-314). */
int ntx_composite_instanced_forward(const float* rgbs, const float* dists, const float* alpha_weight,
                                    const int64_t* ray_ptr, const float* color_last, const float* alpha_last,
                                    const int32_t* ray_index, const float* instance_color, int H,
                                    float patch_scale, float density_scale, int map_exr, int composite_bkgd,
                                    const float* bkgd_color, float* color_map, float* alpha_map, void* stream);

/* ------------------------------------------------------------------------------------------
 * K1 — fused positional encoding + conditioned MLP (bf16 tcgen05 / TMEM, fp32 accumulate)
 * ---------------------------------------------------------------------------------------- */

/* Architecture of network/model.py:58-125 (ParamNerf; Nerf :9-45 is n_geo = n_app = 0,
 * color_depth = 0).  Supported by the sm_100a kernel: n_pos = 3 (or 6, see below), pos_freqs = 10, dir_freqs = 4,
 * param_freqs = 4, width = 256, param_depth = 0, n_geo <= 3, n_app <= 7, depth <= 12. */
typedef struct ntx_mlp_config {
    int32_t n_pos;       /* 3; 6 = IntegratedPositionalEncoding of (mean, diagonal covariance), network/layer.py:25-41 */
    int32_t pos_freqs;   /* 10 */
    int32_t dir_freqs;   /* 4 */
    int32_t param_freqs; /* 4 */
    int32_t n_geo;       /* n_parameters[0] */
    int32_t n_app;       /* n_parameters[1] */
    int32_t depth;       /* 8 */
    int32_t width;       /* 256 */
    int32_t skip_mask;   /* bit i set <=> i in skips (default 1<<4) */
    int32_t color_depth; /* 1 */
} ntx_mlp_config;

typedef struct ntx_mlp ntx_mlp_t; /* opaque host-side handle: layer table only, owns no device memory */

int ntx_mlp_create(const ntx_mlp_config* cfg, ntx_mlp_t** out);
int ntx_mlp_destroy(ntx_mlp_t* mlp);

/* Number of fp32 elements of the flat master weight buffer: for every Dense layer in graph
 * order (trunk0..trunk{depth-1}, alpha, feature, colorhid0.., colorhalf, rgb) the Keras kernel
 * W[in,out] row-major followed by the bias b[out].  Rows of W follow the reference's concat
 * order ([pos_map, h] after a skip, [dir_map, feature]; network/model.py:108,115). */
int64_t ntx_mlp_num_params(const ntx_mlp_t* mlp);
/* Offset (in floats) of layer `i`'s kernel inside the flat buffer; fan_in/fan_out returned. */
int ntx_mlp_num_layers(const ntx_mlp_t* mlp);
int ntx_mlp_layer_info(const ntx_mlp_t* mlp, int i, int64_t* w_offset, int64_t* b_offset, int32_t* fan_in,
                       int32_t* fan_out);
/* Bytes of the packed bf16 weight image consumed by ntx_mlp_forward. */
int64_t ntx_mlp_packed_bytes(const ntx_mlp_t* mlp);
/* fp32 flat master (device) -> packed bf16 UMMA-canonical K-slices + fp32 head block (device). */
int ntx_mlp_pack_weights(const ntx_mlp_t* mlp, const float* master, void* packed, void* stream);

/* Model inputs.  Row r of the batch belongs to group g = r / rows_per_group (rows_per_group = S
 * for the plain renderer where view dir and parameters are per ray, network/renderer.py:152-154;
 * 1 for the instanced renderer where they are per sample, :278-280). */
typedef struct ntx_mlp_inputs {
    const float* pos;        /* [N,3] sample positions, or NULL to use rays_o + rays_d * z */
    const float* rays_o;     /* [G,3] (used when pos == NULL; network/renderer.py:114) */
    const float* rays_d;     /* [G,3] */
    const float* z;          /* [N]   */
    const float* dir;        /* [G,3] view direction fed to the network (already normalised) */
    const float* params;     /* [G,n_geo+n_app] or NULL when the model has no parameters */
    const float* row_scale;  /* [N] or NULL: blur multiplier per row (z, :156) */
    const float* group_scale;/* [G] or NULL: blur multiplier per group (cone_scale, :156) */
    int32_t blur_idx;        /* parameter index scaled by row_scale*group_scale, or -1 (:155-158) */
    int32_t rows_per_group;
    int64_t n_rows;          /* N */
    const float* pos_cov;    /* [N,3] diagonal covariance of pos: required by (and only read for) n_pos = 6 models */
} ntx_mlp_inputs;

/* Forward.  Replaces the Keras model call of network/renderer.py:160-161 / :290-291 including
 * the FourierFeatures of network/layer.py:11-23.  rgbs[N,4] out: raw colour (pre-sigmoid) and
 * raw density (pre-relu), network/model.py:111,123. */
int ntx_mlp_forward(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_inputs* in, float* rgbs, void* stream);

/* Training forward: ntx_mlp_forward that also saves, in bf16, every tensor-core operand the
 * backward pass needs (the reference keeps them on the tf.GradientTape, network/train.py:62-64):
 *   enc_pos[n_tiles][32 KB],      encoded inputs as tile images of 4 blocks (128 columns) in the kernel's
 *   enc_dir[n_tiles][32 KB]       column order (ntx_mlp_enc_columns), constant-1 bias column included;
 *                                 blocks 0..2 are written, block 3 must be zero-filled by the caller once;
 *   acts[n_steps][N,256]          (optional, may be NULL) stored output of every step, row-major
 *                                 (post-activation; the 128-wide last step fills columns 0..127);
 *   acts_img[n_steps][n_tiles][64 KB]  (optional) the same outputs as 128-row tile images: exactly the
 *                                 SWIZZLE_64B operand layout the kernels keep in shared memory
 *                                 (32-column blocks of [128 rows x 64 B], 16-byte chunk index XOR
 *                                 (row >> 1) & 3), written with bulk stores; consumed as-is by
 *                                 ntx_mlp_backward_*; ntx_untile converts to row-major;
 *   masks[n_steps][n_tiles*128][8] (with acts_img) u32 words: bit c of word j set <=> output column
 *                                 32 j + c of that row is > 0 (the ReLU derivative).
 * n_tiles >= ntx_mlp_train_tiles(N) (rows are processed in whole 512-row CTA-pair batches). */
int64_t ntx_mlp_train_tiles(int64_t n_rows);
int ntx_mlp_forward_train(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_inputs* in, float* rgbs,
                          void* enc_pos, void* enc_dir, void* acts, void* acts_img, void* masks, int64_t n_tiles,
                          void* stream);
/* Tile images -> row-major bf16: out[n_rows, cols] from img[n_tiles][tile_bytes] (cols <= 256, multiple of 32;
 * tile_bytes = 65536 for activation / gradient images, 32768 for encoding images). */
int ntx_untile(const void* img, int64_t n_rows, int cols, int64_t tile_bytes, void* out, void* stream);

/* Backward of the MLP on the saved operands (the reference gets it from tape.gradient,
 * network/train.py:66).  Two kernels:
 *   ntx_mlp_backward_data    fused data-gradient chain: from d_rgbs[N,4] (gradient w.r.t. the raw outputs)
 *                            and the ReLU masks it produces dy_img[n_steps][n_tiles][64 KB], the gradient
 *                            w.r.t. every step's pre-activation output, as tile images.  packed_t is the
 *                            transposed packed weight image (ntx_mlp_packed_t_bytes / ntx_mlp_pack_weights_t).
 *   ntx_mlp_backward_weights adds into the flat fp32 gradient buffer `grad` (layout of the master buffer; the caller
 *                            zeroes it): for every step >= 1 X^T dY (rows main_row0.. of that layer's kernel) and the
 *                            bias gradient 1^T dY; the two heads' kernels (h^T d sigma, h^T d rgb).  For the steps that
 *                            read an encoding it adds enc^T dY, in the kernel's column order, into
 *                            enc_scratch[n_steps][128][256] fp32 (caller-zeroed): row c of enc_scratch[s] is the
 *                            gradient of the weight row ntx_mlp_enc_columns maps column c to; the constant-1
 *                            column's row is the bias gradient (needed for step 0).  The host scatters those 96
 *                            rows and adds the heads' bias gradients (sum of d_rgbs). */
int64_t ntx_mlp_packed_t_bytes(const ntx_mlp_t* mlp);
int ntx_mlp_pack_weights_t(const ntx_mlp_t* mlp, const float* master, void* packed_t, void* stream);
int ntx_mlp_backward_data(const ntx_mlp_t* mlp, const void* packed_t, const float* d_rgbs, const void* masks,
                          void* dy_img, int64_t n_rows, int64_t n_tiles, void* stream);
int ntx_mlp_backward_weights(const ntx_mlp_t* mlp, const void* acts_img, const void* dy_img, const void* enc_pos,
                             const void* enc_dir, const float* d_rgbs, float* grad, float* enc_scratch,
                             int64_t n_rows, int64_t n_tiles, void* stream);

/* 8-bit saved operands for the training step (no counterpart in the reference; it exists because the step is bound by the
 * bytes of the saved images): the forward and the data-gradient chain leave their tile images as fp8 e4m3 / scale[step]
 * (32 KB per tile instead of 64 KB) and the weight-gradient GEMM reads them as they are (tcgen05 kind::f8f6f4, MN-major),
 * rescaling the products in its flush.  Scales are per step and per tensor, delayed by one training step:
 * ntx_fp8_update_scales turns the maxima the kernels collected (amax, fp32 bit patterns, cleared by the call) into
 * scale = margin * amax / 448 and its inverse, all on the device.  The encoded-input weight gradients keep bf16: the
 * data-gradient kernel still writes the bf16 dY image of the (three) steps that read an encoding into dy_img.
 * Gradient accuracy: tests/test_gpu_train.py holds this path to the same bars as the bf16 one. */
int ntx_mlp_forward_train_fp8(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_inputs* in, float* rgbs, void* enc_pos,
                              void* enc_dir, void* acts_img8, const float* inv_scale8, void* amax8, void* masks,
                              int64_t n_tiles, void* stream);
int ntx_mlp_backward_data_fp8(const ntx_mlp_t* mlp, const void* packed_t, const float* d_rgbs, const void* masks, void* dy_img,
                              void* dy_img8, const float* inv_scale8, void* amax8, int64_t n_rows, int64_t n_tiles,
                              void* stream);
int ntx_mlp_backward_weights_fp8(const ntx_mlp_t* mlp, const void* acts_img8, const void* dy_img8, const void* dy_img,
                                 const float* x_scale8, const float* dy_scale8, const void* enc_pos, const void* enc_dir,
                                 const float* d_rgbs, float* grad, float* enc_scratch, int64_t n_rows, int64_t n_tiles,
                                 void* stream);
int ntx_fp8_update_scales(void* amax, float* scale, float* inv_scale, int n, float margin, void* stream);

/* Schedule introspection for the host-side backward (nerf_tex_b200/autograd.py).  A "step" is one
 * tensor-core layer: trunk0..trunk{depth-1}, feature, colorhid*, colorhalf. */
int ntx_mlp_num_steps(const ntx_mlp_t* mlp);
int ntx_mlp_step_info(const ntx_mlp_t* mlp, int step, int32_t* layer, int32_t* n_out, int32_t* main_row0,
                      int32_t* enc_data, int32_t* phase, int32_t* relu);
/* ref_index[96] (host): for encoding column c of `phase` (0 pos, 1 dir) the row of the reference's
 * pos_inputs_map / dir_inputs_map it holds (network/model.py:77-101), -1 for the constant-1
 * column, -2 for zero padding. */
int ntx_mlp_enc_columns(const ntx_mlp_t* mlp, int phase, int32_t* ref_index);

/* ------------------------------------------------------------------------------------------
 * Camera rays
 * ---------------------------------------------------------------------------------------- */

/* Replaces network/ray_sampler.py:6-48 (rays_from_camera, Frustum, Proxy) and network/proxy.py:13-35 (AABB).
 *   image_plane_loc[n,2] (row, column) pixel positions, or NULL for the first n pixels of the frame in
 *   row-major order; c2w: HOST pointer to the 4x4 row-major camera-to-world matrix; normalize != 0 divides
 *   rays_d by its norm (Proxy, :34); b0/b1: HOST pointers to the AABB corners (misses give t = inf, inf) or
 *   both NULL for a constant [near, far] (Frustum, :19).
 *   Outputs: rays_o[n,3], rays_d[n,3], t[n,2], cone_scale[n]. */
int ntx_rays_from_camera(const float* image_plane_loc, int64_t n, int height, int width, float focal,
                         const float* c2w, int normalize, const float* b0, const float* b1, float near_, float far_,
                         float* rays_o, float* rays_d, float* t, float* cone_scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * Training step: loss and optimiser (network/loss.py, network/train.py:49-67)
 * ---------------------------------------------------------------------------------------- */

/* AlphaLoss.__call__ (network/loss.py:30-49) with loss_fn / alpha_loss_fn in {0: mse (:51-54), 1: smape (:56-58)}:
 * loss[1] (device) and the gradient of the loss with respect to color_pred[n,3], alpha_pred[n] (and the coarse pair,
 * both NULL or both given) in one launch. */
int ntx_alpha_loss(const float* color_true, const float* alpha_true, const float* color_pred, const float* alpha_pred,
                   const float* color_pred_coarse, const float* alpha_pred_coarse, int64_t n_rays, int color_loss_kind,
                   int alpha_loss_kind, float gamma, int filter_color_loss, int use_hard_mask, float* loss,
                   float* d_color_pred, float* d_alpha_pred, float* d_color_pred_coarse, float* d_alpha_pred_coarse,
                   void* stream);
/* tf.keras.optimizers.Adam.apply_gradients (network/train.py:52,67) on a flat buffer: m, v updated in place,
 * params -= lr_t * m / (sqrt(v) + epsilon) with lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t) computed by the caller. */
int ntx_adam_step(float* params, const float* grad, float* m, float* v, int64_t n, float lr_t, float beta1, float beta2,
                  float epsilon, void* stream);
/* Data-parallel training (not in the reference, which is single-device): the mean over the
 * ranks of the flat fp32 gradient buffer that tape.gradient (network/train.py:66) would have produced on each —
 * one ncclAllReduce(sum) over NVLink on `stream`, then a scale by 1 / world_size.  comm: the caller's ncclComm_t.  NCCL
 * is resolved at run time from the process (or libnccl.so.2); failures return NTX_E_NCCL. */
int ntx_nccl_allreduce_grads(void* comm, float* grad, int64_t n, int world_size, void* stream);

/* ------------------------------------------------------------------------------------------
 * mip variants (network/renderer.py:356-587): cone-segment gaussians + compositing over S + 1 depth edges.
 * The integrated positional encoding itself (network/layer.py:25-41) is the n_pos = 6 mode of the fused MLP kernel.
 * ---------------------------------------------------------------------------------------- */

/* MipRenderer.get_cone_segment_gaussians (renderer.py:411-434): z_edges[R,S+1], radii[R] (blur parameter * cone
 * scale, :384) -> mean[R,S,3], cov[R,S,3] (diagonal). */
int ntx_mip_cone_gaussians(const float* rays_o, const float* rays_d, const float* z_edges, const float* radii, int R,
                           int S, float* mean, float* cov, void* stream);
/* MipInstanceRenderer.get_cone_segment_gaussians (renderer.py:570-587) on packed in-shell samples: dirs[T,3],
 * t[T], radii[T] (blur * cone scale / patch scale, :511), dists[T] -> cov[T,3]. */
int ntx_mip_cone_gaussians_instanced(const float* dirs, const float* t, const float* radii, const float* dists,
                                     int64_t T, float* cov, void* stream);
/* MipRenderer.map_model_output (renderer.py:436-473): like ntx_composite_forward / _backward on packed rgbs[R,S,4],
 * but z_edges[R,S+1] delimits the S segments, so no distance is duplicated (:440). */
int ntx_composite_edges_forward(const float* rgbs, const float* z_edges, const float* rays_d, const float* noise,
                                int R, int S, int map_exr, int composite_bkgd, const float* bkgd_color,
                                float* color_map, float* alpha_map, float* weights, void* stream);
int ntx_composite_edges_backward(const float* rgbs, const float* z_edges, const float* rays_d, const float* noise,
                                 int R, int S, int map_exr, int composite_bkgd, const float* bkgd_color,
                                 const float* d_color_map, const float* d_alpha_map, float* d_rgbs, void* stream);

/* ------------------------------------------------------------------------------------------
 * Image out (the step after a full-frame render)
 * ---------------------------------------------------------------------------------------- */

/* Replaces Logger.render_image's post-processing (network/logger.py:126-137) and the float -> uint8 conversion of
 * write_image (:144): color[H*W,3] and alpha[H*W] (premultiplied) are low-pass filtered and downsampled by `factor`
 * (util/interpolate.py:67-82: depthwise 'SAME' convolution with stride `factor`; taps_host = the n_taps normalised 1-D
 * gaussian weights, HOST pointer, ignored for factor 1), then colour is divided by alpha + 1e-5 if straight_alpha.
 * Outputs (either may be NULL): img[h,w,4] fp32 and img_u8[h,w,4] (values clamped to [0,1], times 255.5, truncated),
 * h = ceil(H / factor), w = ceil(W / factor). */
int ntx_image_finish(const float* color, const float* alpha, int H, int W, int factor, const float* taps_host, int n_taps,
                     int straight_alpha, float* img, uint8_t* img_u8, void* stream);

/* ------------------------------------------------------------------------------------------
 * Shell instancer (the caller in front of the hot path)
 * ---------------------------------------------------------------------------------------- */

/* Replace C_Instancer::GetModelInput (instancer/src/instancer.cpp:751-1037, bound in instancer/instancer.pyx:38-54):
 * patch instances (explicit, or distributed on a base mesh by the host side), optional base mesh whose first hit ends a
 * ray (:804-814,:983,:1019-1023), optional shadow rays (isShadowed :593-603 with the filter of :539-551: a sample whose
 * ray towards the light enters a patch box through its top face, crosses a box's bottom face, hits a mesh triangle on
 * its front side — or the SECOND triangle of any mesh from either side, the filter's `primID == 1` clause at :550 does
 * not look at the geometry — gets the ambient light direction (0, 0, -1), :572-575; looked up at min_shadow_samples /
 * n_shadow_samples points per ray as in :863,:936-951,:992-1000; cast_shadow_rays != 0 adds 1600 B per ray to the
 * workspace), optional parameter textures and auxiliary meshes (ntx_instancer_shading below).  The trace takes the rays'
 * parameters[R,n_param] and light_dir_idx for the shading of auxiliary-mesh hits (NULL / -1 without them) and writes
 * color_last[R,3] (0 except on auxiliary meshes).
 *
 * ntx_instancer_scene — what CommitScene (instancer.cpp:420-424) leaves behind, all DEVICE pointers:
 *   minv[I,3,4]  rows 0..2 of the inverse instance transform (world -> patch), dirt[I,3,3] the row-normalised
 *   transposed rotation, origins[I,3] the instance translation (AddInstance, instancer.cpp:125-141);
 *   spheres[I,4] (optional, 16-byte aligned): world-space bounding sphere of every instanced box as (centre, radius^2),
 *   radius inflated by >= 1 %: a cheap first test that never changes the result;
 *   inst_grid (optional, cell_start == NULL: every ray tests every instance): uniform grid over the instances' world
 *   boxes in place of the reference's Embree BVH — cell_start[nx*ny*nz + 1] / items[] list the instances whose
 *   (slightly inflated) world box overlaps a cell, cell index = (z * ny + y) * nx + x; a ray walks the cells it crosses;
 *   mesh_v[V,3] / mesh_f[F,3] (optional, n_faces == 0: no mesh) the base mesh (DistributeInstancesOnMesh :374-391),
 *   tri_grid the same kind of grid over its triangles (optional).  The triangle test is the watertight one of Woop,
 *   Benthin, Wald (JCGT 2013) in fp32 with the double fallback; ties to the smaller face index.
 * b_0/b_1: HOST pointers to the patch box corners.  Box hits are kept for t in [0, 100] and up to the first mesh hit; at
 * most 200 hits per ray (MAX_TOTAL_HITS, instancer.cpp:22; the smallest t win).  Random draws are a counter hash of
 * (seed, ray id[, step]), not the reference's mt19937; ray_ids[R] (optional) names every ray for those draws — pass the
 * ray's index in the frame and the records do not depend on how the frame is cut into chunks or spread over GPUs.
 * Pass 1 (trace) fills workspace (ntx_instancer_workspace_bytes(R) bytes), n_samples[R] (marching samples of every
 * ray, <= n_pts), hit[R] (ray touched a patch box or the mesh) and alpha_last[R] (1 where the ray ends on the mesh:
 * the opaque last sample of renderer.py:331,339, colour 0).  The caller turns n_samples into ray_ptr[R+1] (exclusive
 * prefix sum, int64; T = ray_ptr[R] samples in total) and sizes the outputs; pass 2 (march) writes the samples of ray r
 * at ray_ptr[r]...:
 *   pts[T,3] patch-local position, dirs[T,3] patch-local unit view direction, params_out[T,n_param] (the ray's
 *   parameters; entries light_dir_idx..+2, if light_dir_idx >= 0, are the normalised light direction mapped into the
 *   patch frame, instancer.cpp:929-952; with light_strength_idx >= 0 the three entries are a point light's POSITION, the
 *   direction becomes position - sample and entry light_strength_idx the received radiance strength / (4 pi d^2 + 1e-6),
 *   instancer.cpp:572-590,954-956), dists[T], t[T], alpha_weight[T] (number of overlapping patches for method 0
 *   'random', 1 for method 1 'nearest', 1 / probability for method 2 'nearest_blend' whose blend_range is the transition
 *   range 0.2 times the patch scale, instancer.cpp:694-713), instance_id[T].  n_param <= 16. */
typedef struct ntx_scene_grid {
    const int32_t* cell_start;
    const int32_t* items;
    float lo[3];
    float cell[3];
    int32_t res[3];
} ntx_scene_grid;
/* One texture channel in the pool: element (i, j) at pool[offset + i * cols + j]; rows follow the image's x, columns
 * its y counted from the bottom (loadTexture, instancer.cpp:34-52). */
typedef struct ntx_texture {
    int64_t offset;
    int32_t rows, cols;
} ntx_texture;
#define NTX_MAX_PARAM_TEXTURES 8
#define NTX_MAX_AUX_MESHES 8
/* Optional per-scene shading data (HOST struct of DEVICE pointers, read at call time):
 *   parameter textures (getParameters, instancer.cpp:640-667): texture file i multiplies parameter param_idx[i] by
 *   param_tex[i] looked up at the UV of the base-mesh point nearest to the sample (within `radius` = the patch extent,
 *   :63,:246), sampled along the ray at min_samples / n_samples points and linearly interpolated (:864,:916-926,:985-991);
 *   auxiliary meshes (AddMesh :393-418, shadeMesh :716-744): faces [aux_face_begin[k], aux_face_begin[k+1]) of the scene's
 *   face list belong to auxiliary mesh k (the base mesh's n_base_faces faces come first); a ray ending there gets the
 *   colour albedo * min(max(n . l, 0) [0 if shadowed] + 0.2, 1), albedo = aux_tex[k][0..2] at the hit's UV (one channel
 *   is used for all three) or 0.8 without a texture.  mesh_uv[V,2] / mesh_n[V,3] are per vertex of the scene's mesh_v. */
typedef struct ntx_instancer_shading {
    const float* pool;
    const float* mesh_uv;
    const float* mesh_n;
    int32_t n_base_faces;
    int32_t n_param_tex;
    int32_t param_idx[NTX_MAX_PARAM_TEXTURES];
    ntx_texture param_tex[NTX_MAX_PARAM_TEXTURES];
    float radius;
    int32_t min_samples, n_samples;
    int32_t n_aux;
    int32_t aux_face_begin[NTX_MAX_AUX_MESHES + 1];
    int32_t aux_channels[NTX_MAX_AUX_MESHES];
    ntx_texture aux_tex[NTX_MAX_AUX_MESHES][3];
} ntx_instancer_shading;
typedef struct ntx_instancer_scene {
    const float* minv;
    const float* dirt;
    const float* origins;
    const float* spheres;
    int32_t n_inst;
    int32_t n_faces;
    const float* mesh_v;
    const int32_t* mesh_f;
    ntx_scene_grid inst_grid;
    ntx_scene_grid tri_grid;
    const ntx_instancer_shading* shading;   /* NULL: no parameter textures, no auxiliary meshes */
} ntx_instancer_scene;
int64_t ntx_instancer_workspace_bytes(int R, int cast_shadow_rays);
/* ray_offsets [R] (may be NULL): the per-ray offset u in [0, 1) the reference draws from its std::mt19937
 * (instancer.cpp:849,858); NULL: a counter hash of (seed, ray name).  A caller that must reproduce a reference run
 * passes the stream (tests/test_gpu_ref_instancer.py does). */
int ntx_instancer_trace(const float* rays_o, const float* rays_d, const int32_t* ray_ids, const float* ray_offsets, int R,
                        const ntx_instancer_scene* scene, const float* b_0, const float* b_1, float step_size, int n_pts,
                        uint32_t seed, int use_mean_distance, int cast_shadow_rays, const float* parameters, int n_param,
                        int light_dir_idx, void* workspace, int32_t* n_samples, int32_t* hit, float* alpha_last,
                        float* color_last, void* stream);
int ntx_instancer_march(const float* rays_o, const float* rays_d, const int32_t* ray_ids, const float* parameters,
                        int n_param, int R, const ntx_instancer_scene* scene, float step_size, int n_pts, uint32_t seed,
                        int light_dir_idx, int light_strength_idx, int method, float blend_range, int use_mean_distance,
                        int cast_shadow_rays, int min_shadow_samples, int n_shadow_samples, const float* b_0,
                        const float* b_1, const void* workspace, const int64_t* ray_ptr, int64_t T, float* pts,
                        float* dirs, float* params_out, float* dists, float* t, float* alpha_weight,
                        int32_t* instance_id, void* stream);

/* ------------------------------------------------------------------------------------------
 * Whole ray chunk: K3 -> K1 -> K2 [-> K3 -> K1 -> K2]
 * ---------------------------------------------------------------------------------------- */

/* View directions fed to the network: out[R,3] = rays_d / |rays_d| (network/renderer.py:98). */
int ntx_normalize_dirs(const float* rays_d, int R, float* out, void* stream);

/* Options of network/renderer.py:34-45 that shape one chunk. */
typedef struct ntx_render_config {
    int32_t n_samples;       /* S */
    int32_t n_importance;    /* N (0: single pass) */
    int32_t perturb;         /* stratified jitter; also sample_pdf's det flag (renderer.py:128) */
    int32_t map_exr;
    int32_t composite_bkgd;
    int32_t blur_idx;        /* -1: none */
    float bkgd_color[3];
} ntx_render_config;

/* Device scratch (bytes) ntx_render_rays needs for R rays: view dirs, depths, raw network outputs
 * and weights of the coarse pass, merged depths and raw outputs of the fine pass. */
int64_t ntx_render_rays_workspace_bytes(int R, int n_samples, int n_importance);

/* Replaces Renderer.render_rays (network/renderer.py:92-143) for one chunk of R rays that all hit
 * the proxy.  rays_o, rays_d[R,3]; t[R,2]; parameters[R,p] or NULL; cone_scale[R] (used iff
 * blur_idx >= 0); z_rand[R,S] iff perturb; u[R,N] iff n_importance > 0 and !perturb; noise[R,S],
 * noise_fine[R,S+N] (already scaled by raw_noise_std) or NULL.  mlp_fine / packed_fine may be NULL
 * (the coarse model is reused, :134).  Outputs: color_pred[R,3], alpha_pred[R]; with
 * n_importance > 0 also color_pred_coarse[R,3], alpha_pred_coarse[R] (:138).
 * workspace: 256-byte aligned device scratch of at least ntx_render_rays_workspace_bytes(). */
int ntx_render_rays(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_t* mlp_fine, const void* packed_fine,
                    const ntx_render_config* cfg, const float* rays_o, const float* rays_d, const float* t,
                    const float* parameters, const float* cone_scale, const float* z_rand, const float* u,
                    const float* noise, const float* noise_fine, int R, float* color_pred, float* alpha_pred,
                    float* color_pred_coarse, float* alpha_pred_coarse, void* workspace, int64_t workspace_bytes,
                    void* stream);

/* Allocation-free twin of ntx_render_rays.  (No counterpart in the reference: it exists for the
 * way this library gets embedded, not for anything the reference computes.)
 *
 * Who needs it: a host that cannot hand the library a device workspace.  The typical case is this
 * library loaded into the reference's own TensorFlow process — TF's allocator claims almost all
 * GPU memory at start-up, so a cudaMalloc made on the side fails, and TF's tensors cannot be sized
 * from C.  The same holds for a render loop that may not tolerate the device-wide stall an
 * allocation causes, and for plug-in hosts that forbid allocation after initialisation.
 *
 * How: the scratch is a fixed array of NTX_STATIC_SCRATCH_BYTES inside the library's device image
 * (every GPU that loads the library holds those bytes until the process exits).  This call never
 * allocates or frees device memory and calls no routine that does.  Inputs and outputs are still
 * the caller's buffers, exactly as for ntx_render_rays (same arguments minus the workspace).
 *
 * Limits and rules:
 *   - a chunk whose ntx_render_rays_workspace_bytes() exceeds NTX_STATIC_SCRATCH_BYTES is refused
 *     with NTX_E_INVALID and a message naming both sizes (4096 rays of 64 + 128 samples fit);
 *     split the chunk, or use ntx_render_rays;
 *   - the first call on each device looks up the array's device address (with that device
 *     current).  Make that first call during set-up, before any stream capture begins;
 *   - "never allocates" holds from that first call on, not for it: the CUDA runtime loads a
 *     library's device image lazily, and it is that load — triggered by the first call into this
 *     file on a device — that reserves the NTX_STATIC_SCRATCH_BYTES.  So load the library and make
 *     the first call (a one-ray chunk will do) BEFORE the host framework takes the rest of the
 *     GPU's memory, or run the process with CUDA_MODULE_LOADING=EAGER so that the reservation
 *     happens when the library is loaded; otherwise that first call can fail with CUDA's
 *     out-of-memory error (reported through ntx_last_error());
 *   - all calls share the one array and it keeps whatever the last call wrote.  Nothing in it
 *     needs clearing — each call overwrites every region before reading it — but two calls that
 *     can overlap on the GPU (different streams, or host threads on the same device) race on
 *     it: keep calls to this entry on one stream, or order them yourself. */
#define NTX_STATIC_SCRATCH_BYTES (24u << 20)
int ntx_render_rays_static(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_t* mlp_fine,
                           const void* packed_fine, const ntx_render_config* cfg, const float* rays_o,
                           const float* rays_d, const float* t, const float* parameters, const float* cone_scale,
                           const float* z_rand, const float* u, const float* noise, const float* noise_fine, int R,
                           float* color_pred, float* alpha_pred, float* color_pred_coarse, float* alpha_pred_coarse,
                           void* stream);

/* ------------------------------------------------------------------------------------------
 * tcgen05 self-test (used by tests/ to validate the UMMA shared-memory descriptors on a GPU)
 * ---------------------------------------------------------------------------------------- */
/* D[128,N] = A[128,K] * B[N,K]^T with bf16 inputs (given as fp32, rounded to bf16 on device),
 * fp32 accumulate.  layout: 0 = SWIZZLE_64B K-slices of 32, 1 = SWIZZLE_128B K-slices of 64,
 * 2 = no swizzle (interleaved core matrices). K multiple of 64, N in {128,256}. */
int ntx_selftest_umma(const float* A, const float* B, int N, int K, int layout, float* D, void* stream);
/* D[128,N] = At^T * Bt, At[K,128] and Bt[K,N] row-major: both operands MN-major, SWIZZLE_64B row images
 * (the operand form of the weight-gradient GEMM, K = sample rows).  K in {64,128}, N in {128,256}. */
int ntx_selftest_umma_mn(const float* At, const float* Bt, int N, int K, float* D, void* stream);
/* The same product with both operands rounded to fp8 e4m3 (kind::f8f6f4, MN-major SWIZZLE_64B images of 64 elements per
 * row, 32 K rows per instruction): probes whether 8-bit saved operands can feed the weight-gradient GEMM as they are. */
int ntx_selftest_umma_mn8(const float* At, const float* Bt, int N, int K, float* D, void* stream);
/* global -> shared through cp.async.bulk + mbarrier, then back out: dst[0:bytes] = src[0:bytes]. */
int ntx_selftest_bulk_copy(const void* src, void* dst, int bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NERFTEX_B200_H_ */

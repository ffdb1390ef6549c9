// The two elementwise pieces of a training step that sit around the hot path (network/train.py:62-67):
//   * AlphaLoss (network/loss.py:20-58) for the shipped combinations of smape / mse: the loss value and its
//     gradient with respect to the renderer's outputs in ONE launch (the reference's tape takes ~35 tiny ops);
//   * the Adam update in tf.keras' form (epsilon added to sqrt(v), bias correction folded into the step size).
// Both are memory-trivial; the point is launch count, not bandwidth.
#include "ntx_common.cuh"

namespace {

// kind 0: mse (loss.py:51-54), kind 1: smape with eps = 1e-2 (loss.py:56-58); returns the term and d term / d pred
__device__ __forceinline__ float loss_term(int kind, float yt, float yp, float& d_pred) {
    if (kind == 0) {
        const float e = yt - yp;
        d_pred = -2.0f * e;
        return e * e;
    }
    const float e = yt - yp, den = yt + yp + 1e-2f;
    const float a = fabsf(e);
    const float sgn = (e > 0.0f) ? 1.0f : (e < 0.0f ? -1.0f : 0.0f);
    d_pred = (-sgn * den - a) / (den * den);
    return a / den;
}

// One block; thread-strided over rays.  out[0] = loss, gradients already divided by the element counts.
__global__ void __launch_bounds__(1024)
alpha_loss_kernel(const float* __restrict__ color_true, const float* __restrict__ alpha_true,
                  const float* __restrict__ color_pred, const float* __restrict__ alpha_pred,
                  const float* __restrict__ color_coarse, const float* __restrict__ alpha_coarse, int64_t n, int color_kind,
                  int alpha_kind, float gamma, int filter_color, int hard_mask, float* __restrict__ loss,
                  float* __restrict__ d_color, float* __restrict__ d_alpha, float* __restrict__ d_color_coarse,
                  float* __restrict__ d_alpha_coarse) {
    __shared__ float partial[32];
    const float inv_nc = 1.0f / (float)(3 * n), inv_na = gamma / (float)n;
    float acc = 0.0f;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float at = alpha_true[i];
        const float mask = filter_color ? (hard_mask ? (at > 0.0f ? 1.0f : 0.0f) : at) : 1.0f;    // loss.py:32-36
        float g;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float ct = color_true[3 * i + c] * mask;
            acc += loss_term(color_kind, ct, color_pred[3 * i + c] * mask, g) * inv_nc;
            d_color[3 * i + c] = g * mask * inv_nc;
            if (color_coarse) {                                                                    // loss.py:43-47
                acc += loss_term(color_kind, ct, color_coarse[3 * i + c] * mask, g) * inv_nc;
                d_color_coarse[3 * i + c] = g * mask * inv_nc;
            }
        }
        acc += loss_term(alpha_kind, at, alpha_pred[i], g) * inv_na;
        d_alpha[i] = g * inv_na;
        if (alpha_coarse) {
            acc += loss_term(alpha_kind, at, alpha_coarse[i], g) * inv_na;
            d_alpha_coarse[i] = g * inv_na;
        }
    }
    acc = ntx_warp_sum(acc);
    if ((threadIdx.x & 31) == 0) partial[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < (blockDim.x >> 5) ? partial[threadIdx.x] : 0.0f;
        v = ntx_warp_sum(v);
        if (threadIdx.x == 0) *loss = v;
    }
}

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            int64_t n, float lr_t, float b1, float b2, float eps) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.0f - b1) * gi;
    const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
}

}  // namespace

NTX_API int ntx_alpha_loss(const float* color_true, const float* alpha_true, const float* color_pred, const float* alpha_pred,
                           const float* color_pred_coarse, const float* alpha_pred_coarse, int64_t n_rays, int color_loss_kind,
                           int alpha_loss_kind, float gamma, int filter_color_loss, int use_hard_mask, float* loss,
                           float* d_color_pred, float* d_alpha_pred, float* d_color_pred_coarse, float* d_alpha_pred_coarse,
                           void* stream) {
    NTX_REQUIRE(n_rays >= 1, "alpha_loss: no rays");
    NTX_REQUIRE(color_true && alpha_true && color_pred && alpha_pred && loss && d_color_pred && d_alpha_pred,
                "alpha_loss: null pointer");
    NTX_REQUIRE((color_pred_coarse != nullptr) == (alpha_pred_coarse != nullptr), "alpha_loss: coarse colour and alpha come together");
    NTX_REQUIRE(!color_pred_coarse || (d_color_pred_coarse && d_alpha_pred_coarse), "alpha_loss: null coarse gradient pointer");
    NTX_REQUIRE((color_loss_kind | alpha_loss_kind) >= 0 && color_loss_kind <= 1 && alpha_loss_kind <= 1,
                "alpha_loss: loss kind must be 0 (mse) or 1 (smape)");
    alpha_loss_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(color_true, alpha_true, color_pred, alpha_pred, color_pred_coarse,
                                                           alpha_pred_coarse, n_rays, color_loss_kind, alpha_loss_kind, gamma,
                                                           filter_color_loss, use_hard_mask, loss, d_color_pred, d_alpha_pred,
                                                           d_color_pred_coarse, d_alpha_pred_coarse);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_adam_step(float* params, const float* grad, float* m, float* v, int64_t n, float lr_t, float beta1, float beta2,
                          float epsilon, void* stream) {
    NTX_REQUIRE(n >= 0, "adam_step: bad size");
    if (n == 0) return NTX_OK;
    NTX_REQUIRE(params && grad && m && v, "adam_step: null pointer");
    adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(params, grad, m, v, n, lr_t, beta1, beta2, epsilon);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

// ---- data-parallel gradient exchange: one ncclAllReduce(sum) of the flat fp32 gradient buffer,
// then 1 / world on the same stream.  NCCL is not linked: the entry resolves ncclAllReduce in the process (the
// communicator the caller passes can only come from an NCCL that is already loaded) or, failing that, in libnccl.so.2.
#include <dlfcn.h>

namespace {
typedef int (*nccl_allreduce_t)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*nccl_errstr_t)(int);
__global__ void scale_kernel(float* __restrict__ x, int64_t n, float s) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] *= s;
}
}  // namespace

NTX_API int ntx_nccl_allreduce_grads(void* comm, float* grad, int64_t n, int world_size, void* stream) {
    NTX_REQUIRE(comm && n >= 0 && world_size >= 1, "nccl_allreduce_grads: bad arguments");
    if (n == 0) return NTX_OK;
    NTX_REQUIRE(grad, "nccl_allreduce_grads: null gradient buffer");
    static nccl_allreduce_t allreduce = nullptr;
    static nccl_errstr_t errstr = nullptr;
    if (!allreduce) {
        void* sym = dlsym(RTLD_DEFAULT, "ncclAllReduce");
        void* h = nullptr;
        if (!sym && (h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL))) sym = dlsym(h, "ncclAllReduce");
        if (!sym) {
            ntx_set_error("nccl_allreduce_grads: ncclAllReduce not found in the process nor in libnccl.so.2");
            return NTX_E_NCCL;
        }
        allreduce = (nccl_allreduce_t)sym;
        errstr = (nccl_errstr_t)(h ? dlsym(h, "ncclGetErrorString") : dlsym(RTLD_DEFAULT, "ncclGetErrorString"));
    }
    const int rc = allreduce(grad, grad, (size_t)n, /*ncclFloat32*/ 7, /*ncclSum*/ 0, comm, (cudaStream_t)stream);
    if (rc != 0) {
        ntx_set_error("nccl_allreduce_grads: ncclAllReduce failed: %s", errstr ? errstr(rc) : "unknown NCCL error");
        return NTX_E_NCCL;
    }
    if (world_size > 1) {
        scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(grad, n, 1.0f / (float)world_size);
        NTX_LAUNCH_CHECK();
    }
    return NTX_OK;
}

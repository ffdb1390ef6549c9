// Image finishing after a full-frame render (Logger.render_image, network/logger.py:121-137):
// [color | alpha] -> gaussian low-pass + strided downsample (util/interpolate.py:67-82, a depthwise 'SAME'
// convolution) -> premultiplied to straight alpha (:132-135) -> optional 8-bit quantisation (:144).
// One thread per output pixel; the separable taps arrive as a kernel parameter.  HBM-bound and tiny.
#include "ntx_common.cuh"

namespace {

constexpr int kMaxTaps = 64;
struct Taps {
    float w[kMaxTaps];
};

__global__ void image_finish_kernel(const float* __restrict__ color, const float* __restrict__ alpha, int H, int W, int factor,
                                    int taps, int pad_top, int pad_left, Taps k, int straight_alpha, int h, int w,
                                    float* __restrict__ img, uint8_t* __restrict__ img_u8) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= h * w) return;
    const int oy = idx / w, ox = idx - oy * w;
    float px[4] = {0.f, 0.f, 0.f, 0.f};
    if (factor == 1) {
        const int64_t p = (int64_t)oy * W + ox;
        px[0] = color[3 * p]; px[1] = color[3 * p + 1]; px[2] = color[3 * p + 2]; px[3] = alpha[p];
    } else {
        for (int i = 0; i < taps; ++i) {
            const int y = oy * factor + i - pad_top;
            if (y < 0 || y >= H) continue;                                   // zero padding
            float row[4] = {0.f, 0.f, 0.f, 0.f};
            for (int j = 0; j < taps; ++j) {
                const int x = ox * factor + j - pad_left;
                if (x < 0 || x >= W) continue;
                const int64_t p = (int64_t)y * W + x;
                const float wj = k.w[j];
                row[0] += wj * color[3 * p]; row[1] += wj * color[3 * p + 1]; row[2] += wj * color[3 * p + 2];
                row[3] += wj * alpha[p];
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) px[c] += k.w[i] * row[c];
        }
    }
    if (straight_alpha) {                                                     // logger.py:133-135
        const float inv = 1.0f / (px[3] + 1e-5f);
        px[0] *= inv; px[1] *= inv; px[2] *= inv;
    }
    if (img) {
#pragma unroll
        for (int c = 0; c < 4; ++c) img[4 * (int64_t)idx + c] = px[c];
    }
    if (img_u8) {                                                             // convert_image_dtype(float -> uint8), saturating
        uchar4 q;
        q.x = (uint8_t)(fminf(fmaxf(px[0], 0.0f), 1.0f) * 255.5f);
        q.y = (uint8_t)(fminf(fmaxf(px[1], 0.0f), 1.0f) * 255.5f);
        q.z = (uint8_t)(fminf(fmaxf(px[2], 0.0f), 1.0f) * 255.5f);
        q.w = (uint8_t)(fminf(fmaxf(px[3], 0.0f), 1.0f) * 255.5f);
        reinterpret_cast<uchar4*>(img_u8)[idx] = q;
    }
}

}  // namespace

NTX_API int ntx_image_finish(const float* color, const float* alpha, int H, int W, int factor, const float* taps_host,
                             int n_taps, int straight_alpha, float* img, uint8_t* img_u8, void* stream) {
    NTX_REQUIRE(H >= 1 && W >= 1 && factor >= 1, "image_finish: bad shape H=%d W=%d factor=%d", H, W, factor);
    NTX_REQUIRE(color && alpha && (img || img_u8), "image_finish: null pointer");
    NTX_REQUIRE(factor == 1 || (taps_host && n_taps >= 1 && n_taps <= kMaxTaps), "image_finish: 1..%d filter taps", kMaxTaps);
    NTX_REQUIRE(!img_u8 || ((uintptr_t)img_u8 & 3) == 0, "image_finish: the 8-bit image must be 4-byte aligned");
    const int h = (H + factor - 1) / factor, w = (W + factor - 1) / factor;   // 'SAME': ceil(in / stride)
    Taps k;
    memset(&k, 0, sizeof(k));
    int pad_top = 0, pad_left = 0;
    if (factor > 1) {
        for (int i = 0; i < n_taps; ++i) k.w[i] = taps_host[i];
        const int pv = (h - 1) * factor + n_taps - H, ph = (w - 1) * factor + n_taps - W;
        pad_top = (pv > 0 ? pv : 0) / 2;
        pad_left = (ph > 0 ? ph : 0) / 2;
    }
    image_finish_kernel<<<(h * w + 255) / 256, 256, 0, (cudaStream_t)stream>>>(color, alpha, H, W, factor, n_taps, pad_top, pad_left,
                                                                             k, straight_alpha, h, w, img, img_u8);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

// Device helpers shared by the fused MLP kernels (forward, data-gradient, weight-gradient).
#pragma once

#include <cuda_bf16.h>
#include <stdint.h>

#include "ntx_mlp.cuh"
#include "ntx_umma.cuh"

namespace {

using namespace ntxmlp;

constexpr uint32_t kMainBytes = kMainBlocks * kBlockBytes;                 // 65536: one [128 x 256] bf16 tile image
constexpr int kEncCols = kEncBlocks * 32;   // 96

__device__ __forceinline__ void st_global_v4(uint4* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// 16-byte chunk `q` (8 bf16 columns) of row `row` inside a SW64 K-major region starting at `base`
__device__ __forceinline__ uint32_t chunk_addr(uint32_t base, int q, int row) {
    return base + (uint32_t)(q >> 2) * kBlockBytes + (uint32_t)row * 64 + (uint32_t)(((q & 3) ^ ((row >> 1) & 3)) << 4);
}

// ---- 8-bit saved operands (training): [128 x 256] tile images of fp8 e4m3, 4 blocks of [128 rows x 64 B] with the
// SWIZZLE_64B pattern of the bf16 images (16-byte chunk index ^= (row >> 1) & 3): 32 KB per tile, consumed MN-major by
// the weight-gradient GEMM (kind::f8f6f4, validated by ntx_selftest_umma_mn8).
constexpr uint32_t kMain8Bytes = 4 * kBlockBytes;                          // 32768
// four floats -> four e4m3 bytes (a in the lowest byte), round to nearest, saturating
__device__ __forceinline__ uint32_t pack_e4m3x4(float a, float b, float c, float d) {
    uint16_t lo, hi;
    asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(b), "f"(a));
    asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(d), "f"(c));
    return (uint32_t)lo | ((uint32_t)hi << 16);
}
__device__ __forceinline__ uint32_t pack_e4m3x4_relu(float a, float b, float c, float d) {
    uint16_t lo, hi;
    asm("cvt.rn.satfinite.relu.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(b), "f"(a));
    asm("cvt.rn.satfinite.relu.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(d), "f"(c));
    return (uint32_t)lo | ((uint32_t)hi << 16);
}
// 32 consecutive columns (chunk j of 8) of this thread's row, scaled, into the tile image at `tile_img`; tracks max |x|
// when `track` (the maxima only steer next step's scales, so a sample of the tiles is enough).  RELU: the conversion clamps
// negative inputs to zero (x is the raw accumulator then).
template <bool RELU = false>
__device__ __forceinline__ void dump_fp8_chunk32(uint8_t* tile_img, int j, int row, const float (&x)[32], float inv_scale,
                                                 float& amax, bool track = true) {
    uint32_t w[8];
    if (track) {
#pragma unroll
        for (int e = 0; e < 32; ++e) amax = fmaxf(amax, RELU ? x[e] : fabsf(x[e]));
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const float a = x[4 * e] * inv_scale, b = x[4 * e + 1] * inv_scale, c = x[4 * e + 2] * inv_scale, d = x[4 * e + 3] * inv_scale;
        w[e] = RELU ? pack_e4m3x4_relu(a, b, c, d) : pack_e4m3x4(a, b, c, d);
    }
    uint8_t* rowp = tile_img + (uint32_t)(j >> 1) * kBlockBytes + (uint32_t)row * 64;
    const int sw = (row >> 1) & 3, c0 = (j & 1) * 2;
    st_global_v4(reinterpret_cast<uint4*>(rowp + (((c0) ^ sw) << 4)), w[0], w[1], w[2], w[3]);
    st_global_v4(reinterpret_cast<uint4*>(rowp + (((c0 + 1) ^ sw) << 4)), w[4], w[5], w[6], w[7]);
}
// block-wide (per role group) maximum into a shared slot: warp shuffle, then one shared-memory atomic per warp
__device__ __forceinline__ void amax_to_shared(float amax, uint32_t* slot) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0) atomicMax(slot, __float_as_uint(amax));       // non-negative floats order like their bits
}

// [sin, cos](2^k x) for k = 0..3 as four bf16x2 words: one accurate sincosf + 3 angle doublings
__device__ __forceinline__ void sincos4(float x, uint32_t (&w)[4]) {
    float s, c;
    sincosf(x, &s, &c);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        w[k] = pack_bf16(s, c);
        const float s2 = 2.0f * s * c;
        c = 1.0f - 2.0f * s * s;
        s = s2;
    }
}


}  // namespace

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

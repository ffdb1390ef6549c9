// Shared host/device helpers for libnerftex_b200.so (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/nerftex_b200.h"

#define NTX_API extern "C" __attribute__((visibility("default")))

// Thread-local message of the last failure; defined in ntx_api.cu.
void ntx_set_error(const char* fmt, ...);

#define NTX_REQUIRE(cond, ...)             \
    do {                                   \
        if (!(cond)) {                     \
            ntx_set_error(__VA_ARGS__);    \
            return NTX_E_INVALID;          \
        }                                  \
    } while (0)

#define NTX_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            ntx_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return NTX_E_CUDA;                                                           \
        }                                                                                \
    } while (0)

#define NTX_LAUNCH_CHECK()                                                               \
    do {                                                                                 \
        cudaError_t e__ = cudaGetLastError();                                            \
        if (e__ != cudaSuccess) {                                                        \
            ntx_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
            return NTX_E_CUDA;                                                           \
        }                                                                                \
    } while (0)

// Number of SMs of the current device (cached per device id).
int ntx_sm_count();

// A packed weight image (this GPU's memory) as a 2-D tensor map over rows of 2 KB (256 eight-byte elements, the widest
// row a box can have; every K-slice is stored pre-swizzled by its packer and starts on a 2 KB boundary), box =
// `box_rows` consecutive rows = one CTA's half of a slice (4 rows = 8 KB for N = 256, 2 rows for N = 128).  No swizzle,
// no interleave: the box lands in shared memory exactly as a linear copy of the same bytes would.  Defined in ntx_api.cu.
int ntx_weight_map(CUtensorMap* map, const void* packed, int64_t packed_bytes, uint32_t box_rows);

// tf.linspace(0, 1, n)[i]: exact endpoints, interior fl(1/(n-1)) * i  (no FMA involved).
__host__ __device__ __forceinline__ float ntx_linspace01(int i, int n) {
    if (n == 1 || i == 0) return 0.0f;
    if (i == n - 1) return 1.0f;
#ifdef __CUDA_ARCH__
    float delta = __fdiv_rn(1.0f, (float)(n - 1));
    return __fmul_rn(delta, (float)i);
#else
    float delta = 1.0f / (float)(n - 1);
    return delta * (float)i;
#endif
}

__device__ __forceinline__ float ntx_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

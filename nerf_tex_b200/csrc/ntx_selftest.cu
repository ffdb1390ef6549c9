// tcgen05 / bulk-TMA self-tests.  One CTA computes D[128,N] = A[128,K] * B[N,K]^T on the
// 5th-gen tensor cores with each of the shared-memory operand layouts the library uses, so a
// wrong descriptor bit shows up as a failing unit test rather than as a wrong render.
#include <cuda_fp8.h>

#include "ntx_common.cuh"
#include "ntx_umma.cuh"

namespace {

using namespace umma;

__device__ __forceinline__ uint32_t layout_offset(int layout, uint32_t row, uint32_t k, uint32_t rows, uint32_t K) {
    if (layout == 0) return sw64_offset(row, k, rows);
    if (layout == 1) return sw128_offset(row, k, rows);
    return interleave_offset(row, k, K);
}

__global__ void __launch_bounds__(128, 1)
selftest_umma_kernel(const float* __restrict__ A, const float* __restrict__ B, int N, int K, int layout,
                     float* __restrict__ D) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar_done;
    __shared__ uint32_t tmem_base_slot;

    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + 128 * K * 2;
    const int tid = threadIdx.x, warp = tid >> 5;

    for (int idx = tid; idx < 128 * K; idx += 128) {
        int r = idx / K, k = idx % K;
        *reinterpret_cast<__nv_bfloat16*>(sA + layout_offset(layout, r, k, 128, K)) = __float2bfloat16(A[idx]);
    }
    for (int idx = tid; idx < N * K; idx += 128) {
        int r = idx / K, k = idx % K;
        *reinterpret_cast<__nv_bfloat16*>(sB + layout_offset(layout, r, k, N, K)) = __float2bfloat16(B[idx]);
    }
    fence_proxy_async_smem();

    if (warp == 0) {
        tmem_alloc(smem_u32(&tmem_base_slot), 256);
        tmem_relinquish();
    }
    if (tid == 0) {
        mbar_init(smem_u32(&bar_done), 1);
        fence_mbar_init();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (tid == 0) {
        const uint32_t idesc = idesc_bf16_f32(128, N);
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        for (int ks = 0; ks < K / 16; ++ks) {
            uint64_t ad, bd;
            if (layout == 0) {          // SW64: 32-wide blocks, two K=16 halves per block
                uint32_t blk = ks >> 1, half = ks & 1;
                ad = smem_desc(a0 + blk * 128 * 64 + half * 32, 16, 512, LAYOUT_SW64);
                bd = smem_desc(b0 + blk * N * 64 + half * 32, 16, 512, LAYOUT_SW64);
            } else if (layout == 1) {   // SW128: 64-wide blocks, four K=16 steps per block
                uint32_t blk = ks >> 2, q = ks & 3;
                ad = smem_desc(a0 + blk * 128 * 128 + q * 32, 16, 1024, LAYOUT_SW128);
                bd = smem_desc(b0 + blk * N * 128 + q * 32, 16, 1024, LAYOUT_SW128);
            } else {                    // interleaved: LBO = 128 B (next K chunk), SBO = K/8*128 B (next 8 rows)
                ad = smem_desc(a0 + ks * 256, 128, (K >> 3) * 128, LAYOUT_NONE);
                bd = smem_desc(b0 + ks * 256, 128, (K >> 3) * 128, LAYOUT_NONE);
            }
            mma_bf16_ss(tmem_base, ad, bd, idesc, ks > 0 ? 1u : 0u);
        }
        mma_commit(smem_u32(&bar_done));
    }
    mbar_wait(smem_u32(&bar_done), 0);
    tc_fence_after();

    // thread t of warp w reads TMEM lane 32*w + t == output row
    const int row = tid;
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) D[(size_t)row * N + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 256);
}

// D[128,N] = At^T * Bt with At[K,128], Bt[K,N] row-major (both operands MN-major): the operand tiles are the
// very same [row][32-column SWIZZLE_64B block] images the forward kernel keeps for its activations, now
// read with the row index as K.  Descriptor: LBO = pitch between 32-element MN blocks, SBO = 8 K-rows.
__global__ void __launch_bounds__(128, 1)
selftest_umma_mn_kernel(const float* __restrict__ At, const float* __restrict__ Bt, int N, int K, float* __restrict__ D) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar_done;
    __shared__ uint32_t tmem_base_slot;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                       // image of At: 4 blocks of [K rows x 64 B]
    uint8_t* sB = smem + 4 * K * 64;          // image of Bt: N/32 blocks of [K rows x 64 B]
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int idx = tid; idx < K * 128; idx += 128) {
        int k = idx / 128, m = idx % 128;
        *reinterpret_cast<__nv_bfloat16*>(sA + sw64_offset(k, m, K)) = __float2bfloat16(At[idx]);
    }
    for (int idx = tid; idx < K * N; idx += 128) {
        int k = idx / N, n = idx % N;
        *reinterpret_cast<__nv_bfloat16*>(sB + sw64_offset(k, n, K)) = __float2bfloat16(Bt[idx]);
    }
    fence_proxy_async_smem();
    if (warp == 0) { tmem_alloc(smem_u32(&tmem_base_slot), 256); tmem_relinquish(); }
    if (tid == 0) { mbar_init(smem_u32(&bar_done), 1); fence_mbar_init(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    if (tid == 0) {
        const uint32_t idesc = idesc_bf16_f32_mn(128, N);
        const uint32_t lbo = K * 64;          // next 32 MN elements: next block
        for (int ks = 0; ks < K / 16; ++ks) { // 16 K rows = 1024 B inside every block
            const uint64_t ad = smem_desc(smem_u32(sA) + ks * 1024, lbo, 512, LAYOUT_SW64);
            const uint64_t bd = smem_desc(smem_u32(sB) + ks * 1024, lbo, 512, LAYOUT_SW64);
            mma_bf16_ss(tmem_base, ad, bd, idesc, ks > 0 ? 1u : 0u);
        }
        mma_commit(smem_u32(&bar_done));
    }
    mbar_wait(smem_u32(&bar_done), 0);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) D[(size_t)tid * N + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 256);
}

// Same as selftest_umma_mn_kernel with 8-bit (e4m3) operands, kind::f8f6f4: D[128,N] = At^T * Bt, At[K,128] and
// Bt[K,N] row-major, both MN-major SWIZZLE_64B row images of 64 one-byte elements per 64-byte row; K = 32 rows per MMA.
__global__ void __launch_bounds__(128, 1)
selftest_umma_mn8_kernel(const float* __restrict__ At, const float* __restrict__ Bt, int N, int K, float* __restrict__ D) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar_done;
    __shared__ uint32_t tmem_base_slot;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                       // image of At: 2 blocks of [K rows x 64 B]
    uint8_t* sB = smem + 2 * K * 64;          // image of Bt: N/64 blocks of [K rows x 64 B]
    const int tid = threadIdx.x, warp = tid >> 5;
    auto off8 = [&](int k, int m) { return (m >> 6) * K * 64 + k * 64 + ((((m & 63) >> 4) ^ ((k >> 1) & 3)) << 4) + (m & 15); };
    for (int idx = tid; idx < K * 128; idx += 128) {
        int k = idx / 128, m = idx % 128;
        sA[off8(k, m)] = (uint8_t)__nv_cvt_float_to_fp8(At[idx], __NV_SATFINITE, __NV_E4M3);
    }
    for (int idx = tid; idx < K * N; idx += 128) {
        int k = idx / N, n = idx % N;
        sB[off8(k, n)] = (uint8_t)__nv_cvt_float_to_fp8(Bt[idx], __NV_SATFINITE, __NV_E4M3);
    }
    fence_proxy_async_smem();
    if (warp == 0) { tmem_alloc(smem_u32(&tmem_base_slot), 256); tmem_relinquish(); }
    if (tid == 0) { mbar_init(smem_u32(&bar_done), 1); fence_mbar_init(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    if (tid == 0) {
        const uint32_t idesc = idesc_e4m3_f32_mn(128, N);
        const uint32_t lbo = K * 64;          // next 64 MN elements: next block
        for (int ks = 0; ks < K / 32; ++ks) { // 32 K rows = 2048 B inside every block
            const uint64_t ad = smem_desc(smem_u32(sA) + ks * 2048, lbo, 512, LAYOUT_SW64);
            const uint64_t bd = smem_desc(smem_u32(sB) + ks * 2048, lbo, 512, LAYOUT_SW64);
            mma_f8_ss(tmem_base, ad, bd, idesc, ks > 0 ? 1u : 0u);
        }
        mma_commit(smem_u32(&bar_done));
    }
    mbar_wait(smem_u32(&bar_done), 0);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) D[(size_t)tid * N + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 256);
}

__global__ void selftest_bulk_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int bytes) {
    extern __shared__ __align__(128) uint8_t buf[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&bar), 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(smem_u32(&bar), bytes);
        tma_bulk_g2s(smem_u32(buf), src, bytes, smem_u32(&bar));
    }
    mbar_wait(smem_u32(&bar), 0);
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = buf[i];
}

}  // namespace

NTX_API int ntx_selftest_umma(const float* A, const float* B, int N, int K, int layout, float* D, void* stream) {
    NTX_REQUIRE(A && B && D, "selftest_umma: null pointer");
    NTX_REQUIRE((N == 128 || N == 256) && K >= 64 && K % 64 == 0 && K <= 256, "selftest_umma: N in {128,256}, K in {64,128,192,256}");
    NTX_REQUIRE(layout >= 0 && layout <= 2, "selftest_umma: layout in {0,1,2}");
    const size_t smem = (size_t)(128 + N) * K * 2 + 1024;
    NTX_CUDA(cudaFuncSetAttribute(selftest_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    selftest_umma_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(A, B, N, K, layout, D);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_selftest_umma_mn(const float* At, const float* Bt, int N, int K, float* D, void* stream) {
    NTX_REQUIRE(At && Bt && D, "selftest_umma_mn: null pointer");
    NTX_REQUIRE((N == 128 || N == 256) && (K == 64 || K == 128), "selftest_umma_mn: N in {128,256}, K in {64,128}");
    const size_t smem = (size_t)(4 + N / 32) * K * 64 + 1024;
    NTX_CUDA(cudaFuncSetAttribute(selftest_umma_mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    selftest_umma_mn_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(At, Bt, N, K, D);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_selftest_umma_mn8(const float* At, const float* Bt, int N, int K, float* D, void* stream) {
    NTX_REQUIRE(At && Bt && D, "selftest_umma_mn8: null pointer");
    NTX_REQUIRE((N == 128 || N == 256) && (K == 64 || K == 128), "selftest_umma_mn8: N in {128,256}, K in {64,128}");
    const size_t smem = (size_t)(2 + N / 64) * K * 64 + 1024;
    NTX_CUDA(cudaFuncSetAttribute(selftest_umma_mn8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    selftest_umma_mn8_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(At, Bt, N, K, D);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

extern "C" __attribute__((visibility("default")))
int ntx_selftest_bulk_copy(const void* src, void* dst, int bytes, void* stream) {
    NTX_REQUIRE(src && dst && bytes > 0 && bytes % 16 == 0 && bytes <= 32768, "selftest_bulk_copy: bytes in (0, 32768], multiple of 16");
    selftest_bulk_kernel<<<1, 128, bytes, (cudaStream_t)stream>>>((const uint8_t*)src, (uint8_t*)dst, bytes);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

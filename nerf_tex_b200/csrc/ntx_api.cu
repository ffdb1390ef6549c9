// Library-level entry points: version, error string, device info.
#include <stdarg.h>

#include "ntx_common.cuh"

static thread_local char g_err[512] = "";

void ntx_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

NTX_API int ntx_version(void) { return NTX_VERSION; }

NTX_API const char* ntx_last_error(void) { return g_err; }

int ntx_sm_count() {
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime (the library does not link libcuda)
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

int ntx_weight_map(CUtensorMap* map, const void* packed, int64_t packed_bytes, uint32_t box_rows) {
    EncodeTiledFn enc = encode_tiled_fn();
    NTX_REQUIRE(enc, "cuTensorMapEncodeTiled is not available from this driver");
    NTX_REQUIRE(packed_bytes > 0 && (packed_bytes & 2047) == 0 && ((uintptr_t)packed & 15) == 0,
                "weight image must be 16-byte aligned and a multiple of 2 KB (%lld bytes)", (long long)packed_bytes);
    const cuuint64_t dims[2] = {256, (cuuint64_t)(packed_bytes / 2048)};
    const cuuint64_t strides[1] = {2048};
    const cuuint32_t box[2] = {256, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<void*>(packed), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    NTX_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return NTX_OK;
}

// K3 — stratified depths, inverse-CDF sampling (sample_pdf) and the hierarchical merge.
//
// Bit-exactness contract: every floating-point op below is an explicit round-to-nearest
// intrinsic (__fadd_rn / __fmul_rn / __fdiv_rn never contract into FMAs) issued in the same
// left-to-right order as oracle/nerftex_oracle.py and oracle/ntx_oracle.c.
//
// Layout: a warp owns `rpw` consecutive rays.  Weights / bins of those rays are staged in
// shared memory with coalesced loads, lane l runs the strictly sequential normalising sum and
// cumsum of ray l (the serial 62-long chain costs nothing when 32 rays run side by side), then
// the whole warp cooperates on each ray: 32 binary searches at a time, coalesced stores, and
// for the hierarchical variant an in-warp bitonic merge of the S coarse and N new depths.
#include <math_constants.h>

#include "ntx_common.cuh"

namespace {

__global__ void stratified_z_kernel(const float* __restrict__ t, const float* __restrict__ z_rand, int64_t total,
                                    int S, int perturb, float* __restrict__ z) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    int64_t r = idx / S;
    int i = (int)(idx - r * S);
    float2 nf = reinterpret_cast<const float2*>(t)[r];
    auto base = [&](int j) {
        float tv = ntx_linspace01(j, S);
        return __fadd_rn(__fmul_rn(nf.x, __fsub_rn(1.0f, tv)), __fmul_rn(nf.y, tv));
    };
    float bi = base(i);
    if (!perturb) {
        z[idx] = bi;
        return;
    }
    float lower = (i == 0) ? bi : __fmul_rn(0.5f, __fadd_rn(bi, base(i - 1)));
    float upper = (i == S - 1) ? bi : __fmul_rn(0.5f, __fadd_rn(base(i + 1), bi));
    z[idx] = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), z_rand[idx]));
}

// HIER = false: bins[R,C], weights[R,C-1]                      -> samples[R,N]
// HIER = true : src = z[R,S] (C = S-1 mid points), weights[R,S] (inner S-2 used)
//               -> z_out[R,S+N] sorted, optional samples / inds
template <bool HIER>
__global__ void sample_pdf_kernel(const float* __restrict__ src, const float* __restrict__ weights,
                                  const float* __restrict__ u_in, int R, int C, int N, int det, int rpw,
                                  int cstride, int bstride, int p2, float* __restrict__ samples,
                                  int32_t* __restrict__ inds, float* __restrict__ z_out) {
    extern __shared__ float smem[];
    const int lane = threadIdx.x & 31;
    const int warp_in_block = threadIdx.x >> 5;
    const int warps_per_block = blockDim.x >> 5;
    const int per_warp = rpw * (cstride + bstride) + p2;
    float* cdf = smem + (size_t)warp_in_block * per_warp;
    float* bins = cdf + rpw * cstride;
    float* sortbuf = bins + rpw * bstride;

    const int W = C - 1;
    const int S = C + 1;  // only meaningful when HIER
    const int64_t ray0 = ((int64_t)blockIdx.x * warps_per_block + warp_in_block) * rpw;
    if (ray0 >= R) return;
    const int nrays = (int)min((int64_t)rpw, (int64_t)R - ray0);

    // Phase A: stage weights (+1e-5, renderer.py:593) and bins, coalesced.
    for (int j = 0; j < nrays; ++j) {
        const int64_t ray = ray0 + j;
        const float* wrow = HIER ? weights + ray * S + 1 : weights + ray * W;
        const float* brow = HIER ? src + ray * S : src + ray * C;
        float* c = cdf + j * cstride;
        float* b = bins + j * bstride;
        for (int i = lane; i < W; i += 32) c[i + 1] = __fadd_rn(wrow[i], 1e-5f);
        for (int i = lane; i < C; i += 32) {
            if (HIER) b[i] = __fmul_rn(0.5f, __fadd_rn(brow[i + 1], brow[i]));   // renderer.py:127
            else b[i] = brow[i];
        }
    }
    __syncwarp();

    // Phase B: lane l owns ray l: sequential sum, pdf, sequential cumsum (renderer.py:594-596).
    if (lane < nrays) {
        float* c = cdf + lane * cstride;
        float total = 0.0f;
        for (int i = 0; i < W; ++i) total = __fadd_rn(total, c[i + 1]);
        float run = 0.0f;
        c[0] = 0.0f;
        for (int i = 0; i < W; ++i) {
            float p = __fdiv_rn(c[i + 1], total);
            run = __fadd_rn(run, p);
            c[i + 1] = run;
        }
    }
    __syncwarp();

    // Phase C: the warp walks its rays; 32 inverse-CDF lookups per step (renderer.py:599-616).
    for (int j = 0; j < nrays; ++j) {
        const int64_t ray = ray0 + j;
        const float* c = cdf + j * cstride;
        const float* b = bins + j * bstride;
        for (int n = lane; n < N; n += 32) {
            float u = det ? ntx_linspace01(n, N) : u_in[ray * N + n];
            int lo = 0, hi = C;  // upper_bound == searchsorted(side='right')
            while (lo < hi) {
                int mid = (lo + hi) >> 1;
                if (c[mid] <= u) lo = mid + 1; else hi = mid;
            }
            int below = max(lo - 1, 0);
            int above = min(lo, C - 1);
            float cb = c[below];
            float denom = __fsub_rn(c[above], cb);
            if (denom < 1e-5f) denom = 1.0f;
            float tt = __fdiv_rn(__fsub_rn(u, cb), denom);
            float bb = b[below];
            float s = __fadd_rn(bb, __fmul_rn(tt, __fsub_rn(b[above], bb)));
            if (samples) samples[ray * N + n] = s;
            if (inds) inds[ray * N + n] = lo;
            if (HIER) sortbuf[S + n] = s;
        }
        if (HIER) {
            // renderer.py:130: sort(concat([z_vals, z_samples])).
            const float* zrow = src + ray * S;
            float* orow = z_out + ray * (int64_t)(S + N);
            for (int i = lane; i < S; i += 32) sortbuf[i] = zrow[i];
            __syncwarp();
            // Fast path (the usual case: stratified depths are increasing, and so are the inverse-CDF samples of
            // increasing u): both runs are already sorted, so every element's final position is its own index
            // plus its rank in the other run (ties: depths first).  Any unsorted pair falls back to the network.
            bool sorted = true;
            for (int i = lane + 1; i < S; i += 32) sorted &= sortbuf[i] >= sortbuf[i - 1];
            for (int n = lane + 1; n < N; n += 32) sorted &= sortbuf[S + n] >= sortbuf[S + n - 1];
            if (__all_sync(0xffffffffu, sorted)) {
                const float* zs = sortbuf;            // [S] sorted depths
                const float* sm = sortbuf + S;        // [N] sorted samples
                for (int i = lane; i < S; i += 32) {  // # samples strictly below z_i
                    const float v = zs[i];
                    int lo = 0, hi = N;
                    while (lo < hi) { const int mid = (lo + hi) >> 1; if (sm[mid] < v) lo = mid + 1; else hi = mid; }
                    orow[i + lo] = v;
                }
                for (int n = lane; n < N; n += 32) {  // # depths at or below s_n
                    const float v = sm[n];
                    int lo = 0, hi = S;
                    while (lo < hi) { const int mid = (lo + hi) >> 1; if (zs[mid] <= v) lo = mid + 1; else hi = mid; }
                    orow[n + lo] = v;
                }
                __syncwarp();
                continue;
            }
            for (int i = S + N + lane; i < p2; i += 32) sortbuf[i] = CUDART_INF_F;
            __syncwarp();
            for (int k = 2; k <= p2; k <<= 1) {
                for (int jj = k >> 1; jj > 0; jj >>= 1) {
                    for (int tix = lane; tix < (p2 >> 1); tix += 32) {
                        int i = ((tix / jj) * 2 * jj) + (tix % jj);
                        int l = i + jj;
                        bool up = ((i & k) == 0);
                        float a = sortbuf[i], bv = sortbuf[l];
                        if ((a > bv) == up) { sortbuf[i] = bv; sortbuf[l] = a; }
                    }
                    __syncwarp();
                }
            }
            for (int i = lane; i < S + N; i += 32) orow[i] = sortbuf[i];
            __syncwarp();
        }
    }
}

int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

template <bool HIER>
int launch_sample(const float* src, const float* weights, const float* u, int R, int C, int N, int det,
                  float* samples, int32_t* inds, float* z_out, cudaStream_t st) {
    if (R == 0 || N == 0) return NTX_OK;
    const int cstride = (C + 1) | 1;   // odd stride: lane-per-ray serial scans hit distinct banks
    const int bstride = C | 1;
    const int p2 = HIER ? next_pow2(C + 1 + N) : 0;
    const int warps = 2;
    const int budget_floats = 46 * 1024 / 4 / warps;   // stay under the 48 KB default limit
    int rpw = 32;
    while (rpw > 1 && rpw * (cstride + bstride) + p2 > budget_floats) rpw >>= 1;
    // small batches: fewer rays per warp so that every SM gets several warps (the per-ray work of a warp is serial)
    while (rpw > 1 && (int64_t)R / (rpw * warps) < 4LL * ntx_sm_count()) rpw >>= 1;
    const size_t smem = (size_t)warps * (rpw * (cstride + bstride) + p2) * sizeof(float);
    NTX_REQUIRE(smem <= 46 * 1024, "sample_pdf: C=%d N=%d needs %zu B of shared memory", C, N, smem);
    const int64_t rays_per_block = (int64_t)rpw * warps;
    const int64_t blocks = (R + rays_per_block - 1) / rays_per_block;
    sample_pdf_kernel<HIER><<<(unsigned)blocks, warps * 32, smem, st>>>(src, weights, u, R, C, N, det, rpw, cstride,
                                                                        bstride, p2, samples, inds, z_out);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

}  // namespace

NTX_API int ntx_stratified_z(const float* t, const float* z_rand, int R, int S, int perturb, float* z, void* stream) {
    NTX_REQUIRE(R >= 0 && S >= 1, "stratified_z: bad shape R=%d S=%d", R, S);
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(t && z, "stratified_z: null pointer");
    NTX_REQUIRE(!perturb || z_rand, "stratified_z: perturb requires z_rand");
    const int64_t total = (int64_t)R * S;
    const int threads = 256;
    const int64_t blocks = (total + threads - 1) / threads;
    stratified_z_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(t, z_rand, total, S, perturb, z);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_sample_pdf(const float* bins, const float* weights, const float* u, int R, int C, int N, int det,
                           float* samples, int32_t* inds, void* stream) {
    NTX_REQUIRE(R >= 0 && C >= 2 && C <= 1025 && N >= 0, "sample_pdf: bad shape R=%d C=%d N=%d", R, C, N);
    if (R == 0 || N == 0) return NTX_OK;
    NTX_REQUIRE(bins && weights && samples, "sample_pdf: null pointer");
    NTX_REQUIRE(det || u, "sample_pdf: random sampling (det=0) requires the uniform draws u");
    return launch_sample<false>(bins, weights, u, R, C, N, det, samples, inds, nullptr, (cudaStream_t)stream);
}

NTX_API int ntx_hierarchical_z(const float* z, const float* weights, const float* u, int R, int S, int N, int det,
                               float* z_out, float* z_samples, int32_t* inds, void* stream) {
    NTX_REQUIRE(R >= 0 && S >= 3 && N >= 1 && S + N <= 1024, "hierarchical_z: bad shape R=%d S=%d N=%d", R, S, N);
    if (R == 0) return NTX_OK;
    NTX_REQUIRE(z && weights && z_out, "hierarchical_z: null pointer");
    NTX_REQUIRE(det || u, "hierarchical_z: random sampling (det=0) requires the uniform draws u");
    return launch_sample<true>(z, weights, u, R, S - 1, N, det, z_samples, inds, z_out, (cudaStream_t)stream);
}

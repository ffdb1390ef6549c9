// K3 — stratified depths, inverse-CDF sampling (sample_pdf) and the hierarchical merge.
//
// Bit-exactness contract: every floating-point op below is an explicit round-to-nearest
// intrinsic (__fadd_rn / __fmul_rn / __fdiv_rn never contract into FMAs) issued in the same
// left-to-right order as oracle/nerftex_oracle.py and oracle/ntx_oracle.c.
//
// Layout: a warp owns `rpw` consecutive rays.  Their weights are staged in shared memory with
// coalesced loads, lane l runs the strictly sequential normalising sum and cumsum of ray l (the
// serial 62-long chains cost little when 32 rays run side by side; the divisions between them are
// spread over the warp), then the whole warp cooperates on each ray: 32 branch-free upper-bound
// searches at a time and, for the hierarchical variant, a rank merge of the S coarse and N new
// depths staged in shared memory and written back with 16-byte stores.
#include <math_constants.h>
#include <stdlib.h>

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
// C_CT / N_CT: compile-time copies of C / N for the common 64 + 128 configuration (0 = run-time sizes).
struct SampleSmem {           // per-warp layout, in floats
    int cdf, zst, zb, mids, sm, merged, end, total;
};
__host__ __device__ inline int pow2_at_least(int v) { int p = 1; while (p < v) p <<= 1; return p; }
__host__ __device__ inline SampleSmem sample_smem(bool hier, int C, int N, int rpw, int cstride) {
    SampleSmem L;
    const int S = C + 1;
    const int zpad = pow2_at_least(S) > S + 2 ? pow2_at_least(S) : S + 2;   // sortable, and two +inf sentinels
    L.cdf = 0;
    L.zst = (rpw * cstride + 3) & ~3;                       // staged z rows of all rays (16-byte stores)
    L.zb = L.zst + (hier ? rpw * S : 0);                    // working copy of one z row, or the bins of one ray
    L.mids = L.zb + (hier ? zpad : C);
    L.sm = L.mids + (hier ? C : 0);                         // samples (sortable)
    L.merged = (L.sm + (hier ? pow2_at_least(N) : 0) + 3) & ~3;   // read back with 16-byte loads
    L.end = L.merged + (hier ? S + N : 0);
    L.total = L.end + (hier ? S + 1 : 0);
    L.total = (L.total + 3) & ~3;                           // keep every warp's block 16-byte aligned
    return L;
}

// In-warp bitonic sort of a power-of-two long shared-memory array.
__device__ __forceinline__ void warp_bitonic(float* a, int p2, int lane) {
    for (int k = 2; k <= p2; k <<= 1) {
        for (int jj = k >> 1; jj > 0; jj >>= 1) {
            for (int tix = lane; tix < (p2 >> 1); tix += 32) {
                const int i = ((tix / jj) * 2 * jj) + (tix % jj);
                const int l = i + jj;
                const bool up = ((i & k) == 0);
                const float x = a[i], y = a[l];
                if ((x > y) == up) { a[i] = y; a[l] = x; }
            }
            __syncwarp();
        }
    }
}

template <bool HIER, int C_CT, int N_CT>
__global__ void __launch_bounds__(128)
sample_pdf_kernel(const float* __restrict__ src, const float* __restrict__ weights, const float* __restrict__ u_in, int R,
                  int C_rt, int N_rt, int det, int rpw, int cstride, float* __restrict__ samples,
                  int32_t* __restrict__ inds, float* __restrict__ z_out) {
    extern __shared__ __align__(16) float smem[];
    const int C = C_CT ? C_CT : C_rt;
    const int N = N_CT ? N_CT : N_rt;
    const int lane = threadIdx.x & 31;
    const int warp_in_block = threadIdx.x >> 5;
    const int warps_per_block = blockDim.x >> 5;
    const SampleSmem L = sample_smem(HIER, C, N, rpw, cstride);
    float* base = smem + (size_t)warp_in_block * L.total;
    float* cdf = base + L.cdf;
    float* zst = base + L.zst;
    float* zb = base + L.zb;
    float* mids = base + L.mids;
    float* sm = base + L.sm;
    float* merged = base + L.merged;
    int* endv = reinterpret_cast<int*>(base + L.end);

    const int W = C - 1;
    const int S = C + 1;  // only meaningful when HIER
    const int64_t ray0 = ((int64_t)blockIdx.x * warps_per_block + warp_in_block) * rpw;
    if (ray0 >= R) return;
    const int nrays = (int)min((int64_t)rpw, (int64_t)R - ray0);

    // Phase A: stage weights + 1e-5 (renderer.py:593) and, for the hierarchical variant, the z rows.
    if (HIER && (S & 3) == 0 && ((reinterpret_cast<uintptr_t>(weights) | reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
        // the rays of a warp are one contiguous block of both arrays: 16-byte loads, many in flight.  Slots 0 and S-1
        // of a cdf row receive the two unused end weights; slot 0 is overwritten below, slot S-1 is never read.
        const float4* w4 = reinterpret_cast<const float4*>(weights + ray0 * S);
        const float4* z4 = reinterpret_cast<const float4*>(src + ray0 * S);
        const int count4 = nrays * (S >> 2);
#pragma unroll 4
        for (int e = lane; e < count4; e += 32) {
            const float4 wv = w4[e];
            const float4 zv = z4[e];
            const int idx = e << 2;
            const int j = idx / S, i = idx - j * S;
            float* c = cdf + j * cstride + i;
            c[0] = __fadd_rn(wv.x, 1e-5f);
            c[1] = __fadd_rn(wv.y, 1e-5f);
            c[2] = __fadd_rn(wv.z, 1e-5f);
            c[3] = __fadd_rn(wv.w, 1e-5f);
            reinterpret_cast<float4*>(zst)[e] = zv;
        }
    } else {
        for (int j = 0; j < nrays; ++j) {
            const int64_t ray = ray0 + j;
            const float* wrow = HIER ? weights + ray * S + 1 : weights + ray * W;
            float* c = cdf + j * cstride;
            for (int i = lane; i < W; i += 32) c[i + 1] = __fadd_rn(wrow[i], 1e-5f);
            if (HIER)
                for (int i = lane; i < S; i += 32) zst[j * S + i] = src[ray * S + i];
        }
    }
    __syncwarp();
    // Phase B (renderer.py:594-596): lane l owns ray l for the two strictly sequential chains (normalising sum,
    // cumsum); the divisions in between are independent, so the whole warp shares them.
    float total = 0.0f;
    if (lane < nrays) {
        const float* c = cdf + lane * cstride;
        for (int i = 0; i < W; ++i) total = __fadd_rn(total, c[i + 1]);
    }
    for (int j = 0; j < nrays; ++j) {
        const float tj = __shfl_sync(0xffffffffu, total, j);
        float* c = cdf + j * cstride;
        for (int i = lane; i < W; i += 32) c[i + 1] = __fdiv_rn(c[i + 1], tj);
    }
    __syncwarp();
    if (lane < nrays) {
        float* c = cdf + lane * cstride;
        float run = 0.0f;
        c[0] = 0.0f;
        for (int i = 0; i < W; ++i) {
            run = __fadd_rn(run, c[i + 1]);
            c[i + 1] = run;
        }
    }
    __syncwarp();

    // Phase C: the warp walks its rays; 32 inverse-CDF lookups per step (renderer.py:599-616).
    int top = 1;                                   // largest power of two <= C
    while (top * 2 <= C) top <<= 1;
    const float delta = (N > 1) ? __fdiv_rn(1.0f, (float)(N - 1)) : 0.0f;       // tf.linspace(0, 1, N) step
    for (int j = 0; j < nrays; ++j) {
        const int64_t ray = ray0 + j;
        const float* c = cdf + j * cstride;
        const float* b;
        bool zs_ok = true;
        if (HIER) {
            const float* zr = zst + j * S;
            for (int i = lane; i < S; i += 32) {
                const float v = zr[i];
                zb[i] = v;
                if (i > 0) zs_ok &= v >= zr[i - 1];
                if (i < C) mids[i] = __fmul_rn(0.5f, __fadd_rn(zr[i + 1], v));   // renderer.py:127
            }
            if (lane < 2) zb[S + lane] = CUDART_INF_F;
            for (int i = lane; i <= S; i += 32) endv[i] = 0;
            zs_ok = __all_sync(0xffffffffu, zs_ok);
            b = mids;
        } else {
            const float* brow = src + ray * C;
            for (int i = lane; i < C; i += 32) zb[i] = brow[i];
            b = zb;
        }
        __syncwarp();
        // renderer.py:130 wants sort(concat([z_vals, z_samples])).  Both runs are normally sorted already (stratified
        // depths increase, and so do the inverse-CDF samples of increasing u), so every element's final position is
        // its own index plus its rank in the other run (ties: depths first).  The sample loop below does that merge on
        // the fly and checks the assumption; if it fails, the slow path after the loop redoes the merge.
        bool sm_ok = true;
        float carry_s = -CUDART_INF_F;
        for (int n0 = 0; n0 < N; n0 += 32) {
            const int n = n0 + lane;
            const bool live = n < N;
            float s = CUDART_INF_F;
            if (live) {
                const float u = det ? ((n == 0) ? 0.0f : (n == N - 1) ? 1.0f : __fmul_rn(delta, (float)n)) : u_in[ray * N + n];
                int lo = 0;                            // upper_bound == searchsorted(side='right'), branch-free
                for (int step = top; step > 0; step >>= 1) {
                    const int t = lo + step;
                    if (t <= C && c[t - 1] <= u) lo = t;
                }
                const int below = max(lo - 1, 0);
                const int above = min(lo, C - 1);
                const float cb = c[below];
                float denom = __fsub_rn(c[above], cb);
                if (denom < 1e-5f) denom = 1.0f;
                const float tt = __fdiv_rn(__fsub_rn(u, cb), denom);
                const float bb = b[below];
                s = __fadd_rn(bb, __fmul_rn(tt, __fsub_rn(b[above], bb)));
                if (samples) samples[ray * N + n] = s;
                if (inds) inds[ray * N + n] = lo;
                if (HIER) {
                    sm[n] = s;
                    if (zs_ok) {
                        // rank among the depths, #{z <= s}: s lies between the mid points around z[lo], so it is
                        // `below` plus at most three (zb carries two +inf sentinels); anything else walks.
                        const float z0 = zb[below], z1 = zb[below + 1], z2 = zb[below + 2];
                        const int cnt = (z0 <= s) + (z1 <= s) + (z2 <= s);
                        int r = below + cnt;
                        if ((cnt == 3 && zb[below + 3] <= s) || (cnt == 0 && below > 0 && zb[below - 1] > s)) {
                            while (r < S && zb[r] <= s) ++r;
                            while (r > 0 && zb[r - 1] > s) --r;
                        }
                        merged[n + r] = s;
                        atomicMax(&endv[r], n + 1);
                    }
                }
            }
            if (HIER) {
                float prev = __shfl_up_sync(0xffffffffu, s, 1);
                if (lane == 0) prev = carry_s;
                sm_ok &= s >= prev;                    // dead lanes hold +inf
                carry_s = __shfl_sync(0xffffffffu, s, 31);
            }
        }
        if (!HIER) { __syncwarp(); continue; }
        sm_ok = __all_sync(0xffffffffu, sm_ok);
        __syncwarp();
        if (!(zs_ok && sm_ok)) {                       // slow path: sort what is unsorted, ranks by binary search
            for (int i = lane; i <= S; i += 32) endv[i] = 0;
            if (!sm_ok) {
                const int p2 = pow2_at_least(N);
                for (int i = N + lane; i < p2; i += 32) sm[i] = CUDART_INF_F;
                __syncwarp();
                warp_bitonic(sm, p2, lane);
            }
            if (!zs_ok) {
                const int p2 = pow2_at_least(S);
                for (int i = S + lane; i < p2; i += 32) zb[i] = CUDART_INF_F;
                __syncwarp();
                warp_bitonic(zb, p2, lane);
            }
            __syncwarp();
            for (int n = lane; n < N; n += 32) {
                const float s = sm[n];
                int lo = 0, hi = S;
                while (lo < hi) { const int mid = (lo + hi) >> 1; if (zb[mid] <= s) lo = mid + 1; else hi = mid; }
                merged[n + lo] = s;
                atomicMax(&endv[lo], n + 1);
            }
            __syncwarp();
        }
        // depth i moves up by #{s < z_i} = #{n : rank_n <= i} = max over r <= i of endv[r] (ranks are non-decreasing)
        int carry = 0;
        for (int i0 = 0; i0 < S; i0 += 32) {
            const int i = i0 + lane;
            int v = (i < S) ? endv[i] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v = max(v, t);
            }
            v = max(v, carry);
            if (i < S) merged[i + v] = zb[i];
            carry = __shfl_sync(0xffffffffu, v, 31);
        }
        __syncwarp();
        float* orow = z_out + ray * (int64_t)(S + N);
        if (((S + N) & 3) == 0 && (reinterpret_cast<uintptr_t>(z_out) & 15) == 0) {
            for (int i = lane; i < ((S + N) >> 2); i += 32)
                reinterpret_cast<float4*>(orow)[i] = reinterpret_cast<const float4*>(merged)[i];
        } else {
            for (int i = lane; i < S + N; i += 32) orow[i] = merged[i];
        }
        __syncwarp();
    }
}

template <bool HIER, int C_CT, int N_CT>
int launch_sample_t(const float* src, const float* weights, const float* u, int R, int C, int N, int det, float* samples,
                    int32_t* inds, float* z_out, cudaStream_t st) {
    const int cstride = (C + 1) | 1;   // odd stride: lane-per-ray serial scans hit distinct banks
    int warps = 4, rpw = HIER ? 8 : 32;          // 8: measured best (occupancy vs idle lanes in phase B)
    if (const char* e = getenv("NTX_K3_RPW")) rpw = max(1, min(32, atoi(e)));   // tuning knob (rays per warp)
    auto bytes = [&](int rpw_, int warps_) { return (size_t)warps_ * sample_smem(HIER, C, N, rpw_, cstride).total * sizeof(float); };
    while (rpw > 1 && bytes(rpw, warps) > 46 * 1024) rpw >>= 1;     // stay under the 48 KB default limit
    while (warps > 1 && bytes(rpw, warps) > 46 * 1024) warps >>= 1;
    // small batches: fewer rays per warp so that every SM gets several warps (the per-ray work of a warp is serial)
    while (rpw > 1 && (int64_t)R / (rpw * warps) < 4LL * ntx_sm_count()) rpw >>= 1;
    const size_t smem = bytes(rpw, warps);
    NTX_REQUIRE(smem <= 46 * 1024, "sample_pdf: C=%d N=%d needs %zu B of shared memory", C, N, smem);
    const int64_t rays_per_block = (int64_t)rpw * warps;
    const int64_t blocks = (R + rays_per_block - 1) / rays_per_block;
    sample_pdf_kernel<HIER, C_CT, N_CT><<<(unsigned)blocks, warps * 32, smem, st>>>(src, weights, u, R, C, N, det, rpw, cstride,
                                                                                   samples, inds, z_out);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

template <bool HIER>
int launch_sample(const float* src, const float* weights, const float* u, int R, int C, int N, int det,
                  float* samples, int32_t* inds, float* z_out, cudaStream_t st) {
    if (R == 0 || N == 0) return NTX_OK;
    if (HIER && C == 63 && N == 128)    // Renderer defaults of every shipped config: n_samples 64, n_importance 128
        return launch_sample_t<HIER, 63, 128>(src, weights, u, R, C, N, det, samples, inds, z_out, st);
    return launch_sample_t<HIER, 0, 0>(src, weights, u, R, C, N, det, samples, inds, z_out, st);
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

// K1-bwd — backward of the fused MLP on the operands saved by ntx_mlp_forward_train.
//
//   mlp_dgrad_kernel  the data-gradient chain, fused like the forward: for every 128-row tile the
//                     gradient w.r.t. each layer's pre-activation output is produced layer by layer
//                     in reverse, dY_{s-1} = relu'(act_{s-1}) * (dY_s W_s^T [+ d sigma w_alpha^T]),
//                     with tcgen05.mma (bf16 in, fp32 accumulate in TMEM), gradients staying in
//                     shared memory between layers and each dY_s leaving once, as a tile image,
//                     through a bulk store.  Same roles / CTA-pair protocol as the forward kernel.
//   mlp_wgrad_kernel  dW_s = X_s^T dY_s and db_s = 1^T dY_s over all rows.  Both operands are the saved
//                     tile images read as MN-major UMMA operands (K = sample rows): no transposes, no
//                     re-layout.  One CTA per SM accumulates a full [256 x 256] fp32 dW in TMEM (all
//                     512 columns) over its share of the tiles, then adds it to the flat gradient
//                     buffer with red.global.add; idle warps sum dY columns for the bias gradient.
//
// No gradient flows into the encodings (nothing trainable upstream; network/renderer.py:129).
// Reference: the tf.GradientTape of network/train.py:62-66 over network/model.py:104-123.
#include <math_constants.h>
#include <stdlib.h>

#include "ntx_common.cuh"
#include "ntx_mlp.cuh"
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include "ntx_mlp_dev.cuh"
#include "ntx_umma.cuh"

namespace {

using namespace umma;
using namespace ntxmlp;

constexpr int kThreadsBwd = 320;
constexpr int kMaxBwdSteps = kMaxSteps;

struct BwdStep {
    int32_t fwd_step;      // forward step s whose weights are applied: dY_{s-1} = f(dY_s W_s^T)
    int32_t n_k;           // K-slices of 32 = n_out(s) / 32
    int32_t relu_prev;     // step s-1 ended in ReLU (mask its gradient)
    int32_t add_alpha;     // step s-1 is the last trunk layer: the density head's gradient joins here
    uint32_t slice_off;    // byte offset of this step's first slice in the packed transposed weights
    int32_t pad_[3];
};

struct BwdProgram {
    int32_t n_bwd;         // MMA steps of the chain = n_steps - 1
    int32_t n_steps;
    int32_t last_n_out;    // width of the last forward step (128)
    int32_t pad_;
    BwdStep steps[kMaxBwdSteps];
};

constexpr uint32_t kBwdHeadFloats = 256 + 3 * 128;      // w_alpha[256], W_rgb as [3][128]
constexpr uint32_t kBwdHeadBytes = 4096;

// shared-memory map of the data-gradient kernel
constexpr uint32_t kDgOffStages = kTilesPerCta * kMainBytes;                  // 131072
constexpr uint32_t kDgOffHeads = kDgOffStages + kStages * kStageBytes;        // +49152
constexpr uint32_t kDgOffBars = kDgOffHeads + kBwdHeadFloats * 4;
constexpr uint32_t kDgNumBars = 2 * kStages + 2 * kTilesPerCta;
constexpr uint32_t kDgOffTmemSlot = kDgOffBars + kDgNumBars * 8;
constexpr uint32_t kDgOffAmax = kDgOffTmemSlot + 16;                          // [16] per-step max |dY| (8-bit images)
constexpr uint32_t kDgSmemBytes = kDgOffAmax + 64;

struct DgradArgs {
    CUtensorMap w_map;           // CTA pairs: packed_t as rows of 2 KB, box = this CTA's 8 KB half of a slice
    const uint8_t* packed_t;     // head block + transposed weight slices
    const float* d_rgbs;         // [N,4]
    const uint4* masks;          // [n_steps][n_tiles*128][2 x uint4]
    uint8_t* dy_img;             // [n_steps][n_tiles][64 KB] out
    int64_t n_rows, n_tiles;
    int32_t depth;
    // 8-bit variant: dy_img8[n_steps][n_tiles][32 KB] e4m3 images of dY / scale8[step]; the bf16 image of a step is then
    // written only if bit `step` of need16 is set (the steps whose encoded-input weight gradient needs it)
    uint32_t need16;
    uint8_t* dy_img8;
    const float* inv_scale8;
    uint32_t* amax8;
};

template <int CTAS, bool F8 = false>
__global__ void __launch_bounds__(kThreadsBwd, 1)
mlp_dgrad_kernel(const __grid_constant__ BwdProgram prog, const __grid_constant__ DgradArgs args) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    if ((sbase & 1023u) != 0) __trap();
    constexpr uint32_t NSTAGE = kStages;
    constexpr uint32_t STAGE_BYTES = kStageBytes;
    constexpr int SPS = CTAS;
    const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar_full = sbase + kDgOffBars;
    const uint32_t bar_empty = bar_full + 8 * kStages;
    const uint32_t bar_act = bar_empty + 8 * kStages;
    const uint32_t bar_acc = bar_act + 8 * kTilesPerCta;
    float* heads = reinterpret_cast<float*>(smem + kDgOffHeads);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kDgOffTmemSlot);

    const int64_t n_rows = args.n_rows;
    const int64_t n_batches = (n_rows + kRowsPerCta * CTAS - 1) / (kRowsPerCta * CTAS);
    const int64_t batch0 = blockIdx.x / CTAS, batch_stride = gridDim.x / CTAS;

    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i < NSTAGE; ++i) {
            mbar_init(bar_full + 8 * i, 1);             // pair: both CTAs' copies credit the issuing CTA's barrier
            mbar_init(bar_empty + 8 * i, 1);
        }
        for (int t = 0; t < kTilesPerCta; ++t) {
            mbar_init(bar_act + 8 * t, 4 * CTAS);
            mbar_init(bar_acc + 8 * t, 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        if (CTAS == 2) { tmem_alloc2(sbase + kDgOffTmemSlot, 512); tmem_relinquish2(); }
        else { tmem_alloc(sbase + kDgOffTmemSlot, 512); tmem_relinquish(); }
    }
    for (int i = threadIdx.x; i < (int)kBwdHeadFloats; i += kThreadsBwd)
        heads[i] = reinterpret_cast<const float*>(args.packed_t)[i];
    if (F8 && threadIdx.x < 16) reinterpret_cast<uint32_t*>(smem + kDgOffAmax)[threadIdx.x] = 0u;
    tc_fence_before();
    __syncthreads();
    if (CTAS == 2) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== weight producer
        if (lane == 0) {
            uint32_t stage = 0, parity = 1;
            const uint32_t bytes = kStageBytes / CTAS;                         // this CTA's rows of one [256 x 32] slice
            const uint32_t leader_full = CTAS == 2 ? mapa_shared(bar_full, 0) : bar_full;
            if (CTAS == 2) prefetch_tensormap(&args.w_map);
            for (int64_t b = batch0; b < n_batches; b += batch_stride) {
                for (int bs = 0; bs < prog.n_bwd; ++bs) {
                    const int n_k = prog.steps[bs].n_k;
                    const uint8_t* src0 = args.packed_t + prog.steps[bs].slice_off + rank * bytes;
                    const int32_t wrow0 = (int32_t)((prog.steps[bs].slice_off + rank * bytes) >> 11);   // rows of 2 KB
                    for (int t = 0; t < kTilesPerCta; ++t) {
                        const uint8_t* src = src0;
                        int32_t wrow = wrow0;
                        for (int left = n_k; left > 0;) {
                            const int n = left < SPS ? left : SPS;
                            mbar_wait(bar_empty + 8 * stage, parity);
                            if (CTAS == 2) {
                                // (see the forward kernel: the bytes of both CTAs' copies go to the issuer's barrier)
                                if (rank == 0) mbar_arrive_expect_tx(bar_full + 8 * stage, 2 * bytes * n);
                                for (int j = 0; j < n; ++j, wrow += (int32_t)(kStageBytes >> 11))
                                    tma_tile2d_g2s_pair(sbase + kDgOffStages + stage * STAGE_BYTES + j * bytes, &args.w_map, 0, wrow,
                                                        leader_full + 8 * stage);
                            } else {
                            mbar_arrive_expect_tx(bar_full + 8 * stage, bytes * n);
                            for (int j = 0; j < n; ++j, src += kStageBytes)
                                tma_bulk_g2s(sbase + kDgOffStages + stage * STAGE_BYTES + j * bytes, src, bytes,
                                             bar_full + 8 * stage);
                            }
                            if (++stage == NSTAGE) { stage = 0; parity ^= 1; }
                            left -= n;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (CTAS == 2 && rank != 0) {
            // peer CTA of a pair: nothing to issue (its copies credit the issuing CTA's barriers themselves)
        } else {
            // ================================================================= MMA issuer
            uint32_t stage = 0, parity = 0, act_phase0 = 0, act_phase1 = 0;
            uint32_t b16 = (sbase + kDgOffStages) >> 4;
            uint32_t full_bar = bar_full, empty_bar = bar_empty;
            const uint32_t idesc = idesc_bf16_f32(kTileRows * CTAS, kWidth);
            const uint32_t b_slice16 = (kStageBytes / CTAS) >> 4;
            for (int64_t b = batch0; b < n_batches; b += batch_stride) {
                for (int bs = 0; bs < prog.n_bwd; ++bs) {
                    const int n_k = prog.steps[bs].n_k;
#pragma unroll
                    for (int t = 0; t < kTilesPerCta; ++t) {
                        if (t == 0) { mbar_wait(bar_act, act_phase0); act_phase0 ^= 1; }
                        else { mbar_wait(bar_act + 8, act_phase1); act_phase1 ^= 1; }
                        tc_fence_after();
                        const uint32_t d_tmem = tmem_base + t * kWidth;
                        const uint32_t a16_tile = (sbase + t * kMainBytes) >> 4;
                        for (int i = 0; i < n_k; i += SPS) {
                            const int n = n_k - i < SPS ? n_k - i : SPS;
                            mbar_wait(full_bar, parity);
                            tc_fence_after();
                            const uint32_t bb = b16, eb = empty_bar;
                            if (++stage == NSTAGE) {
                                stage = 0; parity ^= 1; b16 = (sbase + kDgOffStages) >> 4; full_bar = bar_full; empty_bar = bar_empty;
                            } else {
                                b16 += STAGE_BYTES >> 4; full_bar += 8; empty_bar += 8;
                            }
                            if (elect_one_sync()) {
#pragma unroll
                                for (int j = 0; j < SPS; ++j) {
                                    if (j < n) {
                                        const uint32_t aj = a16_tile + (i + j) * (kBlockBytes >> 4), bj = bb + j * b_slice16;
                                        const uint32_t acc0 = (i + j) > 0 ? 1u : 0u;
                                        if (CTAS == 2) {
                                            mma_bf16_ss_2cta(d_tmem, desc_sw64(aj), desc_sw64(bj), idesc, acc0);
                                            mma_bf16_ss_2cta(d_tmem, desc_sw64(aj + 2), desc_sw64(bj + 2), idesc, 1u);
                                        } else {
                                            mma_bf16_ss(d_tmem, desc_sw64(aj), desc_sw64(bj), idesc, acc0);
                                            mma_bf16_ss(d_tmem, desc_sw64(aj + 2), desc_sw64(bj + 2), idesc, 1u);
                                        }
                                    }
                                }
                                if (CTAS == 2) mma_commit_2cta(eb); else mma_commit(eb);
                            }
                            __syncwarp();
                        }
                        if (elect_one_sync()) {
                            if (CTAS == 2) mma_commit_2cta(bar_acc + 8 * t); else mma_commit(bar_acc + 8 * t);
                        }
                        __syncwarp();
                    }
                }
            }
        }
    } else {
        // ===================================================================== epilogue warpgroups
        const int ew = warp - 2;
        const int tile = ew >> 2;
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const uint32_t main_base = sbase + tile * kMainBytes;
        const uint32_t tmem_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + tile * kWidth;
        const uint32_t my_act = bar_act + 8 * tile, my_acc = bar_acc + 8 * tile;
        const float* w_alpha = heads;              // [256]
        const float* w_rgb = heads + 256;          // [3][128]: w_rgb[j*128 + c] = W_rgb[c][j]
        uint32_t acc_phase = 0;
        const uint32_t act_remote = (CTAS == 2 && rank != 0) ? mapa_shared(my_act, 0) : 0u;
        auto arrive_act = [&]() {
            if (CTAS == 2 && rank != 0) mbar_arrive_remote(act_remote);
            else mbar_arrive(my_act);
        };
        const int last = prog.n_steps - 1;

        for (int64_t b = batch0; b < n_batches; b += batch_stride) {
            const int64_t tile_g = (b * CTAS + rank) * kTilesPerCta + tile;
            const int64_t grow = tile_g * kTileRows + row;
            const bool valid = grow < n_rows;
            float4 dr = make_float4(0.f, 0.f, 0.f, 0.f);
            if (valid) dr = __ldg(reinterpret_cast<const float4*>(args.d_rgbs) + grow);
            auto mask_ptr = [&](int st) { return args.masks + (((int64_t)st * args.n_tiles + tile_g) * kTileRows + row) * 2; };
            auto img8 = [&](int st) { return args.dy_img8 + ((int64_t)st * args.n_tiles + tile_g) * (int64_t)kMain8Bytes; };
            auto store_image = [&](int st, uint32_t bytes) {
                fence_proxy_async_smem();
                named_bar_sync(1 + tile, kTileRows);
                if (row == 0 && (!F8 || ((args.need16 >> st) & 1u))) {
                    tma_bulk_s2g(args.dy_img + ((int64_t)st * args.n_tiles + tile_g) * (int64_t)kMainBytes, main_base, bytes);
                    tma_store_commit();
                }
            };
            // ---- head of the chain: d(colour half output) = relu' * (d rgb W_rgb^T)   (model.py:122-123)
            {
                if (row == 0) tma_store_wait_read();
                named_bar_sync(1 + tile, kTileRows);
                const uint4 m4 = __ldg(mask_ptr(last));
                const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
                float am = 0.0f;
                const float inv8 = F8 ? __ldg(args.inv_scale8 + last) : 0.0f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float x[32];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        uint32_t w[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int col = j * 32 + c * 8 + 2 * e;
                            float v0 = dr.x * w_rgb[col] + dr.y * w_rgb[128 + col] + dr.z * w_rgb[256 + col];
                            float v1 = dr.x * w_rgb[col + 1] + dr.y * w_rgb[128 + col + 1] + dr.z * w_rgb[256 + col + 1];
                            if (!((mw[j] >> (c * 8 + 2 * e)) & 1u)) v0 = 0.0f;
                            if (!((mw[j] >> (c * 8 + 2 * e + 1)) & 1u)) v1 = 0.0f;
                            w[e] = pack_bf16(v0, v1);
                            x[c * 8 + 2 * e] = v0; x[c * 8 + 2 * e + 1] = v1;
                        }
                        st_shared_v4(chunk_addr(main_base, j * 4 + c, row), w[0], w[1], w[2], w[3]);
                    }
                    if (F8) dump_fp8_chunk32<false>(img8(last), j, row, x, inv8, am, (tile_g & 7) == 0);
                }
                if (F8) amax_to_shared(am, reinterpret_cast<uint32_t*>(smem + kDgOffAmax) + last);
                store_image(last, (uint32_t)prog.last_n_out * kTileRows * 2);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) arrive_act();
            }
            // ---- the chain
            for (int bs = 0; bs < prog.n_bwd; ++bs) {
                const int st_prev = prog.steps[bs].fwd_step - 1;           // gradient produced for this forward step
                const bool relu_prev = prog.steps[bs].relu_prev != 0, add_alpha = prog.steps[bs].add_alpha != 0;
                uint32_t mw[8] = {~0u, ~0u, ~0u, ~0u, ~0u, ~0u, ~0u, ~0u};
                if (relu_prev) {
                    const uint4 a = __ldg(mask_ptr(st_prev)), c4 = __ldg(mask_ptr(st_prev) + 1);
                    mw[0] = a.x; mw[1] = a.y; mw[2] = a.z; mw[3] = a.w; mw[4] = c4.x; mw[5] = c4.y; mw[6] = c4.z; mw[7] = c4.w;
                }
                mbar_wait(my_acc, acc_phase);
                acc_phase ^= 1;
                tc_fence_after();
                if (row == 0) tma_store_wait_read();        // the previous image store has read this buffer
                named_bar_sync(1 + tile, kTileRows);
                float am = 0.0f;
                const float inv8 = F8 ? __ldg(args.inv_scale8 + st_prev) : 0.0f;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint32_t v[32];
                    tmem_ld_32x32b_x32(tmem_row + j * 32, v);
                    tmem_ld_wait();
                    float x[32];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        uint32_t w[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int cc = c * 8 + 2 * e;
                            float v0 = __uint_as_float(v[cc]), v1 = __uint_as_float(v[cc + 1]);
                            if (add_alpha) {                                   // model.py:111: sigma = h w_alpha + b
                                v0 += dr.w * w_alpha[j * 32 + cc];
                                v1 += dr.w * w_alpha[j * 32 + cc + 1];
                            }
                            if (!((mw[j] >> cc) & 1u)) v0 = 0.0f;
                            if (!((mw[j] >> (cc + 1)) & 1u)) v1 = 0.0f;
                            w[e] = pack_bf16(v0, v1);
                            x[cc] = v0; x[cc + 1] = v1;
                        }
                        st_shared_v4(chunk_addr(main_base, j * 4 + c, row), w[0], w[1], w[2], w[3]);
                    }
                    if (F8) dump_fp8_chunk32<false>(img8(st_prev), j, row, x, inv8, am, (tile_g & 7) == 0);
                }
                if (F8) amax_to_shared(am, reinterpret_cast<uint32_t*>(smem + kDgOffAmax) + st_prev);
                store_image(st_prev, kMainBytes);
                if (bs + 1 < prog.n_bwd) {
                    fence_proxy_async_smem();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) arrive_act();
                } else {
                    tc_fence_before();
                }
            }
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }

    tc_fence_before();
    __syncthreads();
    if (F8 && threadIdx.x < 16) {
        const uint32_t a = reinterpret_cast<uint32_t*>(smem + kDgOffAmax)[threadIdx.x];
        if (a) atomicMax(args.amax8 + threadIdx.x, a);
    }
    if (CTAS == 2) cluster_sync_all();
    if (warp == 1) {
        __syncwarp();
        if (CTAS == 2) tmem_dealloc2(tmem_base, 512); else tmem_dealloc(tmem_base, 512);
    }
}

// transposed weights for the chain: B[n = input feature (main rows), k = output feature] = W[main_row0 + n][k]
__global__ void mlp_pack_bwd_kernel(const __grid_constant__ Program prog, const __grid_constant__ BwdProgram bprog,
                                    const float* __restrict__ master, uint8_t* __restrict__ packed_t) {
    const int bs = blockIdx.y;
    if (bs == bprog.n_bwd) {             // head block
        if (blockIdx.x != 0) return;
        float* h = reinterpret_cast<float*>(packed_t);
        const Layer& la = prog.layers[prog.alpha_layer];
        const Layer& lr = prog.layers[prog.rgb_layer];
        for (int i = threadIdx.x; i < (int)(kBwdHeadBytes / 4); i += blockDim.x) {
            float v = 0.0f;
            if (i < 256) v = master[la.w_off + i];
            else if (i < 256 + 384) { const int j = (i - 256) / 128, c = (i - 256) % 128; v = master[lr.w_off + c * 3 + j]; }
            h[i] = v;
        }
        return;
    }
    const BwdStep& B = bprog.steps[bs];
    const int ks = blockIdx.x;
    if (ks >= B.n_k) return;
    const Step& step = prog.steps[B.fwd_step];
    const Layer& L = prog.layers[step.layer];
    uint8_t* dst = packed_t + B.slice_off + (size_t)ks * kStageBytes;
    for (int idx = threadIdx.x; idx < 256 * 32; idx += blockDim.x) {
        const int n = idx >> 5, kk = idx & 31;
        const float v = master[L.w_off + (int64_t)(step.main_row0 + n) * L.fan_out + ks * 32 + kk];
        *reinterpret_cast<__nv_bfloat16*>(dst + sw64_offset(n, kk, 256)) = __float2bfloat16(v);
    }
}

int64_t build_bwd_program(const ntx_mlp* m, BwdProgram* bp) {
    memset(bp, 0, sizeof(*bp));
    const Program& p = m->prog;
    bp->n_steps = p.n_steps;
    bp->n_bwd = p.n_steps - 1;
    bp->last_n_out = p.steps[p.n_steps - 1].n_out;
    uint32_t off = kBwdHeadBytes;
    for (int bs = 0; bs < bp->n_bwd; ++bs) {
        const int st = p.n_steps - 1 - bs;
        BwdStep& B = bp->steps[bs];
        B.fwd_step = st;
        B.n_k = p.steps[st].n_out / 32;
        B.relu_prev = (p.steps[st - 1].epi & EPI_RELU) ? 1 : 0;
        B.add_alpha = (st - 1 == m->cfg.depth - 1) ? 1 : 0;
        B.slice_off = off;
        off += (uint32_t)B.n_k * kStageBytes;
    }
    return off;
}

// ----------------------------------------------------------------------------------------------
// weight gradient
// ----------------------------------------------------------------------------------------------
constexpr int kWgThreads = 192;                      // warp 0 producer, warp 1 MMA issuer, warps 2-5 bias sums + flush
constexpr int kWgStages = 3;
constexpr uint32_t kWgHalfRows = 64;                 // K rows per stage (half a tile)
constexpr uint32_t kWgOperandBytes = 8 * kWgHalfRows * 64;      // 32 KB: 8 blocks of [64 rows x 64 B]
constexpr uint32_t kWgStageBytes = 2 * kWgOperandBytes;         // X half + dY half
constexpr uint32_t kWgOffBias = kWgStages * kWgStageBytes;      // 196608: [4][256] fp32 scratch for the bias reduction
constexpr uint32_t kWgOffBars = kWgOffBias + 8 * 256 * 4;          // (8 row groups in the 8-bit path, 4 in the bf16 path)
constexpr uint32_t kWgOffTmemSlot = kWgOffBars + (2 * kWgStages + 2) * 8;
constexpr uint32_t kWgSmemBytes = kWgOffTmemSlot + 16;
static_assert(kWgSmemBytes <= 232448, "weight-gradient kernel exceeds shared memory");

// one GEMM of the weight-gradient kernel: out[m][n] += sum_rows X[row][m] * dY[row][n]
struct WgJob {
    const uint8_t* x;        // X tile images
    const uint8_t* dy;       // dY tile images
    float* out;              // row m of the result goes to out + m * out_ld
    float* bias;             // 1^T dY goes here (NULL: skip)
    int64_t x_tile_bytes;    // 65536 (activation images) or 32768 (encoding images)
    int32_t x_blocks;        // 64-byte-wide column blocks of X.  bf16: 8 = 256 columns (two accumulators), 4 = 128;
                             // e4m3: 4 = 256 columns (two accumulators), 2 = 128
    int32_t n_out;           // N (256 or 128)
    int32_t out_ld;
    int32_t fp8;             // both operands are e4m3 tile images (32 KB per tile); `scale` rescales the product
    int32_t pad_;
    const float* x_scale;    // device scalars: value = stored e4m3 * scale (NULL: 1)
    const float* dy_scale;
    int64_t dy_tile_bytes;   // 65536 (bf16) or 32768 (e4m3)
};

struct WgradArgs {
    WgJob jobs[kMaxSteps + 4];
    int32_t n_jobs;
    int32_t pad_;
    int64_t n_tiles;             // tiles that hold real rows
};

__global__ void __launch_bounds__(kWgThreads, 1)
mlp_wgrad_kernel(const __grid_constant__ WgradArgs args) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    if ((sbase & 1023u) != 0) __trap();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar_full = sbase + kWgOffBars;               // [3]
    const uint32_t bar_empty = bar_full + 8 * kWgStages;        // [3]  1 commit + 4 bias warps
    const uint32_t bar_acc = bar_empty + 8 * kWgStages;         // accumulators of a job complete
    const uint32_t bar_free = bar_acc + 8;                      // accumulators flushed (4 warps)
    float* bias_scratch = reinterpret_cast<float*>(smem + kWgOffBias);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kWgOffTmemSlot);

    if (threadIdx.x == 0) {
        for (int i = 0; i < kWgStages; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, 5);
        }
        mbar_init(bar_acc, 1);
        mbar_init(bar_free, 4);
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(sbase + kWgOffTmemSlot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // my tiles: blockIdx.x, += gridDim.x; each tile = 2 stages (rows 0..63, 64..127)
    const int64_t my_tiles = args.n_tiles > blockIdx.x ? (args.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    constexpr uint32_t kHalf = kWgHalfRows * 64;                         // bytes of one block's half (64 rows)

    if (warp == 0) {
        if (lane == 0) {
            uint32_t stage = 0, parity = 1;
            for (int jb = 0; jb < args.n_jobs; ++jb) {
                const WgJob& J = args.jobs[jb];
                const int xb = J.x_blocks, n_blk = J.fp8 ? J.n_out / 64 : J.n_out / 32;
                for (int64_t i = 0; i < my_tiles; ++i) {
                    const int64_t tile = blockIdx.x + i * gridDim.x;
                    for (int half = 0; half < 2; ++half) {
                        mbar_wait(bar_empty + 8 * stage, parity);
                        mbar_arrive_expect_tx(bar_full + 8 * stage, (xb + n_blk) * kHalf);
                        const uint32_t dst = sbase + stage * kWgStageBytes;
                        const uint8_t* xt = J.x + tile * J.x_tile_bytes + half * kHalf;
                        const uint8_t* dt = J.dy + tile * J.dy_tile_bytes + half * kHalf;
                        for (int blk = 0; blk < xb; ++blk)
                            tma_bulk_g2s(dst + blk * kHalf, xt + blk * kBlockBytes, kHalf, bar_full + 8 * stage);
                        for (int blk = 0; blk < n_blk; ++blk)
                            tma_bulk_g2s(dst + kWgOperandBytes + blk * kHalf, dt + blk * kBlockBytes, kHalf, bar_full + 8 * stage);
                        if (++stage == kWgStages) { stage = 0; parity ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        uint32_t stage = 0, parity = 0, free_phase = 0;
        for (int jb = 0; jb < args.n_jobs; ++jb) {
            const bool f8 = args.jobs[jb].fp8 != 0;
            const uint32_t idesc = f8 ? idesc_e4m3_f32_mn(128, args.jobs[jb].n_out) : idesc_bf16_f32_mn(128, args.jobs[jb].n_out);
            const bool two = args.jobs[jb].x_blocks == (f8 ? 4 : 8);
            if (jb > 0) { mbar_wait(bar_free, free_phase); free_phase ^= 1; tc_fence_after(); }
            uint32_t accumulate = 0;
            for (int64_t i = 0; i < my_tiles * 2; ++i) {
                mbar_wait(bar_full + 8 * stage, parity);
                tc_fence_after();
                const uint32_t xa = sbase + stage * kWgStageBytes, da = xa + kWgOperandBytes;
                if (elect_one_sync()) {
                    if (f8) {
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {                 // 32 K rows per MMA (kind::f8f6f4)
                            const uint64_t bd = smem_desc(da + ks * 2048, kHalf, 512, LAYOUT_SW64);
                            mma_f8_ss(tmem_base, smem_desc(xa + ks * 2048, kHalf, 512, LAYOUT_SW64), bd, idesc, accumulate);
                            if (two)
                                mma_f8_ss(tmem_base + 256, smem_desc(xa + 2 * kHalf + ks * 2048, kHalf, 512, LAYOUT_SW64), bd, idesc,
                                          accumulate);
                            accumulate = 1;
                        }
                    } else {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {                 // 16 K rows per MMA
                            const uint64_t bd = smem_desc(da + ks * 1024, kHalf, 512, LAYOUT_SW64);
                            mma_bf16_ss(tmem_base, smem_desc(xa + ks * 1024, kHalf, 512, LAYOUT_SW64), bd, idesc, accumulate);
                            if (two)
                                mma_bf16_ss(tmem_base + 256, smem_desc(xa + 4 * kHalf + ks * 1024, kHalf, 512, LAYOUT_SW64), bd, idesc,
                                            accumulate);
                            accumulate = 1;
                        }
                    }
                    mma_commit(bar_empty + 8 * stage);
                }
                __syncwarp();
                accumulate = 1;
                if (++stage == kWgStages) { stage = 0; parity ^= 1; }
            }
            if (elect_one_sync()) mma_commit(bar_acc);
            __syncwarp();
        }
    } else {
        // column sums of the staged dY halves (bias gradient), then the flush of the accumulators
        const int tid = threadIdx.x - 64;                 // 0..127
        const int quarter = warp & 3;                      // TMEM lane quarter of this warp
        const int cc = tid & 31, rg = tid >> 5;            // chunk column (8 cols), row group (16 rows)
        uint32_t stage = 0, parity = 0, acc_phase = 0;
        for (int jb = 0; jb < args.n_jobs; ++jb) {
            const WgJob& J = args.jobs[jb];
            const int n_out = J.n_out;
            float bsum[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) bsum[e] = 0.0f;
            const bool f8 = J.fp8 != 0;
            const int c16 = tid & 15, rg8 = tid >> 4;                  // 8-bit path: 16-column chunk, group of 8 rows
            const bool do_bias = J.bias != nullptr && (f8 ? c16 * 16 < n_out : cc * 8 < n_out);
            for (int64_t i = 0; i < my_tiles * 2; ++i) {
                mbar_wait(bar_full + 8 * stage, parity);
                if (do_bias && f8) {
                    const uint8_t* dy = smem + stage * kWgStageBytes + kWgOperandBytes + (c16 >> 2) * kHalf;
#pragma unroll 4
                    for (int r = rg8 * 8; r < rg8 * 8 + 8; ++r) {
                        const uint4 v = *reinterpret_cast<const uint4*>(dy + r * 64 + (((c16 & 3) ^ ((r >> 1) & 3)) << 4));
                        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const __half2_raw lo = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(w[e] & 0xFFFFu), __NV_E4M3);
                            const __half2_raw hi = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(w[e] >> 16), __NV_E4M3);
                            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo));
                            const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi));
                            bsum[4 * e] += a.x; bsum[4 * e + 1] += a.y; bsum[4 * e + 2] += b.x; bsum[4 * e + 3] += b.y;
                        }
                    }
                } else if (do_bias) {
                    const uint8_t* dy = smem + stage * kWgStageBytes + kWgOperandBytes + (cc >> 2) * kHalf;
#pragma unroll 4
                    for (int r = rg * 16; r < rg * 16 + 16; ++r) {
                        const uint4 v = *reinterpret_cast<const uint4*>(dy + r * 64 + (((cc & 3) ^ ((r >> 1) & 3)) << 4));
                        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            bsum[2 * e] += __uint_as_float(w[e] << 16);
                            bsum[2 * e + 1] += __uint_as_float(w[e] & 0xFFFF0000u);
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_empty + 8 * stage);
                if (++stage == kWgStages) { stage = 0; parity ^= 1; }
            }
            if (J.bias != nullptr && f8) {
#pragma unroll
                for (int e = 0; e < 16; ++e) bias_scratch[rg8 * 256 + c16 * 16 + e] = bsum[e];
                named_bar_sync(1, 128);
                for (int c = tid; c < n_out; c += 128) {
                    float sm = 0.0f;
#pragma unroll
                    for (int g8 = 0; g8 < 8; ++g8) sm += bias_scratch[g8 * 256 + c];
                    if (my_tiles > 0) atomicAdd(J.bias + c, sm * (J.dy_scale ? __ldg(J.dy_scale) : 1.0f));
                }
                named_bar_sync(1, 128);
            } else if (J.bias != nullptr) {
#pragma unroll
                for (int e = 0; e < 8; ++e) bias_scratch[rg * 256 + cc * 8 + e] = bsum[e];
                named_bar_sync(1, 128);
                for (int c = tid; c < n_out; c += 128) {
                    const float sm = bias_scratch[c] + bias_scratch[256 + c] + bias_scratch[512 + c] + bias_scratch[768 + c];
                    if (my_tiles > 0) atomicAdd(J.bias + c, sm);
                }
                named_bar_sync(1, 128);
            }
            // weight gradient: thread = result row m of accumulator 0 (and m + 128 of accumulator 1)
            mbar_wait(bar_acc, acc_phase);
            acc_phase ^= 1;
            tc_fence_after();
            if (my_tiles > 0) {
                const int m = quarter * 32 + lane;
                const float oscale = (J.x_scale ? __ldg(J.x_scale) : 1.0f) * (J.dy_scale ? __ldg(J.dy_scale) : 1.0f);
                for (int mh = 0; mh < (f8 ? J.x_blocks / 2 : J.x_blocks / 4); ++mh) {
                    float* orow = J.out + (int64_t)(mh * 128 + m) * J.out_ld;
                    for (int j = 0; j < n_out / 32; ++j) {
                        uint32_t v[32];
                        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + mh * 256 + j * 32, v);
                        tmem_ld_wait();
#pragma unroll
                        for (int c = 0; c < 32; ++c) atomicAdd(orow + j * 32 + c, __uint_as_float(v[c]) * oscale);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_free);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

// The two heads' kernels on CUDA cores, straight from the tile images (model.py:111,123):
//   out[c * n_d + j] += sum_rows X[row][c] * d_rgbs[row][d0 + j]       (n_d = 1: density head, 3: colour head)
// A warp reads 32 consecutive rows of one 32-column block (2 KB contiguous); partial sums stay in registers
// over all tiles of the CTA and are reduced across the warp once at the end.
// CH: 16-byte chunks (8 columns) of a 32-column block a warp owns: 4 (density head: 8 warps = 8 blocks) or 2 (colour
// head: 8 warps = 4 blocks x 2 halves), which keeps the accumulators at <= 48 registers and two blocks resident per SM
// (with 96 the kernel ran at one block per SM and 11 % occupancy, latency-bound at 3 TB/s).
template <int ND, int CH>
__device__ __forceinline__ void heads_wgrad_body(const uint8_t* __restrict__ x_img, const float* __restrict__ d_rgbs,
                                                 int d0, float* __restrict__ out, int64_t n_rows, int64_t n_tiles) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int blk = CH == 4 ? warp : warp >> 1, c0 = CH == 4 ? 0 : (warp & 1) * 2;
    float acc[8 * CH * ND];
#pragma unroll
    for (int i = 0; i < 8 * CH * ND; ++i) acc[i] = 0.0f;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint8_t* tb = x_img + tile * (int64_t)kMainBytes + blk * kBlockBytes;
#pragma unroll
        for (int rgp = 0; rgp < 4; ++rgp) {
            const int row = rgp * 32 + lane;
            const int64_t grow = tile * kTileRows + row;
            float d[ND];
#pragma unroll
            for (int j = 0; j < ND; ++j) d[j] = grow < n_rows ? __ldg(d_rgbs + 4 * grow + d0 + j) : 0.0f;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(tb + row * 64 + (((c0 + c) ^ ((row >> 1) & 3)) << 4)));
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float x0 = __uint_as_float(w[e] << 16), x1 = __uint_as_float(w[e] & 0xFFFF0000u);
#pragma unroll
                    for (int j = 0; j < ND; ++j) {
                        acc[(c * 8 + 2 * e) * ND + j] += x0 * d[j];
                        acc[(c * 8 + 2 * e + 1) * ND + j] += x1 * d[j];
                    }
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 8 * CH * ND; ++i) {
        const float v = ntx_warp_sum(acc[i]);
        if (lane == 0) atomicAdd(out + (blk * 32 + c0 * 8 + i / ND) * ND + (i % ND), v);
    }
}

// the same on e4m3 tile images (64 columns per 64-byte row): a warp owns CH chunks of 16 columns
template <int ND, int CH>
__device__ __forceinline__ void heads_wgrad_body8(const uint8_t* __restrict__ x_img, const float* __restrict__ x_scale,
                                                  const float* __restrict__ d_rgbs, int d0, float* __restrict__ out, int64_t n_rows,
                                                  int64_t n_tiles) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int blk = CH == 2 ? warp >> 1 : warp >> 2, c0 = CH == 2 ? (warp & 1) * 2 : (warp & 3);
    float acc[16 * CH * ND];
#pragma unroll
    for (int i = 0; i < 16 * CH * ND; ++i) acc[i] = 0.0f;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint8_t* tb = x_img + tile * (int64_t)kMain8Bytes + blk * kBlockBytes;
#pragma unroll
        for (int rgp = 0; rgp < 4; ++rgp) {
            const int row = rgp * 32 + lane;
            const int64_t grow = tile * kTileRows + row;
            float d[ND];
#pragma unroll
            for (int j = 0; j < ND; ++j) d[j] = grow < n_rows ? __ldg(d_rgbs + 4 * grow + d0 + j) : 0.0f;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(tb + row * 64 + (((c0 + c) ^ ((row >> 1) & 3)) << 4)));
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const __half2_raw lo = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(w[e] & 0xFFFFu), __NV_E4M3);
                    const __half2_raw hi = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(w[e] >> 16), __NV_E4M3);
                    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo));
                    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi));
                    const float x[4] = {a.x, a.y, b.x, b.y};
#pragma unroll
                    for (int q = 0; q < 4; ++q)
#pragma unroll
                        for (int j = 0; j < ND; ++j) acc[(c * 16 + 4 * e + q) * ND + j] += x[q] * d[j];
                }
            }
        }
    }
    const float sx = __ldg(x_scale);
#pragma unroll
    for (int i = 0; i < 16 * CH * ND; ++i) {
        const float v = ntx_warp_sum(acc[i]);
        if (lane == 0) atomicAdd(out + (blk * 64 + c0 * 16 + i / ND) * ND + (i % ND), v * sx);
    }
}

__global__ void __launch_bounds__(256, 2)
mlp_heads_wgrad8_kernel(const uint8_t* __restrict__ x_alpha, const float* __restrict__ sx_alpha, const uint8_t* __restrict__ x_rgb,
                        const float* __restrict__ sx_rgb, const float* __restrict__ d_rgbs, float* __restrict__ out_alpha,
                        float* __restrict__ out_rgb, int64_t n_rows, int64_t n_tiles) {
    if (blockIdx.y == 0) heads_wgrad_body8<1, 2>(x_alpha, sx_alpha, d_rgbs, 3, out_alpha, n_rows, n_tiles);
    else heads_wgrad_body8<3, 1>(x_rgb, sx_rgb, d_rgbs, 0, out_rgb, n_rows, n_tiles);
}

// per-step scales of the 8-bit images from the maxima the previous step saw: scale = margin * amax / 448 (e4m3's largest
// finite value), kept where nothing was seen; the maxima are cleared for the step to come
__global__ void fp8_update_scales_kernel(uint32_t* __restrict__ amax, float* __restrict__ scale, float* __restrict__ inv_scale, int n,
                                         float margin) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float a = __uint_as_float(amax[i]);
    if (a > 0.0f && a < CUDART_INF_F) {
        const float sc = fmaxf(a * margin / 448.0f, 1e-30f);
        scale[i] = sc;
        inv_scale[i] = 1.0f / sc;
    }
    amax[i] = 0u;
}

// both heads in one launch (blockIdx.y = 0: density head, 1: colour head)
__global__ void __launch_bounds__(256, 2)
mlp_heads_wgrad_kernel(const uint8_t* __restrict__ x_alpha, const uint8_t* __restrict__ x_rgb, const float* __restrict__ d_rgbs,
                       float* __restrict__ out_alpha, float* __restrict__ out_rgb, int64_t n_rows, int64_t n_tiles) {
    if (blockIdx.y == 0) heads_wgrad_body<1, 4>(x_alpha, d_rgbs, 3, out_alpha, n_rows, n_tiles);
    else heads_wgrad_body<3, 2>(x_rgb, d_rgbs, 0, out_rgb, n_rows, n_tiles);
}

}  // namespace

NTX_API int64_t ntx_mlp_packed_t_bytes(const ntx_mlp_t* mlp) {
    if (!mlp) return -1;
    BwdProgram bp;
    return build_bwd_program(mlp, &bp);
}

NTX_API int ntx_mlp_pack_weights_t(const ntx_mlp_t* mlp, const float* master, void* packed_t, void* stream) {
    NTX_REQUIRE(mlp && master && packed_t, "mlp_pack_weights_t: null pointer");
    NTX_REQUIRE(((uintptr_t)packed_t & 15) == 0, "mlp_pack_weights_t: buffer must be 16-byte aligned");
    BwdProgram bp;
    build_bwd_program(mlp, &bp);
    dim3 grid(8, bp.n_bwd + 1);
    mlp_pack_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(mlp->prog, bp, master, (uint8_t*)packed_t);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

static int backward_data_impl(const ntx_mlp_t* mlp, const void* packed_t, const float* d_rgbs, const void* masks,
                              void* dy_img, int64_t n_rows, int64_t n_tiles, void* stream, void* dy_img8,
                              const float* inv_scale8, uint32_t* amax8) {
    NTX_REQUIRE(mlp && packed_t && d_rgbs && masks && dy_img, "mlp_backward_data: null pointer");
    NTX_REQUIRE(!dy_img8 || (inv_scale8 && amax8 && ((uintptr_t)dy_img8 & 15) == 0 && mlp->prog.n_steps <= 16),
                "mlp_backward_data: 8-bit images need inv_scale8[n_steps], amax8[n_steps] and 16-byte alignment");
    NTX_REQUIRE(n_rows >= 0 && n_tiles >= ((n_rows + 511) / 512) * 4, "mlp_backward_data: %lld tiles for %lld rows",
                (long long)n_tiles, (long long)n_rows);
    if (n_rows == 0) return NTX_OK;
    NTX_REQUIRE((((uintptr_t)packed_t | (uintptr_t)d_rgbs | (uintptr_t)masks | (uintptr_t)dy_img) & 15) == 0,
                "mlp_backward_data: buffers must be 16-byte aligned");
    static const int ctas = [] { const char* e = getenv("NTX_MLP_CTAS"); return (e && e[0] == '1') ? 1 : 2; }();
    using KernelFn = void (*)(const BwdProgram, const DgradArgs);
    const int f8 = dy_img8 ? 1 : 0;
    KernelFn fn = ctas == 2 ? (f8 ? mlp_dgrad_kernel<2, true> : mlp_dgrad_kernel<2>) : (f8 ? mlp_dgrad_kernel<1, true> : mlp_dgrad_kernel<1>);
    static bool attr_set[64][2][2] = {};
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && !attr_set[dev][ctas - 1][f8]) {
        NTX_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDgSmemBytes));
        attr_set[dev][ctas - 1][f8] = true;
    }
    BwdProgram bp;
    build_bwd_program(mlp, &bp);
    DgradArgs a;
    if (ctas == 2) {
        if (int rc = ntx_weight_map(&a.w_map, packed_t, build_bwd_program(mlp, &bp), 4)) return rc;
    } else {
        memset(&a.w_map, 0, sizeof(a.w_map));
    }
    a.packed_t = (const uint8_t*)packed_t;
    a.d_rgbs = d_rgbs;
    a.masks = (const uint4*)masks;
    a.dy_img = (uint8_t*)dy_img;
    a.n_rows = n_rows;
    a.n_tiles = n_tiles;
    a.depth = mlp->cfg.depth;
    a.dy_img8 = (uint8_t*)dy_img8;
    a.inv_scale8 = inv_scale8;
    a.amax8 = amax8;
    a.need16 = 0;
    for (int st = 0; st < mlp->prog.n_steps; ++st)
        if (mlp->prog.steps[st].enc_data) a.need16 |= 1u << st;       // their encoded-input weight gradient reads bf16 dY
    const int64_t rows_per_unit = (int64_t)kRowsPerCta * ctas;
    const int64_t n_units = (n_rows + rows_per_unit - 1) / rows_per_unit;
    const int64_t max_units = ntx_sm_count() / ctas;
    const int grid = (int)(n_units < max_units ? n_units : max_units) * ctas;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreadsBwd);
    cfg.dynamicSmemBytes = kDgSmemBytes;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ctas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    NTX_CUDA(cudaLaunchKernelEx(&cfg, fn, bp, a));
    return NTX_OK;
}

NTX_API int ntx_mlp_backward_data(const ntx_mlp_t* mlp, const void* packed_t, const float* d_rgbs, const void* masks,
                                  void* dy_img, int64_t n_rows, int64_t n_tiles, void* stream) {
    return backward_data_impl(mlp, packed_t, d_rgbs, masks, dy_img, n_rows, n_tiles, stream, nullptr, nullptr, nullptr);
}

NTX_API int ntx_mlp_backward_data_fp8(const ntx_mlp_t* mlp, const void* packed_t, const float* d_rgbs, const void* masks,
                                      void* dy_img, void* dy_img8, const float* inv_scale8, void* amax8, int64_t n_rows,
                                      int64_t n_tiles, void* stream) {
    NTX_REQUIRE(dy_img8, "mlp_backward_data_fp8: null image buffer");
    return backward_data_impl(mlp, packed_t, d_rgbs, masks, dy_img, n_rows, n_tiles, stream, dy_img8, inv_scale8, (uint32_t*)amax8);
}

static int backward_weights_impl(const ntx_mlp_t* mlp, const void* acts_img, const void* dy_img, const void* enc_pos,
                                 const void* enc_dir, const float* d_rgbs, float* grad, float* enc_scratch,
                                 int64_t n_rows, int64_t n_tiles, void* stream, const void* acts_img8, const void* dy_img8,
                                 const float* x_scale8, const float* dy_scale8) {
    const bool f8 = acts_img8 != nullptr;
    NTX_REQUIRE(mlp && (acts_img || f8) && dy_img && enc_pos && enc_dir && d_rgbs && grad && enc_scratch,
                "mlp_backward_weights: null pointer");
    NTX_REQUIRE(!f8 || (dy_img8 && x_scale8 && dy_scale8), "mlp_backward_weights_fp8: null 8-bit image / scale pointer");
    NTX_REQUIRE(n_rows >= 0 && n_tiles >= (n_rows + 127) / 128, "mlp_backward_weights: bad tile count");
    if (n_rows == 0) return NTX_OK;
    static bool attr_set[64] = {};
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && !attr_set[dev]) {
        NTX_CUDA(cudaFuncSetAttribute(mlp_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWgSmemBytes));
        attr_set[dev] = true;
    }
    const Program& p = mlp->prog;
    const int64_t step_bytes = n_tiles * (int64_t)kMainBytes;
    const int64_t step8_bytes = n_tiles * (int64_t)kMain8Bytes;
    WgradArgs a;
    memset(&a, 0, sizeof(a));
    a.n_tiles = (n_rows + 127) / 128;
    int nj = 0;
    for (int st = 0; st < p.n_steps; ++st) {
        const Step& S = p.steps[st];
        const Layer& L = p.layers[S.layer];
        if (st > 0) {                                   // main operand: the previous step's stored output
            WgJob& J = a.jobs[nj++];
            J.out = grad + L.w_off + (int64_t)S.main_row0 * L.fan_out;
            J.bias = grad + L.b_off;
            J.n_out = S.n_out; J.out_ld = L.fan_out;
            if (f8) {                                   // both operands as e4m3 images, rescaled in the flush
                J.x = (const uint8_t*)acts_img8 + (st - 1) * step8_bytes;
                J.dy = (const uint8_t*)dy_img8 + st * step8_bytes;
                J.x_tile_bytes = kMain8Bytes; J.dy_tile_bytes = kMain8Bytes; J.x_blocks = 4; J.fp8 = 1;
                J.x_scale = x_scale8 + (st - 1); J.dy_scale = dy_scale8 + st;
            } else {
                J.x = (const uint8_t*)acts_img + (st - 1) * step_bytes;
                J.dy = (const uint8_t*)dy_img + st * step_bytes;
                J.x_tile_bytes = kMainBytes; J.dy_tile_bytes = kMainBytes; J.x_blocks = 8;
            }
        }
        if (S.enc_data) {                               // encoded inputs: rows in kernel column order -> scratch
            WgJob& J = a.jobs[nj++];
            J.x = (const uint8_t*)(S.phase == 0 ? enc_pos : enc_dir);
            J.dy = (const uint8_t*)dy_img + st * step_bytes;
            J.out = enc_scratch + (int64_t)st * 128 * 256;
            J.bias = nullptr;
            J.x_tile_bytes = 4 * kBlockBytes; J.dy_tile_bytes = kMainBytes; J.x_blocks = 4; J.n_out = S.n_out; J.out_ld = 256;
        }
    }
    a.n_jobs = nj;
    cudaStream_t stc = (cudaStream_t)stream;
    const int64_t g = a.n_tiles < ntx_sm_count() ? a.n_tiles : ntx_sm_count();
    mlp_wgrad_kernel<<<(unsigned)g, kWgThreads, kWgSmemBytes, stc>>>(a);
    NTX_LAUNCH_CHECK();
    // heads: density (input = last trunk layer's output), colour (input = the 128-wide last step's output)
    const Layer& la = p.layers[p.alpha_layer];
    const Layer& lr = p.layers[p.rgb_layer];
    const int64_t gh = a.n_tiles < 4 * ntx_sm_count() ? a.n_tiles : 4 * ntx_sm_count();
    if (f8)
        mlp_heads_wgrad8_kernel<<<dim3((unsigned)gh, 2), 256, 0, stc>>>(
            (const uint8_t*)acts_img8 + (int64_t)(mlp->cfg.depth - 1) * step8_bytes, x_scale8 + (mlp->cfg.depth - 1),
            (const uint8_t*)acts_img8 + (int64_t)(p.n_steps - 1) * step8_bytes, x_scale8 + (p.n_steps - 1), d_rgbs, grad + la.w_off,
            grad + lr.w_off, n_rows, a.n_tiles);
    else
        mlp_heads_wgrad_kernel<<<dim3((unsigned)gh, 2), 256, 0, stc>>>(
            (const uint8_t*)acts_img + (int64_t)(mlp->cfg.depth - 1) * step_bytes,
            (const uint8_t*)acts_img + (int64_t)(p.n_steps - 1) * step_bytes, d_rgbs, grad + la.w_off, grad + lr.w_off, n_rows, a.n_tiles);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_mlp_backward_weights(const ntx_mlp_t* mlp, const void* acts_img, const void* dy_img, const void* enc_pos,
                                     const void* enc_dir, const float* d_rgbs, float* grad, float* enc_scratch,
                                     int64_t n_rows, int64_t n_tiles, void* stream) {
    NTX_REQUIRE(acts_img, "mlp_backward_weights: null pointer");
    return backward_weights_impl(mlp, acts_img, dy_img, enc_pos, enc_dir, d_rgbs, grad, enc_scratch, n_rows, n_tiles, stream,
                                 nullptr, nullptr, nullptr, nullptr);
}

NTX_API int ntx_mlp_backward_weights_fp8(const ntx_mlp_t* mlp, const void* acts_img8, const void* dy_img8, const void* dy_img,
                                         const float* x_scale8, const float* dy_scale8, const void* enc_pos, const void* enc_dir,
                                         const float* d_rgbs, float* grad, float* enc_scratch, int64_t n_rows, int64_t n_tiles,
                                         void* stream) {
    NTX_REQUIRE(acts_img8, "mlp_backward_weights_fp8: null pointer");
    return backward_weights_impl(mlp, nullptr, dy_img, enc_pos, enc_dir, d_rgbs, grad, enc_scratch, n_rows, n_tiles, stream, acts_img8,
                                 dy_img8, x_scale8, dy_scale8);
}

NTX_API int ntx_fp8_update_scales(void* amax, float* scale, float* inv_scale, int n, float margin, void* stream) {
    NTX_REQUIRE(amax && scale && inv_scale && n >= 0 && n <= 64 && margin > 0.0f, "fp8_update_scales: bad arguments");
    if (n == 0) return NTX_OK;
    fp8_update_scales_kernel<<<1, 64, 0, (cudaStream_t)stream>>>((uint32_t*)amax, scale, inv_scale, n, margin);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

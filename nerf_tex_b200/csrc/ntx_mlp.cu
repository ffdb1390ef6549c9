// K1 — fused Fourier-feature encoding + conditioned NeRF-Tex MLP forward for sm_100a.
//
// One persistent CTA per SM (320 threads) walks 256-row batches of samples:
//
//   warp 0      weight producer: streams the packed bf16 weight image, one [N x 32] K-slice per
//               bulk-TMA copy (cp.async.bulk -> UBLKCP), into a 3-deep shared-memory ring.  In a CTA
//               pair each CTA copies its N/2 rows of every slice as a tensor-map box
//               (cp.async.bulk.tensor ... cta_group::2) whose bytes are credited straight to the
//               issuing CTA's barrier: the ring's turnaround has no hop through the peer.
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma (M=128, N=256|128, K=16, bf16 in,
//               fp32 accumulate in TMEM).  The CTA's two 128-row tiles alternate layer by layer:
//               while the tensor core runs tile 1's layer, tile 0's epilogue warps turn their
//               accumulators into the next layer's operand, and vice versa.
//   warps 2-9   two epilogue warpgroups (one per tile, one thread per row): compute the encodings,
//               and per layer pull the fp32 accumulator row out of TMEM (tcgen05.ld), apply ReLU,
//               round to bf16 and write it back as next layer's A operand in the UMMA-canonical
//               SWIZZLE_64B K-major layout.  Activations never leave the SM.
//
// Biases ride the tensor core: every encoding carries a constant-1 column, and the packer puts
// the bias in the weight row that multiplies it.  Skip / direction concats are never
// materialised: they are extra K-slices accumulated into the same TMEM tile.  The two tiny
// heads (density 256->1, colour 128->3) run on CUDA cores in fp32 inside the epilogues.
//
// Reference: network/layer.py:11-23, network/model.py:58-125, network/renderer.py:145-168.
#include <cuda.h>
#include <cuda_bf16.h>

#include <stdlib.h>
#include <string.h>

#include <new>

#include "ntx_common.cuh"
#include "ntx_mlp.cuh"
#include "ntx_umma.cuh"
#include "ntx_mlp_dev.cuh"

namespace {

using namespace umma;
using namespace ntxmlp;

// ----------------------------------------------------------------------------------------------
// shared-memory map (bytes, from a 1024-aligned base)
// ----------------------------------------------------------------------------------------------
constexpr uint32_t kEncBytes = kEncBlocks * kBlockBytes;                   // 24576
constexpr uint32_t kTileBytes = kMainBytes + kEncBytes;                    // 90112
#ifdef NTX_EXP_STAGES
constexpr int kStagesX = NTX_EXP_STAGES;      // timing experiment only: a 4th stage overlaps tile 1's encoding buffer
constexpr uint32_t kOffStages = kTilesPerCta * kTileBytes - (kStagesX > 3 ? (kStagesX - 3) * kStageBytes : 0);
constexpr uint32_t kRingBytes = kStagesX * kStageBytes;
#define kStages kStagesX
#else
constexpr uint32_t kOffStages = kTilesPerCta * kTileBytes;                 // 180224
constexpr uint32_t kRingBytes = kStages * kStageBytes;                     // 49152: 3 x 16 KB, or 6 x 8 KB per CTA of a pair
#endif
constexpr uint32_t kOffHeads = kOffStages + kRingBytes;                    // 229376
constexpr uint32_t kOffBars = kOffHeads + kHeadFloats * 4;                 // +2592
constexpr uint32_t kMaxStages = 2 * kStages;
constexpr uint32_t kNumBars = 2 * kMaxStages + 2 * kTilesPerCta;
constexpr uint32_t kOffTmemSlot = kOffBars + kNumBars * 8;
constexpr uint32_t kOffAmax = kOffTmemSlot + 16;                           // [16] per-step max |activation| (8-bit dumps)
constexpr uint32_t kSmemBytes = kOffAmax + 64;
static_assert(kSmemBytes <= 232448, "exceeds the 227 KB per-CTA shared memory limit");

constexpr int kThreads = 320;
#ifndef NTX_MMA_PREFETCH
#define NTX_MMA_PREFETCH 0
#endif
constexpr int kEpiWarp0 = 2;

struct KernelArgs {
    // CTA pairs only: the packed weight image as rows of 2 KB; one box = this CTA's half of a K-slice
    // (8 KB for N = 256, 4 KB for N = 128)
    CUtensorMap w_map128, w_map64;
    ntx_mlp_inputs in;
    const uint8_t* packed;
    float* rgbs;
    // training only: bf16 copies of every tensor-core operand the backward needs
    uint8_t* dump_enc_pos; // [n_tiles][32 KB] tile images of the encoded pos-phase inputs (kernel column order,
                           // constant-1 column included): blocks 0..2 = 96 columns are written, block 3 stays as allocated
    uint8_t* dump_enc_dir; // [n_tiles][32 KB]
    uint4* dump_acts;      // [n_steps][N, 256] stored outputs of every step (the 128-wide last step uses cols 0..127); may be NULL
    uint8_t* dump_img;     // [n_steps][n_tiles][64 KB]: the same outputs as 128-row SWIZZLE_64B tile images (the A
                           // operand buffer as it sits in shared memory), written by bulk stores; may be NULL
    uint4* dump_masks;     // [n_steps][n_tiles*128][8 x u32]: bit c of word j = output column 32 j + c is > 0
    int64_t n_tiles;       // tiles allocated per step in dump_img / dump_masks
    // 8-bit variant of dump_img (the bf16 images are then not written): [n_steps][n_tiles][32 KB] e4m3 tile images of
    // output / scale8[step]; amax8[step] receives max |output| (for the next training step's scales)
    uint8_t* dump_img8;
    const float* inv_scale8;
    uint32_t* amax8;
};

// ----------------------------------------------------------------------------------------------
// the fused forward kernel
// ----------------------------------------------------------------------------------------------

// bf16x2 word {lo, hi} with ReLU folded into the conversion (one F2FP per two activations)
__device__ __forceinline__ uint32_t pack_bf16_relu(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// 32 fp32 accumulator columns (chunk j of this thread's row) -> bf16 -> next layer's A operand
// bits {2e, 2e+1} of the result: the two bf16 halves of packed word e are non-zero (i.e. > 0 after ReLU)
__device__ __forceinline__ uint32_t nonzero_bits4(const uint32_t (&w)[4]) {
    uint32_t m = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        m |= ((w[e] & 0xFFFFu) != 0u ? 1u : 0u) << (2 * e);
        m |= ((w[e] >> 16) != 0u ? 1u : 0u) << (2 * e + 1);
    }
    return m;
}

struct Dump8 { uint8_t* tile_img; float inv_scale; float amax; bool track; };

template <bool RELU, bool DUMP, bool F8>
__device__ __forceinline__ uint32_t store_chunk32(uint32_t main_base, int j, int row, const uint32_t (&v)[32], uint4* grow_ptr,
                                                  Dump8& d8) {
    uint32_t mask = 0;
    if (F8) {
        float x[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) x[c] = __uint_as_float(v[c]);
        dump_fp8_chunk32<RELU>(d8.tile_img, j, row, x, d8.inv_scale, d8.amax, d8.track);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float lo = __uint_as_float(v[8 * c + 2 * e]), hi = __uint_as_float(v[8 * c + 2 * e + 1]);
            w[e] = RELU ? pack_bf16_relu(lo, hi) : pack_bf16(lo, hi);
        }
        st_shared_v4(chunk_addr(main_base, j * 4 + c, row), w[0], w[1], w[2], w[3]);
        if (DUMP && grow_ptr) st_global_v4(grow_ptr + j * 4 + c, w[0], w[1], w[2], w[3]);
        if (DUMP) mask |= nonzero_bits4(w) << (8 * c);
    }
    return mask;
}

// plain hidden layer epilogue: TMEM loads software-pipelined against the convert + store
template <bool RELU, bool DUMP, bool F8>
__device__ __forceinline__ void epilogue_store_layer(uint32_t tmem_row, uint32_t main_base, int row, int n_chunks,
                                                     uint4* grow_ptr, uint32_t (&masks)[8], Dump8& d8) {
    uint32_t va[32], vb[32];
    tmem_ld_32x32b_x32(tmem_row, va);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        if (j < n_chunks) {
            tmem_ld_32x32b_x32(tmem_row + (j + 1) * 32, vb);
            masks[j] = store_chunk32<RELU, DUMP, F8>(main_base, j, row, va, grow_ptr, d8);
            tmem_ld_wait();
            if (j + 2 < n_chunks) tmem_ld_32x32b_x32(tmem_row + (j + 2) * 32, va);
            masks[j + 1] = store_chunk32<RELU, DUMP, F8>(main_base, j + 1, row, vb, grow_ptr, d8);
            tmem_ld_wait();
        }
    }
}

// CTAS == 1: every CTA is on its own (cta_group::1, M = 128 per instruction).
// CTAS == 2: CTA pairs (cluster of 2, cta_group::2): one tcgen05.mma issued by the even CTA drives
//            both SMs (M = 256), each CTA stages only its half of every weight slice (N/2 rows), so
//            the L2->SM weight traffic per SM halves and the 48 KB ring holds 6 slices instead of 3.
template <bool DUMP, int CTAS, bool F8 = false>
__global__ void __launch_bounds__(kThreads, 1)
mlp_forward_kernel(const __grid_constant__ Program prog, const __grid_constant__ KernelArgs args) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    if ((sbase & 1023u) != 0) __trap();

    // A ring stage is 16 KB: one whole [N x 32] slice for a lone CTA; for a pair, this CTA's halves
    // ([N/2 x 32], 8 KB) of TWO consecutive slices, so that one barrier round trip feeds four MMAs.
    constexpr uint32_t NSTAGE = kStages;                        // ring stages
    constexpr uint32_t STAGE_BYTES = kStageBytes;               // bytes of one ring stage
    constexpr int SPS = CTAS;                                   // slices per stage
    const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;   // 0: issues the MMAs
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar_full = sbase + kOffBars;                 // [NSTAGE]  weight slice landed
    const uint32_t bar_empty = bar_full + 8 * kMaxStages;       // [NSTAGE]  weight slice consumed
    const uint32_t bar_act = bar_empty + 8 * kMaxStages;        // [2] tile's A operand (activations) ready
    const uint32_t bar_acc = bar_act + 8 * kTilesPerCta;        // [2] tile's accumulators ready
    float* heads = reinterpret_cast<float*>(smem + kOffHeads);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOffTmemSlot);

    const int64_t n_rows = args.in.n_rows;
    // a "batch" is what one CTA (or CTA pair) processes per trip: 256 rows per CTA
    const int64_t n_batches = (n_rows + kRowsPerCta * CTAS - 1) / (kRowsPerCta * CTAS);
    const int64_t batch0 = blockIdx.x / CTAS, batch_stride = gridDim.x / CTAS;

    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i < NSTAGE; ++i) {
            // pair: the issuing CTA's "landed" barrier also collects the peer's relay arrive
            // pair: only the issuing CTA's "landed" barrier is used; both CTAs' copies credit their bytes to it
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, 1);
        }
        for (int t = 0; t < kTilesPerCta; ++t) {
            mbar_init(bar_act + 8 * t, 4 * CTAS);   // one arrive per epilogue warp of the tile (both CTAs)
            mbar_init(bar_acc + 8 * t, 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        if (CTAS == 2) { tmem_alloc2(sbase + kOffTmemSlot, 512); tmem_relinquish2(); }
        else { tmem_alloc(sbase + kOffTmemSlot, 512); tmem_relinquish(); }
    }
    // fp32 head weights -> shared (broadcast reads in the epilogues)
    for (int i = threadIdx.x; i < kHeadFloats; i += kThreads)
        heads[i] = reinterpret_cast<const float*>(args.packed)[i];
    if (F8 && threadIdx.x < 16) reinterpret_cast<uint32_t*>(smem + kOffAmax)[threadIdx.x] = 0u;
    tc_fence_before();
    __syncthreads();
    if (CTAS == 2) cluster_sync_all();          // peer barriers initialised before any remote arrive / multicast
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== weight producer
        // The two tiles run half a layer out of phase (tile 1's MMAs overlap tile 0's epilogue and
        // vice versa), so every layer's slices are streamed once per tile, in MMA issue order.
        if (lane == 0) {
            uint32_t stage = 0, parity = 1;      // a fresh barrier passes a wait on parity 1
            const uint32_t leader_full = CTAS == 2 ? mapa_shared(bar_full, 0) : bar_full;
            if (CTAS == 2) { prefetch_tensormap(&args.w_map128); prefetch_tensormap(&args.w_map64); }
            for (int64_t b = batch0; b < n_batches; b += batch_stride) {
                for (int st = 0; st < prog.n_steps; ++st) {
                    const int n_main = prog.steps[st].n_main, enc_n = prog.steps[st].enc_n;
                    const uint32_t slice_bytes = (uint32_t)prog.steps[st].n_out * 64u;
                    const uint32_t bytes = slice_bytes / CTAS;               // pair: this CTA's N/2 rows of a slice
                    const uint8_t* src0 = args.packed + prog.slice_off[prog.steps[st].slice_begin] + rank * bytes;
                    // pair: box coordinates in rows of 2 KB of the packed image
                    const int32_t wrow0 = (int32_t)((prog.slice_off[prog.steps[st].slice_begin] + rank * bytes) >> 11);
                    const int32_t wrows = (int32_t)(slice_bytes >> 11);
                    const void* w_map = bytes == 8192u ? (const void*)&args.w_map128 : (const void*)&args.w_map64;
                    for (int t = 0; t < kTilesPerCta; ++t) {
                        const uint8_t* src = src0;
                        int32_t wrow = wrow0;
                        // stage groups: main slices first, then encoding slices, SPS slices per stage
                        for (int part = 0; part < 2; ++part) {
                            int left = part == 0 ? n_main : enc_n;
                            while (left > 0) {
                                const int n = left < SPS ? left : SPS;
                                mbar_wait(bar_empty + 8 * stage, parity);
                                if (CTAS == 2) {
                                    if (rank == 0) mbar_arrive_expect_tx(bar_full + 8 * stage, 2 * bytes * n);
                                    for (int j = 0; j < n; ++j, wrow += wrows)
                                        tma_tile2d_g2s_pair(sbase + kOffStages + stage * STAGE_BYTES + j * bytes, w_map, 0, wrow,
                                                            leader_full + 8 * stage);
                                } else {
                                    mbar_arrive_expect_tx(bar_full + 8 * stage, bytes * n);
                                    for (int j = 0; j < n; ++j, src += slice_bytes)
                                        tma_bulk_g2s(sbase + kOffStages + stage * STAGE_BYTES + j * bytes, src, bytes,
                                                     bar_full + 8 * stage);
                                }
                                if (++stage == NSTAGE) { stage = 0; parity ^= 1; }
                                left -= n;
                            }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer
        // The whole warp walks the schedule (so every operand is warp-uniform); one elected lane
        // issues tcgen05.mma / tcgen05.commit.
        if (CTAS == 2 && rank != 0) {
            // peer CTA of a pair: nothing to issue.  Its weight copies credit the issuing CTA's barriers themselves
            // (cp.async.bulk.tensor ... cta_group::2), its stages are freed by the issuer's multicast commits.
        } else {
        uint32_t stage = 0, parity = 0, act_phase0 = 0, act_phase1 = 0;
        uint32_t b16 = (sbase + kOffStages) >> 4;         // ring stage base >> 4
        uint32_t full_bar = bar_full, empty_bar = bar_empty;
        // one ring stage: wait until it landed, issue its MMAs (K=16 each; up to 2 per slice, SPS slices),
        // release the stage.  a16: A operand of the first slice (>>4); consecutive slices read consecutive
        // 32-column A blocks.  nh_last: K=16 halves of the stage's last slice (the others use 2).
        auto issue_stage = [&](uint32_t d_tmem, uint32_t a16, uint32_t half0, int n_slices, uint32_t nh_last,
                               uint32_t b_slice16, uint32_t idesc, uint32_t accumulate) {
            mbar_wait(full_bar, parity);
            tc_fence_after();
            const uint32_t bb = b16 + half0 * 2, eb = empty_bar;
            if (++stage == NSTAGE) {
                stage = 0; parity ^= 1; b16 = (sbase + kOffStages) >> 4; full_bar = bar_full; empty_bar = bar_empty;
            } else {
                b16 += STAGE_BYTES >> 4; full_bar += 8; empty_bar += 8;
            }
            if (elect_one_sync()) {
#pragma unroll
                for (int j = 0; j < SPS; ++j) {
                    if (j < n_slices) {
                        const uint32_t aj = a16 + j * (kBlockBytes >> 4), bj = bb + j * b_slice16;
                        const uint32_t nh = (j == n_slices - 1) ? nh_last : 2u;
                        if (CTAS == 2) {
                            mma_bf16_ss_2cta(d_tmem, desc_sw64(aj), desc_sw64(bj), idesc, (j > 0) ? 1u : accumulate);
                            if (nh == 2) mma_bf16_ss_2cta(d_tmem, desc_sw64(aj + 2), desc_sw64(bj + 2), idesc, 1u);
                        } else {
                            mma_bf16_ss(d_tmem, desc_sw64(aj), desc_sw64(bj), idesc, (j > 0) ? 1u : accumulate);
                            if (nh == 2) mma_bf16_ss(d_tmem, desc_sw64(aj + 2), desc_sw64(bj + 2), idesc, 1u);
                        }
                    }
                }
                if (CTAS == 2) mma_commit_2cta(eb);         // stage free (in both CTAs) once these MMAs retire
                else mma_commit(eb);
            }
            __syncwarp();
        };
        for (int64_t b = batch0; b < n_batches; b += batch_stride) {
            for (int st = 0; st < prog.n_steps; ++st) {
                // copy the step into registers once (asm memory clobbers would otherwise reload it per slice)
                const uint32_t idesc = idesc_bf16_f32(kTileRows * CTAS, prog.steps[st].n_out);
                const int n_main = prog.steps[st].n_main, enc_n = prog.steps[st].enc_n;
                const uint32_t enc_half0 = prog.steps[st].enc_half0, enc_last_nh = prog.steps[st].enc_last_nh;
                const uint32_t enc16 = ((uint32_t)kMainBytes + prog.steps[st].enc_blk0 * kBlockBytes + enc_half0 * 32u) >> 4;
                const uint32_t b_slice16 = ((uint32_t)prog.steps[st].n_out * 64u / CTAS) >> 4;   // this CTA's bytes of one slice
#pragma unroll
                for (int t = 0; t < kTilesPerCta; ++t) {
                    if (t == 0) { mbar_wait(bar_act, act_phase0); act_phase0 ^= 1; }
                    else { mbar_wait(bar_act + 8, act_phase1); act_phase1 ^= 1; }
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + t * kWidth;
                    const uint32_t a16_tile = (sbase + t * kTileBytes) >> 4;
                    for (int i = 0; i < n_main; i += SPS)
                        issue_stage(d_tmem, a16_tile + i * (kBlockBytes >> 4), 0u, SPS, 2u, b_slice16, idesc, i > 0 ? 1u : 0u);
                    for (int e = 0; e < enc_n; e += SPS) {
                        const int n = enc_n - e < SPS ? enc_n - e : SPS;
                        issue_stage(d_tmem, a16_tile + enc16 + e * (kBlockBytes >> 4), enc_half0, n,
                                    e + n == enc_n ? enc_last_nh : 2u, b_slice16, idesc, (n_main > 0 || e > 0) ? 1u : 0u);
                    }
                    if (elect_one_sync()) {                 // this tile's accumulators complete
                        if (CTAS == 2) mma_commit_2cta(bar_acc + 8 * t);
                        else mma_commit(bar_acc + 8 * t);
                    }
                    __syncwarp();
                }
            }
        }
        }
    } else {
        // ===================================================================== epilogue warpgroups
        const int ew = warp - kEpiWarp0;            // 0..7
        const int tile = ew >> 2;                   // 0: rows 0..127, 1: rows 128..255 of the batch
        const int quarter = warp & 3;               // TMEM lane quarter this warp may access
        const int row = quarter * 32 + lane;        // row inside the tile == TMEM lane
        const uint32_t main_base = sbase + tile * kTileBytes;
        const uint32_t enc_base = main_base + kMainBytes;
        const uint32_t tmem_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + tile * kWidth;
        const uint32_t my_act = bar_act + 8 * tile, my_acc = bar_acc + 8 * tile;
        const int g = prog.n_geo, a = prog.n_app;
        const float* w_alpha = heads;               // [256], b at [256]
        const float* w_rgb = heads + 260;           // [3][128], b at [260 + 384 ..]
        uint32_t acc_phase = 0;
        // pair: the tile's "operand ready" barrier lives in the issuing CTA
        const uint32_t act_remote = (CTAS == 2 && rank != 0) ? mapa_shared(my_act, 0) : 0u;
        auto arrive_act = [&]() {
            if (CTAS == 2 && rank != 0) mbar_arrive_remote(act_remote);
            else mbar_arrive(my_act);
        };

        for (int64_t b = batch0; b < n_batches; b += batch_stride) {
            const int64_t grow = (b * CTAS + rank) * kRowsPerCta + tile * kTileRows + row;
            const bool valid = grow < n_rows;
            // ------------------------------------------------------------- load this row's inputs
            float x[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 0.f}, pg[3] = {0.f, 0.f, 0.f};
            float xv[3] = {0.f, 0.f, 0.f};          // diagonal covariance of the position (integrated encoding only)
            float pa[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (valid) {
                const int64_t grp = grow / args.in.rows_per_group;
                if (args.in.pos) {
#pragma unroll
                    for (int i = 0; i < 3; ++i) x[i] = __ldg(args.in.pos + 3 * grow + i);
                    if (prog.ipe) {
#pragma unroll
                        for (int i = 0; i < 3; ++i) xv[i] = __ldg(args.in.pos_cov + 3 * grow + i);
                    }
                } else {
                    const float zz = __ldg(args.in.z + grow);                   // renderer.py:114
#pragma unroll
                    for (int i = 0; i < 3; ++i)
                        x[i] = __ldg(args.in.rays_o + 3 * grp + i) + __ldg(args.in.rays_d + 3 * grp + i) * zz;
                }
#pragma unroll
                for (int i = 0; i < 3; ++i) d[i] = __ldg(args.in.dir + 3 * grp + i);
                if (args.in.params) {
                    const float* pr = args.in.params + grp * (g + a);
                    float blur = 1.0f;                                          // renderer.py:155-158
                    if (args.in.blur_idx >= 0) {
                        if (args.in.group_scale) blur = __ldg(args.in.group_scale + grp);
                        if (args.in.row_scale) blur *= __ldg(args.in.row_scale + grow);
                    }
#pragma unroll
                    for (int j = 0; j < 3; ++j)
                        if (j < g) pg[j] = __ldg(pr + j) * (j == args.in.blur_idx ? blur : 1.0f);
#pragma unroll
                    for (int j = 0; j < 7; ++j)
                        if (j < a) pa[j] = __ldg(pr + g + j) * (g + j == args.in.blur_idx ? blur : 1.0f);
                }
            }
            // ------------------------------------------------------------- pos-phase encoding
            if (DUMP) {
                if (row == 0) tma_store_wait_read();         // last batch's dir-encoding image has left the buffer
                named_bar_sync(1 + tile, kTileRows);
            }
            {
                uint32_t w[32];
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    float s, c;
#pragma unroll
                    for (int k = 0; k < 10; ++k) {
                        if (k == 0) sincosf(x[i], &s, &c);
                        else if (k == 5) sincosf(32.0f * x[i], &s, &c);     // restart: bounds the doubling error
                        else { const float s2 = 2.0f * s * c; c = 1.0f - 2.0f * s * s; s = s2; }
                        if (prog.ipe) {
                            // E[sin(2^k x)] for x ~ N(mean, var): sin(2^k mean) * exp(-var * 4^k / 2)   (layer.py:33-41)
                            const float att = __expf(-0.5f * (float)(1 << (2 * k)) * xv[i]);
                            w[i * 10 + k] = pack_bf16(s * att, c * att);
                        } else {
                            w[i * 10 + k] = pack_bf16(s, c);
                        }
                    }
                }
                w[30] = pack_bf16(1.0f, prog.ipe ? 0.0f : x[0]);            // no identity term in the integrated encoding
                w[31] = prog.ipe ? 0u : pack_bf16(x[1], x[2]);
                auto put = [&](int q, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
                    st_shared_v4(chunk_addr(enc_base, q, row), w0, w1, w2, w3);
                };
#pragma unroll
                for (int q = 0; q < 8; ++q) put(q, w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
                if (g > 0) {
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        if (j < g) {
                            uint32_t p4[4];
                            sincos4(pg[j], p4);
                            put(8 + j, p4[0], p4[1], p4[2], p4[3]);
                        }
                    }
                    put(8 + g, pack_bf16(pg[0], pg[1]), pack_bf16(pg[2], 0.f), 0u, 0u);
                    if ((g & 1) == 0) put(9 + g, 0u, 0u, 0u, 0u);
                }
                if (DUMP) {
                    for (int q = pos_chunks(g) + (pos_chunks(g) & 1); q < kEncCols / 8; ++q) put(q, 0u, 0u, 0u, 0u);
                    fence_proxy_async_smem();
                    named_bar_sync(1 + tile, kTileRows);
                    if (row == 0) {
                        const int64_t tile_g = (b * CTAS + rank) * kTilesPerCta + tile;
                        tma_bulk_s2g(args.dump_enc_pos + tile_g * (int64_t)(4 * kBlockBytes), enc_base, kEncBytes);
                        tma_store_commit();
                    }
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) arrive_act();

            float alpha_raw = 0.0f;
            // ------------------------------------------------------------- layer epilogues
            for (int st = 0; st < prog.n_steps; ++st) {
                const Step& step = prog.steps[st];
                const uint32_t epi = step.epi;
                const int n_chunks = step.n_out >> 5;
                mbar_wait(my_acc, acc_phase);
                acc_phase ^= 1;
                tc_fence_after();
                float o0 = 0.f, o1 = 0.f, o2 = 0.f;
                uint4* gp = (DUMP && valid && args.dump_acts) ? args.dump_acts + ((int64_t)st * n_rows + grow) * (kWidth / 8) : nullptr;
                uint32_t masks[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
                Dump8 d8 = {nullptr, 0.0f, 0.0f, false};
                if (F8) {
                    const int64_t tile_g8 = (b * CTAS + rank) * kTilesPerCta + tile;
                    d8.tile_img = args.dump_img8 + ((int64_t)st * args.n_tiles + tile_g8) * (int64_t)kMain8Bytes;
                    d8.inv_scale = __ldg(args.inv_scale8 + st);
                    d8.track = (tile_g8 & 7) == 0;          // maxima from every eighth tile
                }
                if (DUMP && args.dump_img) {
                    // the previous step's image store must have finished reading this tile's buffer
                    if (row == 0) tma_store_wait_read();
                    named_bar_sync(1 + tile, kTileRows);
                }
                if ((epi & (EPI_ALPHA | EPI_FINAL)) == 0) {
                    // hidden layer: (ReLU) -> bf16 -> A operand of the next layer
                    if (epi & EPI_RELU) epilogue_store_layer<true, DUMP, F8>(tmem_row, main_base, row, n_chunks, gp, masks, d8);
                    else epilogue_store_layer<false, DUMP, F8>(tmem_row, main_base, row, n_chunks, gp, masks, d8);
                } else {
                    // layers that also feed an fp32 head (density after the trunk, colour at the end)
                    for (int j = 0; j < n_chunks; ++j) {
                        uint32_t v[32];
                        tmem_ld_32x32b_x32(tmem_row + j * 32, v);
                        tmem_ld_wait();
                        float f[32];
#pragma unroll
                        for (int c = 0; c < 32; ++c) f[c] = fmaxf(__uint_as_float(v[c]), 0.0f);   // both are ReLU layers
                        if (epi & EPI_ALPHA) {                                   // model.py:111
                            const float4* wa = reinterpret_cast<const float4*>(w_alpha + j * 32);
#pragma unroll
                            for (int c = 0; c < 8; ++c) {
                                const float4 ww = wa[c];
                                alpha_raw += f[4 * c] * ww.x + f[4 * c + 1] * ww.y + f[4 * c + 2] * ww.z + f[4 * c + 3] * ww.w;
                            }
                        }
                        if (epi & EPI_FINAL) {                                   // model.py:123
                            const float4* w0 = reinterpret_cast<const float4*>(w_rgb + j * 32);
                            const float4* w1 = reinterpret_cast<const float4*>(w_rgb + 128 + j * 32);
                            const float4* w2 = reinterpret_cast<const float4*>(w_rgb + 256 + j * 32);
#pragma unroll
                            for (int c = 0; c < 8; ++c) {
                                const float4 a0 = w0[c], a1 = w1[c], a2 = w2[c];
                                o0 += f[4 * c] * a0.x + f[4 * c + 1] * a0.y + f[4 * c + 2] * a0.z + f[4 * c + 3] * a0.w;
                                o1 += f[4 * c] * a1.x + f[4 * c + 1] * a1.y + f[4 * c + 2] * a1.z + f[4 * c + 3] * a1.w;
                                o2 += f[4 * c] * a2.x + f[4 * c + 1] * a2.y + f[4 * c + 2] * a2.z + f[4 * c + 3] * a2.w;
                            }
                        }
                        if (F8) dump_fp8_chunk32<false>(d8.tile_img, j, row, f, d8.inv_scale, d8.amax, d8.track);
                        if (!(epi & EPI_FINAL) || DUMP) {
                            uint32_t mj = 0;
#pragma unroll
                            for (int c = 0; c < 4; ++c) {
                                uint32_t w[4] = {pack_bf16(f[8 * c], f[8 * c + 1]), pack_bf16(f[8 * c + 2], f[8 * c + 3]),
                                                 pack_bf16(f[8 * c + 4], f[8 * c + 5]), pack_bf16(f[8 * c + 6], f[8 * c + 7])};
                                // (the last layer's output only enters shared memory to be dumped as an image)
                                if (!(epi & EPI_FINAL) || DUMP) st_shared_v4(chunk_addr(main_base, j * 4 + c, row), w[0], w[1], w[2], w[3]);
                                if (DUMP && gp) st_global_v4(gp + j * 4 + c, w[0], w[1], w[2], w[3]);
                                if (DUMP) mj |= nonzero_bits4(w) << (8 * c);
                            }
                            if (DUMP) {
#pragma unroll
                                for (int q = 0; q < 8; ++q) if (q == j) masks[q] = mj;
                            }
                        }
                    }
                    if (epi & EPI_ALPHA) alpha_raw += w_alpha[256];
                }
                if (epi & EPI_SWITCH) {
                    // ----------------------------------------------------- dir-phase encoding
                    auto put = [&](int q, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
                        st_shared_v4(chunk_addr(enc_base, q, row), w0, w1, w2, w3);
                    };
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        uint32_t p4[4];
                        sincos4(d[i], p4);
                        put(i, p4[0], p4[1], p4[2], p4[3]);
                    }
#pragma unroll
                    for (int j = 0; j < 7; ++j) {
                        if (j < a) {
                            uint32_t p4[4];
                            sincos4(pa[j], p4);
                            put(3 + j, p4[0], p4[1], p4[2], p4[3]);
                        }
                    }
                    put(3 + a, pack_bf16(1.0f, d[0]), pack_bf16(d[1], d[2]), pack_bf16(pa[0], pa[1]), pack_bf16(pa[2], pa[3]));
                    if (a > 4) put(4 + a, pack_bf16(pa[4], pa[5]), pack_bf16(pa[6], 0.f), 0u, 0u);
                    if (dir_chunks(a) & 1) put(dir_chunks(a), 0u, 0u, 0u, 0u);
                    if (DUMP) {
                        for (int q = dir_chunks(a) + (dir_chunks(a) & 1); q < kEncCols / 8; ++q) put(q, 0u, 0u, 0u, 0u);
                        fence_proxy_async_smem();
                        named_bar_sync(1 + tile, kTileRows);
                        if (row == 0) {
                            const int64_t tile_g = (b * CTAS + rank) * kTilesPerCta + tile;
                            tma_bulk_s2g(args.dump_enc_dir + tile_g * (int64_t)(4 * kBlockBytes), enc_base, kEncBytes);
                            tma_store_commit();
                        }
                    }
                }
                if (F8) {
                    // relu sign bits; the 8-bit image left through the register stores above
                    const int64_t tile_g = (b * CTAS + rank) * kTilesPerCta + tile;
                    uint4* mp = args.dump_masks + (((int64_t)st * args.n_tiles + tile_g) * kTileRows + row) * 2;
                    st_global_v4(mp, masks[0], masks[1], masks[2], masks[3]);
                    st_global_v4(mp + 1, masks[4], masks[5], masks[6], masks[7]);
                    amax_to_shared(d8.amax, reinterpret_cast<uint32_t*>(smem + kOffAmax) + st);
                }
                if (DUMP && args.dump_img) {
                    // relu sign bits (for the backward's masks) and the tile image of this step's stored output
                    const int64_t tile_g = (b * CTAS + rank) * kTilesPerCta + tile;
                    uint4* mp = args.dump_masks + (((int64_t)st * args.n_tiles + tile_g) * kTileRows + row) * 2;
                    st_global_v4(mp, masks[0], masks[1], masks[2], masks[3]);
                    st_global_v4(mp + 1, masks[4], masks[5], masks[6], masks[7]);
                    fence_proxy_async_smem();
                    named_bar_sync(1 + tile, kTileRows);
                    if (row == 0) {
                        tma_bulk_s2g(args.dump_img + ((int64_t)st * args.n_tiles + tile_g) * (int64_t)kMainBytes, main_base,
                                     (uint32_t)step.n_out * kTileRows * 2);
                        tma_store_commit();
                    }
                }
                if (epi & EPI_FINAL) {
                    if (valid) {
                        const float* brgb = w_rgb + 384;
                        reinterpret_cast<float4*>(args.rgbs)[grow] =
                            make_float4(o0 + brgb[0], o1 + brgb[1], o2 + brgb[2], alpha_raw);
                    }
                    tc_fence_before();      // TMEM reads done before the next batch's MMAs overwrite it
                } else {
                    fence_proxy_async_smem();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) arrive_act();
                }
            }
        }
    }

    if (DUMP) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // image stores fully written
    tc_fence_before();
    __syncthreads();
    if (F8 && threadIdx.x < 16) {
        const uint32_t a = reinterpret_cast<uint32_t*>(smem + kOffAmax)[threadIdx.x];
        if (a) atomicMax(args.amax8 + threadIdx.x, a);
    }
    if (CTAS == 2) cluster_sync_all();          // no CTA leaves while its peer may still signal it
    if (warp == 1) {
        __syncwarp();
        if (CTAS == 2) tmem_dealloc2(tmem_base, 512);
        else tmem_dealloc(tmem_base, 512);
    }
}

// tile images -> row-major (test / host-side helper)
__global__ void untile_kernel(const uint8_t* __restrict__ img, int64_t n_rows, int cols, int64_t tile_bytes,
                              uint4* __restrict__ out) {
    const int chunks_per_row = cols / 8;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_rows * chunks_per_row) return;
    const int64_t r = idx / chunks_per_row;
    const int q = (int)(idx % chunks_per_row);
    const int64_t tile = r / kTileRows;
    const int row = (int)(r % kTileRows);
    const uint32_t off = (uint32_t)(q >> 2) * kBlockBytes + (uint32_t)row * 64 + (uint32_t)(((q & 3) ^ ((row >> 1) & 3)) << 4);
    out[idx] = *reinterpret_cast<const uint4*>(img + tile * tile_bytes + off);
}

// ----------------------------------------------------------------------------------------------
// weight packer: flat fp32 master -> packed image (fp32 head block + bf16 SW64 K-slices)
// ----------------------------------------------------------------------------------------------
__global__ void mlp_pack_kernel(const __grid_constant__ Program prog, const float* __restrict__ master,
                                uint8_t* __restrict__ packed) {
    const int s = blockIdx.x;
    if (s == prog.n_slices) {            // head block
        float* h = reinterpret_cast<float*>(packed);
        const Layer& la = prog.layers[prog.alpha_layer];
        const Layer& lr = prog.layers[prog.rgb_layer];
        for (int i = threadIdx.x; i < kHeadBytes / 4; i += blockDim.x) {
            float v = 0.0f;
            if (i < 256) v = master[la.w_off + i];                               // W_alpha[in=256, out=1]
            else if (i == 256) v = master[la.b_off];
            else if (i >= 260 && i < 260 + 384) { int c = (i - 260) / 128, k = (i - 260) % 128; v = master[lr.w_off + k * 3 + c]; }
            else if (i >= 644 && i < 647) v = master[lr.b_off + (i - 644)];
            h[i] = v;
        }
        return;
    }
    const uint32_t sl = prog.slices[s];
    const Step& step = prog.steps[slice_step(sl)];
    const Layer& L = prog.layers[step.layer];
    const int N = step.n_out;
    uint8_t* dst = packed + prog.slice_off[s];
    const int h0 = slice_half0(sl), nh = slice_nh(sl);
    for (int idx = threadIdx.x; idx < N * 32; idx += blockDim.x) {
        const int n = idx >> 5, kk = idx & 31, half = kk >> 4;
        float v = 0.0f;
        if (half >= h0 && half < h0 + nh) {
            if (slice_is_enc(sl)) {
                const int col = slice_blk(sl) * 32 + kk;
                const int ref = step.phase == 0 ? pos_col_to_ref(col, prog.n_geo, prog.ipe) : dir_col_to_ref(col, prog.n_app);
                if (ref == kRefOnes) v = master[L.b_off + n];
                else if (ref >= 0 && step.enc_data) v = master[L.w_off + (int64_t)ref * L.fan_out + n];
            } else {
                const int c = slice_blk(sl) * 32 + kk;
                v = master[L.w_off + (int64_t)(step.main_row0 + c) * L.fan_out + n];
            }
        }
        *reinterpret_cast<__nv_bfloat16*>(dst + sw64_offset(n, kk, N)) = __float2bfloat16(v);
    }
}

int build_program(ntx_mlp* m) {
    const ntx_mlp_config& c = m->cfg;
    Program& p = m->prog;
    memset(&p, 0, sizeof(p));
    const int g = c.n_geo, a = c.n_app, W = c.width;
    p.n_geo = g; p.n_app = a;
    p.ipe = c.n_pos == 6;
    p.P = (p.ipe ? 6 * c.pos_freqs : 3 * (1 + 2 * c.pos_freqs)) + g * (1 + 2 * c.param_freqs);   // layer.py:21-23 / :33-37
    p.D = 3 * (1 + 2 * c.dir_freqs) + a * (1 + 2 * c.param_freqs);

    // ---- Dense layer table in graph order (model.py:104-123)
    int64_t off = 0;
    int nl = 0;
    auto add_layer = [&](int fin, int fout) {
        Layer& L = p.layers[nl];
        L.fan_in = fin; L.fan_out = fout; L.w_off = off; off += (int64_t)fin * fout; L.b_off = off; off += fout;
        return nl++;
    };
    int trunk[16];
    int k = p.P;
    for (int i = 0; i < c.depth; ++i) {
        trunk[i] = add_layer(k, W);
        k = ((c.skip_mask >> i) & 1) ? W + p.P : W;
    }
    p.alpha_layer = add_layer(k, 1);
    const int feature = add_layer(k, W);
    int colorhid[8];
    k = W + p.D;
    for (int i = 0; i < c.color_depth; ++i) { colorhid[i] = add_layer(k, W); k = W; }
    const int colorhalf = add_layer(k, W / 2);
    p.rgb_layer = add_layer(W / 2, 3);
    p.n_layers = nl;
    m->n_params = off;

    // ---- MMA steps
    int last_pos_step = 0;   // last step that consumes pos-encoding features
    for (int i = 0; i < c.depth; ++i) if (i > 0 && ((c.skip_mask >> (i - 1)) & 1)) last_pos_step = i;
    int ns = 0, nsl = 0;
    uint32_t boff = kHeadBytes;
    auto add_step = [&](int layer, int n_out, bool has_main, int enc_data, int phase, int main_row0, uint32_t epi) {
        Step& s = p.steps[ns];
        s.layer = layer; s.n_out = n_out; s.main_row0 = main_row0; s.enc_data = enc_data; s.phase = phase; s.epi = epi;
        s.slice_begin = nsl;
        s.n_main = has_main ? kMainBlocks : 0;
        if (has_main)
            for (int b = 0; b < kMainBlocks; ++b) p.slices[nsl++] = make_slice(0, b, 0, 2, ns);
        if (enc_data) {
            const int halves = ((phase == 0 ? pos_chunks(g) : dir_chunks(a)) + 1) / 2;
            for (int h = 0; h < halves; h += 2) p.slices[nsl++] = make_slice(1, h / 2, 0, halves - h >= 2 ? 2 : 1, ns);
            s.enc_n = (halves + 1) / 2; s.enc_blk0 = 0; s.enc_half0 = 0; s.enc_last_nh = (halves & 1) ? 1 : 2;
        } else {
            const int oc = phase == 0 ? pos_ones_col() : dir_ones_col(a);
            p.slices[nsl++] = make_slice(1, oc / 32, (oc % 32) / 16, 1, ns);
            s.enc_n = 1; s.enc_blk0 = oc / 32; s.enc_half0 = (oc % 32) / 16; s.enc_last_nh = 1;
        }
        s.slice_end = nsl;
        for (int q = s.slice_begin; q < s.slice_end; ++q) { p.slice_off[q] = boff; boff += (uint32_t)n_out * 64; }
        return ns++;
    };
    for (int i = 0; i < c.depth; ++i) {
        const bool skip_in = i > 0 && ((c.skip_mask >> (i - 1)) & 1);
        uint32_t epi = EPI_RELU;
        if (i == c.depth - 1) epi |= EPI_ALPHA;
        if (i == last_pos_step) epi |= EPI_SWITCH;
        add_step(trunk[i], W, i > 0, (i == 0 || skip_in) ? 1 : 0, i <= last_pos_step ? 0 : 1, skip_in ? p.P : 0, epi);
    }
    add_step(feature, W, true, 0, 1, 0, 0u);                                    // model.py:114 (linear)
    for (int i = 0; i < c.color_depth; ++i)
        add_step(colorhid[i], W, true, i == 0 ? 1 : 0, 1, i == 0 ? p.D : 0, EPI_RELU);
    add_step(colorhalf, W / 2, true, c.color_depth == 0 ? 1 : 0, 1, c.color_depth == 0 ? p.D : 0, EPI_RELU | EPI_FINAL);
    p.n_steps = ns;
    p.n_slices = nsl;
    p.slice_off[nsl] = boff;
    m->packed_bytes = boff;
    return NTX_OK;
}

}  // namespace

NTX_API int ntx_mlp_create(const ntx_mlp_config* cfg, ntx_mlp_t** out) {
    NTX_REQUIRE(cfg && out, "mlp_create: null pointer");
    const ntx_mlp_config& c = *cfg;
    if (!((c.n_pos == 3 || c.n_pos == 6) && c.pos_freqs == 10 && c.dir_freqs == 4 && c.param_freqs == 4 && c.width == 256)) {
        ntx_set_error("mlp_create: the sm_100a kernel implements n_pos=3 (or 6: integrated encoding), pos_freqs=10, dir_freqs=4, param_freqs=4, "
                      "width=256 (got %d,%d,%d,%d,%d)", c.n_pos, c.pos_freqs, c.dir_freqs, c.param_freqs, c.width);
        return NTX_E_UNSUPPORTED;
    }
    if (!(c.n_geo >= 0 && c.n_geo <= 3 && c.n_app >= 0 && c.n_app <= 7 && c.depth >= 1 && c.depth <= 12 &&
          c.color_depth >= 0 && c.color_depth <= 2 && ((c.skip_mask >> (c.depth - 1)) == 0))) {
        ntx_set_error("mlp_create: unsupported n_parameters=[%d,%d] depth=%d color_depth=%d skip_mask=0x%x "
                      "(need n_geo<=3, n_app<=7, depth<=12, color_depth<=2, no skip on the last trunk layer)",
                      c.n_geo, c.n_app, c.depth, c.color_depth, c.skip_mask);
        return NTX_E_UNSUPPORTED;
    }
    ntx_mlp* m = new (std::nothrow) ntx_mlp();
    NTX_REQUIRE(m, "mlp_create: out of host memory");
    m->cfg = c;
    build_program(m);
    if (m->prog.n_steps > kMaxSteps || m->prog.n_slices > kMaxSlices) {
        delete m;
        ntx_set_error("mlp_create: network too deep for the step table");
        return NTX_E_UNSUPPORTED;
    }
    *out = m;
    return NTX_OK;
}

NTX_API int ntx_mlp_destroy(ntx_mlp_t* mlp) {
    delete mlp;
    return NTX_OK;
}

NTX_API int64_t ntx_mlp_num_params(const ntx_mlp_t* mlp) { return mlp ? mlp->n_params : -1; }
NTX_API int ntx_mlp_num_layers(const ntx_mlp_t* mlp) { return mlp ? mlp->prog.n_layers : -1; }
NTX_API int64_t ntx_mlp_packed_bytes(const ntx_mlp_t* mlp) { return mlp ? mlp->packed_bytes : -1; }

NTX_API int ntx_mlp_layer_info(const ntx_mlp_t* mlp, int i, int64_t* w_offset, int64_t* b_offset, int32_t* fan_in,
                               int32_t* fan_out) {
    NTX_REQUIRE(mlp && i >= 0 && i < mlp->prog.n_layers, "mlp_layer_info: bad layer index %d", i);
    const Layer& L = mlp->prog.layers[i];
    if (w_offset) *w_offset = L.w_off;
    if (b_offset) *b_offset = L.b_off;
    if (fan_in) *fan_in = L.fan_in;
    if (fan_out) *fan_out = L.fan_out;
    return NTX_OK;
}

NTX_API int ntx_mlp_pack_weights(const ntx_mlp_t* mlp, const float* master, void* packed, void* stream) {
    NTX_REQUIRE(mlp && master && packed, "mlp_pack_weights: null pointer");
    NTX_REQUIRE(((uintptr_t)packed & 15) == 0, "mlp_pack_weights: packed image must be 16-byte aligned");
    mlp_pack_kernel<<<mlp->prog.n_slices + 1, 256, 0, (cudaStream_t)stream>>>(mlp->prog, master, (uint8_t*)packed);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

// 128-row tiles the training kernels touch for n_rows rows (whole 512-row CTA-pair batches)
NTX_API int64_t ntx_mlp_train_tiles(int64_t n_rows) {
    return ((n_rows + 511) / 512) * 4;
}

static int launch_forward(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_inputs* in, float* rgbs,
                          void* enc_pos, void* enc_dir, void* acts, void* acts_img, void* masks, int64_t n_tiles,
                          bool dump, void* stream, void* acts_img8 = nullptr, const float* inv_scale8 = nullptr,
                          uint32_t* amax8 = nullptr) {
    NTX_REQUIRE(mlp && packed && in, "mlp_forward: null pointer");
    NTX_REQUIRE(in->n_rows >= 0, "mlp_forward: negative row count");
    if (in->n_rows == 0) return NTX_OK;
    NTX_REQUIRE(rgbs && in->dir, "mlp_forward: null rgbs / dir");
    NTX_REQUIRE(in->pos || (in->rays_o && in->rays_d && in->z), "mlp_forward: need pos, or rays_o + rays_d + z");
    NTX_REQUIRE(!mlp->prog.ipe || (in->pos && in->pos_cov), "mlp_forward: an n_pos = 6 model needs pos (means) and pos_cov (variances)");
    NTX_REQUIRE(in->rows_per_group >= 1, "mlp_forward: rows_per_group must be >= 1");
    const int np = mlp->cfg.n_geo + mlp->cfg.n_app;
    NTX_REQUIRE(np == 0 || in->params, "mlp_forward: the model takes %d parameters but params is NULL", np);
    NTX_REQUIRE(in->blur_idx < np, "mlp_forward: blur_idx %d out of range", in->blur_idx);
    NTX_REQUIRE((((uintptr_t)packed | (uintptr_t)rgbs) & 15) == 0, "mlp_forward: packed / rgbs must be 16-byte aligned");
    if (dump) {
        NTX_REQUIRE(enc_pos && enc_dir && (acts_img || acts_img8), "mlp_forward_train: null save buffer");
        NTX_REQUIRE(!acts_img8 || (inv_scale8 && amax8 && ((uintptr_t)acts_img8 & 15) == 0 && mlp->prog.n_steps <= 16),
                    "mlp_forward_train: 8-bit images need inv_scale8[n_steps], amax8[n_steps] and 16-byte alignment");
        NTX_REQUIRE((((uintptr_t)enc_pos | (uintptr_t)enc_dir | (uintptr_t)acts | (uintptr_t)acts_img | (uintptr_t)masks) & 15) == 0,
                    "mlp_forward_train: save buffers must be 16-byte aligned");
        {
            NTX_REQUIRE(masks, "mlp_forward_train: tile images need the mask buffer too");
            NTX_REQUIRE(n_tiles >= ntx_mlp_train_tiles(in->n_rows), "mlp_forward_train: %lld tiles given, %lld needed",
                        (long long)n_tiles, (long long)ntx_mlp_train_tiles(in->n_rows));
        }
    }
    // NTX_MLP_CTAS=1 selects the single-CTA kernel (cta_group::1); default is CTA pairs (cta_group::2)
    static const int ctas = [] { const char* e = getenv("NTX_MLP_CTAS"); return (e && e[0] == '1') ? 1 : 2; }();
    using KernelFn = void (*)(const Program, const KernelArgs);
    const int mode = acts_img8 ? 2 : (dump ? 1 : 0);
    KernelFn fn = ctas == 2 ? (mode == 2 ? mlp_forward_kernel<true, 2, true> : mode == 1 ? mlp_forward_kernel<true, 2> : mlp_forward_kernel<false, 2>)
                            : (mode == 2 ? mlp_forward_kernel<true, 1, true> : mode == 1 ? mlp_forward_kernel<true, 1> : mlp_forward_kernel<false, 1>);
    static bool attr_set[64][3][2] = {};
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && !attr_set[dev][mode][ctas - 1]) {
        NTX_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
        attr_set[dev][mode][ctas - 1] = true;
    }
    KernelArgs ka;
    if (ctas == 2) {
        NTX_REQUIRE((mlp->packed_bytes & 2047) == 0, "mlp_forward: packed image is not a multiple of 2 KB");
        if (int rc = ntx_weight_map(&ka.w_map128, packed, mlp->packed_bytes, 4)) return rc;
        if (int rc = ntx_weight_map(&ka.w_map64, packed, mlp->packed_bytes, 2)) return rc;
    } else {
        memset(&ka.w_map128, 0, sizeof(CUtensorMap));
        memset(&ka.w_map64, 0, sizeof(CUtensorMap));
    }
    ka.in = *in;
    ka.packed = (const uint8_t*)packed;
    ka.rgbs = rgbs;
    ka.dump_enc_pos = (uint8_t*)enc_pos;
    ka.dump_enc_dir = (uint8_t*)enc_dir;
    ka.dump_acts = (uint4*)acts;
    ka.dump_img = (uint8_t*)acts_img;
    ka.dump_masks = (uint4*)masks;
    ka.n_tiles = n_tiles;
    ka.dump_img8 = (uint8_t*)acts_img8;
    ka.inv_scale8 = inv_scale8;
    ka.amax8 = amax8;
    const int64_t rows_per_unit = (int64_t)kRowsPerCta * ctas;
    const int64_t n_units = (in->n_rows + rows_per_unit - 1) / rows_per_unit;
    const int64_t max_units = ntx_sm_count() / ctas;
    const int grid = (int)(n_units < max_units ? n_units : max_units) * ctas;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kSmemBytes;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ctas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    NTX_CUDA(cudaLaunchKernelEx(&cfg, fn, mlp->prog, ka));
    return NTX_OK;
}

NTX_API int ntx_mlp_forward(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_inputs* in, float* rgbs,
                            void* stream) {
    return launch_forward(mlp, packed, in, rgbs, nullptr, nullptr, nullptr, nullptr, nullptr, 0, false, stream);
}

NTX_API int ntx_mlp_forward_train(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_inputs* in, float* rgbs,
                                  void* enc_pos, void* enc_dir, void* acts, void* acts_img, void* masks,
                                  int64_t n_tiles, void* stream) {
    return launch_forward(mlp, packed, in, rgbs, enc_pos, enc_dir, acts, acts_img, masks, n_tiles, true, stream);
}

NTX_API int ntx_mlp_forward_train_fp8(const ntx_mlp_t* mlp, const void* packed, const ntx_mlp_inputs* in, float* rgbs,
                                      void* enc_pos, void* enc_dir, void* acts_img8, const float* inv_scale8, void* amax8,
                                      void* masks, int64_t n_tiles, void* stream) {
    NTX_REQUIRE(acts_img8, "mlp_forward_train_fp8: null image buffer");
    return launch_forward(mlp, packed, in, rgbs, enc_pos, enc_dir, nullptr, nullptr, masks, n_tiles, true, stream, acts_img8,
                          inv_scale8, (uint32_t*)amax8);
}

NTX_API int ntx_untile(const void* img, int64_t n_rows, int cols, int64_t tile_bytes, void* out, void* stream) {
    NTX_REQUIRE(n_rows >= 0 && cols >= 32 && cols <= 256 && cols % 32 == 0 && tile_bytes >= (int64_t)cols * 256,
                "untile: bad shape");
    if (n_rows == 0) return NTX_OK;
    NTX_REQUIRE(img && out, "untile: null pointer");
    const int64_t total = n_rows * (cols / 8);
    untile_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const uint8_t*)img, n_rows, cols, tile_bytes, (uint4*)out);
    NTX_LAUNCH_CHECK();
    return NTX_OK;
}

NTX_API int ntx_mlp_num_steps(const ntx_mlp_t* mlp) { return mlp ? mlp->prog.n_steps : -1; }

NTX_API int ntx_mlp_step_info(const ntx_mlp_t* mlp, int st, int32_t* layer, int32_t* n_out, int32_t* main_row0,
                              int32_t* enc_data, int32_t* phase, int32_t* relu) {
    NTX_REQUIRE(mlp && st >= 0 && st < mlp->prog.n_steps, "mlp_step_info: bad step %d", st);
    const Step& s = mlp->prog.steps[st];
    if (layer) *layer = s.layer;
    if (n_out) *n_out = s.n_out;
    if (main_row0) *main_row0 = s.main_row0;
    if (enc_data) *enc_data = s.enc_data;
    if (phase) *phase = s.phase;
    if (relu) *relu = (s.epi & EPI_RELU) ? 1 : 0;
    return NTX_OK;
}

NTX_API int ntx_mlp_enc_columns(const ntx_mlp_t* mlp, int phase, int32_t* ref_index) {
    NTX_REQUIRE(mlp && ref_index && (phase == 0 || phase == 1), "mlp_enc_columns: bad argument");
    for (int c = 0; c < kEncCols; ++c)
        ref_index[c] = phase == 0 ? pos_col_to_ref(c, mlp->prog.n_geo, mlp->prog.ipe) : dir_col_to_ref(c, mlp->prog.n_app);
    return NTX_OK;
}

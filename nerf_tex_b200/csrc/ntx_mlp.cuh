// Host/device description of the fused MLP "program" shared by the forward kernel, the weight
// packer and (later) the backward kernels.
//
// Encoded-input column order (ours; the packer permutes weight rows to match, so results equal
// the reference's concat order of network/layer.py:23 and network/model.py:93,101):
//
//   pos phase (trunk0 .. last skip consumer)           dir phase (afterwards)
//   cols  0..59 : x_i, i<3: [sin 2^k x_i, cos 2^k x_i]  cols 8i..8i+7, i<3 : [sin,cos](2^k d_i), k<4
//                 at 20 i + 2 k (k < 10)                cols 8(3+j)..    : [sin,cos](2^k pa_j), j<a
//   col   60    : 1.0  (bias column)                    col  8(3+a)      : 1.0 (bias column)
//   cols 61..63 : x_0, x_1, x_2                         cols 8(3+a)+1..3 : d_0, d_1, d_2
//   cols 64+8j..: [sin,cos](2^k pg_j), k<4, j<g         cols 8(3+a)+4..7 : pa_0..pa_3
//   cols 64+8g..: pg_0..pg_{g-1}, 0...                  cols 8(4+a)..    : pa_4.., 0...
#pragma once

#include <stdint.h>

namespace ntxmlp {

constexpr int kMaxSteps = 16;
constexpr int kMaxSlices = 176;
constexpr int kWidth = 256;
constexpr int kTileRows = 128;        // rows per UMMA tile (M)
constexpr int kTilesPerCta = 2;       // two tiles share every staged weight slice
constexpr int kRowsPerCta = kTileRows * kTilesPerCta;
constexpr int kBlockBytes = kTileRows * 64;           // one 32-column K block of a 128-row tile (SW64)
constexpr int kMainBlocks = kWidth / 32;              // 8
constexpr int kEncBlocks = 3;                         // up to 96 encoded columns
constexpr int kStages = 3;
constexpr int kStageBytes = 256 * 64;                 // one [N=256 x K=32] bf16 weight slice
constexpr int kHeadFloats = 256 + 4 + 3 * 128 + 4;    // w_alpha, b_alpha(+pad), w_rgb[3][128], b_rgb(+pad)
constexpr int kHeadBytes = 4096;                      // head block is padded to 4 KB in the packed image

enum : uint32_t {
    EPI_RELU = 1u,        // relu before storing
    EPI_ALPHA = 2u,       // also evaluate the density head on this layer's fp32 output
    EPI_SWITCH = 4u,      // after this layer, overwrite the pos encoding with the dir encoding
    EPI_FINAL = 8u        // colour half layer: relu, rgb head, write rgbs
};

struct Step {
    int32_t layer;          // index into the Dense layer table (flat master order)
    int32_t n_out;          // 256 or 128
    int32_t slice_begin;    // [slice_begin, slice_end) in Program::slices
    int32_t slice_end;
    int32_t main_row0;      // row of W that multiplies main-buffer column 0 (P or D after a concat, else 0)
    int32_t enc_data;       // 1: this layer consumes the resident encoding's features, 0: bias column only
    int32_t phase;          // 0: pos encoding resident, 1: dir encoding resident
    uint32_t epi;           // EPI_* flags
    // compact form of the slice list for the MMA issuer (no per-slice table lookups)
    int32_t n_main;         // leading slices that read main-buffer blocks 0..n_main-1 (two K=16 halves each)
    int32_t enc_n;          // trailing slices that read encoding blocks enc_blk0..enc_blk0+enc_n-1
    int32_t enc_blk0;
    int32_t enc_half0;      // first K=16 half used inside the encoding block(s)
    int32_t enc_last_nh;    // halves issued for the last encoding slice (the others use 2)
    int32_t pad_[3];
};

// slice word: bit0 a_is_enc | bits1-4 a_blk | bit5 first half | bits6-7 n_halves | bits8-15 step
//             | bits16-29 (byte offset of the A operand inside its tile) >> 4, ready for the descriptor
__host__ __device__ inline uint32_t make_slice(int is_enc, int blk, int half0, int nh, int step) {
    const uint32_t a_rel = (is_enc ? (uint32_t)(kMainBlocks * kBlockBytes) : 0u) + (uint32_t)blk * kBlockBytes + (uint32_t)half0 * 32u;
    return (uint32_t)is_enc | ((uint32_t)blk << 1) | ((uint32_t)half0 << 5) | ((uint32_t)nh << 6) | ((uint32_t)step << 8) |
           ((a_rel >> 4) << 16);
}
__host__ __device__ inline uint32_t slice_a16(uint32_t s) { return (s >> 16) & 0x3FFFu; }
__host__ __device__ inline int slice_is_enc(uint32_t s) { return s & 1; }
__host__ __device__ inline int slice_blk(uint32_t s) { return (s >> 1) & 15; }
__host__ __device__ inline int slice_half0(uint32_t s) { return (s >> 5) & 1; }
__host__ __device__ inline int slice_nh(uint32_t s) { return (s >> 6) & 3; }
__host__ __device__ inline int slice_step(uint32_t s) { return (s >> 8) & 255; }

struct Layer {
    int64_t w_off, b_off;
    int32_t fan_in, fan_out;
};

struct Program {
    int32_t n_steps, n_slices;
    int32_t n_geo, n_app;
    int32_t ipe;                  // 1: integrated positional encoding of (mean, variance) inputs (layer.py:25-41)
    int32_t P, D;                 // reference encoded widths (72 / 63 for n_parameters [1,4])
    int32_t alpha_layer, rgb_layer;
    Step steps[kMaxSteps];
    uint32_t slices[kMaxSlices];
    uint32_t slice_off[kMaxSlices + 1];   // byte offset of every slice in the packed image
    Layer layers[kMaxSteps + 2];
    int32_t n_layers;
};

constexpr int kRefOnes = -1, kRefZero = -2;

// our pos-phase column -> index into the reference's pos_inputs_map (model.py:77,88-93).  With the integrated
// encoding (layer.py:33-37) the map has no identity term and is ordered [sin block, cos block], each frequency-major.
__host__ __device__ inline int pos_col_to_ref(int col, int g, int ipe = 0) {
    const int base = ipe ? 60 : 63;
    if (col < 60) {
        int i = col / 20, r = col % 20, k = r >> 1, c = r & 1;
        return ipe ? 30 * c + 3 * k + i : 3 + 6 * k + 3 * c + i;
    }
    if (col == 60) return kRefOnes;
    if (col < 64) return ipe ? kRefZero : col - 61;
    int q = (col - 64) >> 3, e = (col - 64) & 7;
    if (q < g) { int k = e >> 1, c = e & 1; return base + g + 2 * g * k + g * c + q; }
    if (q == g && e < g) return base + e;
    return kRefZero;
}
// our dir-phase column -> index into the reference's dir_inputs_map (model.py:78,96-101)
__host__ __device__ inline int dir_col_to_ref(int col, int a) {
    int q = col >> 3, e = col & 7;
    if (q < 3) { int k = e >> 1, c = e & 1; return 3 + 6 * k + 3 * c + q; }
    if (q < 3 + a) { int j = q - 3, k = e >> 1, c = e & 1; return 27 + a + 2 * a * k + a * c + j; }
    if (q == 3 + a) {
        if (e == 0) return kRefOnes;
        if (e < 4) return e - 1;
        return (e - 4 < a) ? 27 + (e - 4) : kRefZero;
    }
    if (q == 4 + a) return (4 + e < a) ? 27 + 4 + e : kRefZero;
    return kRefZero;
}
__host__ __device__ inline int pos_chunks(int g) { return 8 + (g > 0 ? g + 1 : 0); }
__host__ __device__ inline int dir_chunks(int a) { return 3 + a + 1 + (a > 4 ? 1 : 0); }
__host__ __device__ inline int pos_ones_col() { return 60; }
__host__ __device__ inline int dir_ones_col(int a) { return 8 * (3 + a); }

}  // namespace ntxmlp

struct ntx_mlp {
    ntx_mlp_config cfg;
    ntxmlp::Program prog;
    int64_t n_params;
    int64_t packed_bytes;
};

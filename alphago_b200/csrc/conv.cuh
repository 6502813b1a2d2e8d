// Shifted-window implicit-GEMM convolution on tcgen05 (see conv.cu).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ag {

// Activation layout in HBM ("padded pixel rows"): a bf16 matrix [rows][C] whose row index is
//   m = n*img_stride + x*row_stride + y      (x, y in [0, size); n = image)
// with at least `pad` all-zero slots after every board row and `pad` all-zero board rows after
// every image, so a KxK 'same' convolution tap (dx, dy) is the plain row shift dx*row_stride + dy:
// the halo is read from the zero slots (and, before the first / after the last image, from TMA's
// out-of-bounds zero fill).  3x3 layers use (row_stride, img_stride) = (20, 400); the 5x5 first
// layer reads (21, 441).  Pad slots are never written after the one-time zero fill.
struct PixelLayout {
    int size;        // board edge (19)
    int row_stride;  // slots per board row
    int img_stride;  // slots per image
};

constexpr int kConvMaxTaps = 25;
constexpr int kConvMaskWords = 8;  // 256 channels / 32: one 32-byte mask row per pixel row

struct ConvArgs {
    int m_total;          // rows of the A matrix the GEMM covers (= images * in.img_stride)
    int n;                // output channels of this GEMM (UMMA N): multiple of 8... 16, <= 256
    int taps;
    int kchunks;          // K per tap / 64
    int off[kConvMaxTaps];  // A row shift per tap
    int b_tap_rows;       // rows of the B tensor per tap (tap t's matrix starts at row t*b_tap_rows)
    PixelLayout in, dst;  // row decode of the GEMM's M index, and where the result row goes
    const float* bias;    // [n] or nullptr
    // ReLU bookkeeping as bitmasks, kConvMaskWords uint32 per pixel row (bit c%32 of word c/32 = channel c was > 0):
    uint32_t* mask_out;         // forward (relu): written for every output row, indexed like `out` (dst layout rows); may be nullptr
    const uint32_t* gate_mask;  // ReLU backward: nullptr, or rows in gate_layout; the result is zeroed where the bit is clear
    PixelLayout gate_layout;
    __nv_bfloat16* out;   // rows in dst layout, out_ld wide
    int out_ld;
    int relu;
    int halo, win_rows;   // filled by conv_gemm_launch (window kernel): largest |off|, rows of the activation window box
    int ksteps_last;      // chain kernel: K=16 steps worth issuing in the LAST 64-channel slice (0 = all 4): the channels behind
                          // the real ones are zero padding (48 input planes fill 3 of the first layer's 4 steps)
    int debug;            // measurement only (AG_CONV_DEBUG): 1 = skip A loads, 2 = skip B loads, 4 = skip epilogue memory traffic
};

// cuTensorMapEncodeTiled through the runtime's driver entry point (no libcuda link dependency).
// 2D bf16 matrix [rows][cols], box [box_rows][64 columns], 128-byte swizzle, zero OOB fill.
int make_tmap_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows);

// rows of the activation window box (128 + 2 * largest |off|) the window kernel wants tmap_a_win built with; 0 = too tall
int conv_window_rows(const int* off, int taps);

// D[m, :] = sum_t sum_k A[m + off[t], k] * B[t][:, k]   (+bias, ReLU / gate), bf16 in, fp32 accumulate
// tmap_b_half (optional): the same B tensor with a box of N/2 rows; when given (and N % 32 == 0) a CTA-pair kernel runs
// tmap_a_win (optional): the A tensor with a box of conv_window_rows() rows: the pair kernel then fetches each activation
// row once per channel slice instead of once per tap
int conv_gemm_launch(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap* tmap_b_half, const CUtensorMap* tmap_a_win,
                     const ConvArgs& args, cudaStream_t stream);

// One entry of a layer chain run by a single persistent launch (conv.cu, conv_chain_kernel)
struct __align__(64) ChainLayer {
    CUtensorMap tmap_aw;   // A: the layer's input with a box of win_rows rows
    CUtensorMap tmap_bh;   // B: its weights with a box of N/2 rows
    ConvArgs a;
    int n_pairs;           // 256-row tiles of this layer
    int tile0;             // g of its first tile
    int flag0;             // index of its first flag
    int dep;               // 0: input comes from outside the chain; 1: from the layer before, same row space; 2: from the layer
                           // before, whose GEMM rows live in another layout than its output (first layer 5x5): wait by image
    int prev_in_img_stride;  // img_stride of the previous layer's GEMM rows
    int prev_rows;           // GEMM rows of the previous layer (bounds its flags)
    // Image-aligned tiling: 0 = half-tile h covers GEMM rows [128h, 128h+128); k > 0 = every image has k half-tiles of its
    // own, starting at the image's first row (3 x 128 = 384 rows cover the 19 x 20 = 380 rows of an image that produce
    // output; the 20-slot zero row behind it is never computed: 4 % fewer tiles than 400 / 128 per image)
    int tiles_img;
    int prev_tiles_img;      // the same of the layer before (maps its rows to its half-tile flags)
};


constexpr int kChainMaxLayers = 16;   // per launch (the table is a kernel parameter: 16 x ~0.6 KB of the 32 KB limit)
struct ChainParams {
    ChainLayer layers[kChainMaxLayers];
    int n_layers;
    int total_tiles;
};

bool conv_chain_enabled();
// runs P.layers[0..n_layers) in one launch; flags: one unsigned per 128-row tile of every layer (layer l's start at flag0),
// epoch: a value no earlier launch on these flags used; win_rows_max / n_max: largest window box and N over the layers
int conv_chain_launch(const ChainParams& P, unsigned int* d_flags, unsigned int epoch, int win_rows_max, int n_max, cudaStream_t stream);

}  // namespace ag

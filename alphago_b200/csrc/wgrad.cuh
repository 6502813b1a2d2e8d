// Split-K weight-gradient GEMM (see wgrad.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "conv.cuh"

namespace ag {

constexpr int kWgradMaxSplits = 32;

struct WgradArgs {
    int m_total;   // pixel rows reduced over (= images * img_stride of the layer's INPUT layout)
    int taps;
    int off[kConvMaxTaps];  // forward row shift per tap
    int ldout;     // row length of G (output channels padded to 64): dW rows per tap
    int kin;       // row length of X (input channels padded to 64): dW columns, UMMA N
    int splits;    // split-K factor
    int tiles;     // filled by wgrad_launch
    int tpg;       // filled by wgrad_launch: taps per B operand (1 = shifted-G tiling, >= 2 = taps-in-N tiling, see wgrad.cu)
    int groups;    // filled by wgrad_launch: tap groups (taps-in-N tiling)
    float* scratch;  // [splits][taps*ldout][kin] fp32
    float* bias_scratch;  // nullptr, or [splits][ldout] fp32: column sums of G (the bias gradient) ride along as one more B column
};

size_t wgrad_smem_bytes(int taps, int kin, bool with_bias);
int wgrad_plan_splits(int taps, int ldout, int kin, int m_total);
// tmap_g: G [m_total rows][ldout], tmap_x: X [m_total rows][kin], both with box [64][64] and EXACTLY
// m_total rows (rows past the batch must read as zeros).  dw: [taps][ldout][kin] fp32.
// db (optional, needs args.bias_scratch and kin <= 240): [ldout] fp32 bias gradient = sum over all rows of G
int wgrad_launch(const CUtensorMap& tmap_g, const CUtensorMap& tmap_x, WgradArgs args, float* dw, float* db, cudaStream_t stream);

}  // namespace ag

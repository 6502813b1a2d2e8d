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
    float* scratch;  // [splits][taps*ldout][kin] fp32
};

size_t wgrad_smem_bytes(int kin);
int wgrad_plan_splits(int taps, int ldout, int m_total);
// tmap_g: G [m_total rows][ldout], tmap_x: X [m_total rows][kin], both with box [64][64] and EXACTLY
// m_total rows (rows past the batch must read as zeros).  dw: [taps][ldout][kin] fp32.
int wgrad_launch(const CUtensorMap& tmap_g, const CUtensorMap& tmap_x, WgradArgs args, float* dw, cudaStream_t stream);

}  // namespace ag

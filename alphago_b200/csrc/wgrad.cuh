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
    int kb_img;    // 0: K blocks tile the rows linearly; k > 0: every image has k K blocks of its own starting at its first row
                   // (6 x 64 = 384 rows cover the 380 rows of a 19 x 20 image that can be non-zero; the zero row behind them is
                   // skipped: 4 % fewer UMMAs).  Set by the caller together with img_stride; both operands share the layout.
    int img_stride;
    int win_rows;  // filled by wgrad_launch from the maps
    int krows;     // filled by wgrad_launch from the maps: pixel rows per pipeline stage (32 or 64)
    int debug;     // experiments only (AG_WGRAD_DEBUG)
    float* scratch;  // splits * wgrad_partial_floats() fp32, laid out by the kernel
    float* bias_scratch;  // nullptr, or [splits][ldout] fp32: column sums of G (the bias gradient) ride along as one more B column
};

// floats of one split's partials in `scratch` (>= taps * ldout * kin: whole accumulators per CTA tile, see wgrad.cu)
size_t wgrad_partial_floats(int taps, int ldout, int kin);
// sm_budget: SMs the launch may fill (0 = all): data-parallel training leaves a few to the collective running beside it
int wgrad_plan_splits(int taps, int ldout, int kin, int m_total, int sm_budget);
// Tensor maps of one weight-gradient problem: G [rows][ldout] and X [rows][kin] with box [krows][64] and EXACTLY `rows`
// rows (rows past the batch must read as zeros), plus G with a box of win_rows rows (the shared window, wgrad.cu).
struct WgradMaps {
    CUtensorMap g, gw, x;
    int win_rows, krows;
};
int wgrad_make_maps(WgradMaps* m, const void* g, const void* x, uint64_t rows, int ldout, int kin, int taps, const int* off);
// dw: [taps][ldout][kin] fp32.
// db (optional, needs args.bias_scratch and kin <= 240): [ldout] fp32 bias gradient = sum over all rows of G
int wgrad_launch(const WgradMaps& maps, WgradArgs args, float* dw, float* db, cudaStream_t stream);

// One layer of a chained weight-gradient launch (wgrad.cu, wgrad_chain_kernel): same geometry for all, own buffers each.
struct WgradChainEntry {
    const WgradMaps* maps;
    float* scratch;            // splits * tiles * 256 * kin floats (tiles = ceil(taps * ldout / 256)), laid out by the kernel;
                               // consecutive entries must alternate between two buffers
    float* bias_scratch;       // [splits][ldout] or nullptr (then db is not produced)
    float* dw;
    float* db;
    unsigned int* counters;    // [tiles] arrival counters (running totals over launches)
    unsigned int target;       // their value once this launch's partials are all in: (launches so far, this one included) * splits
    unsigned int* done;        // nullptr, or a running count of CTAs whose share of this layer's gradient is final: it reaches
                               // (launches so far) * tiles * splits when dw / db of the layer are complete
};
// CTAs of one chained launch for this geometry (tiles * splits)
int wgrad_chain_ctas(const WgradArgs& args);
// can this problem run chained (taps not folded into N, all CTAs resident at once, ...)?
bool wgrad_chain_ok(const WgradArgs& args, const WgradMaps& maps, bool with_bias);
// weight (+bias) gradients of n_layers layers in one launch, split reduction included; args: m_total, taps, off, ldout, kin, splits
int wgrad_chain_launch(WgradArgs args, const WgradChainEntry* entries, int n_layers, cudaStream_t stream);
constexpr int kWgradChainMaxLayers = 16;

}  // namespace ag

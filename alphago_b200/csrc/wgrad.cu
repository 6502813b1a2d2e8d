// Weight gradient of a KxK 'same' convolution as a split-K tcgen05 GEMM over the pixel dimension
// (the backward of Keras' Convolution2D inside fit_generator, supervised_policy_trainer.py:207-213):
//
//     dW[tap][co][ci] = sum_m G[m - shift(tap)][co] * X[m][ci]
//
// G = gradient w.r.t. the layer's pre-activation output, X = the layer's input, both padded-pixel-
// row matrices (conv.cuh) whose zero pad slots make the halo terms vanish.  The reduction index m
// is the ROW index of both matrices, so both operands are MN-major: TMA brings [64 pixel rows][64
// channels] boxes (128-byte swizzle) that the UMMA reads transposed.  A CTA owns 128 rows of the
// flattened (tap, co) axis = two 64-channel chunks, each loaded with its own tap shift, and a
// contiguous slice of the pixel rows; its fp32 partial goes to scratch [split][tap*co][ci] and
// wgrad_reduce sums the splits in a fixed order (deterministic, no atomics).
//
// Narrow inputs (kin <= 128: the first layer, 48 planes -> kin = 64) would leave most of the UMMA's N = 256 unused
// that way (an M = 128 UMMA costs the same ~128 cycles for any N <= 256), so they use the mirrored tiling
// ("taps in N"): dW[tap][co][ci] = sum_m G[m][co] * X[m + shift(tap)][ci] with the CTA's 128 rows = two co chunks of
// G (unshifted) and its N = up to 256/kin taps side by side, each X chunk loaded with its own tap shift.
#include <string.h>

#include "common.cuh"
#include "conv.cuh"
#include "tcgen.cuh"
#include "wgrad.cuh"

namespace ag {

namespace {

constexpr int kStages = 4;
constexpr int kChunkBytes = 64 * 128;  // [64 pixel rows][64 bf16]
constexpr int kThreads = 192;
constexpr int kTmemCols = 256;

struct __align__(8) WgradBarriers {
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t acc_full;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap tmap_g, const __grid_constant__ CUtensorMap tmap_x, const WgradArgs p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    const int xchunks = p.kin / 64;
    const int cochunks = p.ldout / 64;
    const int nchunks = p.taps * cochunks;
    const int tile = blockIdx.x % p.tiles, split = blockIdx.x / p.tiles;
    const bool taps_in_n = p.tpg > 1;
    const int cotiles = (cochunks + 1) / 2;
    const int grp = taps_in_n ? tile / cotiles : 0, cpair = taps_in_n ? tile - grp * cotiles : 0;
    const int nb = taps_in_n ? min(p.tpg, p.taps - grp * p.tpg) : 1;          // taps side by side in this CTA's B operand
    const bool with_bias = p.bias_scratch != nullptr && (!taps_in_n || grp == p.groups - 1);  // the last group has room for the ones
    const int b_chunks = nb * xchunks;
    const int tx_bytes = (2 + b_chunks) * kChunkBytes;                     // what TMA brings per stage
    // + a chunk of ones behind the X chunks: one more B column that is 1 for every pixel row
    const int stage_bytes = (2 + (taps_in_n ? p.tpg * xchunks : xchunks + (p.bias_scratch ? 1 : 0))) * kChunkBytes;
    const int acc_cols = b_chunks * 64;
    const int umma_n = acc_cols + (with_bias ? 16 : 0);
    WgradBarriers* bars = reinterpret_cast<WgradBarriers*>(base + kStages * stage_bytes);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;  // warp-uniform by construction

    const int kb_total = (p.m_total + 63) / 64;
    const int kb0 = (int)((long long)kb_total * split / p.splits), kb1 = (int)((long long)kb_total * (split + 1) / p.splits);

    pdl_trigger();
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { tc::mbar_init(&bars->full[s], 1); tc::mbar_init(&bars->empty[s], 1); }
        tc::mbar_init(&bars->acc_full, 1);
        tc::fence_mbar_init();
        tc::tma_prefetch_desc(&tmap_g);
        tc::tma_prefetch_desc(&tmap_x);
    }
    if (warp == 1) tc::tmem_alloc(&bars->tmem_base, kTmemCols);
    if (with_bias) {  // the ones chunk of every stage (all elements equal, so the swizzle does not matter)
        for (int s = 0; s < kStages; s++) {
            uint32_t* ones = reinterpret_cast<uint32_t*>(base + s * stage_bytes + tx_bytes);
            for (int i = threadIdx.x; i < kChunkBytes / 4; i += kThreads) ones[i] = 0x3F803F80u;  // bf16 1.0 x2
        }
        tc::fence_proxy_async();  // generic-proxy writes, read by the tensor core
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = bars->tmem_base;
    pdl_wait();  // everything above overlapped the previous kernel's tail; G, X and the scratch are touched only from here on

    if (warp == 0) {
        // producer: warp-wide loop, one elected lane issues the TMA (uniform control flow, see conv.cu)
        int shift[2], col[2];
        bool live[2];
        for (int h = 0; h < 2; h++) {  // the two halves of this CTA's 128 dW rows: (tap, co-chunk), or co-chunk alone
            if (taps_in_n) {
                int c = cpair * 2 + h;
                live[h] = c < cochunks;
                shift[h] = 0;
                col[h] = live[h] ? c * 64 : 0;
            } else {
                int c = tile * 2 + h;
                live[h] = c < nchunks;
                int t = live[h] ? c / cochunks : 0;
                shift[h] = p.off[t];
                col[h] = live[h] ? (c - t * cochunks) * 64 : 0;
            }
        }
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = kb0; kb < kb1; kb++) {
            tc::mbar_wait(&bars->empty[stage], phase ^ 1);
            if (tc::elect_one()) {
                uint8_t* s = base + stage * stage_bytes;
                tc::mbar_expect_tx(&bars->full[stage], (uint32_t)tx_bytes);
                const int m0 = kb * 64;
                for (int h = 0; h < 2; h++)  // a dead half reads far out of bounds -> zeros
                    tc::tma_load_2d(s + h * kChunkBytes, &tmap_g, &bars->full[stage], col[h], live[h] ? m0 - shift[h] : 0x40000000);
                for (int j = 0; j < nb; j++) {
                    const int xrow = taps_in_n ? m0 + p.off[grp * p.tpg + j] : m0;  // rows outside the tensor read as zeros
                    for (int c = 0; c < xchunks; c++)
                        tc::tma_load_2d(s + (2 + j * xchunks + c) * kChunkBytes, &tmap_x, &bars->full[stage], c * 64, xrow);
                }
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        const uint32_t idesc = tc::idesc_bf16(128, umma_n, 1, 1);
        int stage = 0;
        uint32_t phase = 0;
        bool ready = false;
        for (int kb = kb0; kb < kb1; kb++) {
            if (!ready) tc::mbar_wait(&bars->full[stage], phase);
            tc::fence_after_sync();
            if (tc::elect_one()) {
                const uint32_t s = tc::smem_u32(base + stage * stage_bytes);
                const uint64_t adesc = tc::desc_mnmajor_sw128(s, kChunkBytes);
                const uint64_t bdesc = tc::desc_mnmajor_sw128(s + 2 * kChunkBytes, kChunkBytes);
#pragma unroll
                for (int k = 0; k < 4; k++)  // 16 pixel rows per UMMA: +2048 B
                    tc::umma_bf16(tmem_base, adesc + (2048 >> 4) * k, bdesc + (2048 >> 4) * k, idesc, (kb > kb0) || (k > 0));
                tc::umma_commit(&bars->empty[stage]);
                if (kb == kb1 - 1) tc::umma_commit(&bars->acc_full);
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
            ready = tc::mbar_try_wait(&bars->full[stage], phase);  // peek at the next slice while these UMMAs run
        }
    } else {
        const int q = warp & 3;
        const int r = q * 32 + lane;           // row of the 128-row tile
        const int chunk = (taps_in_n ? cpair : tile) * 2 + (r >> 6);
        const bool valid = chunk < (taps_in_n ? cochunks : nchunks);
        // scratch row = tap*ldout + co  (shifted-G tiling: chunk*64 + r%64 by construction; taps in N: the tap comes from the column)
        float* dst = p.scratch + ((size_t)split * p.taps * p.ldout + (size_t)chunk * 64 + (r & 63)) * p.kin;
        if (kb1 > kb0) {
            tc::mbar_wait(&bars->acc_full, 0);
            tc::fence_after_sync();
        }
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        for (int c0 = 0; c0 < acc_cols; c0 += 32) {
            uint32_t v[32];
            if (kb1 > kb0) {
                tc::tmem_ld32(taddr + c0, v);
                tc::tmem_ld_wait();
            } else {
#pragma unroll
                for (int j = 0; j < 32; j++) v[j] = 0;
            }
            if (valid) {
                float* d = dst + c0;
                if (taps_in_n) {  // column c0 = (tap within the group) * kin + ci
                    const int j = c0 / p.kin;
                    d = dst + (size_t)(grp * p.tpg + j) * p.ldout * p.kin + (c0 - j * p.kin);
                }
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<uint4*>(d + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
        }
        if (with_bias) {
            // the column behind the last X chunk is sum_m G[m][co]: the bias gradient.  Shifted-G tiling: taken from the centre
            // tap's rows (no shift); taps in N: G is never shifted.  (warp-uniform branch: a warp's 32 rows share a chunk)
            const int t = taps_in_n ? 0 : (valid ? chunk / cochunks : -1);
            if (valid && (taps_in_n || t == p.taps / 2)) {
                uint32_t v[32];
                if (kb1 > kb0) { tc::tmem_ld32(taddr + acc_cols, v); tc::tmem_ld_wait(); } else v[0] = 0;
                p.bias_scratch[(size_t)split * p.ldout + (chunk - t * cochunks) * 64 + (r & 63)] = __uint_as_float(v[0]);
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_base, kTmemCols);
    }
}

// dW[i] = sum_s scratch[s][i], fixed order
__global__ void wgrad_reduce_kernel(const float* __restrict__ scratch, float* __restrict__ dw, size_t count4, int splits,
                                    const float* __restrict__ bias_scratch, float* __restrict__ db, int ldout) {
    pdl_trigger();
    pdl_wait();
    if (db && blockIdx.x == 0) {
        for (int c = threadIdx.x; c < ldout; c += blockDim.x) {
            float t = 0.f;
            for (int s = 0; s < splits; s++) t += bias_scratch[(size_t)s * ldout + c];
            db[c] = t;
        }
    }
    const float4* s4 = reinterpret_cast<const float4*>(scratch);
    float4* d4 = reinterpret_cast<float4*>(dw);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count4; i += (size_t)gridDim.x * blockDim.x) {
        float4 a = s4[i];
        for (int s = 1; s < splits; s++) {
            float4 b = s4[(size_t)s * count4 + i];
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        d4[i] = a;
    }
}

}  // namespace

// Tiling choice: taps per B operand (1 = shifted-G tiling) and the CTA tiles of one split.  Taps in N needs a spare
// 16 columns behind the last group's taps for the ones column, which every odd tap count (1x1, 3x3, 5x5) leaves.
static int wgrad_tiling(int taps, int ldout, int kin, int* tpg, int* groups) {
    int t = 256 / kin;
    if (t >= 2 && taps >= 2 && taps % t != 0) {
        *tpg = t;
        *groups = (taps + t - 1) / t;
        return ((ldout / 64 + 1) / 2) * *groups;
    }
    *tpg = 1;
    *groups = taps;
    return (taps * (ldout / 64) + 1) / 2;
}

size_t wgrad_smem_bytes(int taps, int kin, bool with_bias) {
    int tpg, groups;
    wgrad_tiling(taps, 64, kin, &tpg, &groups);
    int chunks = 2 + (tpg > 1 ? tpg * (kin / 64) : kin / 64 + (with_bias ? 1 : 0));
    return 1024 + (size_t)kStages * chunks * kChunkBytes + sizeof(WgradBarriers) + 64;
}

int wgrad_plan_splits(int taps, int ldout, int kin, int m_total) {
    // CTAs = tiles * splits should fill whole waves of the SMs: one wave if that keeps >= 85 % of them busy,
    // else two.  Fewer, longer K slices also mean fewer fp32 partials to write and reduce.
    int tpg, groups;
    int tiles = wgrad_tiling(taps, ldout, kin, &tpg, &groups);
    int kb = (m_total + 63) / 64;
    int sms = sm_count();
    int s1 = sms / tiles, s2 = 2 * sms / tiles;
    int s = s1;
    if (s1 < 1 || ((double)tiles * s1 / sms < 0.85 && (double)tiles * s2 / (2.0 * sms) > (double)tiles * s1 / sms + 0.05)) s = s2;
    if (s > kb) s = kb;
    if (s < 1) s = 1;
    if (s > kWgradMaxSplits) s = kWgradMaxSplits;
    return s;
}

int wgrad_launch(const CUtensorMap& tmap_g, const CUtensorMap& tmap_x, WgradArgs args, float* dw, float* db, cudaStream_t stream) {
    if (args.kin % 64 != 0 || args.kin < 64 || args.kin > 256 || args.ldout % 64 != 0) {
        set_error("wgrad: channel counts must be padded to 64 (kin=%d <= 256, ldout=%d)", args.kin, args.ldout);
        return AG_EINVAL;
    }
    args.tiles = wgrad_tiling(args.taps, args.ldout, args.kin, &args.tpg, &args.groups);
    if (args.splits < 1) args.splits = 1;
    static size_t configured = 0;
    if (!db || args.kin + 16 > 256) args.bias_scratch = nullptr;
    if (db && !args.bias_scratch) { set_error("wgrad: bias gradient needs bias_scratch and kin <= 240"); return AG_EINVAL; }
    size_t smem = wgrad_smem_bytes(args.taps, args.kin, args.bias_scratch != nullptr);
    if (smem > 227 * 1024) { set_error("wgrad: %zu bytes of shared memory needed", smem); return AG_EINVAL; }
    if (smem > configured) {
        AG_CUDA(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    launch_k(wgrad_kernel, dim3(args.tiles * args.splits), dim3(kThreads), smem, stream, tmap_g, tmap_x, args);
    AG_LAUNCH_CHECK("wgrad_kernel");
    size_t count4 = (size_t)args.taps * args.ldout * args.kin / 4;
    int grid = (int)((count4 + 255) / 256);
    if (grid > sm_count() * 8) grid = sm_count() * 8;
    launch_k(wgrad_reduce_kernel, dim3(grid), dim3(256), 0, stream, args.scratch, dw, count4, args.splits, args.bias_scratch, args.bias_scratch ? db : nullptr, args.ldout);
    AG_LAUNCH_CHECK("wgrad_reduce_kernel");
    return AG_OK;
}

}  // namespace ag

// ------------------------------------------------------------------------------------ standalone op
namespace ag {
static int wgrad_offsets(int width, int row_stride, int* off) {
    if (width < 1 || width > 5 || width % 2 == 0) { set_error("wgrad: filter width %d must be odd and <= 5", width); return AG_EINVAL; }
    int pad = width / 2, t = 0;
    for (int kx = 0; kx < width; kx++)
        for (int ky = 0; ky < width; ky++, t++) off[t] = (kx - pad) * row_stride + (ky - pad);
    return AG_OK;
}
static int wgrad_check(long long rows, int ldout, int kin, int row_stride) {
    if (rows < 1 || rows > 0x7fffffffLL - 4096 || row_stride < 1) { set_error("wgrad: bad rows/row_stride"); return AG_EINVAL; }
    if (ldout < 64 || ldout % 64 || kin < 64 || kin % 64 || kin > 256) { set_error("wgrad: channels must be padded to 64 (ldout=%d, kin=%d <= 256)", ldout, kin); return AG_EINVAL; }
    return AG_OK;
}
}  // namespace ag

extern "C" {

// bytes of scratch ag_wgrad wants for this problem (the split-K partials)
size_t ag_wgrad_scratch_bytes(int width, int ldout, int kin, long long rows) {
    if (ag::wgrad_check(rows, ldout, kin, 1) || width < 1 || width > 5) return 0;
    return (size_t)ag::wgrad_plan_splits(width * width, ldout, kin, (int)rows) * width * width * ldout * kin * sizeof(float);
}

int ag_wgrad(const void* d_g, const void* d_x, long long rows, int ldout, int kin, int width, int row_stride, float* d_scratch,
             size_t scratch_bytes, float* d_dw, void* stream) {
    int rc = ag::wgrad_check(rows, ldout, kin, row_stride);
    if (rc) return rc;
    ag::WgradArgs a;
    memset(&a, 0, sizeof(a));
    rc = ag::wgrad_offsets(width, row_stride, a.off);
    if (rc) return rc;
    a.m_total = (int)rows; a.taps = width * width; a.ldout = ldout; a.kin = kin;
    const size_t partial = (size_t)a.taps * ldout * kin * sizeof(float);
    a.splits = ag::wgrad_plan_splits(a.taps, ldout, kin, (int)rows);
    if ((size_t)a.splits * partial > scratch_bytes) a.splits = (int)(scratch_bytes / partial);
    if (a.splits < 1 || !d_scratch) { ag::set_error("ag_wgrad: scratch of %zu bytes is smaller than one partial (%zu bytes)", scratch_bytes, partial); return AG_EINVAL; }
    a.scratch = d_scratch;
    CUtensorMap tg, tx;
    rc = ag::make_tmap_2d(&tg, d_g, (uint64_t)rows, ldout, 64);
    if (!rc) rc = ag::make_tmap_2d(&tx, d_x, (uint64_t)rows, kin, 64);
    if (rc) return rc;
    return ag::wgrad_launch(tg, tx, a, d_dw, nullptr, (cudaStream_t)stream);
}

}  // extern "C"

// ------------------------------------------------------------------------------------ no-allocation variant
// Same computation as ag_wgrad, for callers that are not allowed to allocate device memory: code running inside
// another framework's memory plan (a pool sized ahead of time that must not be disturbed), inside a region where
// the allocator would synchronise the device or fail outright (a latency-critical serving loop), or in a process
// whose memory budget was fixed at start-up.  The split-K partials live in an array that is part of this library's
// image: every GPU that loads libalphago_b200.so sets aside AG_WGRAD_STATIC_SCRATCH_BYTES for it until the process
// exits, whether or not this entry point is ever called.
#define AG_WGRAD_STATIC_SCRATCH_BYTES (16u << 20)
static __device__ __align__(256) float g_wgrad_static_scratch[AG_WGRAD_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace ag {
// In host code `g_wgrad_static_scratch` is only a handle for the symbol, never a device address; the address a
// kernel can use has to be asked for, per device, with that device current.
static int wgrad_static_scratch_ptr(float** out) {
    constexpr int kMaxDevices = 64;
    static float* cached[kMaxDevices] = {nullptr};
    int dev = 0;
    AG_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { set_error("ag_wgrad_static: device ordinal %d not supported", dev); return AG_EINVAL; }
    if (!cached[dev]) {
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_wgrad_static_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_wgrad_static_scratch)");
        if (!p) { set_error("ag_wgrad_static: no device address for the built-in scratch"); return AG_ECUDA; }
        cached[dev] = p;
    }
    *out = cached[dev];
    return AG_OK;
}
}  // namespace ag

extern "C" {

size_t ag_wgrad_static_max_bytes(void) { return AG_WGRAD_STATIC_SCRATCH_BYTES; }

int ag_wgrad_static(const void* d_g, const void* d_x, long long rows, int ldout, int kin, int width, int row_stride, float* d_dw,
                    void* stream) {
    int rc = ag::wgrad_check(rows, ldout, kin, row_stride);
    if (rc) return rc;
    ag::WgradArgs a;
    memset(&a, 0, sizeof(a));
    rc = ag::wgrad_offsets(width, row_stride, a.off);
    if (rc) return rc;
    a.m_total = (int)rows; a.taps = width * width; a.ldout = ldout; a.kin = kin;
    const size_t partial = (size_t)a.taps * ldout * kin * sizeof(float);
    if (partial > AG_WGRAD_STATIC_SCRATCH_BYTES) {
        ag::set_error("ag_wgrad_static: one %dx%d x %d x %d gradient (%zu bytes) exceeds the built-in scratch of %u bytes; "
                      "use ag_wgrad with your own scratch", width, width, ldout, kin, partial, AG_WGRAD_STATIC_SCRATCH_BYTES);
        return AG_EINVAL;
    }
    a.splits = ag::wgrad_plan_splits(a.taps, ldout, kin, (int)rows);
    if ((size_t)a.splits * partial > AG_WGRAD_STATIC_SCRATCH_BYTES) a.splits = (int)(AG_WGRAD_STATIC_SCRATCH_BYTES / partial);
    rc = ag::wgrad_static_scratch_ptr(&a.scratch);
    if (rc) return rc;
    CUtensorMap tg, tx;
    rc = ag::make_tmap_2d(&tg, d_g, (uint64_t)rows, ldout, 64);
    if (!rc) rc = ag::make_tmap_2d(&tx, d_x, (uint64_t)rows, kin, 64);
    if (rc) return rc;
    return ag::wgrad_launch(tg, tx, a, d_dw, nullptr, (cudaStream_t)stream);
}

}  // extern "C"

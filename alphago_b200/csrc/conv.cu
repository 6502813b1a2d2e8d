// Implicit-GEMM convolution for the policy/value conv stack (AlphaGo/models/policy.py:119-150,
// AlphaGo/models/value.py:14-25; in the reference these are Keras Convolution2D layers executed by
// Theano).  One persistent, warp-specialised tcgen05 kernel serves every KxK 'same' layer, forward
// and data-gradient:
//
//     D[m, co] = sum_taps sum_ci A[m + shift(tap), ci] * W[tap][co][ci]
//
// A is the padded-pixel-row activation matrix (conv.cuh), so each tap is a row-shifted TMA box of
// the same 2D tensor; W[tap] is a K-major [co][ci] matrix.  Per CTA: 128 output rows x N channels,
// accumulated in TMEM (two accumulator buffers so the epilogue of tile i overlaps the MMAs of
// tile i+1), K streamed in 64-channel slices through a 4-stage TMA->smem ring.
//   warp 0    TMA producer (one elected lane)
//   warp 1    TMEM allocation + UMMA issue (one elected lane)
//   warps 2-5 epilogue: tcgen05.ld -> bias / ReLU (+bitmask) / ReLU-gate (bitmask) -> bf16 -> global (pad rows skipped)
#include <stdlib.h>

#include "common.cuh"
#include "conv.cuh"
#include "tcgen.cuh"

namespace ag {

namespace {

constexpr int kTileM = 128;
constexpr int kABytes = kTileM * 128;  // 128 rows x 64 bf16
constexpr int kThreads = 192;
constexpr int kAccStride = 256;  // TMEM columns between the two accumulator buffers
constexpr int kTmemCols = 512;

constexpr int kMaxStages = 6;
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {  // read-only 16-byte load (L1-cached: a row's four loads share sectors)
    uint4 r;
    asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

struct __align__(8) ConvBarriers {
    uint64_t full[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint32_t tmem_base;
};

// Epilogue of one accumulator tile, one GEMM row per thread: tcgen05.ld -> bias / ReLU (+ its bitmask) / ReLU gate -> bf16
// -> global.  Rows that decode to pad slots produce no output.  The gate is the forward pass's ReLU bitmask of the
// layer below (32 bytes per row instead of the bf16 activations' 2*ld), fetched before waiting for the accumulator.
__device__ __forceinline__ void epilogue_tile(const ConvArgs& p, const float* s_bias, const int m, const uint32_t taddr,
                                              uint64_t* acc_full, const uint32_t parity) {
    const int img = m / p.in.img_stride;
    const int rem = m - img * p.in.img_stride;
    const int x = rem / p.in.row_stride, y = rem - x * p.in.row_stride;
    const bool valid = m < p.m_total && x < p.in.size && y < p.in.size && !(p.debug & 4);
    const size_t orow = (size_t)img * p.dst.img_stride + (size_t)x * p.dst.row_stride + y;
    __nv_bfloat16* dst = p.out + orow * p.out_ld;
    uint32_t* mrow = p.mask_out ? p.mask_out + orow * kConvMaskWords : nullptr;
    uint32_t mk[kConvMaskWords];
#pragma unroll
    for (int i = 0; i < kConvMaskWords; i++) mk[i] = 0xffffffffu;
    if (p.gate_mask && valid) {
        const uint32_t* g = p.gate_mask + ((size_t)img * p.gate_layout.img_stride + (size_t)x * p.gate_layout.row_stride + y) * kConvMaskWords;
        const uint4 lo = ld_nc_v4(g);
        mk[0] = lo.x; mk[1] = lo.y; mk[2] = lo.z; mk[3] = lo.w;
        if (p.n > 128) {
            const uint4 hi = ld_nc_v4(g + 4);
            mk[4] = hi.x; mk[5] = hi.y; mk[6] = hi.z; mk[7] = hi.w;
        }
    }
    tc::mbar_wait(acc_full, parity);
    tc::fence_after_sync();
    for (int c0 = 0; c0 < p.n; c0 += 32) {
        uint32_t r[32];
        tc::tmem_ld32(taddr + c0, r);
        const uint32_t mcur = mk[0];  // this chunk's gate word; the rest shift down (keeps the array in registers)
#pragma unroll
        for (int i = 0; i + 1 < kConvMaskWords; i++) mk[i] = mk[i + 1];
        tc::tmem_ld_wait();
        if (valid) {
            uint32_t mbits = 0;
#pragma unroll
            for (int g = 0; g < 4; g++) {  // 8 channels -> one 16-byte store
                if (c0 + g * 8 < p.n) {
                    float v[8];
#pragma unroll
                    for (int j = 0; j < 8; j++) v[j] = __uint_as_float(r[g * 8 + j]) + s_bias[c0 + g * 8 + j];
                    if (p.relu) {
#pragma unroll
                        for (int j = 0; j < 8; j++) v[j] = fmaxf(v[j], 0.f);
                    }
#pragma unroll
                    for (int j = 0; j < 8; j++) v[j] = ((mcur >> (g * 8 + j)) & 1u) ? v[j] : 0.f;
                    uint4 o;
                    __nv_bfloat162* ob = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
                    for (int j = 0; j < 4; j++) ob[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                    *reinterpret_cast<uint4*>(dst + c0 + g * 8) = o;
                    // the mask records what was STORED (bf16 > 0), so the backward gate agrees with the activations bit for bit
                    if (mrow) {
                        const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const uint32_t w = ow[j] & 0x7fff7fffu;
                            mbits |= ((w & 0xffffu) ? 1u : 0u) << (g * 8 + 2 * j);
                            mbits |= ((w >> 16) ? 1u : 0u) << (g * 8 + 2 * j + 1);
                        }
                    }
                }
            }
            if (mrow) mrow[c0 >> 5] = mbits;
        }
    }
}

__global__ void __launch_bounds__(kThreads, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const ConvArgs p, const int kStages) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // carve: stages x (A | B), then barriers, then bias
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    const int b_bytes = p.n * 128;
    const int stage_bytes = kABytes + b_bytes;
    ConvBarriers* bars = reinterpret_cast<ConvBarriers*>(base + kStages * stage_bytes);
    float* s_bias = reinterpret_cast<float*>(bars + 1);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;  // warp-uniform by construction
    const int n_tiles = (p.m_total + kTileM - 1) / kTileM;
    const int kblocks = p.taps * p.kchunks;

    pdl_trigger();
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { tc::mbar_init(&bars->full[s], 1); tc::mbar_init(&bars->empty[s], 1); }
        for (int b = 0; b < 2; b++) { tc::mbar_init(&bars->acc_full[b], 1); tc::mbar_init(&bars->acc_empty[b], 128); }
        tc::fence_mbar_init();
        tc::tma_prefetch_desc(&tmap_a);
        tc::tma_prefetch_desc(&tmap_b);
    }
    if (warp == 1) tc::tmem_alloc(&bars->tmem_base, kTmemCols);
    pdl_wait();  // set-up above overlaps the previous kernel; global memory (bias included) only from here on
    for (int i = threadIdx.x; i < p.n; i += kThreads) s_bias[i] = p.bias ? p.bias[i] : 0.f;
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ===== TMA producer: the whole warp walks the loop (uniform control flow), one elected lane issues
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int m0 = tile * kTileM;
            for (int t = 0; t < p.taps; t++) {
                const int arow = m0 + p.off[t];
                for (int kc = 0; kc < p.kchunks; kc++) {
                    tc::mbar_wait(&bars->empty[stage], phase ^ 1);
                    if (tc::elect_one()) {
                        uint8_t* sa = base + stage * stage_bytes;
                        const uint32_t tx = ((p.debug & 1) ? 0 : kABytes) + ((p.debug & 2) ? 0 : b_bytes);
                        if (tx) tc::mbar_expect_tx(&bars->full[stage], tx); else tc::mbar_arrive(&bars->full[stage]);
                        if (!(p.debug & 1)) tc::tma_load_2d(sa, &tmap_a, &bars->full[stage], kc * 64, arow);
                        if (!(p.debug & 2)) tc::tma_load_2d(sa + kABytes, &tmap_b, &bars->full[stage], kc * 64, t * p.b_tap_rows);
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== UMMA issuer: warp-wide loop, the UMMAs and commits come from one elected lane (keeps every operand
        // in uniform registers: a divergent single-lane loop makes the compiler wrap each UMMA in a uniformity loop)
        const uint32_t idesc = tc::idesc_bf16(kTileM, p.n, 0, 0);
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        bool ready = false;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
            const int buf = it & 1;
            tc::mbar_wait(&bars->acc_empty[buf], ((it >> 1) & 1) ^ 1);
            tc::fence_after_sync();
            const uint32_t d_tmem = tmem_base + buf * kAccStride;
            for (int kb = 0; kb < kblocks; kb++) {
                if (!ready) tc::mbar_wait(&bars->full[stage], phase);
                tc::fence_after_sync();
                if (tc::elect_one()) {
                    const uint32_t sa = tc::smem_u32(base + stage * stage_bytes);
                    const uint64_t adesc = tc::desc_kmajor_sw128(sa);
                    const uint64_t bdesc = tc::desc_kmajor_sw128(sa + kABytes);
#pragma unroll
                    for (int k = 0; k < 4; k++)  // 4 x K=16 inside the 64-channel slice: +32 B each
                        tc::umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    tc::umma_commit(&bars->empty[stage]);  // frees the smem slot when these MMAs retire
                    if (kb == kblocks - 1) tc::umma_commit(&bars->acc_full[buf]);
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
                ready = tc::mbar_try_wait(&bars->full[stage], phase);  // peek at the next slice while these UMMAs run
            }
        }
    } else {
        // ===== epilogue: warp w may only touch TMEM lanes [32*(w%4), +32)
        const int q = warp & 3;
        int it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
            const int buf = it & 1;
            epilogue_tile(p, s_bias, tile * kTileM + q * 32 + lane, tmem_base + ((uint32_t)(q * 32) << 16) + buf * kAccStride,
                          &bars->acc_full[buf], (it >> 1) & 1);
            tc::fence_before_sync();
            tc::mbar_arrive(&bars->acc_empty[buf]);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---- CTA-pair variant: one UMMA of M = 256 per cluster of two CTAs (cta_group::2).  Each CTA stages its own 128 pixel
// rows of A and HALF of the weight slice (N/2 rows); the tensor cores of both SMs read the two halves of B from each
// other's shared memory, so per output row the B traffic through shared memory halves.  The leader CTA owns the
// smem-full and accumulator-empty barriers and issues the UMMAs; commits are multicast to both CTAs.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_bh, const ConvArgs p,
                      const int kStages) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    const int half_n = p.n / 2;
    const int b_bytes = half_n * 128;
    const int stage_bytes = kABytes + b_bytes;
    ConvBarriers* bars = reinterpret_cast<ConvBarriers*>(base + kStages * stage_bytes);
    float* s_bias = reinterpret_cast<float*>(bars + 1);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int n_pairs = (p.m_total + 2 * kTileM - 1) / (2 * kTileM);
    const int n_clusters = gridDim.x / 2, cluster_id = blockIdx.x / 2;
    const int kblocks = p.taps * p.kchunks;

    pdl_trigger();
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { tc::mbar_init(&bars->full[s], 1); tc::mbar_init(&bars->empty[s], 1); }
        for (int b = 0; b < 2; b++) { tc::mbar_init(&bars->acc_full[b], 1); tc::mbar_init(&bars->acc_empty[b], 256); }
        tc::fence_mbar_init();
        tc::tma_prefetch_desc(&tmap_a);
        tc::tma_prefetch_desc(&tmap_bh);
    }
    if (warp == 1) tc::tmem_alloc_pair(&bars->tmem_base, kTmemCols);
    pdl_wait();
    for (int i = threadIdx.x; i < p.n; i += kThreads) s_bias[i] = p.bias ? p.bias[i] : 0.f;
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync_all();  // both CTAs' barriers exist before anyone signals across the pair
    tc::fence_after_sync();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        int stage = 0;
        uint32_t phase = 0;
        for (int pair = cluster_id; pair < n_pairs; pair += n_clusters) {
            const int m0 = pair * 2 * kTileM + (int)rank * kTileM;
            for (int t = 0; t < p.taps; t++) {
                const int arow = m0 + p.off[t];
                for (int kc = 0; kc < p.kchunks; kc++) {
                    tc::mbar_wait(&bars->empty[stage], phase ^ 1);
                    if (tc::elect_one()) {
                        uint8_t* sa = base + stage * stage_bytes;
                        const uint32_t tx = ((p.debug & 1) ? 0 : kABytes) + ((p.debug & 2) ? 0 : b_bytes);
                        if (leader) { if (tx) tc::mbar_expect_tx(&bars->full[stage], 2 * tx); else tc::mbar_arrive(&bars->full[stage]); }
                        if (!(p.debug & 1)) tc::tma_load_2d_pair(sa, &tmap_a, &bars->full[stage], kc * 64, arow);
                        if (!(p.debug & 2)) tc::tma_load_2d_pair(sa + kABytes, &tmap_bh, &bars->full[stage], kc * 64, t * p.b_tap_rows + (int)rank * half_n);
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (leader) {
            const uint32_t idesc = tc::idesc_bf16(2 * kTileM, p.n, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            bool ready = false;
            for (int pair = cluster_id; pair < n_pairs; pair += n_clusters, it++) {
                const int buf = it & 1;
                tc::mbar_wait(&bars->acc_empty[buf], ((it >> 1) & 1) ^ 1);
                tc::fence_after_sync();
                const uint32_t d_tmem = tmem_base + buf * kAccStride;
                for (int kb = 0; kb < kblocks; kb++) {
                    if (!ready) tc::mbar_wait(&bars->full[stage], phase);
                    tc::fence_after_sync();
                    if (tc::elect_one()) {
                        const uint32_t sa = tc::smem_u32(base + stage * stage_bytes);
                        const uint64_t adesc = tc::desc_kmajor_sw128(sa);
                        const uint64_t bdesc = tc::desc_kmajor_sw128(sa + kABytes);
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            tc::umma_bf16_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                        tc::umma_commit_pair(&bars->empty[stage]);
                        if (kb == kblocks - 1) tc::umma_commit_pair(&bars->acc_full[buf]);
                    }
                    __syncwarp();
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                    ready = tc::mbar_try_wait(&bars->full[stage], phase);
                }
            }
        }
    } else {
        const int q = warp & 3;
        const uint32_t empty_remote0 = tc::map_to_cta(&bars->acc_empty[0], 0), empty_remote1 = tc::map_to_cta(&bars->acc_empty[1], 0);
        int it = 0;
        for (int pair = cluster_id; pair < n_pairs; pair += n_clusters, it++) {
            const int buf = it & 1;
            epilogue_tile(p, s_bias, pair * 2 * kTileM + (int)rank * kTileM + q * 32 + lane,
                          tmem_base + ((uint32_t)(q * 32) << 16) + buf * kAccStride, &bars->acc_full[buf], (it >> 1) & 1);
            tc::fence_before_sync();
            tc::mbar_arrive_cluster(buf ? empty_remote1 : empty_remote0);  // the leader's accumulator-empty barrier
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync_all();  // the peer's smem and barriers stay alive until the leader's last UMMA / arrival is done
    if (warp == 1) {
        tc::fence_after_sync();
        tc::tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

// ---- CTA-pair variant with a shared activation window.  The pair kernel above is bound by shared-memory bandwidth, not by
// the tensor cores: per 64-channel slice a CTA's smem takes 28 KB of TMA writes and 28 KB of UMMA reads for 4 UMMAs whose
// tensor-core floor is 4 x 96 cycles (N = 192) -- 146 B/cycle against the SM's 128.  But the nine (or 25) taps of one
// channel slice read the SAME activation rows shifted by a few pixel rows, and the 128-byte swizzle is a function of the
// shared-memory address, so a K-major UMMA descriptor may start at any 128-byte row of a larger box.  Hence: per channel
// slice ONE box of 128 + 2*halo rows (the "window", halo = largest |tap shift|) is fetched, and every tap's UMMAs start at
// row halo + shift(tap) of it.  Activation traffic (L2 -> smem) drops by the number of taps; weights stream through their
// own ring.  Loop order: channel slice outer, tap inner.
constexpr int kWinSlots = 2;
constexpr int kMaxBStages = 12;
constexpr int kTapsPerStage = 3;   // weight slices (taps) per ring stage = UMMAs / 4 per barrier round trip

struct __align__(8) WinBarriers {
    uint64_t a_full[kWinSlots];
    uint64_t a_empty[kWinSlots];
    uint64_t b_full[kMaxBStages];
    uint64_t b_empty[kMaxBStages];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint32_t tmem_base;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_gemm_win_kernel(const __grid_constant__ CUtensorMap tmap_aw, const __grid_constant__ CUtensorMap tmap_bh, const ConvArgs p,
                     const int kBStages) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    const int half_n = p.n / 2;
    const int b_bytes = half_n * 128;
    const int win_box = p.win_rows * 128;                 // bytes one window box brings
    const int win_bytes = (win_box + 1023) & ~1023;       // its slot
    uint8_t* b_base = base + kWinSlots * win_bytes;
    WinBarriers* bars = reinterpret_cast<WinBarriers*>(b_base + kBStages * kTapsPerStage * b_bytes);
    float* s_bias = reinterpret_cast<float*>(bars + 1);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int n_pairs = (p.m_total + 2 * kTileM - 1) / (2 * kTileM);
    const int n_clusters = gridDim.x / 2, cluster_id = blockIdx.x / 2;

    pdl_trigger();
    if (threadIdx.x == 0) {
        for (int s = 0; s < kWinSlots; s++) { tc::mbar_init(&bars->a_full[s], 1); tc::mbar_init(&bars->a_empty[s], 1); }
        for (int s = 0; s < kBStages; s++) { tc::mbar_init(&bars->b_full[s], 1); tc::mbar_init(&bars->b_empty[s], 1); }
        for (int b = 0; b < 2; b++) { tc::mbar_init(&bars->acc_full[b], 1); tc::mbar_init(&bars->acc_empty[b], 256); }
        tc::fence_mbar_init();
        tc::tma_prefetch_desc(&tmap_aw);
        tc::tma_prefetch_desc(&tmap_bh);
    }
    if (warp == 1) tc::tmem_alloc_pair(&bars->tmem_base, kTmemCols);
    pdl_wait();
    for (int i = threadIdx.x; i < p.n; i += kThreads) s_bias[i] = p.bias ? p.bias[i] : 0.f;
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync_all();  // both CTAs' barriers exist before anyone signals across the pair
    tc::fence_after_sync();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ===== TMA producer (each CTA: its own 128 rows' window, its half of the weight rows; the leader's barriers count both)
        int ai = 0, bi = 0;
        uint32_t aph = 0, bph = 0;
        for (int pair = cluster_id; pair < n_pairs; pair += n_clusters) {
            const int m0 = pair * 2 * kTileM + (int)rank * kTileM;
            for (int kc = 0; kc < p.kchunks; kc++) {
                tc::mbar_wait(&bars->a_empty[ai], aph ^ 1);
                if (tc::elect_one()) {
                    if (p.debug & 1) { if (leader) tc::mbar_arrive(&bars->a_full[ai]); }
                    else {
                        if (leader) tc::mbar_expect_tx(&bars->a_full[ai], 2 * (uint32_t)win_box);
                        tc::tma_load_2d_pair(base + ai * win_bytes, &tmap_aw, &bars->a_full[ai], kc * 64, m0 - p.halo);
                    }
                }
                __syncwarp();
                if (++ai == kWinSlots) { ai = 0; aph ^= 1; }
                for (int t0 = 0; t0 < p.taps; t0 += kTapsPerStage) {
                    const int nt = min(kTapsPerStage, p.taps - t0);
                    tc::mbar_wait(&bars->b_empty[bi], bph ^ 1);
                    if (tc::elect_one()) {
                        if (p.debug & 2) { if (leader) tc::mbar_arrive(&bars->b_full[bi]); }
                        else {
                            if (leader) tc::mbar_expect_tx(&bars->b_full[bi], 2 * (uint32_t)(nt * b_bytes));
                            for (int j = 0; j < nt; j++)
                                tc::tma_load_2d_pair(b_base + (bi * kTapsPerStage + j) * b_bytes, &tmap_bh, &bars->b_full[bi], kc * 64,
                                                     (t0 + j) * p.b_tap_rows + (int)rank * half_n);
                        }
                    }
                    __syncwarp();
                    if (++bi == kBStages) { bi = 0; bph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (leader) {
            const uint32_t idesc = tc::idesc_bf16(2 * kTileM, p.n, 0, 0);
            int ai = 0, bi = 0, it = 0;
            uint32_t aph = 0, bph = 0;
            bool ready = false;
            for (int pair = cluster_id; pair < n_pairs; pair += n_clusters, it++) {
                const int buf = it & 1;
                tc::mbar_wait(&bars->acc_empty[buf], ((it >> 1) & 1) ^ 1);
                tc::fence_after_sync();
                const uint32_t d_tmem = tmem_base + buf * kAccStride;
                for (int kc = 0; kc < p.kchunks; kc++) {
                    tc::mbar_wait(&bars->a_full[ai], aph);
                    const uint64_t a_desc0 = tc::desc_kmajor_sw128(tc::smem_u32(base + ai * win_bytes)) + (uint64_t)(p.halo * 8);  // window row `halo`
                    for (int t0 = 0; t0 < p.taps; t0 += kTapsPerStage) {
                        const int nt = min(kTapsPerStage, p.taps - t0);
                        if (!ready) tc::mbar_wait(&bars->b_full[bi], bph);
                        tc::fence_after_sync();
                        if (tc::elect_one()) {
                            // one barrier round trip per kTapsPerStage taps: the issuing thread, not the tensor core, was the bottleneck
                            const uint64_t b_desc0 = tc::desc_kmajor_sw128(tc::smem_u32(b_base + bi * kTapsPerStage * b_bytes));
#pragma unroll
                            for (int j = 0; j < kTapsPerStage; j++) {
                                if (j < nt && !(p.debug & 16)) {
                                    const uint64_t adesc = a_desc0 + (int64_t)(p.off[t0 + j] * 8);   // a pixel row = 128 B = 8 descriptor units
                                    const uint64_t bdesc = b_desc0 + (uint64_t)(j * (b_bytes >> 4));
#pragma unroll
                                    for (int k = 0; k < 4; k++)  // 4 x K=16 inside the 64-channel slice: +32 B each
                                        tc::umma_bf16_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kc | t0 | j | k) != 0);
                                }
                            }
                            tc::umma_commit_pair(&bars->b_empty[bi]);
                            if (t0 + nt == p.taps) {
                                tc::umma_commit_pair(&bars->a_empty[ai]);
                                if (kc == p.kchunks - 1) tc::umma_commit_pair(&bars->acc_full[buf]);
                            }
                        }
                        __syncwarp();
                        if (++bi == kBStages) { bi = 0; bph ^= 1; }
                        ready = tc::mbar_try_wait(&bars->b_full[bi], bph);  // peek at the next weight slices while these UMMAs run
                    }
                    if (++ai == kWinSlots) { ai = 0; aph ^= 1; }
                }
            }
        }
    } else {
        const int q = warp & 3;
        const uint32_t empty_remote0 = tc::map_to_cta(&bars->acc_empty[0], 0), empty_remote1 = tc::map_to_cta(&bars->acc_empty[1], 0);
        int it = 0;
        for (int pair = cluster_id; pair < n_pairs; pair += n_clusters, it++) {
            const int buf = it & 1;
            epilogue_tile(p, s_bias, pair * 2 * kTileM + (int)rank * kTileM + q * 32 + lane,
                          tmem_base + ((uint32_t)(q * 32) << 16) + buf * kAccStride, &bars->acc_full[buf], (it >> 1) & 1);
            tc::fence_before_sync();
            tc::mbar_arrive_cluster(buf ? empty_remote1 : empty_remote0);  // the leader's accumulator-empty barrier
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync_all();  // the peer's smem and barriers stay alive until the leader's last UMMA / arrival is done
    if (warp == 1) {
        tc::fence_after_sync();
        tc::tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

// ---- A whole chain of layers in ONE persistent launch (the forward conv stack, or the data-gradient chain of backward).
// One kernel per layer leaves ~10 % of the machine idle: 400 tiles of 256 rows over 74 CTA pairs is 5.4 rounds, run as 6,
// and every launch pays its own prologue and drain.  Here the tiles of all layers form one sequence g = 0, 1, 2 ...
// (layer-major), pair c takes g = c, c + pairs, c + 2*pairs ..., and a tile starts as soon as the 128-row tiles of the layer
// before it that cover its rows (+ halo) are finished: every CTA publishes a tile with a release store of the launch's
// epoch to flags[layer][tile128], and the TMA producer acquires the flags it needs before fetching the window.  All CTAs are
// resident (grid <= SM count, 1 CTA/SM) and each walks its tiles in increasing g, while a tile only waits for tiles with
// smaller g: no cycle, so no deadlock.  The kernel body is the window kernel's; the layer table travels as a kernel
// parameter (constant bank: the tensor maps and per-layer arguments are read like any other kernel argument).
__device__ __forceinline__ void chain_wait_flag(const unsigned int* flag, unsigned int epoch) {
    unsigned int v;
    long long t0 = 0;
    for (;;) {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v == epoch) return;
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > 4000000000LL) {
            printf("alphago_b200: conv chain dependency wait timed out (block %d flag %p has %u, want %u)\n", blockIdx.x, (const void*)flag, v, epoch);
            __trap();
        }
        __nanosleep(64);
    }
}

// layer of global tile g (layers are few: a linear scan over the parameter table)
__device__ __forceinline__ int chain_layer_of(const ChainParams& P, int g, int l) {
    while (l + 1 < P.n_layers && P.layers[l + 1].tile0 <= g) l++;
    return l;
}

// first GEMM row of half-tile h of a layer, and the half-tile that holds GEMM row r (see ChainLayer::tiles_img)
__device__ __forceinline__ int chain_tile_row(int h, int tiles_img, int img_stride) {
    if (tiles_img == 0) return h * kTileM;
    const int img = h / tiles_img;
    return img * img_stride + (h - img * tiles_img) * kTileM;
}
__device__ __forceinline__ int chain_row_tile(int r, int tiles_img, int img_stride) {
    if (tiles_img == 0) return r / kTileM;
    const int img = r / img_stride;
    const int part = (r - img * img_stride) / kTileM;
    return img * tiles_img + (part < tiles_img ? part : tiles_img - 1);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_chain_kernel(const __grid_constant__ ChainParams P, unsigned int* __restrict__ flags, const unsigned int epoch, const int win_bytes,
                  const int b_bytes_max, const int kBStages) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    uint8_t* b_base = base + kWinSlots * win_bytes;
    const int stage_bytes = kTapsPerStage * b_bytes_max;
    WinBarriers* bars = reinterpret_cast<WinBarriers*>(b_base + kBStages * stage_bytes);
    float* s_bias = reinterpret_cast<float*>(bars + 1);  // [256]

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const bool leader = rank == 0;
    const int n_clusters = gridDim.x / 2, cluster_id = blockIdx.x / 2;
    const int total_tiles = P.total_tiles;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kWinSlots; s++) { tc::mbar_init(&bars->a_full[s], 1); tc::mbar_init(&bars->a_empty[s], 1); }
        for (int s = 0; s < kBStages; s++) { tc::mbar_init(&bars->b_full[s], 1); tc::mbar_init(&bars->b_empty[s], 1); }
        for (int b = 0; b < 2; b++) { tc::mbar_init(&bars->acc_full[b], 1); tc::mbar_init(&bars->acc_empty[b], 256); }
        tc::fence_mbar_init();
    }
    if (warp == 1) tc::tmem_alloc_pair(&bars->tmem_base, kTmemCols);
    pdl_wait();
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync_all();  // both CTAs' barriers exist before anyone signals across the pair
    tc::fence_after_sync();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ===== TMA producer
        int ai = 0, bi = 0, l = 0;
        uint32_t aph = 0, bph = 0;
        for (int g = cluster_id; g < total_tiles; g += n_clusters) {
            l = chain_layer_of(P, g, l);
            const ChainLayer& L = P.layers[l];
            const ConvArgs& p = L.a;
            const int half_n = p.n / 2, b_bytes = half_n * 128, win_box = p.win_rows * 128;
            const int m0 = chain_tile_row((g - L.tile0) * 2 + (int)rank, L.tiles_img, p.in.img_stride);
            if (L.dep) {
                // rows [m0 - halo, m0 + 128 + halo) of the input must be final: the 128-row tiles of the layer before that wrote them
                int lo = m0 - p.halo, hi = m0 + kTileM - 1 + p.halo;
                if (L.dep == 2) {  // rows -> images -> GEMM rows of the previous layer
                    lo = (lo < 0 ? 0 : lo / p.in.img_stride) * L.prev_in_img_stride;
                    hi = (hi / p.in.img_stride + 1) * L.prev_in_img_stride - 1;
                }
                if (lo < 0) lo = 0;
                if (hi > L.prev_rows - 1) hi = L.prev_rows - 1;
                const unsigned int* f = flags + P.layers[l - 1].flag0;
                const int j0 = chain_row_tile(lo, L.prev_tiles_img, L.prev_in_img_stride), j1 = chain_row_tile(hi, L.prev_tiles_img, L.prev_in_img_stride);
                for (int j = j0 + lane; j <= j1; j += 32) chain_wait_flag(f + j, epoch);
                __syncwarp();
                asm volatile("fence.proxy.async.global;" ::: "memory");  // the generic-proxy stores just acquired, before TMA reads them
            }
            for (int kc = 0; kc < p.kchunks; kc++) {
                tc::mbar_wait(&bars->a_empty[ai], aph ^ 1);
                if (tc::elect_one()) {
                    if (leader) tc::mbar_expect_tx(&bars->a_full[ai], 2 * (uint32_t)win_box);
                    tc::tma_load_2d_pair(base + ai * win_bytes, &L.tmap_aw, &bars->a_full[ai], kc * 64, m0 - p.halo);
                }
                __syncwarp();
                if (++ai == kWinSlots) { ai = 0; aph ^= 1; }
                for (int t0 = 0; t0 < p.taps; t0 += kTapsPerStage) {
                    const int nt = min(kTapsPerStage, p.taps - t0);
                    tc::mbar_wait(&bars->b_empty[bi], bph ^ 1);
                    if (tc::elect_one()) {
                        if (leader) tc::mbar_expect_tx(&bars->b_full[bi], 2 * (uint32_t)(nt * b_bytes));
                        for (int j = 0; j < nt; j++)
                            tc::tma_load_2d_pair(b_base + bi * stage_bytes + j * b_bytes, &L.tmap_bh, &bars->b_full[bi], kc * 64,
                                                 (t0 + j) * p.b_tap_rows + (int)rank * half_n);
                    }
                    __syncwarp();
                    if (++bi == kBStages) { bi = 0; bph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (leader) {
            int ai = 0, bi = 0, it = 0, l = 0;
            uint32_t aph = 0, bph = 0;
            bool ready = false;
            for (int g = cluster_id; g < total_tiles; g += n_clusters, it++) {
                l = chain_layer_of(P, g, l);
                const ConvArgs& p = P.layers[l].a;
                const uint32_t idesc = tc::idesc_bf16(2 * kTileM, p.n, 0, 0);
                const int b_bytes = (p.n / 2) * 128;
                const int buf = it & 1;
                tc::mbar_wait(&bars->acc_empty[buf], ((it >> 1) & 1) ^ 1);
                tc::fence_after_sync();
                const uint32_t d_tmem = tmem_base + buf * kAccStride;
                for (int kc = 0; kc < p.kchunks; kc++) {
                    tc::mbar_wait(&bars->a_full[ai], aph);
                    const uint64_t a_desc0 = tc::desc_kmajor_sw128(tc::smem_u32(base + ai * win_bytes)) + (uint64_t)(p.halo * 8);  // window row `halo`
                    const int ksteps = (kc == p.kchunks - 1 && p.ksteps_last > 0) ? p.ksteps_last : 4;
                    for (int t0 = 0; t0 < p.taps; t0 += kTapsPerStage) {
                        const int nt = min(kTapsPerStage, p.taps - t0);
                        if (!ready) tc::mbar_wait(&bars->b_full[bi], bph);
                        tc::fence_after_sync();
                        if (tc::elect_one()) {
                            const uint64_t b_desc0 = tc::desc_kmajor_sw128(tc::smem_u32(b_base + bi * stage_bytes));
#pragma unroll
                            for (int j = 0; j < kTapsPerStage; j++) {
                                if (j < nt) {
                                    const uint64_t adesc = a_desc0 + (int64_t)(p.off[t0 + j] * 8);   // a pixel row = 128 B = 8 descriptor units
                                    const uint64_t bdesc = b_desc0 + (uint64_t)(j * (b_bytes >> 4));
#pragma unroll
                                    for (int k = 0; k < 4; k++)
                                        if (k < ksteps) tc::umma_bf16_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kc | t0 | j | k) != 0);
                                }
                            }
                            tc::umma_commit_pair(&bars->b_empty[bi]);
                            if (t0 + nt == p.taps) {
                                tc::umma_commit_pair(&bars->a_empty[ai]);
                                if (kc == p.kchunks - 1) tc::umma_commit_pair(&bars->acc_full[buf]);
                            }
                        }
                        __syncwarp();
                        if (++bi == kBStages) { bi = 0; bph ^= 1; }
                        ready = tc::mbar_try_wait(&bars->b_full[bi], bph);
                    }
                    if (++ai == kWinSlots) { ai = 0; aph ^= 1; }
                }
            }
        }
    } else {
        // ===== epilogue warps (128 threads, named barrier 1)
        const int q = warp & 3, et = threadIdx.x - 64;
        const uint32_t empty_remote0 = tc::map_to_cta(&bars->acc_empty[0], 0), empty_remote1 = tc::map_to_cta(&bars->acc_empty[1], 0);
        int it = 0, l = 0, bias_l = -1;
        for (int g = cluster_id; g < total_tiles; g += n_clusters, it++) {
            l = chain_layer_of(P, g, l);
            const ChainLayer& L = P.layers[l];
            if (l != bias_l) {
                bias_l = l;
                asm volatile("bar.sync 1, 128;" ::: "memory");  // everyone is done with the previous layer's bias
                const float* bias = L.a.bias;
                for (int i = et; i < 256; i += 128) s_bias[i] = (bias && i < L.a.n) ? __ldg(bias + i) : 0.f;
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            const int buf = it & 1;
            const int tile128 = (g - L.tile0) * 2 + (int)rank;
            epilogue_tile(L.a, s_bias, chain_tile_row(tile128, L.tiles_img, L.a.in.img_stride) + q * 32 + lane,
                          tmem_base + ((uint32_t)(q * 32) << 16) + buf * kAccStride,
                          &bars->acc_full[buf], (it >> 1) & 1);
            tc::fence_before_sync();
            tc::mbar_arrive_cluster(buf ? empty_remote1 : empty_remote0);  // the leader's accumulator-empty barrier
            // publish the tile: every thread's stores, made visible device-wide, then one release store of the epoch
            // (the barrier orders the other 127 threads' stores before thread 0's release, which is cumulative; the consumer
            // adds the generic -> async proxy fence on its side before TMA reads the rows)
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flags + L.flag0 + tile128), "r"(epoch) : "memory");
            }
        }
    }
    // Only now may the kernels behind this one start: a dependent that launched while some of OUR pairs were still waiting for
    // an SM would fill the SMs with CTAs that wait for us, and the tiles those pairs own would never be computed (seen as
    // a dependency timeout with the trigger at the top of the kernel).
    pdl_trigger();
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync_all();  // the peer's smem and barriers stay alive until the leader's last UMMA / arrival is done
    if (warp == 1) {
        tc::fence_after_sync();
        tc::tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

}  // namespace

int make_tmap_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return AG_ECUDA; }
    if (cols % 64 != 0 || box_rows == 0 || box_rows > 256) { set_error("make_tmap_2d: cols %llu / box_rows %u unsupported", (unsigned long long)cols, box_rows); return AG_EINVAL; }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d) rows=%llu cols=%llu box=%u", (int)r, (unsigned long long)rows, (unsigned long long)cols, box_rows); return AG_ECUDA; }
    return AG_OK;
}

static size_t conv_smem_bytes(int n, int stages) { return 1024 + (size_t)stages * (kABytes + n * 128) + sizeof(ConvBarriers) + 256 * sizeof(float); }

// pipeline depth: as many stages as fit in 227 KB, at most kMaxStages (AG_CONV_STAGES overrides, for experiments)
static int conv_stages(int n) {
    static int forced = -1;
    if (forced < 0) { const char* e = getenv("AG_CONV_STAGES"); forced = e ? atoi(e) : 0; }
    int s = kMaxStages;
    while (s > 2 && conv_smem_bytes(n, s) > 227 * 1024) s--;
    if (forced >= 2 && forced <= s) s = forced;
    return s;
}

static size_t conv_pair_smem_bytes(int n, int stages) { return 1024 + (size_t)stages * (kABytes + (n / 2) * 128) + sizeof(ConvBarriers) + 256 * sizeof(float); }

static bool conv_pair_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("AG_CONV_PAIR"); on = e ? atoi(e) : 1; }
    return on != 0;
}

static bool conv_window_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("AG_CONV_WINDOW"); on = e ? atoi(e) : 1; }
    return on != 0;
}

int conv_window_rows(const int* off, int taps) {
    int halo = 0;
    for (int t = 0; t < taps; t++) { int a = off[t] < 0 ? -off[t] : off[t]; if (a > halo) halo = a; }
    int rows = kTileM + 2 * halo;
    return rows <= 256 ? rows : 0;
}

static size_t conv_win_smem_bytes(int n, int win_rows, int bstages) {
    return 1024 + (size_t)kWinSlots * ((win_rows * 128 + 1023) & ~1023) + (size_t)bstages * kTapsPerStage * (n / 2) * 128 + sizeof(WinBarriers) +
           256 * sizeof(float);
}

int conv_gemm_launch(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap* tmap_b_half, const CUtensorMap* tmap_a_win,
                     const ConvArgs& args, cudaStream_t stream) {
    if (args.n < 16 || args.n > 256 || args.n % 16 != 0) { set_error("conv: N=%d must be a multiple of 16 in [16,256]", args.n); return AG_EINVAL; }
    if (args.taps < 1 || args.taps > kConvMaxTaps || args.kchunks < 1) { set_error("conv: bad taps/kchunks"); return AG_EINVAL; }
    static int debug = -1;
    if (debug < 0) { const char* e = getenv("AG_CONV_DEBUG"); debug = e ? atoi(e) : 0; }
    ConvArgs dargs = args;
    dargs.debug = debug;
    const int tiles = (args.m_total + kTileM - 1) / kTileM;
    if (tiles < 1) return AG_OK;
    // CTA pairs: N/2 rows of weights per CTA must be whole 8-row swizzle groups, and there must be work for a pair
    if (tmap_b_half && conv_pair_enabled() && args.n % 32 == 0 && tiles >= 2) {
        const int win_rows = tmap_a_win && conv_window_enabled() ? conv_window_rows(args.off, args.taps) : 0;
        if (win_rows) {
            // one activation window per channel slice, taps as row-shifted descriptor starts
            static PerDevice<size_t> configured3_dev;
            size_t& configured3 = configured3_dev.cur();
            int bstages = kMaxBStages;
            while (bstages > 2 && conv_win_smem_bytes(args.n, win_rows, bstages) > 227 * 1024) bstages--;
            size_t smem = conv_win_smem_bytes(args.n, win_rows, bstages);
            if (smem <= 227 * 1024) {
                if (smem > configured3) {
                    AG_CUDA(cudaFuncSetAttribute(conv_gemm_win_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                    configured3 = smem;
                }
                dargs.win_rows = win_rows;
                dargs.halo = (win_rows - kTileM) / 2;
                int pairs = (tiles + 1) / 2;
                int clusters = sm_count() / 2;
                if (clusters > pairs) clusters = pairs;
                launch_k(conv_gemm_win_kernel, dim3(2 * clusters), dim3(kThreads), smem, stream, *tmap_a_win, *tmap_b_half, dargs, bstages);
                AG_LAUNCH_CHECK("conv_gemm_win_kernel");
                return AG_OK;
            }
        }
        static PerDevice<size_t> configured2_dev;
        size_t& configured2 = configured2_dev.cur();
        int stages = kMaxStages;
        while (stages > 2 && conv_pair_smem_bytes(args.n, stages) > 227 * 1024) stages--;
        size_t smem = conv_pair_smem_bytes(args.n, stages);
        if (smem > configured2) {
            AG_CUDA(cudaFuncSetAttribute(conv_gemm_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured2 = smem;
        }
        int pairs = (tiles + 1) / 2;
        int clusters = sm_count() / 2;
        if (clusters > pairs) clusters = pairs;
        launch_k(conv_gemm_pair_kernel, dim3(2 * clusters), dim3(kThreads), smem, stream, tmap_a, *tmap_b_half, dargs, stages);
        AG_LAUNCH_CHECK("conv_gemm_pair_kernel");
        return AG_OK;
    }
    static PerDevice<size_t> configured_dev;
    size_t& configured = configured_dev.cur();
    const int stages = conv_stages(args.n);
    size_t smem = conv_smem_bytes(args.n, stages);
    if (smem > 227 * 1024) { set_error("conv: %zu bytes of shared memory needed", smem); return AG_EINVAL; }
    if (smem > configured) {
        AG_CUDA(cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    int grid = tiles < sm_count() ? tiles : sm_count();
    launch_k(conv_gemm_kernel, dim3(grid), dim3(kThreads), smem, stream, tmap_a, tmap_b, dargs, stages);
    AG_LAUNCH_CHECK("conv_gemm_kernel");
    return AG_OK;
}

bool conv_chain_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("AG_CONV_CHAIN"); on = e ? atoi(e) : 1; }
    return on != 0 && conv_pair_enabled() && conv_window_enabled();
}

static size_t conv_chain_smem_bytes(int win_bytes, int b_bytes, int bstages) {
    return 1024 + (size_t)kWinSlots * win_bytes + (size_t)bstages * kTapsPerStage * b_bytes + sizeof(WinBarriers) + 256 * sizeof(float) + 64;
}

int conv_chain_launch(const ChainParams& P, unsigned int* d_flags, unsigned int epoch, int win_rows_max, int n_max, cudaStream_t stream) {
    if (P.n_layers < 1 || P.total_tiles < 1) return AG_OK;
    if (P.n_layers > kChainMaxLayers) { set_error("conv chain: %d layers in one launch (max %d)", P.n_layers, kChainMaxLayers); return AG_EINVAL; }
    const int win_bytes = (win_rows_max * 128 + 1023) & ~1023, b_bytes = (n_max / 2) * 128;
    int bstages = kMaxBStages;
    while (bstages > 2 && conv_chain_smem_bytes(win_bytes, b_bytes, bstages) > 227 * 1024) bstages--;
    size_t smem = conv_chain_smem_bytes(win_bytes, b_bytes, bstages);
    // one CTA per SM, always: two CTAs of this kernel on one SM would deadlock (the second blocks in the TMEM allocation while
    // the first waits for tiles of the second's pair), so ask for more than half of the SM's shared memory
    if (smem < 120 * 1024) smem = 120 * 1024;
    if (smem > 227 * 1024) { set_error("conv chain: %zu bytes of shared memory needed", smem); return AG_EINVAL; }
    static PerDevice<size_t> configured_dev;
    static PerDevice<int> max_clusters_dev;
    size_t& configured = configured_dev.cur();
    int& max_clusters = max_clusters_dev.cur();
    if (smem > configured) {
        AG_CUDA(cudaFuncSetAttribute(conv_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
        // every CTA pair must be resident at once (tiles wait for each other): ask how many the device can hold
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(sm_count() & ~1); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem;
        int n = 0;
        AG_CUDA(cudaOccupancyMaxActiveClusters(&n, conv_chain_kernel, &cfg));
        max_clusters = n;
    }
    int clusters = sm_count() / 2;
    if (clusters > max_clusters) clusters = max_clusters;
    if (clusters > P.total_tiles) clusters = P.total_tiles;
    if (clusters < 1) { set_error("conv chain: no CTA pair fits on this device"); return AG_ESTATE; }
    cudaError_t e = launch_k_whole(conv_chain_kernel, dim3(2 * clusters), dim3(kThreads), smem, stream, P, d_flags, epoch, win_bytes, b_bytes, bstages);
    if (e == cudaErrorCooperativeLaunchTooLarge || e == cudaErrorNotSupported) {   // the grid cannot be co-resident here: not an error of the stream
        (void)cudaGetLastError();
        set_error("conv chain: %d CTA pairs cannot be co-resident on this device (%s)", clusters, cudaGetErrorString(e));
        return AG_ESTATE;
    }
    AG_LAUNCH_CHECK("conv_chain_kernel");
    return AG_OK;
}

}  // namespace ag

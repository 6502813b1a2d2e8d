// Weight gradient of a KxK 'same' convolution as a split-K tcgen05 GEMM over the pixel dimension
// (the backward of Keras' Convolution2D inside fit_generator, supervised_policy_trainer.py:207-213):
//
//     dW[tap][co][ci] = sum_m G[m - shift(tap)][co] * X[m][ci]
//
// G = gradient w.r.t. the layer's pre-activation output, X = the layer's input, both padded-pixel-
// row matrices (conv.cuh) whose zero pad slots make the halo terms vanish.  The reduction index m
// is the ROW index of both matrices, so both operands are MN-major: TMA brings [64 pixel rows][64
// channels] boxes (128-byte swizzle) that the UMMA reads transposed.  A CTA owns two 128-row tiles of the
// flattened (tap, co) axis = four 64-channel chunks, each loaded with its own tap shift, and a
// contiguous slice of the pixel rows; its fp32 partial goes to scratch [split][tile][...] in an order that makes the
// epilogue's stores contiguous (see there) and wgrad_reduce sums the splits in a fixed order (deterministic, no atomics).
//
// Narrow inputs (kin <= 128: the first layer, 48 planes -> kin = 64) would leave most of the UMMA's N = 256 unused
// that way (an M = 128 UMMA costs the same ~128 cycles for any N <= 256), so they use the mirrored tiling
// ("taps in N"): dW[tap][co][ci] = sum_m G[m][co] * X[m + shift(tap)][ci] with the CTA's 128 rows = two co chunks of
// G (unshifted) and its N = up to 256/kin taps side by side, each X chunk loaded with its own tap shift.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "conv.cuh"
#include "tcgen.cuh"
#include "wgrad.cuh"

namespace ag {

namespace {

constexpr int kMaxStages = 8;
constexpr int kThreads = 192;
constexpr int kTmemCols = 512;
constexpr int kAccStride = 256;  // TMEM columns between the CTA's two accumulators
constexpr int kASlots = 4;       // 64-row slots of dW rows per CTA: two accumulators x two halves

struct __align__(8) WgradBarriers {
    uint64_t full[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t acc_full;
    uint32_t tmem_base;
};

// Per-CTA constants, worked out once by thread 0.
struct WgradPlan {
    int n_win;               // G boxes fetched per stage
    int win_off[kASlots];    // byte offset of the box inside the stage
    int win_col[kASlots];    // its G column (co chunk * 64)
    int win_shift[kASlots];  // its first row is G row m0 - win_shift
    int slot_off[kASlots];   // byte offset, inside the stage, of the slot's first pixel row; -1 = dead slot
    int slot_row[kASlots];   // scratch row of the slot's first channel: tap*ldout + co chunk*64 (taps in N: co chunk*64)
    int slot_bias[kASlots];  // co chunk*64 when the slot's ones column is the bias gradient, else -1
    int lo[2], hi[2];        // slots behind rows 0-63 / 64-127 of each accumulator (hi = -1: dead half); lo has the lower address
};

// float4 elements of one CTA's partial: two accumulators of 128 rows x (the widest B operand of the tiling) columns
__host__ __device__ __forceinline__ int wgrad_tile_units(const WgradArgs& p) { return 2 * ((p.tpg > 1 ? p.tpg : 1) * p.kin / 4) * 128; }

// Per-CTA constants (see WgradPlan), worked out once by one thread.
__device__ __forceinline__ void wgrad_make_plan(const WgradArgs& p, WgradPlan* plan, const bool taps_in_n, const bool windowed,
                                                const int chunk_bytes, const int a_first, const int a_live) {
    const int win_bytes = windowed ? ((p.win_rows * 128 + 1023) & ~1023) : chunk_bytes;
    int cc[kASlots], shift[kASlots], win_of[kASlots];
    int n_win = 0;
    for (int i = 0; i < kASlots; i++) {
        plan->slot_off[i] = -1; plan->slot_row[i] = 0; plan->slot_bias[i] = -1;
        if (i >= a_live) continue;
        const int c = a_first + i;
        int tap = 0;
        if (taps_in_n) { cc[i] = c; shift[i] = 0; }
        else { cc[i] = c / p.taps; tap = c - cc[i] * p.taps; shift[i] = p.off[tap]; }
        plan->slot_row[i] = tap * p.ldout + cc[i] * 64;
        if (taps_in_n || tap == p.taps / 2) plan->slot_bias[i] = cc[i] * 64;  // unshifted G: its column sums are the bias gradient
        if (windowed && i > 0 && cc[i] == cc[i - 1]) {
            win_of[i] = win_of[i - 1];
        } else {
            win_of[i] = n_win;
            plan->win_off[n_win] = n_win * win_bytes;
            plan->win_col[n_win] = cc[i] * 64;
            n_win++;
        }
        plan->win_shift[win_of[i]] = shift[i];  // shifts grow with the tap: the last member's is the window's first row
    }
    plan->n_win = n_win;
    for (int i = 0; i < a_live; i++) plan->slot_off[i] = plan->win_off[win_of[i]] + (plan->win_shift[win_of[i]] - shift[i]) * 128;
    for (int a = 0; a < 2; a++) {
        const int s0 = 2 * a, s1 = 2 * a + 1;
        if (s1 >= a_live) { plan->lo[a] = s0; plan->hi[a] = -1; }
        else if (plan->slot_off[s0] < plan->slot_off[s1]) { plan->lo[a] = s0; plan->hi[a] = s1; }
        else { plan->lo[a] = s1; plan->hi[a] = s0; }
    }
}

// The kernel is bound by L2 -> shared-memory bandwidth, not by the tensor cores (a 128-row tile alone would need 16 KB
// of G and 24 KB of X per 4 UMMAs; the chip's L2 delivers ~43 B per SM per cycle), so
//   * a CTA owns TWO 128-row tiles that share the X operand: two accumulators in TMEM, 8 UMMAs per stage;
//   * its four 64-row slots are consecutive TAPS of one co chunk (co-major order), and taps of one co chunk are the same
//     G rows shifted by a few pixel rows: ONE box of 64 + span rows is fetched per co chunk ("window") and every slot's
//     UMMA descriptor starts at its own row of it (128 B per row; the 128-byte swizzle is a function of the address, so
//     a row-shifted start reads the right bytes).
__global__ void __launch_bounds__(kThreads, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap tmap_g, const __grid_constant__ CUtensorMap tmap_gw, const __grid_constant__ CUtensorMap tmap_x,
             const WgradArgs p, const int kStages) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    const int xchunks = p.kin / 64;
    const int cochunks = p.ldout / 64;
    const int chunk_bytes = p.krows * 128;          // one TMA box: [krows pixel rows][64 bf16]
    const bool windowed = p.win_rows > 0;
    // bytes of a stage set aside for G: four slots' boxes, or the (at most two) shared windows
    const int a_region = windowed ? 2 * ((p.win_rows * 128 + 1023) & ~1023) : kASlots * chunk_bytes;
    const int tile = blockIdx.x % p.tiles, split = blockIdx.x / p.tiles;
    const bool taps_in_n = p.tpg > 1;
    const int coquads = (cochunks + kASlots - 1) / kASlots;
    const int grp = taps_in_n ? tile / coquads : 0;
    // this CTA's slots are a_first .. a_first+3 of a_count: (co chunk, tap) pairs co-major, or co chunks alone (taps in N)
    const int a_first = (taps_in_n ? tile - grp * coquads : tile) * kASlots;
    const int a_count = taps_in_n ? cochunks : p.taps * cochunks;
    const int a_live = min(kASlots, a_count - a_first);
    const int nb = taps_in_n ? min(p.tpg, p.taps - grp * p.tpg) : 1;          // taps side by side in this CTA's B operand
    const bool with_bias = p.bias_scratch != nullptr && (!taps_in_n || grp == p.groups - 1);  // the last group has room for the ones
    const int b_chunks = nb * xchunks;
    // stage = G region (4 slots' worth) | X chunks | a chunk of ones: one more B column that is 1 for every pixel row
    const int stage_bytes = a_region + (taps_in_n ? p.tpg * xchunks : xchunks + (p.bias_scratch ? 1 : 0)) * chunk_bytes;
    const int ones_off = a_region + b_chunks * chunk_bytes;
    const int acc_cols = b_chunks * 64;
    const int umma_n = acc_cols + (with_bias ? 16 : 0);
    WgradBarriers* bars = reinterpret_cast<WgradBarriers*>(base + kStages * stage_bytes);
    WgradPlan* plan = reinterpret_cast<WgradPlan*>(bars + 1);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;  // warp-uniform by construction

    const int kb_total = p.kb_img ? (p.m_total / p.img_stride) * p.kb_img : (p.m_total + p.krows - 1) / p.krows;
    const int kb0 = (int)((long long)kb_total * split / p.splits), kb1 = (int)((long long)kb_total * (split + 1) / p.splits);

    pdl_trigger();
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { tc::mbar_init(&bars->full[s], 1); tc::mbar_init(&bars->empty[s], 1); }
        tc::mbar_init(&bars->acc_full, 1);
        tc::fence_mbar_init();
        tc::tma_prefetch_desc(&tmap_g);
        tc::tma_prefetch_desc(&tmap_gw);
        tc::tma_prefetch_desc(&tmap_x);
        wgrad_make_plan(p, plan, taps_in_n, windowed, chunk_bytes, a_first, a_live);
    }
    if (warp == 1) tc::tmem_alloc(&bars->tmem_base, kTmemCols);
    if (with_bias) {  // the ones chunk of every stage (all elements equal, so the swizzle does not matter)
        for (int s = 0; s < kStages; s++) {
            uint32_t* ones = reinterpret_cast<uint32_t*>(base + s * stage_bytes + ones_off);
            for (int i = threadIdx.x; i < chunk_bytes / 4; i += kThreads) ones[i] = 0x3F803F80u;  // bf16 1.0 x2
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
        const int n_win = plan->n_win;
        const CUtensorMap* gmap = windowed ? &tmap_gw : &tmap_g;
        // measurement only (AG_WGRAD_DEBUG & 4): fetch just the X chunks a CTA of a pair would (half of N, rounded up to chunks)
        const int xload = (p.debug & 4) ? (xchunks * 32 + 63) / 64 : xchunks;
        const uint32_t tx_bytes = (uint32_t)(n_win * (windowed ? p.win_rows * 128 : chunk_bytes) + (b_chunks - nb * (xchunks - xload)) * chunk_bytes);
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = kb0; kb < kb1; kb++) {
            tc::mbar_wait(&bars->empty[stage], phase ^ 1);
            if (tc::elect_one()) {
                uint8_t* s = base + stage * stage_bytes;
                tc::mbar_expect_tx(&bars->full[stage], tx_bytes);
                const int m0 = p.kb_img ? (kb / p.kb_img) * p.img_stride + (kb % p.kb_img) * p.krows : kb * p.krows;
                for (int w = 0; w < n_win; w++)  // rows outside the tensor read as zeros
                    tc::tma_load_2d(s + plan->win_off[w], gmap, &bars->full[stage], plan->win_col[w], m0 - plan->win_shift[w]);
                for (int j = 0; j < nb; j++) {
                    const int xrow = taps_in_n ? m0 + p.off[grp * p.tpg + j] : m0;
                    for (int c = 0; c < xload; c++)
                        tc::tma_load_2d(s + a_region + (j * xchunks + c) * chunk_bytes, &tmap_x, &bars->full[stage], c * 64, xrow);
                }
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        // the ones column (N + 16) only where an accumulator holds a slot whose column sums are wanted (an unshifted slot):
        // the other accumulators run with N = the X columns alone (3 of 14 accumulators of a 3x3 192-filter layer need it)
        uint32_t idesc_acc[2];
        for (int a = 0; a < 2; a++) {
            const int lo = plan->lo[a], hi = plan->hi[a];
            const bool ones = with_bias && ((plan->slot_off[lo] >= 0 && plan->slot_bias[lo] >= 0) || (hi >= 0 && plan->slot_bias[hi] >= 0));
            idesc_acc[a] = tc::idesc_bf16(128, ones ? umma_n : acc_cols, 1, 1);
            // measurement only (AG_WGRAD_DEBUG & 2): UMMAs of half the width = the operand bytes one CTA of a pair would read
            // (& 8: each issued twice = the tensor work of the full width on top); results are wrong in these modes
            if (p.debug & 2) idesc_acc[a] = tc::idesc_bf16(128, acc_cols / 2, 1, 1);
            // & 8 / & 16: read A / B as if it were K-major (same bytes fetched, different order); & 32: M = 64
            if (p.debug & (8 | 16 | 32))
                idesc_acc[a] = tc::idesc_bf16((p.debug & 32) ? 64 : 128, (p.debug & 2) ? acc_cols / 2 : acc_cols, (p.debug & 8) ? 0 : 1, (p.debug & 16) ? 0 : 1);
        }
        const bool second = a_live > 2;  // the second accumulator has at least one live half
        // A operand of accumulator a: rows 0-63 from slot lo, rows 64-127 from slot hi (LBO = their distance)
        uint32_t a_off[2], a_lbo[2];
        for (int a = 0; a < 2; a++) {
            const int lo = plan->lo[a], hi = plan->hi[a];
            a_off[a] = (uint32_t)max(plan->slot_off[lo], 0);
            a_lbo[a] = hi >= 0 ? (uint32_t)(plan->slot_off[hi] - plan->slot_off[lo]) : (uint32_t)chunk_bytes;
        }
        const int ksteps = p.krows / 16;
        int stage = 0;
        uint32_t phase = 0;
        bool ready = false;
        for (int kb = kb0; kb < kb1; kb++) {
            if (!ready) tc::mbar_wait(&bars->full[stage], phase);
            tc::fence_after_sync();
            if (tc::elect_one()) {
                const uint32_t s = tc::smem_u32(base + stage * stage_bytes);
                uint64_t adesc0 = tc::desc_mnmajor_sw128(s + a_off[0], a_lbo[0]);
                uint64_t adesc1 = tc::desc_mnmajor_sw128(s + a_off[1], a_lbo[1]);
                if (p.debug & 1) {  // experiment: descriptor base-offset field = position of the start inside the 1024-byte swizzle pattern
                    adesc0 |= (uint64_t)(((s + a_off[0]) >> 7) & 7) << 49;
                    adesc1 |= (uint64_t)(((s + a_off[1]) >> 7) & 7) << 49;
                }
                uint64_t bdesc = tc::desc_mnmajor_sw128(s + a_region, chunk_bytes);
                if (p.debug & 8) { adesc0 = tc::desc_kmajor_sw128(s + a_off[0]); adesc1 = tc::desc_kmajor_sw128(s + a_off[1]); }
                if (p.debug & 16) bdesc = tc::desc_kmajor_sw128(s + a_region);
#pragma unroll 4
                for (int k = 0; k < ksteps; k++) {  // 16 pixel rows per UMMA: +2048 B
                    tc::umma_bf16(tmem_base, adesc0 + (2048 >> 4) * k, bdesc + (2048 >> 4) * k, idesc_acc[0], (kb > kb0) || (k > 0));
                    if (second)
                        tc::umma_bf16(tmem_base + kAccStride, adesc1 + (2048 >> 4) * k, bdesc + (2048 >> 4) * k, idesc_acc[1], (kb > kb0) || (k > 0));
                }
                tc::umma_commit(&bars->empty[stage]);
                if (kb == kb1 - 1) tc::umma_commit(&bars->acc_full);
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
            ready = tc::mbar_try_wait(&bars->full[stage], phase);  // peek at the next slice while these UMMAs run
        }
    } else {
        const int q = warp & 3;
        const int r = q * 32 + lane;           // row of a 128-row accumulator
        if (kb1 > kb0) {
            tc::mbar_wait(&bars->acc_full, 0);
            tc::fence_after_sync();
        }
        for (int a = 0; a < 2; a++) {
            if (a * 2 >= a_live) break;  // warp-uniform
            const int slot = (r >> 6) ? plan->hi[a] : plan->lo[a];  // warp-uniform: a warp's 32 rows share a half
            const bool valid = slot >= 0;
            // partials as [4-column group][slot * 64 + row of the slot][4] (wgrad_tile_units): a warp's 32 rows of one column
            // group are 512 contiguous bytes (row-major partials would scatter every store over 32 lines); wgrad_reduce_kernel
            // puts them where they belong in dW
            uint4* dst = reinterpret_cast<uint4*>(p.scratch) + ((size_t)split * p.tiles + tile) * wgrad_tile_units(p) +
                         (valid ? slot : 0) * 64 + (r & 63);
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * kAccStride;
            for (int c0 = 0; c0 < acc_cols; c0 += 32) {
                uint32_t v[32];
                if (kb1 > kb0) {
                    tc::tmem_ld32(taddr + c0, v);
                    tc::tmem_ld_wait();
                } else {
#pragma unroll
                    for (int j = 0; j < 32; j++) v[j] = 0;
                }
                if (valid && !(p.debug & 64)) {  // & 64 (measurement only): leave the partial stores out
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        dst[(size_t)((c0 + j) >> 2) * 256] = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                }
            }
            if (with_bias && valid && plan->slot_bias[slot] >= 0) {
                // the column behind the last X chunk is sum_m G[m][co] for an unshifted slot: the bias gradient
                uint32_t v[32];
                if (kb1 > kb0) { tc::tmem_ld32(taddr + acc_cols, v); tc::tmem_ld_wait(); } else v[0] = 0;
                p.bias_scratch[(size_t)split * p.ldout + plan->slot_bias[slot] + (r & 63)] = __uint_as_float(v[0]);
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

// dW = sum over the splits of the partials, in a fixed association (deterministic, and the chained kernel's): the block's 4
// thread rows each sum every 4th split (independent loads, all in flight together: the kernel lives on memory-level
// parallelism), then the four sums are added in order through shared memory.
// The partials are [split][tile][4-column group][slot * 64 + row][4] (see the kernel's epilogue); a block takes 16 rows of one
// slot for four neighbouring column groups -- reads of 256-byte runs, writes of 64 contiguous bytes per dW row -- and sends
// them to their place in dW [tap][co][ci].  Block (0, 0) also sums the bias column.
constexpr int kReduceWays = 4;
constexpr int kReduceElems = 64;   // float4 elements per block: 16 rows x 4 column groups
__global__ void __launch_bounds__(kReduceElems * kReduceWays)
wgrad_reduce_kernel(const WgradArgs p, float* __restrict__ dw, float* __restrict__ db) {
    pdl_trigger();
    pdl_wait();
    __shared__ float4 part[kReduceWays][kReduceElems];
    const int tile = blockIdx.x;
    const bool taps_in_n = p.tpg > 1;
    const int cochunks = p.ldout / 64;
    const int coquads = (cochunks + kASlots - 1) / kASlots;
    const int grp = taps_in_n ? tile / coquads : 0;
    const int a_first = (taps_in_n ? tile - grp * coquads : tile) * kASlots;
    const int a_count = taps_in_n ? cochunks : p.taps * cochunks;
    const int nb = taps_in_n ? min(p.tpg, p.taps - grp * p.tpg) : 1;
    if (db && tile == 0 && blockIdx.y == 0) {
        for (int c = threadIdx.x; c < p.ldout; c += blockDim.x) {
            float t = 0.f;
            for (int s = 0; s < p.splits; s++) t += p.bias_scratch[(size_t)s * p.ldout + c];
            db[c] = t;
        }
    }
    const int kin4 = p.kin / 4;
    const int tile_units = wgrad_tile_units(p);
    const int e = threadIdx.x % kReduceElems, way = threadIdx.x / kReduceElems;
    // blockIdx.y = (group of 4 column groups, slot, group of 16 rows); kin is a multiple of 64, so 4 | kin4
    const int c4 = (blockIdx.y >> 4) * 4 + (e >> 4), slot = (blockIdx.y >> 2) & 3, row = (blockIdx.y & 3) * 16 + (e & 15);
    const int c = a_first + slot;                                   // (co chunk, tap) pair co-major, or the co chunk (taps in N)
    if (c >= a_count || (blockIdx.y >> 4) * 4 >= nb * kin4) return; // dead slot / columns of a short last tap group (block-uniform)
    const int cc = taps_in_n ? c : c / p.taps;
    const int j = c4 / kin4;                                        // tap within the group (taps in N), else 0
    const int tap = taps_in_n ? grp * p.tpg + j : c - cc * p.taps;
    const size_t split_units = (size_t)p.tiles * tile_units;
    const float4* s4 = reinterpret_cast<const float4*>(p.scratch) + (size_t)tile * tile_units + (size_t)c4 * 256 + slot * 64 + row;
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int s = way; s < p.splits; s += kReduceWays) {
        const float4 b = __ldg(s4 + (size_t)s * split_units);
        t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
    }
    part[way][e] = t;
    __syncthreads();
    if (way == 0) {
#pragma unroll
        for (int w = 1; w < kReduceWays; w++) {
            const float4 b = part[w][e];
            t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
        }
        reinterpret_cast<float4*>(dw)[((size_t)tap * p.ldout + cc * 64 + row) * kin4 + (c4 - j * kin4)] = t;
    }
}

// ---- Weight gradients of a run of layers with identical geometry in ONE launch, split reduction included.
// Per layer a CTA does the same (tile, split) work item as wgrad_kernel; what changes is what happens around it:
//   * no launch / prologue / pipeline fill per layer: the TMA ring simply continues with the next layer's operands while the
//     epilogue warps are still busy with this one's accumulators;
//   * the split-K reduction is done by the same CTAs: after writing its partial a CTA bumps the tile's counter, waits until all
//     `splits` partials of the tile are there, and sums its 1/splits share of the tile's rows (same association as
//     wgrad_reduce_kernel: bit-identical results) -- all on the epilogue warps, which would otherwise idle during the next
//     layer's main loop.  Needs every CTA of the grid resident at once (checked on the host).
// Partials alternate between two scratch buffers by layer parity (a CTA cannot reach layer l+2 before every CTA of its tile
// has finished reducing layer l, see the counter wait), so they stay in L2.
constexpr int kWgradChainMax = 16;
struct WgradChainLayer {
    CUtensorMap g, gw, x;
    float* scratch;        // this layer's partial buffer [splits][tiles][2 accumulators][kin/4][128 rows][4] (TMEM order)
    float* bias_scratch;   // [splits][ldout]
    float* dw;             // [taps][ldout][kin]
    float* db;             // [ldout]
    unsigned int* counters;  // [tiles], running totals over launches
    unsigned int target;     // counter value that means "all splits of this launch have arrived"
    unsigned int* done;      // running total of CTAs that have finished their share of this layer's dW / db (may be null)
};
struct WgradChainParams {
    WgradArgs a;           // shared geometry (scratch pointers unused)
    WgradChainLayer layers[kWgradChainMax];
    int n_layers;
};

__device__ __forceinline__ float4 ldcg4(const float4* p) {  // L2 only: other CTAs wrote these lines during this launch
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

__global__ void __launch_bounds__(kThreads, 1)
wgrad_chain_kernel(const __grid_constant__ WgradChainParams P, const int kStages) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    const WgradArgs& p = P.a;
    const int xchunks = p.kin / 64;
    const int cochunks = p.ldout / 64;
    const int chunk_bytes = p.krows * 128;
    const bool windowed = p.win_rows > 0;
    const int a_region = windowed ? 2 * ((p.win_rows * 128 + 1023) & ~1023) : kASlots * chunk_bytes;
    const int tile = blockIdx.x % p.tiles, split = blockIdx.x / p.tiles;
    const int a_first = tile * kASlots;
    const int a_count = p.taps * cochunks;
    const int a_live = min(kASlots, a_count - a_first);
    const bool with_bias = P.layers[0].bias_scratch != nullptr;
    const int stage_bytes = a_region + (xchunks + (with_bias ? 1 : 0)) * chunk_bytes;
    const int ones_off = a_region + xchunks * chunk_bytes;
    const int acc_cols = xchunks * 64;
    const int umma_n = acc_cols + (with_bias ? 16 : 0);
    WgradBarriers* bars = reinterpret_cast<WgradBarriers*>(base + kStages * stage_bytes);
    WgradPlan* plan = reinterpret_cast<WgradPlan*>(bars + 1);
    uint64_t* acc_empty = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(plan + 1) + 7) & ~uintptr_t(7));
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;

    const int kb_total = p.kb_img ? (p.m_total / p.img_stride) * p.kb_img : (p.m_total + p.krows - 1) / p.krows;
    const int kb0 = (int)((long long)kb_total * split / p.splits), kb1 = (int)((long long)kb_total * (split + 1) / p.splits);

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { tc::mbar_init(&bars->full[s], 1); tc::mbar_init(&bars->empty[s], 1); }
        tc::mbar_init(&bars->acc_full, 1);
        tc::mbar_init(acc_empty, 128);
        tc::fence_mbar_init();
        wgrad_make_plan(p, plan, false, windowed, chunk_bytes, a_first, a_live);
    }
    if (warp == 1) tc::tmem_alloc(&bars->tmem_base, kTmemCols);
    if (with_bias) {  // the ones chunk of every stage (all elements equal, so the swizzle does not matter)
        for (int s = 0; s < kStages; s++) {
            uint32_t* ones = reinterpret_cast<uint32_t*>(base + s * stage_bytes + ones_off);
            for (int i = threadIdx.x; i < chunk_bytes / 4; i += kThreads) ones[i] = 0x3F803F80u;  // bf16 1.0 x2
        }
        tc::fence_proxy_async();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = bars->tmem_base;
    pdl_wait();

    if (warp == 0) {
        // producer: one ring over all layers
        const int n_win = plan->n_win;
        const uint32_t tx_bytes = (uint32_t)(n_win * (windowed ? p.win_rows * 128 : chunk_bytes) + xchunks * chunk_bytes);
        int stage = 0;
        uint32_t phase = 0;
        for (int li = 0; li < P.n_layers; li++) {
            const WgradChainLayer& L = P.layers[li];
            const CUtensorMap* gmap = windowed ? &L.gw : &L.g;
            for (int kb = kb0; kb < kb1; kb++) {
                tc::mbar_wait(&bars->empty[stage], phase ^ 1);
                if (tc::elect_one()) {
                    uint8_t* s = base + stage * stage_bytes;
                    tc::mbar_expect_tx(&bars->full[stage], tx_bytes);
                    const int m0 = p.kb_img ? (kb / p.kb_img) * p.img_stride + (kb % p.kb_img) * p.krows : kb * p.krows;
                    for (int w = 0; w < n_win; w++)  // rows outside the tensor read as zeros
                        tc::tma_load_2d(s + plan->win_off[w], gmap, &bars->full[stage], plan->win_col[w], m0 - plan->win_shift[w]);
                    for (int c = 0; c < xchunks; c++)
                        tc::tma_load_2d(s + a_region + c * chunk_bytes, &L.x, &bars->full[stage], c * 64, m0);
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // the ones column (N + 16) only where an accumulator holds a slot whose column sums are wanted (an unshifted slot):
        // the other accumulators run with N = the X columns alone (3 of 14 accumulators of a 3x3 192-filter layer need it)
        uint32_t idesc_acc[2];
        for (int a = 0; a < 2; a++) {
            const int lo = plan->lo[a], hi = plan->hi[a];
            const bool ones = with_bias && ((plan->slot_off[lo] >= 0 && plan->slot_bias[lo] >= 0) || (hi >= 0 && plan->slot_bias[hi] >= 0));
            idesc_acc[a] = tc::idesc_bf16(128, ones ? umma_n : acc_cols, 1, 1);
        }
        const bool second = a_live > 2;
        uint32_t a_off[2], a_lbo[2];
        for (int a = 0; a < 2; a++) {
            const int lo = plan->lo[a], hi = plan->hi[a];
            a_off[a] = (uint32_t)max(plan->slot_off[lo], 0);
            a_lbo[a] = hi >= 0 ? (uint32_t)(plan->slot_off[hi] - plan->slot_off[lo]) : (uint32_t)chunk_bytes;
        }
        const int ksteps = p.krows / 16;
        int stage = 0;
        uint32_t phase = 0;
        bool ready = false;
        for (int li = 0; li < P.n_layers; li++) {
            tc::mbar_wait(acc_empty, (li & 1) ^ 1);  // the epilogue warps have read the previous layer's accumulators out of TMEM
            tc::fence_after_sync();
            for (int kb = kb0; kb < kb1; kb++) {
                if (!ready) tc::mbar_wait(&bars->full[stage], phase);
                tc::fence_after_sync();
                if (tc::elect_one()) {
                    const uint32_t s = tc::smem_u32(base + stage * stage_bytes);
                    const uint64_t adesc0 = tc::desc_mnmajor_sw128(s + a_off[0], a_lbo[0]);
                    const uint64_t adesc1 = tc::desc_mnmajor_sw128(s + a_off[1], a_lbo[1]);
                    const uint64_t bdesc = tc::desc_mnmajor_sw128(s + a_region, chunk_bytes);
#pragma unroll 4
                    for (int k = 0; k < ksteps; k++) {  // 16 pixel rows per UMMA: +2048 B
                        tc::umma_bf16(tmem_base, adesc0 + (2048 >> 4) * k, bdesc + (2048 >> 4) * k, idesc_acc[0], (kb > kb0) || (k > 0));
                        if (second)
                            tc::umma_bf16(tmem_base + kAccStride, adesc1 + (2048 >> 4) * k, bdesc + (2048 >> 4) * k, idesc_acc[1], (kb > kb0) || (k > 0));
                    }
                    tc::umma_commit(&bars->empty[stage]);
                    if (kb == kb1 - 1) tc::umma_commit(&bars->acc_full);
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
                ready = tc::mbar_try_wait(&bars->full[stage], phase);
            }
        }
    } else {
        // epilogue warps: partial out of TMEM, then this CTA's share of the tile's split reduction
        const int q = warp & 3, et = threadIdx.x - 64;
        const int r = q * 32 + lane;
        const int kin4 = p.kin / 4;
        const int tile_units = 2 * kin4 * 128;                    // float4 elements of one CTA's partial (both accumulators)
        const size_t split_units = (size_t)p.tiles * tile_units;  // ... of one split's partials
        const int units = (a_live > 2 ? 2 : 1) * kin4 * 128;      // ... of this tile's live accumulators
        const int u0 = (int)((long long)units * split / p.splits), u1 = (int)((long long)units * (split + 1) / p.splits);
        for (int li = 0; li < P.n_layers; li++) {
            const WgradChainLayer& L = P.layers[li];
            tc::mbar_wait(&bars->acc_full, li & 1);
            tc::fence_after_sync();
            for (int a = 0; a < 2; a++) {
                if (a * 2 >= a_live) break;  // warp-uniform
                const int slot = (r >> 6) ? plan->hi[a] : plan->lo[a];
                const bool valid = slot >= 0;
                // partials are kept in the order the accumulator comes out of TMEM, [accumulator][4-column group][row][4]:
                // a warp's 32 rows of one column group are 512 contiguous bytes (row-major partials would scatter every
                // store over 32 lines: 5.3 us of this epilogue per layer, during which the next layer's UMMAs wait for TMEM)
                uint4* dst = reinterpret_cast<uint4*>(L.scratch) + ((size_t)split * p.tiles + tile) * tile_units + (size_t)a * kin4 * 128 + r;
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * kAccStride;
                for (int c0 = 0; c0 < acc_cols; c0 += 32) {
                    uint32_t v[32];
                    tc::tmem_ld32(taddr + c0, v);
                    tc::tmem_ld_wait();
                    if (valid) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            dst[(size_t)((c0 + j) >> 2) * 128] = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                    }
                }
                if (with_bias && valid && plan->slot_bias[slot] >= 0) {
                    uint32_t v[32];
                    tc::tmem_ld32(taddr + acc_cols, v);
                    tc::tmem_ld_wait();
                    L.bias_scratch[(size_t)split * p.ldout + plan->slot_bias[slot] + (r & 63)] = __uint_as_float(v[0]);
                }
            }
            tc::fence_before_sync();
            tc::mbar_arrive(acc_empty);  // TMEM is free for the next layer's UMMAs
            // ---- publish the partial, wait for the tile's other splits
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
                asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(L.counters + tile) : "memory");
                unsigned int v;
                long long t0 = 0;
                for (;;) {
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(L.counters + tile) : "memory");
                    if (v >= L.target) break;
                    if (t0 == 0) t0 = clock64();
                    else if (clock64() - t0 > 4000000000LL) {
                        printf("alphago_b200: wgrad chain split wait timed out (block %d layer %d counter %u, want %u)\n", blockIdx.x, li, v, L.target);
                        __trap();
                    }
                    __nanosleep(64);
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            // ---- this CTA's share of the tile: units [u0, u1), four interleaved partial sums added in order
            const float4* s4 = reinterpret_cast<const float4*>(L.scratch);
            float4* d4 = reinterpret_cast<float4*>(L.dw);
            for (int u = u0 + et; u < u1; u += 128) {   // u = (accumulator, column group, row): the partials' own order
                const int ua = u / (kin4 * 128);
                const int c4 = (u >> 7) - ua * kin4, row = u & 127;
                const int sl = (row >> 6) ? plan->hi[ua] : plan->lo[ua];
                if (sl < 0) continue;                    // dead half of the last tile
                const size_t i = (size_t)(plan->slot_row[sl] + (row & 63)) * kin4 + c4;   // dW is [tap][co][ci]
                const float4* src = s4 + (size_t)tile * tile_units + u;
                float4 acc[kReduceWays];
#pragma unroll
                for (int w = 0; w < kReduceWays; w++) {
                    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int s = w; s < p.splits; s += kReduceWays) {
                        const float4 b = ldcg4(src + (size_t)s * split_units);
                        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
                    }
                    acc[w] = a;
                }
                float4 a = acc[0];
#pragma unroll
                for (int w = 1; w < kReduceWays; w++) { a.x += acc[w].x; a.y += acc[w].y; a.z += acc[w].z; a.w += acc[w].w; }
                d4[i] = a;
            }
            if (with_bias && split == 0 && et < 64) {
                for (int sl = 0; sl < a_live; sl++) {
                    if (plan->slot_bias[sl] < 0) continue;
                    const int c = plan->slot_bias[sl] + et;
                    float t = 0.f;
                    for (int s = 0; s < p.splits; s++) t += __ldcg(L.bias_scratch + (size_t)s * p.ldout + c);
                    L.db[c] = t;
                }
            }
            // ---- this CTA's part of the layer's gradient is final: count it.  A stream that waits for the count to reach
            // (launches so far) x (CTAs per launch) may start exchanging this layer while the layers below are still running.
            if (L.done) {
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (et == 0) {
                    asm volatile("fence.acq_rel.gpu;" ::: "memory");
                    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(L.done) : "memory");
                }
            }
        }
    }
    pdl_trigger();  // not earlier: every CTA of this grid must be resident before anything queued behind it may take SMs
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_base, kTmemCols);
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
        return ((ldout / 64 + kASlots - 1) / kASlots) * *groups;
    }
    *tpg = 1;
    *groups = taps;
    return (taps * (ldout / 64) + kASlots - 1) / kASlots;
}

// floats of ONE split's partials: every CTA tile keeps two whole accumulators (128 rows x the tiling's widest B operand), which
// is a little more than the taps * ldout * kin elements of dW when the tiles do not divide them
size_t wgrad_partial_floats(int taps, int ldout, int kin) {
    WgradArgs a;
    a.kin = kin;
    const int tiles = wgrad_tiling(taps, ldout, kin, &a.tpg, &a.groups);
    return (size_t)tiles * wgrad_tile_units(a) * 4;
}

// pixel rows per pipeline stage (one TMA box per operand chunk).  Measured (3x3, 192 channels, batch 256): 64 rows x 3
// stages 65.9 us, 32 rows x 6 stages 83 us: the smaller boxes cost more than the deeper pipeline gains.
static int wgrad_krows() {
    static int v = 0;
    if (!v) { const char* e = getenv("AG_WGRAD_KROWS"); v = (e && atoi(e) == 32) ? 32 : 64; }
    return v;
}

static int wgrad_stage_bytes(int taps, int kin, bool with_bias, int krows, int win_rows) {
    int tpg, groups;
    wgrad_tiling(taps, 64, kin, &tpg, &groups);
    const int a_region = win_rows > 0 ? 2 * ((win_rows * 128 + 1023) & ~1023) : kASlots * krows * 128;  // as in the kernels
    return a_region + (tpg > 1 ? tpg * (kin / 64) : kin / 64 + (with_bias ? 1 : 0)) * krows * 128;
}

static int wgrad_stages(int taps, int kin, bool with_bias, int krows, int win_rows) {
    int s = (227 * 1024 - 1024 - (int)sizeof(WgradBarriers) - (int)sizeof(WgradPlan) - 64) / wgrad_stage_bytes(taps, kin, with_bias, krows, win_rows);
    return s > kMaxStages ? kMaxStages : s;
}

static size_t wgrad_smem_bytes(int taps, int kin, bool with_bias, int krows, int win_rows) {
    return 1024 + (size_t)wgrad_stages(taps, kin, with_bias, krows, win_rows) * wgrad_stage_bytes(taps, kin, with_bias, krows, win_rows) +
           sizeof(WgradBarriers) + sizeof(WgradPlan) + 64;
}

// Rows of the G box shared by the taps of one co chunk (0 = every slot fetches its own krows rows): krows + the largest
// shift distance inside any four consecutive taps.  A CTA holds at most two such boxes (taps >= 3), which must fit
// the G region of a stage.
static int wgrad_window_rows(int taps, const int* off, int kin, int krows) {
    int tpg, groups;
    wgrad_tiling(taps, 64, kin, &tpg, &groups);
    static int enabled = -1;
    if (enabled < 0) { const char* e = getenv("AG_WGRAD_WINDOW"); enabled = (e && e[0] == '0') ? 0 : 1; }
    if (tpg > 1 || taps < 3 || !enabled) return 0;
    int span = 0;
    for (int t = 0; t < taps; t++) {
        int u = t + kASlots - 1 < taps ? t + kASlots - 1 : taps - 1;
        if (off[u] < off[t]) return 0;  // shifts must grow with the tap
        if (off[u] - off[t] > span) span = off[u] - off[t];
    }
    int rows = (krows + span + 7) & ~7;
    if (rows > 256 || 2 * ((rows * 128 + 1023) & ~1023) > kASlots * krows * 128) return 0;
    return rows;
}

int wgrad_make_maps(WgradMaps* m, const void* g, const void* x, uint64_t rows, int ldout, int kin, int taps, const int* off) {
    m->krows = wgrad_krows();
    int rc = make_tmap_2d(&m->g, g, rows, ldout, m->krows);
    if (!rc) rc = make_tmap_2d(&m->x, x, rows, kin, m->krows);
    m->win_rows = wgrad_window_rows(taps, off, kin, m->krows);
    if (!rc) rc = make_tmap_2d(&m->gw, g, rows, ldout, m->win_rows ? m->win_rows : m->krows);
    return rc;
}

int wgrad_plan_splits(int taps, int ldout, int kin, int m_total, int sm_budget) {
    // CTAs = tiles * splits should fill whole waves of the SMs: one wave if that keeps >= 85 % of them busy,
    // else two.  Fewer, longer K slices also mean fewer fp32 partials to write and reduce.
    int tpg, groups;
    int tiles = wgrad_tiling(taps, ldout, kin, &tpg, &groups);
    int kb = (m_total + 63) / 64;
    int sms = sm_budget > 0 && sm_budget < sm_count() ? sm_budget : sm_count();  // SMs the grid may occupy
    int s1 = sms / tiles, s2 = 2 * sms / tiles;
    int s = s1;
    if (s1 < 1 || ((double)tiles * s1 / sms < 0.85 && (double)tiles * s2 / (2.0 * sms) > (double)tiles * s1 / sms + 0.05)) s = s2;
    if (s > kb) s = kb;
    if (s < 1) s = 1;
    if (s > kWgradMaxSplits) s = kWgradMaxSplits;
    return s;
}

int wgrad_launch(const WgradMaps& maps, WgradArgs args, float* dw, float* db, cudaStream_t stream) {
    if (args.kin % 64 != 0 || args.kin < 64 || args.kin > 256 || args.ldout % 64 != 0) {
        set_error("wgrad: channel counts must be padded to 64 (kin=%d <= 256, ldout=%d)", args.kin, args.ldout);
        return AG_EINVAL;
    }
    args.tiles = wgrad_tiling(args.taps, args.ldout, args.kin, &args.tpg, &args.groups);
    if (args.splits < 1) args.splits = 1;
    if (args.kb_img && (args.img_stride < 1 || args.m_total % args.img_stride != 0 || args.kb_img * maps.krows > args.img_stride)) args.kb_img = 0;
    static PerDevice<size_t> configured_dev;
    size_t& configured = configured_dev.cur();
    if (!db || args.kin + 16 > 256) args.bias_scratch = nullptr;
    if (db && !args.bias_scratch) { set_error("wgrad: bias gradient needs bias_scratch and kin <= 240"); return AG_EINVAL; }
    args.krows = maps.krows;
    size_t smem = wgrad_smem_bytes(args.taps, args.kin, args.bias_scratch != nullptr, args.krows, maps.win_rows);
    if (smem > 227 * 1024) { set_error("wgrad: %zu bytes of shared memory needed", smem); return AG_EINVAL; }
    if (smem > configured) {
        AG_CUDA(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    { static int dbg = -1; if (dbg < 0) { const char* e = getenv("AG_WGRAD_DEBUG"); dbg = e ? atoi(e) : 0; } args.debug = dbg; }
    const int stages = wgrad_stages(args.taps, args.kin, args.bias_scratch != nullptr, args.krows, maps.win_rows);
    if (stages < 2) { set_error("wgrad: shared memory fits only %d stage(s)", stages); return AG_EINVAL; }
    args.win_rows = maps.win_rows;
    launch_k(wgrad_kernel, dim3(args.tiles * args.splits), dim3(kThreads), smem, stream, maps.g, maps.gw, maps.x, args, stages);
    AG_LAUNCH_CHECK("wgrad_kernel");
    launch_k(wgrad_reduce_kernel, dim3(args.tiles, wgrad_tile_units(args) / kReduceElems), dim3(kReduceElems * kReduceWays), 0, stream, args, dw,
             args.bias_scratch ? db : nullptr);
    AG_LAUNCH_CHECK("wgrad_reduce_kernel");
    return AG_OK;
}

// ---- chained launch (wgrad_chain_kernel): layers share args' geometry; each entry brings its maps, buffers and counters
bool wgrad_chain_ok(const WgradArgs& a_in, const WgradMaps& maps, bool with_bias) {
    static int enabled = -1;
    if (enabled < 0) { const char* e = getenv("AG_WGRAD_CHAIN"); enabled = (e && e[0] == '0') ? 0 : 1; }
    if (!enabled) return false;
    WgradArgs a = a_in;
    a.tiles = wgrad_tiling(a.taps, a.ldout, a.kin, &a.tpg, &a.groups);
    if (a.tpg != 1 || a.kin % 64 || a.ldout % 64 || a.kin > 256) return false;
    if (with_bias && a.kin + 16 > 256) return false;
    const int kb_total = a.kb_img ? (a.m_total / a.img_stride) * a.kb_img : (a.m_total + maps.krows - 1) / maps.krows;
    if (a.splits < 1 || kb_total < a.splits) return false;
    // every CTA must be resident at once (they wait for each other's partials): one CTA per SM with this much shared memory
    return a.tiles * a.splits <= sm_count() && wgrad_stages(a.taps, a.kin, with_bias, maps.krows, maps.win_rows) >= 2;
}

int wgrad_chain_ctas(const WgradArgs& a_in) {
    WgradArgs a = a_in;
    return wgrad_tiling(a.taps, a.ldout, a.kin, &a.tpg, &a.groups) * (a.splits < 1 ? 1 : a.splits);
}

int wgrad_chain_launch(WgradArgs args, const WgradChainEntry* entries, int n_layers, cudaStream_t stream) {
    if (n_layers < 1) return AG_OK;
    if (n_layers > kWgradChainMax) { set_error("wgrad chain: %d layers in one launch (max %d)", n_layers, kWgradChainMax); return AG_EINVAL; }
    WgradChainParams P;
    memset(&P, 0, sizeof(P));
    args.tiles = wgrad_tiling(args.taps, args.ldout, args.kin, &args.tpg, &args.groups);
    args.win_rows = entries[0].maps->win_rows;
    args.krows = entries[0].maps->krows;
    if (args.kb_img && (args.img_stride < 1 || args.m_total % args.img_stride != 0 || args.kb_img * args.krows > args.img_stride)) args.kb_img = 0;
    args.debug = 0;
    const bool with_bias = entries[0].bias_scratch != nullptr;
    for (int i = 0; i < n_layers; i++) {
        WgradChainLayer& L = P.layers[i];
        L.g = entries[i].maps->g; L.gw = entries[i].maps->gw; L.x = entries[i].maps->x;
        L.scratch = entries[i].scratch; L.bias_scratch = entries[i].bias_scratch;
        L.dw = entries[i].dw; L.db = entries[i].db; L.counters = entries[i].counters; L.target = entries[i].target; L.done = entries[i].done;
        if ((entries[i].bias_scratch != nullptr) != with_bias || entries[i].maps->win_rows != args.win_rows) {
            set_error("wgrad chain: layers of one launch must share their geometry");
            return AG_EINVAL;
        }
    }
    P.a = args; P.n_layers = n_layers;
    const int stages = wgrad_stages(args.taps, args.kin, with_bias, args.krows, args.win_rows);
    size_t smem = wgrad_smem_bytes(args.taps, args.kin, with_bias, args.krows, args.win_rows);
    if (smem < 120 * 1024) smem = 120 * 1024;  // one CTA per SM (see wgrad_chain_ok)
    static PerDevice<size_t> configured_dev;
    size_t& configured = configured_dev.cur();
    if (smem > configured) {
        AG_CUDA(cudaFuncSetAttribute(wgrad_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    cudaError_t e = launch_k_whole(wgrad_chain_kernel, dim3(args.tiles * args.splits), dim3(kThreads), smem, stream, P, stages);
    if (e == cudaErrorCooperativeLaunchTooLarge || e == cudaErrorNotSupported) {   // the grid cannot be co-resident here: not an error of the stream
        (void)cudaGetLastError();
        set_error("wgrad chain: %d CTAs cannot be co-resident on this device (%s)", args.tiles * args.splits, cudaGetErrorString(e));
        return AG_ESTATE;
    }
    AG_LAUNCH_CHECK("wgrad_chain_kernel");
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
    return (size_t)ag::wgrad_plan_splits(width * width, ldout, kin, (int)rows, 0) * ag::wgrad_partial_floats(width * width, ldout, kin) * sizeof(float);
}

// bytes of one split's partials: the least scratch ag_wgrad can work with
size_t ag_wgrad_partial_bytes(int width, int ldout, int kin) {
    if (ag::wgrad_check(1, ldout, kin, 1) || width < 1 || width > 5) return 0;
    return ag::wgrad_partial_floats(width * width, ldout, kin) * sizeof(float);
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
    const size_t partial = ag::wgrad_partial_floats(a.taps, ldout, kin) * sizeof(float);
    a.splits = ag::wgrad_plan_splits(a.taps, ldout, kin, (int)rows, 0);
    if ((size_t)a.splits * partial > scratch_bytes) a.splits = (int)(scratch_bytes / partial);
    if (a.splits < 1 || !d_scratch) { ag::set_error("ag_wgrad: scratch of %zu bytes is smaller than one partial (%zu bytes)", scratch_bytes, partial); return AG_EINVAL; }
    a.scratch = d_scratch;
    ag::WgradMaps maps;
    rc = ag::wgrad_make_maps(&maps, d_g, d_x, (uint64_t)rows, ldout, kin, a.taps, a.off);
    if (rc) return rc;
    return ag::wgrad_launch(maps, a, d_dw, nullptr, (cudaStream_t)stream);
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
    const size_t partial = ag::wgrad_partial_floats(a.taps, ldout, kin) * sizeof(float);
    if (partial > AG_WGRAD_STATIC_SCRATCH_BYTES) {
        ag::set_error("ag_wgrad_static: one %dx%d x %d x %d gradient (%zu bytes) exceeds the built-in scratch of %u bytes; "
                      "use ag_wgrad with your own scratch", width, width, ldout, kin, partial, AG_WGRAD_STATIC_SCRATCH_BYTES);
        return AG_EINVAL;
    }
    a.splits = ag::wgrad_plan_splits(a.taps, ldout, kin, (int)rows, 0);
    if ((size_t)a.splits * partial > AG_WGRAD_STATIC_SCRATCH_BYTES) a.splits = (int)(AG_WGRAD_STATIC_SCRATCH_BYTES / partial);
    rc = ag::wgrad_static_scratch_ptr(&a.scratch);
    if (rc) return rc;
    ag::WgradMaps maps;
    rc = ag::wgrad_make_maps(&maps, d_g, d_x, (uint64_t)rows, ldout, kin, a.taps, a.off);
    if (rc) return rc;
    return ag::wgrad_launch(maps, a, d_dw, nullptr, (cudaStream_t)stream);
}

}  // extern "C"

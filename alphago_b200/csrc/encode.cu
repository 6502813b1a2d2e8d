// 48-plane AlphaGo feature encoder: one warp per position, integer work in shared memory.
//
// Replaces the reference's Preprocess.state_to_tensor hot loop
// (AlphaGo/preprocessing/preprocessing.py:9-245 on top of AlphaGo/go.py:63-267) for batches of
// positions.  Derived rules state (groups, liberty sets) is a pure function of the board, so it
// is rebuilt per position by a warp-wide union-find instead of porting the reference's
// incremental Python sets (go.py:104-157).
//
// Per position (flat point p = x*size+y, planes [f][x][y], AlphaGo/util.py:6-8); lane l owns points l, l+32, ...:
//   1. board / age -> shared memory; colour and turns_since codes           (coalesced byte loads)
//   2. groups: horizontal runs by a segmented max-scan over the lanes, then ONE pass of lock-free unions (atomicMin hooking)
//      over the vertical edges that join two runs for the first time, then a flatten to the root point
//   3. per group: stones and liberties (shared-memory atomics); the liberties of the side to move's groups, up to 8
//      each, in one pool laid out by a warp prefix sum (sum of min(libs, 8) <= 964 on any 19x19 board)
//   4. per point: liberties plane; for empty points legality (go.py:181-216), capture_size, and |L| for
//      self_atari_size / liberties_after (preprocessing.py:91-144, incl. the negative-index wrap of :143 and the
//      "captures ignored" quirk) counted WITHOUT a set: the first own neighbour group gives libs-1, every further
//      liberty counts if none of ITS neighbours belongs to a group taken before (labels compared, no bit sets);
//      sensibleness via an iterative restatement of the path-dependent is_eye recursion (go.py:230-259)
//   5. ladder candidates (moves touching an opponent group with two liberties / an own group in atari) go to a work
//      list in global memory; ladder_resolve_kernel reads them out one query per warp, dynamically balanced
//   6. nibble codes -> one-hot uint8 planes, 16-byte coalesced stores
//
// HBM-bound by construction (18,054 algorithmic bytes per position); in practice the integer work of
// steps 2-5 is what the kernel spends its time on, so the design goals are warps per SM (7.3 KB of shared memory per
// warp: 30 warps/SM) and short dependent chains.
#include <stdlib.h>

#include <atomic>

#include "common.cuh"
#include "encode.cuh"
#include "ladder.cuh"
#include "ladder_warp.cuh"

namespace ag {
namespace enc {

__device__ __forceinline__ bool eyeish(const int8_t* col, int size, int p, int owner) {  // go.py:218-228
    if (col[p] != 0) return false;
    int x = p / size, y = p - x * size;
    if (x > 0 && col[p - size] != owner) return false;
    if (x < size - 1 && col[p + size] != owner) return false;
    if (y > 0 && col[p - 1] != owner) return false;
    if (y < size - 1 && col[p + 1] != owner) return false;
    return true;
}

// go.py:230-259 as an explicit-stack DFS.  The reference's result for inner calls depends on the
// current recursion path (`d not in stack`) and on the early exit once bad > allowable, so the
// visit order (go.py:102) and both of those are reproduced literally.  Frame = pos | next<<9 | bad<<12.
__device__ __noinline__ bool is_eye(const int8_t* col, int size, int p0, int owner) {
    if (!eyeish(col, size, p0, owner)) return false;
    uint16_t stk[184];  // path nodes are distinct same-parity points: <= 181 on 19x19
    int sp = 0;
    stk[sp++] = (uint16_t)p0;
    bool ret = false, have_ret = false;
    while (true) {
        uint32_t f = stk[sp - 1];
        int pos = f & 511, j = (f >> 9) & 7, bad = (f >> 12) & 3;
        int x = pos / size, y = pos - x * size;
        int allow = (x > 0 && x < size - 1 && y > 0 && y < size - 1) ? 1 : 0;  // go.py:242
        bool failed = false, descended = false;
        if (have_ret) {
            have_ret = false;
            if (!ret) bad++;
            if (bad > allow) failed = true;
        }
        while (!failed && j < 4) {
            int dx = (j == 0 || j == 3) ? -1 : 1, dy = (j == 0 || j == 2) ? -1 : 1;
            j++;
            int xx = x + dx, yy = y + dy;
            if (xx < 0 || yy < 0 || xx >= size || yy >= size) continue;
            int d = xx * size + yy, c = col[d];
            if (c == -owner) bad++;
            else if (c == 0) {
                bool on = false;
                for (int t = 0; t < sp - 1; t++) on |= ((stk[t] & 511) == d);
                if (!on) {
                    if (!eyeish(col, size, d, owner)) bad++;
                    else if (sp < 184) {
                        stk[sp - 1] = (uint16_t)(pos | (j << 9) | (bad << 12));
                        stk[sp++] = (uint16_t)d;
                        descended = true;
                        break;
                    }
                }
            }
            if (bad > allow) failed = true;
        }
        if (descended) continue;
        ret = !failed;
        have_ret = true;
        if (--sp == 0) return ret;
    }
}

#ifdef AG_ENC_PROFILE
__device__ unsigned long long g_enc_phase_cycles[8];
#define ENC_T(i) do { long long _t = clock64(); if (lane == 0) atomicAdd(&g_enc_phase_cycles[i], (unsigned long long)(_t - t_phase)); t_phase = _t; } while (0)
#else
#define ENC_T(i) do {} while (0)
#endif

constexpr int COLPAD = 384;   // 12 x 32 points; col[] is zero beyond the board, so p + size may be read unguarded
constexpr int POOL = 968;     // >= max over boards of sum over one colour's groups of min(liberties, 8)  (= 964 on 19x19)
constexpr int NCODE = 4;      // nibble-packed code arrays: BOARD|TURNS, LIBS|CAPTURE, SELFATARI|LIBSAFTER, SENSIBLE|ladder marks

struct WarpMem {
    int8_t col[COLPAD];           // colour of the point
    uint16_t lab[NPAD];           // root point of the stone's group; 0xFFFF where empty
    uint32_t cnt[NPAD];           // by root: stones << 16 | liberties
    uint32_t off[NPAD];           // union-find parents (step 2); then, by root of an own group: filled << 16 | pool offset
    uint16_t pool[POOL];          // liberty lists (<= 8 each) of the side to move's groups
    uint8_t codes[NCODE][NPAD];   // low / high nibble = plane index within a family, 0xF = none
};

// plane descriptor in shared memory: bits 0-11 byte offset of the code array, 12-15 nibble shift, 16-19 value,
// 24-25 kind (0 compare, 1 all zeros, 2 all ones, 3 ones when black is to move)
__device__ __forceinline__ uint32_t plane_desc(int arr, int val) {
    switch (arr) {
        case C_BOARD: return 0u * NPAD | 0u << 12 | (uint32_t)val << 16;
        case C_TURNS: return 0u * NPAD | 4u << 12 | (uint32_t)val << 16;
        case C_LIBS: return 1u * NPAD | 0u << 12 | (uint32_t)val << 16;
        case C_CAPTURE: return 1u * NPAD | 4u << 12 | (uint32_t)val << 16;
        case C_SELFATARI: return 2u * NPAD | 0u << 12 | (uint32_t)val << 16;
        case C_LIBSAFTER: return 2u * NPAD | 4u << 12 | (uint32_t)val << 16;
        case C_SENSIBLE: return 3u * NPAD | 0u << 12 | (uint32_t)val << 16;
        case C_CONST: return (val == 0 ? 2u : 1u) << 24;
        case C_COLOR: return 3u << 24;
        default: return 1u << 24;   // ladder planes: zeros here, ladder_resolve_kernel writes the ones
    }
}

// work list of ladder queries: header {appended, taken} then entries position << 16 | point << 5 | kind << 4 | prey mask
// (bit j of the mask: neighbour j of the point, in go.py:96 order, is the first to touch a prey group)
struct LadderWork {
    unsigned int* head;      // [0] capture queries appended by encode_kernel, [2] escape queries, [1] next query to take
    uint32_t* entries;
    uint32_t* bits;          // per position: 12 words of black stones, 12 words of white stones
    unsigned int esc_base;   // where the escape queries start in `entries` (head[2] counts them)
};

__device__ __forceinline__ uint32_t uf_find(volatile uint32_t* parent, uint32_t a) {
    while (true) {
        const uint32_t b = parent[a];
        if (b == a) return a;
        a = b;
    }
}
// lock-free union: hook the larger root under the smaller one; a lost race continues from what the winner left
__device__ __forceinline__ void uf_unite(volatile uint32_t* parent, uint32_t a, uint32_t b) {
    while (true) {
        a = uf_find(parent, a);
        b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { const uint32_t t = a; a = b; b = t; }
        const uint32_t old = atomicMin((uint32_t*)&parent[a], b);
        if (old == a) return;
        a = old;
    }
}
// does point e touch (4-neighbourhood) a stone whose group root is one of the first m roots packed in r01 / r23
// (16 bits each)?  Not inlined: the kernel must stay small enough for the instruction cache (see the header of the kernel).
__device__ __noinline__ bool touches(const uint16_t* lab, const uint16_t* s_xy, int size, int e, uint32_t r01, uint32_t r23, int m) {
    const int x = s_xy[e] & 255, y = s_xy[e] >> 8;
    const uint32_t l0 = x > 0 ? lab[e - size] : 0xFFFFu, l1 = x < size - 1 ? lab[e + size] : 0xFFFFu;
    const uint32_t l2 = y > 0 ? lab[e - 1] : 0xFFFFu, l3 = y < size - 1 ? lab[e + 1] : 0xFFFFu;
    bool hit = false;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint32_t r = ((i < 2 ? r01 : r23) >> (16 * (i & 1))) & 0xFFFFu;
        if (i < m) hit |= (l0 == r) | (l1 == r) | (l2 == r) | (l3 == r);
    }
    return hit;
}
// one output byte: plane descriptor d at point p
__device__ __forceinline__ uint32_t point_value(const uint8_t* codes, uint32_t d, int p, uint32_t black) {
    const uint32_t kind = d >> 24;
    if (kind) return kind == 2 ? 1u : (kind == 3 ? (black & 1u) : 0u);
    return (((uint32_t)codes[(d & 0xFFFu) + p] >> ((d >> 12) & 0xFu)) & 0xFu) == ((d >> 16) & 0xFu) ? 1u : 0u;
}
// four consecutive output bytes of a plane starting at point p (bytes past the plane's end are garbage: the caller masks)
__device__ __forceinline__ uint32_t plane_word(const uint8_t* codes, uint32_t d, int p, uint32_t black) {
    const uint32_t kind = d >> 24;
    if (kind) return kind == 2 ? 0x01010101u : (kind == 3 ? black : 0u);
    const uint32_t* ca = reinterpret_cast<const uint32_t*>(codes + (d & 0xFFFu) + (p & ~3));
    const uint32_t cv = __funnelshift_r(ca[0], ca[1], (p & 3) * 8);
    const uint32_t nib = (cv >> ((d >> 12) & 0xFu)) & 0x0F0F0F0Fu;
    return __vcmpeq4(nib, ((d >> 16) & 0xFu) * 0x01010101u) & 0x01010101u;
}

__global__ void __launch_bounds__(WARPS * 32, 3) encode_kernel(
    const int8_t* __restrict__ g_board, const uint8_t* __restrict__ g_age, const int8_t* __restrict__ g_player,
    const int16_t* __restrict__ g_ko, int npos, int size, PlaneTable tab, uint8_t* __restrict__ g_out, LadderWork work) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_tab[MAX_PLANES];
    __shared__ uint16_t s_xy[COLPAD];  // x | y << 8 of every point: replaces p / size in the hot loops
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpMem& W = reinterpret_cast<WarpMem*>(smem_raw)[warp];
    const int n = size * size;
    const int F = tab.n_planes;
    for (int i = threadIdx.x; i < F; i += blockDim.x) s_tab[i] = plane_desc(tab.arr[i], tab.val[i]);
    for (int i = threadIdx.x; i < COLPAD; i += blockDim.x) { int x = i / size; s_xy[i] = (uint16_t)(x | ((i - x * size) << 8)); }
    __syncthreads();
    volatile uint32_t* parent = W.off;
    const float inv_n = 1.0f / (float)n;
    const int nwarps = blockDim.x >> 5;

    for (int pos = blockIdx.x * nwarps + warp; pos < npos; pos += gridDim.x * nwarps) {
        const int8_t* board = g_board + (size_t)pos * n;
        const uint8_t* age = g_age + (size_t)pos * n;
        const int me = g_player[pos];
        const int ko = g_ko[pos];
#ifdef AG_ENC_PROFILE
        long long t_phase = clock64();
#endif
        // ---- 1. load; board / turns_since codes; the position's bit boards for the ladder read-out
        // The EMPTY points are listed at the end of the liberty pool (its front is filled in step 3): steps 3 and 4 then run on
        // full warps of empty points instead of on whatever mix of stones and empty points 32 consecutive points hold.
        uint32_t my_blk = 0, my_wht = 0;
        int n_empty = 0;
        const uint32_t lt = (1u << lane) - 1u;
#pragma unroll 1
        for (int k = 0; k < 12; k++) {
            const int p = k * 32 + lane;
            int c = 0, a = 255;
            if (p < n) { c = board[p]; a = age[p]; }
            W.col[p] = (int8_t)c;
            const uint32_t bb = __ballot_sync(0xffffffffu, c > 0), bw = __ballot_sync(0xffffffffu, c < 0);
            const uint32_t be = __ballot_sync(0xffffffffu, p < n && c == 0);
            if (p < n && c == 0) W.pool[POOL - 1 - (n_empty + __popc(be & lt))] = (uint16_t)p;
            n_empty += __popc(be);
            if (lane == k) { my_blk = bb; my_wht = bw; }
            if (p < NPAD) {
                W.cnt[p] = 0;
                const int cb = c == 0 ? 2 : (c == me ? 0 : 1);           // preprocessing.py:9-17
                const int ct = (c != 0 && a < 8) ? a : 0xF;              // preprocessing.py:20-43
                W.codes[0][p] = (uint8_t)(cb | ct << 4);
            }
        }
        if (work.bits && lane < 12) { work.bits[(size_t)pos * 24 + lane] = my_blk; work.bits[(size_t)pos * 24 + 12 + lane] = my_wht; }
        __syncwarp();
        ENC_T(0);

        // ---- 2. groups.  2a: every stone points at the left end of its horizontal run.  Run starts grow with the point index, so
        // an inclusive max-scan of "my index if I start a run, else 0" over the lanes, carried from block to block, finds them
        {
            uint32_t carry = 0;
#pragma unroll 1
            for (int k = 0; k < 12; k++) {
                const int p = k * 32 + lane;
                const int c = p < n ? W.col[p] : 0;
                const bool cont = c != 0 && (s_xy[p] >> 8) > 0 && W.col[p - 1] == c;
                uint32_t r = cont ? 0u : (uint32_t)p;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, r, d); if (lane >= d) r = max(r, v); }
                r = max(r, carry);
                if (p < NPAD) W.off[p] = r;
                carry = __shfl_sync(0xffffffffu, r, 31);
            }
        }
        __syncwarp();
        // 2b: one union per pair of vertically touching runs (the edge whose left neighbour does not already join them)
#pragma unroll 1
        for (int k = 0; k < 12; k++) {
            const int p = k * 32 + lane;
            if (p >= n) continue;
            const int c = W.col[p];
            if (c == 0 || W.col[p + size] != c || p + size >= n) continue;
            const int y = s_xy[p] >> 8;
            if (y > 0 && W.col[p - 1] == c && W.col[p + size - 1] == c) continue;
            uf_unite(parent, (uint32_t)p, (uint32_t)(p + size));
        }
        __syncwarp();
        // 2c: flatten to the root point
#pragma unroll 1
        for (int k = 0; k < 12; k++) {
            const int p = k * 32 + lane;
            if (p < NPAD) W.lab[p] = (p < n && W.col[p] != 0) ? (uint16_t)uf_find(parent, (uint32_t)p) : (uint16_t)0xFFFF;
        }
        __syncwarp();
        ENC_T(1);

        // ---- 3. group sizes and liberty counts; then the liberty lists of the side to move's groups
#pragma unroll 1
        for (int k = 0; k < 12; k++) {
            const int p = k * 32 + lane;
            if (p < n && W.col[p] != 0) atomicAdd(&W.cnt[W.lab[p]], 1u << 16);
        }
        const int e_iters = (n_empty + 31) >> 5;
#pragma unroll 1
        for (int k = 0; k < e_iters; k++) {
            const int slot = k * 32 + lane;
            if (slot >= n_empty) continue;
            const int p = W.pool[POOL - 1 - slot];
            const int x = s_xy[p] & 255, y = s_xy[p] >> 8;
            const uint32_t r0 = x > 0 ? W.lab[p - size] : 0xFFFFu, r1 = x < size - 1 ? W.lab[p + size] : 0xFFFFu;
            const uint32_t r2 = y > 0 ? W.lab[p - 1] : 0xFFFFu, r3 = y < size - 1 ? W.lab[p + 1] : 0xFFFFu;
            if (r0 != 0xFFFFu) atomicAdd(&W.cnt[r0], 1u);
            if (r1 != 0xFFFFu && r1 != r0) atomicAdd(&W.cnt[r1], 1u);
            if (r2 != 0xFFFFu && r2 != r0 && r2 != r1) atomicAdd(&W.cnt[r2], 1u);
            if (r3 != 0xFFFFu && r3 != r0 && r3 != r1 && r3 != r2) atomicAdd(&W.cnt[r3], 1u);
        }
        __syncwarp();
        uint32_t pool_used = 0;
        {   // pool offsets: exclusive prefix sum of min(liberties, 8) over the root points of own groups
            uint32_t base = 0;
#pragma unroll 1
            for (int k = 0; k < 12; k++) {
                const int p = k * 32 + lane;
                uint32_t m = 0;
                if (p < n && W.col[p] == me && W.lab[p] == p) m = min(W.cnt[p] & 0xFFFFu, 8u);
                uint32_t incl = m;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
                if (p < NPAD) W.off[p] = base + incl - m;
                base += __shfl_sync(0xffffffffu, incl, 31);
            }
            pool_used = base;
        }
        __syncwarp();
        // the list of empty points survives only if the liberty lists written now stay below it (always, on positions from play)
        const bool compact = pool_used + (uint32_t)n_empty <= (uint32_t)POOL;
        const int iters = compact ? e_iters : 12;
#pragma unroll 1
        for (int k = 0; k < iters; k++) {
            const int slot = k * 32 + lane;
            int p = slot;
            if (compact) p = slot < n_empty ? (int)W.pool[POOL - 1 - slot] : n;
            if (p >= n || W.col[p] != 0) continue;
            const int x = s_xy[p] & 255, y = s_xy[p] >> 8;
            uint32_t r[4];
            r[0] = (x > 0 && W.col[p - size] == me) ? W.lab[p - size] : 0xFFFFu;
            r[1] = (x < size - 1 && W.col[p + size] == me) ? W.lab[p + size] : 0xFFFFu;
            r[2] = (y > 0 && W.col[p - 1] == me) ? W.lab[p - 1] : 0xFFFFu;
            r[3] = (y < size - 1 && W.col[p + 1] == me) ? W.lab[p + 1] : 0xFFFFu;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                bool dup = r[j] == 0xFFFFu;
#pragma unroll
                for (int t = 0; t < j; t++) dup |= r[t] == r[j];
                if (dup) continue;
                const uint32_t o = atomicAdd(&W.off[r[j]], 1u << 16);
                const uint32_t idx = o >> 16, at = (o & 0xFFFFu) + idx;
                if (idx < 8 && at < POOL) W.pool[at] = (uint16_t)p;
            }
        }
        __syncwarp();
        ENC_T(2);

        // ---- 4. per-point codes.  4a: stones (liberties plane) and pad points; 4b: the empty points, by the list when it is intact
#pragma unroll 1
        for (int k = 0; k < 12; k++) {
            const int p = k * 32 + lane;
            const int c = p < n ? W.col[p] : 1;
            if (c != 0 && p < NPAD) {
                uint32_t c_libs = 0xF;
                if (p < n) {
                    const int l = W.cnt[W.lab[p]] & 0xFFFF;                   // preprocessing.py:46-61
                    if (l >= 1) c_libs = (uint32_t)(l >= 8 ? 7 : l - 1);
                }
                W.codes[1][p] = (uint8_t)(c_libs | 0xF0u);
                W.codes[2][p] = 0xFF;
                W.codes[3][p] = 0x0F;
            }
        }
        __syncwarp();
        // 4b: the empty points
#pragma unroll 1
        for (int k = 0; k < iters; k++) {
            const int slot = k * 32 + lane;
            int p = slot;
            if (compact) p = slot < n_empty ? (int)W.pool[POOL - 1 - slot] : -1;
            const bool valid = p >= 0 && p < n && W.col[p < 0 ? 0 : p] == 0;
            uint32_t c_cap = 0xF, c_sa = 0xF, c_la = 0xF, c_sens = 0xF, lad = 0;
            if (valid) {
                const int x = s_xy[p] & 255, y = s_xy[p] >> 8;
                int nb[4];                                                    // go.py:96 order; -1 = off the board
                nb[0] = x > 0 ? p - size : -1; nb[1] = x < size - 1 ? p + size : -1;
                nb[2] = y > 0 ? p - 1 : -1;    nb[3] = y < size - 1 ? p + 1 : -1;
                int cq[4];
                uint32_t rt[4], cn[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {                                 // all loads first: they are independent
                    const int q = nb[j] < 0 ? p : nb[j];
                    cq[j] = W.col[q];
                    rt[j] = W.lab[q];
                }
#pragma unroll
                for (int j = 0; j < 4; j++) cn[j] = (nb[j] >= 0 && cq[j] != 0) ? W.cnt[rt[j]] : 0;
                bool has_empty = false, ok = false, all_mine = true;
                uint32_t mcap = 0, mesc = 0;   // neighbours (go.py:96 order) that name a ladder prey group for the first time
                int cap = 0, gsz = 1, ko_n = 0;
                uint32_t ro[4] = {0xFFFEu, 0xFFFEu, 0xFFFEu, 0xFFFEu}, lo[4] = {0, 0, 0, 0};
                bool big = false;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (nb[j] < 0) continue;
                    if (cq[j] != me) all_mine = false;
                    if (cq[j] == 0) { has_empty = true; continue; }
                    const int libs = cn[j] & 0xFFFF;
                    bool dup = false;
#pragma unroll
                    for (int t = 0; t < j; t++) dup |= (nb[t] >= 0 && cq[t] != 0 && rt[t] == rt[j]);
                    if (cq[j] == me) {
                        if (libs > 1) ok = true;                              // go.py:192-195
                        if (libs == 1 && !dup) mesc |= 1u << j;
                        if (!dup) {
                            gsz += cn[j] >> 16;
                            if (libs > 8) big = true;                         // >= 9 liberties: |L| >= 8 for certain
#pragma unroll
                            for (int t = 0; t < 4; t++) if (t == ko_n) { ro[t] = rt[j]; lo[t] = (uint32_t)libs; }
                            ko_n++;
                        }
                    } else {
                        if (libs == 1) { ok = true; if (!dup) cap += cn[j] >> 16; }   // go.py:197-199, pp:79-86
                        if (libs == 2 && !dup) mcap |= 1u << j;
                    }
                }
                const bool legal = (p != ko) && (has_empty || ok);            // go.py:204-216
                if (legal) {
                    c_cap = (uint32_t)min(cap, 7);                            // preprocessing.py:87
                    // |L|, L = empty neighbours U liberties of adjacent own groups, minus p (pp:98-109, 131-142)
                    int cntL = 8;
                    if (!big) {
                        const uint32_t r01 = ro[0] | ro[1] << 16, r23 = ro[2] | ro[3] << 16;
                        const uint32_t l123 = lo[1] | lo[2] << 8 | lo[3] << 16;
                        cntL = ko_n > 0 ? (int)lo[0] - 1 : 0;
                        if (ko_n == 0) {
#pragma unroll
                            for (int j = 0; j < 4; j++) cntL += (nb[j] >= 0 && cq[j] == 0) ? 1 : 0;
                        } else {
                            uint32_t emp = 0;   // the (<= 3) empty neighbours, 9 bits each, 0x1FF = none
#pragma unroll
                            for (int j = 3; j >= 0; j--) if (nb[j] >= 0 && cq[j] == 0) emp = emp << 9 | (uint32_t)nb[j];
                            int nemp = 0;
#pragma unroll
                            for (int j = 0; j < 4; j++) nemp += (nb[j] >= 0 && cq[j] == 0) ? 1 : 0;
#pragma unroll 1
                            for (int t = 0; t < nemp; t++, emp >>= 9)
                                if (!touches(W.lab, s_xy, size, (int)(emp & 0x1FFu), r01, r23, 4)) cntL++;
#pragma unroll 1
                            for (int j = 1; j < ko_n; j++) {
                                const uint32_t rj = ((j < 2 ? r01 : r23) >> (16 * (j & 1))) & 0xFFFFu, lj = (l123 >> (8 * (j - 1))) & 0xFFu;
                                const uint32_t o = W.off[rj] & 0xFFFFu;
#pragma unroll 1
                                for (uint32_t t = 0; t < lj; t++) {
                                    const int e = W.pool[o + t];
                                    if (e != p && !touches(W.lab, s_xy, size, e, r01, r23, j)) cntL++;
                                }
                            }
                        }
                    }
                    if (cntL == 1) c_sa = (uint32_t)min(gsz - 1, 7);          // preprocessing.py:111-114
                    c_la = cntL == 0 ? 7u : (uint32_t)min(7, cntL - 1);       // preprocessing.py:143 (-1 wraps to plane 7)
                    // go.py:261-267: legal and not is_eye(p, current_player)
                    c_sens = 0;
                    if (tab.need_eye && all_mine && is_eye(W.col, size, p, me)) c_sens = 0xF;
                    if (tab.need_ladder) lad = mcap | mesc << 4;
                }
            }
            if (valid) {
                W.codes[1][p] = (uint8_t)(0xFu | c_cap << 4);   // no liberties plane on an empty point
                W.codes[2][p] = (uint8_t)(c_sa | c_la << 4);
                W.codes[3][p] = (uint8_t)c_sens;
            }
            // ---- 5. ladder candidates (moves touching an opponent group with two liberties / an own group in atari) -> work
            // list, with the prey neighbours the read-out starts from; everything else on the two planes is 0 for certain
            if (tab.need_ladder && work.entries) {
                __syncwarp();
                const uint32_t bc = __ballot_sync(0xffffffffu, (lad & 0xFu) != 0), be = __ballot_sync(0xffffffffu, (lad >> 4) != 0);
                if (bc | be) {   // two lists: capture queries from the front of the area, escape queries from its middle
                    unsigned int bcap = 0, besc = 0;
                    if (lane == 0) {
                        if (bc) bcap = atomicAdd(work.head, (unsigned int)__popc(bc));
                        if (be) besc = atomicAdd(work.head + 2, (unsigned int)__popc(be));
                    }
                    bcap = __shfl_sync(0xffffffffu, bcap, 0);
                    besc = __shfl_sync(0xffffffffu, besc, 0);
                    const uint32_t lt = (1u << lane) - 1u;
                    const uint32_t e = (uint32_t)pos << 16 | (uint32_t)p << 5;
                    if (lad & 0xFu) work.entries[bcap + __popc(bc & lt)] = e | (lad & 0xFu);
                    if (lad >> 4) work.entries[work.esc_base + besc + __popc(be & lt)] = e | 16u | (lad >> 4);
                }
            }
        }
        __syncwarp();
        ENC_T(3);

        ENC_T(4);

        // ---- 6. one-hot expansion, coalesced stores.  out[pos][plane][p] = (nibble code of the plane's family == its value).
        // A position's planes are one contiguous run of F*n bytes: a few head bytes up to the next
        // 16-byte boundary, 16-byte vectors, a few tail bytes.
        const uint8_t* codes = &W.codes[0][0];
        const int total = F * n;
        uint8_t* out = g_out + (size_t)pos * total;
        int lead = (int)((16 - (reinterpret_cast<uintptr_t>(out) & 15)) & 15);
        if (lead > total) lead = total;
        const int nvec = (total - lead) >> 4;
        const int tail0 = lead + (nvec << 4);
        const uint32_t black = me == 1 ? 0x01010101u : 0u;
        for (int i = lane; i < lead + (total - tail0); i += 32) {
            const int b = i < lead ? i : tail0 + (i - lead);
            const int plane = b / n, p = b - plane * n;
            out[b] = (uint8_t)point_value(codes, s_tab[plane], p, black);
        }
        for (int v = lane; v < nvec; v += 32) {
            const int i0 = lead + (v << 4);
            int plane = (int)((float)i0 * inv_n);
            int p = i0 - plane * n;
            if (p < 0) { plane--; p += n; }
            if (p >= n) { plane++; p -= n; }
            uint32_t wds[4];
#pragma unroll
            for (int w = 0; w < 4; w++) {
                // four points of this plane from p on; where the plane ends inside the word, the rest comes from the next
                // plane's first points (n >= 4: a word never spans more than two planes)
                uint32_t word = plane_word(codes, s_tab[plane], p, black);
                const int left = n - p;
                if (left < 4) {
                    const uint32_t nxt = plane + 1 < F ? plane_word(codes, s_tab[plane + 1], 0, black) : 0u;
                    word = (word & (0xFFFFFFFFu >> (8 * (4 - left)))) | nxt << (8 * left);
                    p = 4 - left; plane++;
                } else {
                    p += 4;
                    if (p == n) { p = 0; plane++; }
                }
                wds[w] = word;
            }
            *reinterpret_cast<uint4*>(out + i0) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
        }
        __syncwarp();
        ENC_T(5);
    }
}

// Ladder planes, second half.  encode_kernel has queued every candidate move (position, kind, point) and stored each
// position's stones as bit boards.  Warps take queries from the list one at a time (a global counter: the launch ends at
// most one query after the list runs dry, whatever the spread of query costs) and read each out with the whole warp
// (ladder_warp.cuh: the board's bit sets are spread over the lanes, so the search has one control path).  Only "yes"
// answers are written: encode_kernel has already written zeros on both planes.
constexpr int kResolveWarps = 8;
#ifndef AG_LADDER_PICK
#define AG_LADDER_PICK(idx) ((idx) < nesc ? work.entries[work.esc_base + nesc - 1u - (idx)] : work.entries[ncap - 1u - ((idx) - nesc)])
#endif

__global__ void __launch_bounds__(kResolveWarps * 32)
ladder_resolve_kernel(const int8_t* __restrict__ g_player, const int16_t* __restrict__ g_ko, int size, int n_planes,
                      int cap_plane, int esc_plane, uint8_t* __restrict__ g_out, LadderWork work) {
    __shared__ LadderFrame s_frames[kResolveWarps][kLadderDepth + 2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = size * size;
    WBoard B;
    B.m = wmasks_make(size, lane);
    B.size = size; B.lane = lane;
    const unsigned int ncap = work.head[0], nesc = work.head[2], total = ncap + nesc;
    unsigned int cur = 0xFFFFFFFFu;
    while (true) {
        unsigned int idx = 0;
        if (lane == 0) idx = atomicAdd(&work.head[1], 1u);
        idx = __shfl_sync(0xffffffffu, idx, 0);
        if (idx >= total) break;
        // escape queries first, each list from its END: the queries of the positions that took the encoder longest
        // come last in a list and are, on average, the long read-outs -- started first, they do not end up as the launch's tail
        const uint32_t e = AG_LADDER_PICK(idx);
        const unsigned int pos = e >> 16;           // position << 16 | point << 5 | kind << 4 | prey-neighbour mask
        const int kind = (e >> 4) & 1, a = (e >> 5) & 511;
        const int pl = kind ? esc_plane : cap_plane;
        if (pl < 0) continue;
        if (pos != cur) {   // wl_query leaves the board as it found it: consecutive queries of one position share the load
            B.blk = lane < 12 ? work.bits[(size_t)pos * 24 + lane] : 0u;
            B.wht = lane < 12 ? work.bits[(size_t)pos * 24 + 12 + lane] : 0u;
            cur = pos;
        }
        B.player = g_player[pos]; B.ko = g_ko[pos]; B.undo_top = 0;
        const bool yes = wl_query(B, s_frames[warp], kind, a, (int)(e & 15u));
        if (yes && lane == 0) g_out[((size_t)pos * n_planes + pl) * n + a] = 1;
        __syncwarp();
    }
}

// copies the resolved ladder plane into further planes of the same kind (a feature listed more than once)
__global__ void ladder_copy_plane_kernel(uint8_t* out, int npos, int n, int n_planes, int src, int dst) {
    size_t total = (size_t)npos * n;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t pos = i / n, p = i - pos * n;
        out[(pos * n_planes + dst) * n + p] = out[(pos * n_planes + src) * n + p];
    }
}

}  // namespace enc

size_t encode_smem_bytes() { return sizeof(enc::WarpMem) * enc::WARPS; }

#ifdef AG_ENC_PROFILE
extern "C" int ag_debug_enc_phases(unsigned long long* h_out, int reset) {
    cudaMemcpyFromSymbol(h_out, enc::g_enc_phase_cycles, sizeof(unsigned long long) * 8);
    if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(enc::g_enc_phase_cycles, z, sizeof(z)); }
    return 0;
}
#endif

// Work-list memory of the ladder read-out: worst case 2 queries per point, so 722 entries + 24 bit-board words per position.
// It belongs to the library, per device, allocated on the first encode that asks for ladder planes and grown (never
// shrunk) when a call brings more positions than any before it; calls in between allocate nothing.  Four slots are
// handed out in turn, so up to four encodes may be in flight on different streams of one device.
namespace {
constexpr int kWorkSlots = 4;
constexpr int kWorkChunk = 65536;          // positions per launch pair (190 MB of work list at most)
constexpr int kWorkEntriesPerPos = 722;
struct WorkSlot { unsigned char* mem; size_t cap_pos; };
struct WorkSlots { WorkSlot s[kWorkSlots]; };
PerDevice<WorkSlots> g_work;
std::atomic<unsigned> g_next_slot{0};

int work_get(int npos, enc::LadderWork* w) {
    WorkSlot& s = g_work.cur().s[g_next_slot.fetch_add(1) % kWorkSlots];
    if ((size_t)npos > s.cap_pos) {
        size_t cap = 1024;
        while (cap < (size_t)npos) cap <<= 1;
        if (s.mem) AG_CUDA(cudaFree(s.mem));   // waits for the kernels that may still read it
        s.mem = nullptr; s.cap_pos = 0;
        AG_CUDA(cudaMalloc(&s.mem, 256 + cap * (24 + kWorkEntriesPerPos) * sizeof(uint32_t)));
        s.cap_pos = cap;
    }
    w->head = reinterpret_cast<unsigned int*>(s.mem);
    w->bits = reinterpret_cast<uint32_t*>(s.mem + 256);
    w->entries = w->bits + s.cap_pos * 24;
    w->esc_base = (unsigned int)(s.cap_pos * (kWorkEntriesPerPos / 2));
    return AG_OK;
}
}  // namespace

int encode_launch(const int8_t* board, const uint8_t* age, const int8_t* player, const int16_t* ko, int npos, int size,
                  const enc::PlaneTable& tab, uint8_t* out, cudaStream_t stream) {
    static PerDevice<bool> configured_dev;
    bool& configured = configured_dev.cur();
    const size_t smem = encode_smem_bytes();
    if (!configured) {
        AG_CUDA(cudaFuncSetAttribute(enc::encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    int cap_plane = -1, esc_plane = -1;
    for (int j = 0; j < tab.n_planes; j++) {
        if (tab.arr[j] == enc::C_LADCAP && cap_plane < 0) cap_plane = j;
        if (tab.arr[j] == enc::C_LADESC && esc_plane < 0) esc_plane = j;
    }
    const int n = size * size;
    for (int start = 0; start < npos; start += kWorkChunk) {
        const int cnt = npos - start < kWorkChunk ? npos - start : kWorkChunk;
        enc::LadderWork work = {nullptr, nullptr, nullptr, 0};
        if (tab.need_ladder) {
            int rc = work_get(cnt, &work);
            if (rc) return rc;
            AG_CUDA(cudaMemsetAsync(work.head, 0, 4 * sizeof(unsigned int), stream));
        }
        // small batches: fewer warps per CTA so that every SM gets positions
        int warps = cnt / (sm_count() * 3);
        warps = warps < 1 ? 1 : (warps > enc::WARPS ? enc::WARPS : warps);
        const int blocks_needed = (cnt + warps - 1) / warps;
        const int max_blocks = sm_count() * 24;  // grid-stride beyond that
        const int grid = blocks_needed < max_blocks ? blocks_needed : max_blocks;
        uint8_t* o = out + (size_t)start * tab.n_planes * n;
        enc::encode_kernel<<<grid, warps * 32, sizeof(enc::WarpMem) * warps, stream>>>(
            board + (size_t)start * n, age + (size_t)start * n, player + start, ko + start, cnt, size, tab, o, work);
        AG_LAUNCH_CHECK("encode_kernel");
        if (!tab.need_ladder) continue;
        enc::ladder_resolve_kernel<<<sm_count() * 6, enc::kResolveWarps * 32, 0, stream>>>(player + start, ko + start, size, tab.n_planes,
                                                                                          cap_plane, esc_plane, o, work);
        AG_LAUNCH_CHECK("ladder_resolve_kernel");
    }
    if (tab.need_ladder) {
        for (int j = 0; j < tab.n_planes; j++) {  // repeated ladder features
            int src = tab.arr[j] == enc::C_LADCAP ? cap_plane : (tab.arr[j] == enc::C_LADESC ? esc_plane : -1);
            if (src < 0 || src == j) continue;
            enc::ladder_copy_plane_kernel<<<sm_count() * 4, 256, 0, stream>>>(out, npos, n, tab.n_planes, src, j);
            AG_LAUNCH_CHECK("ladder_copy_plane_kernel");
        }
    }
    return AG_OK;
}

}  // namespace ag

// 48-plane AlphaGo feature encoder: one warp per position, integer work in shared memory.
//
// Replaces the reference's Preprocess.state_to_tensor hot loop
// (AlphaGo/preprocessing/preprocessing.py:9-245 on top of AlphaGo/go.py:63-267) for batches of
// positions.  Derived rules state (groups, liberty sets) is a pure function of the board, so it
// is rebuilt per position by a warp-wide union-find instead of porting the reference's
// incremental Python sets (go.py:104-157).
//
// Per position (flat point index p = x*size+y, planes [f][x][y], AlphaGo/util.py:6-8):
//   1. board/age -> shared memory                      (coalesced byte loads)
//   2. union-find over same-colour edges               (atomicMin hooking + path walk)
//   3. per-group size / liberty count / first 8 liberties (shared-memory atomics)
//   4. per-point feature codes: liberties, legality (go.py:181-216), capture_size, the
//      liberty union for self_atari_size / liberties_after (preprocessing.py:91-144, incl. the
//      negative-index wrap of :143 and the "captures ignored" quirk), sensibleness via an
//      iterative restatement of the path-dependent is_eye recursion (go.py:230-259)
//   5. ladder planes (see ladder.cuh)
//   6. codes -> one-hot uint8 planes, 16-byte coalesced stores
//
// HBM-bound by construction (18,054 algorithmic bytes per position); in practice
// the integer work in steps 2-5 is what the kernel spends its time on.
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
__device__ bool is_eye(const int8_t* col, int size, int p0, int owner) {
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

struct WarpMem {
    uint32_t label[NPAD];         // union-find parent, then root per point
    uint32_t cnt[NPAD];           // by root: group size << 16 | liberty count
    uint16_t liblist[MAXN][8];    // by root: first 8 liberties
    uint32_t lanebits[LW][32];    // lane-private 361-bit scratch set, [word][lane] (conflict-free)
    uint32_t codes[N_CODE][CODE_STRIDE / 4];  // per-point plane index per feature family, 0xFF = none
    int8_t col[NPAD];
    uint32_t ncand;               // ladder candidates queued in liblist (reused once the codes are final)
};

__device__ __forceinline__ uint32_t find_root(volatile uint32_t* label, uint32_t p) {
    uint32_t r = label[p];
    while (true) {
        uint32_t n = label[r];
        if (n == r) return r;
        r = n;
    }
}

__global__ void __launch_bounds__(WARPS * 32) encode_kernel(
    const int8_t* __restrict__ g_board, const uint8_t* __restrict__ g_age, const int8_t* __restrict__ g_player,
    const int16_t* __restrict__ g_ko, int npos, int size, PlaneTable tab, uint8_t* __restrict__ g_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint8_t s_arr[MAX_PLANES], s_val[MAX_PLANES];
    __shared__ uint16_t s_xy[NPAD];  // x | y << 8 of every point: replaces p / size in the hot loops
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpMem& W = reinterpret_cast<WarpMem*>(smem_raw)[warp];
    const int n = size * size;
    const int F = tab.n_planes;
    for (int i = threadIdx.x; i < F; i += blockDim.x) { s_arr[i] = tab.arr[i]; s_val[i] = tab.val[i]; }
    for (int i = threadIdx.x; i < NPAD; i += blockDim.x) { int x = i / size; s_xy[i] = (uint16_t)(x | ((i - x * size) << 8)); }
    __syncthreads();
    uint8_t* const codes = reinterpret_cast<uint8_t*>(W.codes);  // [N_CODE][CODE_STRIDE]
#define CODE(a, p) codes[(a) * CODE_STRIDE + (p)]
    volatile uint32_t* label = W.label;

    for (int pos = blockIdx.x * WARPS + warp; pos < npos; pos += gridDim.x * WARPS) {
        const int8_t* board = g_board + (size_t)pos * n;
        const uint8_t* age = g_age + (size_t)pos * n;
        const int me = g_player[pos];
        const int ko = g_ko[pos];

        // ---- 1. load; board / turns_since / constant codes
        for (int p = lane; p < NPAD; p += 32) {
            int c = 0, a = 255;
            if (p < n) { c = board[p]; a = age[p]; }
            W.col[p] = (int8_t)c;
            label[p] = p;
            W.cnt[p] = 0;
            CODE(C_BOARD, p) = c == 0 ? 2 : (c == me ? 0 : 1);          // preprocessing.py:9-17
            CODE(C_TURNS, p) = (c != 0 && a < 8) ? (uint8_t)a : 0xFF;    // preprocessing.py:20-43
            CODE(C_CONST, p) = 0;
            CODE(C_COLOR, p) = me == 1 ? 0 : 0xFF;
            CODE(C_LADCAP, p) = 0xFF;
            CODE(C_LADESC, p) = 0xFF;
        }
        for (int w = 0; w < LW; w++) W.lanebits[w][lane] = 0;
        if (lane == 0) W.ncand = 0;
        __syncwarp();

        // ---- 2. union-find over same-colour edges (+x and +y direction of every stone)
        while (true) {
            bool changed = false;
            for (int p = lane; p < n; p += 32) {
                int c = W.col[p];
                if (c == 0) continue;
                const int x = s_xy[p] & 255, y = s_xy[p] >> 8;
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    int q = e == 0 ? p + 1 : p + size;
                    bool ok = e == 0 ? (y < size - 1) : (x < size - 1);
                    if (!ok || W.col[q] != c) continue;
                    uint32_t rp = find_root(label, p), rq = find_root(label, q);
                    if (rp != rq) {
                        uint32_t hi = max(rp, rq), lo = min(rp, rq);
                        atomicMin((uint32_t*)&label[hi], lo);
                        changed = true;
                    }
                }
            }
            __syncwarp();
            if (!__any_sync(0xffffffffu, changed)) break;
        }
        // flatten: concurrent walks stay correct, a store only ever shortens a path to the same root
        for (int p = lane; p < n; p += 32) label[p] = find_root(label, p);
        __syncwarp();

        // ---- 3. group sizes, liberty counts, first 8 liberties
        for (int p = lane; p < n; p += 32) {
            int c = W.col[p];
            if (c != 0) { atomicAdd(&W.cnt[label[p]], 1u << 16); continue; }
            const int x = s_xy[p] & 255, y = s_xy[p] >> 8;
            uint32_t r[4]; int k = 0;
            if (x > 0 && W.col[p - size] != 0) r[k++] = label[p - size];
            if (x < size - 1 && W.col[p + size] != 0) r[k++] = label[p + size];
            if (y > 0 && W.col[p - 1] != 0) r[k++] = label[p - 1];
            if (y < size - 1 && W.col[p + 1] != 0) r[k++] = label[p + 1];
            for (int j = 0; j < k; j++) {
                bool dup = false;
                for (int t = 0; t < j; t++) dup |= (r[t] == r[j]);
                if (dup) continue;
                uint32_t idx = atomicAdd(&W.cnt[r[j]], 1u) & 0xFFFFu;
                if (idx < 8) W.liblist[r[j]][idx] = (uint16_t)p;
            }
        }
        __syncwarp();

        // ---- 4. per-point codes
        for (int p = lane; p < NPAD; p += 32) {
            uint8_t c_libs = 0xFF, c_cap = 0xFF, c_sa = 0xFF, c_la = 0xFF, c_sens = 0xFF;
            int c = p < n ? W.col[p] : 1;
            if (p < n && c != 0) {
                int l = W.cnt[label[p]] & 0xFFFF;                         // preprocessing.py:46-61
                if (l >= 1) c_libs = (uint8_t)(l >= 8 ? 7 : l - 1);
            } else if (p < n) {
                const int x = s_xy[p] & 255, y = s_xy[p] >> 8;
                int nb[4], k = 0;                                        // go.py:96 order
                if (x > 0) nb[k++] = p - size;
                if (x < size - 1) nb[k++] = p + size;
                if (y > 0) nb[k++] = p - 1;
                if (y < size - 1) nb[k++] = p + 1;
                uint32_t roots[4];
                bool has_empty = false, ok = false, all_mine = true;
                int cap = 0;
                for (int j = 0; j < k; j++) {
                    int q = nb[j], cq = W.col[q];
                    roots[j] = 0xFFFFFFFFu;
                    if (cq != me) all_mine = false;
                    if (cq == 0) { has_empty = true; continue; }
                    uint32_t r = label[q], cn = W.cnt[r];
                    int libs = cn & 0xFFFF;
                    roots[j] = r;
                    bool dup = false;
                    for (int t = 0; t < j; t++) dup |= (roots[t] == r);
                    if (cq == me) { if (libs > 1) ok = true; }             // go.py:192-195
                    else if (libs == 1) { ok = true; if (!dup) cap += cn >> 16; }  // go.py:197-199, pp:79-86
                }
                bool legal = (p != ko) && (has_empty || ok);              // go.py:204-216
                if (legal) {
                    c_cap = (uint8_t)min(cap, 7);                         // preprocessing.py:87
                    // L = empty neighbours U liberties of adjacent own groups, minus p (pp:98-109, 131-142)
                    int gsz = 1, cntL = 0;
                    bool big = false, any_own = false;
                    for (int j = 0; j < k; j++) {
                        int q = nb[j], cq = W.col[q];
                        if (cq == 0) { W.lanebits[q >> 5][lane] |= 1u << (q & 31); continue; }
                        if (cq != me) continue;
                        bool dup = false;
                        for (int t = 0; t < j; t++) dup |= (roots[t] == roots[j]);
                        if (dup) continue;
                        any_own = true;
                        uint32_t cn = W.cnt[roots[j]];
                        int libs = cn & 0xFFFF;
                        gsz += cn >> 16;
                        if (libs > 8) { big = true; continue; }          // >= 9 liberties: |L| >= 8 for certain
                        for (int t = 0; t < libs; t++) {
                            int e = W.liblist[roots[j]][t];
                            W.lanebits[e >> 5][lane] |= 1u << (e & 31);
                        }
                    }
                    (void)any_own;
                    W.lanebits[p >> 5][lane] &= ~(1u << (p & 31));
#pragma unroll
                    for (int w = 0; w < LW; w++) { cntL += __popc(W.lanebits[w][lane]); W.lanebits[w][lane] = 0; }
                    if (big) cntL = max(cntL, 8);
                    if (cntL == 1) c_sa = (uint8_t)min(gsz - 1, 7);        // preprocessing.py:111-114
                    c_la = cntL == 0 ? 7 : (uint8_t)min(7, cntL - 1);     // preprocessing.py:143 (-1 wraps to plane 7)
                    // go.py:261-267: legal and not is_eye(p, current_player)
                    c_sens = 0;
                    if (tab.need_eye && all_mine && is_eye(W.col, size, p, me)) c_sens = 0xFF;
                }
            }
            CODE(C_LIBS, p) = c_libs;
            CODE(C_CAPTURE, p) = c_cap;
            CODE(C_SELFATARI, p) = c_sa;
            CODE(C_LIBSAFTER, p) = c_la;
            CODE(C_SENSIBLE, p) = c_sens;
        }
        __syncwarp();

        // ---- 5. ladder planes, first half: mark the moves that touch a prey group (an adjacent opponent group with
        // two liberties / an adjacent own group in atari).  Everything else is 0 for certain; the marked moves are
        // resolved by ladder_resolve_kernel, which runs at full occupancy with dynamic work distribution.
        if (tab.need_ladder) {
            for (int p = lane; p < n; p += 32) {
                if (CODE(C_CAPTURE, p) == 0xFF) continue;  // not a legal move
                const int x = s_xy[p] & 255, y = s_xy[p] >> 8;
                bool cap = false, esc = false;
                int nb[4], k = 0;
                if (x > 0) nb[k++] = p - size;
                if (x < size - 1) nb[k++] = p + size;
                if (y > 0) nb[k++] = p - 1;
                if (y < size - 1) nb[k++] = p + 1;
                for (int j = 0; j < k; j++) {
                    int cq = W.col[nb[j]];
                    if (cq == 0) continue;
                    int libs = W.cnt[label[nb[j]]] & 0xFFFF;
                    if (cq == -me && libs == 2) cap = true;
                    if (cq == me && libs == 1) esc = true;
                }
                if (cap) CODE(C_LADCAP, p) = 0;
                if (esc) CODE(C_LADESC, p) = 0;
            }
            __syncwarp();
        }

        // ---- 6. one-hot expansion, coalesced stores.  out[pos][plane][p] = (code[arr][p] == val).
        // A position's planes are one contiguous run of F*n bytes: a few head bytes up to the next
        // 16-byte boundary, 16-byte vectors, a few tail bytes.
        const int total = F * n;
        uint8_t* out = g_out + (size_t)pos * total;
        int lead = (int)((16 - (reinterpret_cast<uintptr_t>(out) & 15)) & 15);
        if (lead > total) lead = total;
        const int nvec = (total - lead) >> 4;
        const int tail0 = lead + (nvec << 4);
        for (int i = lane; i < lead + (total - tail0); i += 32) {
            int b = i < lead ? i : tail0 + (i - lead);
            int plane = b / n, p = b - plane * n;
            out[b] = CODE(s_arr[plane], p) == s_val[plane];
        }
        for (int v = lane; v < nvec; v += 32) {
            int i0 = lead + (v << 4);
            int plane = i0 / n, p = i0 - plane * n;
            uint32_t wds[4];
#pragma unroll
            for (int w = 0; w < 4; w++) {
                uint32_t word;
                if (p + 3 < n) {
                    const uint32_t* ca = W.codes[s_arr[plane]];
                    uint32_t lo = ca[p >> 2], hi = ca[(p >> 2) + 1];
                    uint32_t cv = __funnelshift_r(lo, hi, (p & 3) * 8);
                    word = __vcmpeq4(cv, s_val[plane] * 0x01010101u) & 0x01010101u;
                    p += 4;
                    if (p == n) { p = 0; plane++; }
                } else {
                    word = 0;
#pragma unroll
                    for (int b = 0; b < 4; b++) {
                        uint32_t bit = (plane < F) && (CODE(s_arr[plane], p) == s_val[plane]);
                        word |= bit << (8 * b);
                        if (++p == n) { p = 0; plane++; }
                    }
                }
                wds[w] = word;
            }
            *reinterpret_cast<uint4*>(out + i0) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
        }
        __syncwarp();
    }
}

#undef CODE

// Ladder planes, second half.  Input: the planes encode_kernel wrote, whose ladder planes hold 1 on every candidate
// move.  Each warp claims positions one at a time from a global counter, lists the position's candidates (ascending, by
// ballot) in shared memory and reads them out one after the other with the whole warp (ladder_warp.cuh: the board's bit
// sets are spread over the lanes, so the search has one control path); each 0/1 answer overwrites the candidate mark.
constexpr int kResolveWarps = 8;

__global__ void __launch_bounds__(kResolveWarps * 32)
ladder_resolve_kernel(const int8_t* __restrict__ g_board, const int8_t* __restrict__ g_player, const int16_t* __restrict__ g_ko,
                      int npos, int size, int n_planes, int cap_plane, int esc_plane, uint8_t* __restrict__ g_out,
                      unsigned int* __restrict__ counter, int slices) {
    __shared__ LadderFrame s_frames[kResolveWarps][kLadderDepth + 2];
    __shared__ uint16_t s_list[kResolveWarps][2 * MAXN];  // kind << 9 | point
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = size * size;
    WBoard B;
    B.m = wmasks_make(size, lane);
    B.size = size; B.lane = lane;
    while (true) {
        // work unit = (position, slice): a small batch is cut into `slices` shares of each position's candidates (by point),
        // so that one position with many deep read-outs does not become the whole launch's critical path
        unsigned int unit = 0;
        if (lane == 0) unit = atomicAdd(counter, 1u);
        unit = __shfl_sync(0xffffffffu, unit, 0);
        if (unit >= (unsigned)npos * (unsigned)slices) break;
        const unsigned int pos = unit / (unsigned)slices;
        const int slice = (int)(unit - pos * (unsigned)slices);
        uint8_t* planes = g_out + (size_t)pos * n_planes * n;
        int count = 0;
        for (int kind = 0; kind < 2; kind++) {
            const int pl = kind ? esc_plane : cap_plane;
            if (pl < 0) continue;
            for (int p0 = 0; p0 < n; p0 += 32) {
                const int p = p0 + lane;
                // a share = the candidates at points p with p % slices == slice: other warps are overwriting the marks of their
                // own shares with answers at the same time, so membership must not depend on the other marks
                const bool mark = p < n && (p % slices) == slice && planes[(size_t)pl * n + p] != 0;
                const unsigned bal = __ballot_sync(0xffffffffu, mark);
                if (mark) s_list[warp][count + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)(kind << 9 | p);
                count += __popc(bal);
            }
        }
        __syncwarp();
        if (count == 0) continue;
        wl_load(B, g_board + (size_t)pos * n, size, g_player[pos], g_ko[pos], lane);
        for (int i = 0; i < count; i++) {
            const int c = s_list[warp][i];
            const int kind = c >> 9, a = c & 0x1FF;
            const bool yes = wl_query(B, s_frames[warp], kind, a);
            if (lane == 0) planes[(size_t)(kind ? esc_plane : cap_plane) * n + a] = yes ? 1 : 0;
        }
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

// work counters of ladder_resolve_kernel: a small ring so that encodes in flight on different streams do not share one
static __device__ unsigned int g_ladder_counters[16];

size_t encode_smem_bytes() { return sizeof(enc::WarpMem) * enc::WARPS; }

int encode_launch(const int8_t* board, const uint8_t* age, const int8_t* player, const int16_t* ko, int npos, int size,
                  const enc::PlaneTable& tab, uint8_t* out, cudaStream_t stream) {
    static PerDevice<bool> configured_dev;
    bool& configured = configured_dev.cur();
    size_t smem = encode_smem_bytes();
    if (!configured) {
        AG_CUDA(cudaFuncSetAttribute(enc::encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    int blocks_needed = (npos + enc::WARPS - 1) / enc::WARPS;
    int max_blocks = sm_count() * 16;  // grid-stride beyond that
    int grid = blocks_needed < max_blocks ? blocks_needed : max_blocks;
    if (grid < 1) grid = 1;
    enc::encode_kernel<<<grid, enc::WARPS * 32, smem, stream>>>(board, age, player, ko, npos, size, tab, out);
    AG_LAUNCH_CHECK("encode_kernel");
    if (!tab.need_ladder) return AG_OK;
    int cap_plane = -1, esc_plane = -1;
    for (int j = 0; j < tab.n_planes; j++) {
        if (tab.arr[j] == enc::C_LADCAP && cap_plane < 0) cap_plane = j;
        if (tab.arr[j] == enc::C_LADESC && esc_plane < 0) esc_plane = j;
    }
    // the symbol's address differs per device; the ring position is shared by host threads
    static PerDevice<unsigned int*> counters_dev;
    static std::atomic<unsigned> next_counter{0};
    unsigned int*& counters = counters_dev.cur();
    if (!counters) AG_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&counters), g_ladder_counters));
    unsigned int* ctr = counters + (next_counter.fetch_add(1) & 15);
    AG_CUDA(cudaMemsetAsync(ctr, 0, sizeof(unsigned int), stream));
    // enough work units to fill the machine even for a handful of positions; and never whole positions: the launch ends with
    // the slowest unit, and one position can hold several deep read-outs
    int slices = 65536 / (npos > 0 ? npos : 1);
    slices = slices < 8 ? 8 : (slices > 16 ? 16 : slices);   // measured at 16384 positions: 4 -> 4.37 ms, 8 -> 4.19 ms, 16 -> 4.23 ms
    { static int forced = -1; if (forced < 0) { const char* e = getenv("AG_LADDER_SLICES"); forced = e ? atoi(e) : 0; } if (forced > 0) slices = forced; }
    int rgrid = sm_count() * 8;
    int need = (npos * slices + enc::kResolveWarps - 1) / enc::kResolveWarps;
    if (rgrid > need) rgrid = need;
    enc::ladder_resolve_kernel<<<rgrid, enc::kResolveWarps * 32, 0, stream>>>(board, player, ko, npos, size, tab.n_planes, cap_plane, esc_plane, out, ctr,
                                                                              slices);
    AG_LAUNCH_CHECK("ladder_resolve_kernel");
    for (int j = 0; j < tab.n_planes; j++) {  // repeated ladder features
        int src = tab.arr[j] == enc::C_LADCAP ? cap_plane : (tab.arr[j] == enc::C_LADESC ? esc_plane : -1);
        if (src < 0 || src == j) continue;
        enc::ladder_copy_plane_kernel<<<sm_count() * 4, 256, 0, stream>>>(out, npos, size * size, tab.n_planes, src, j);
        AG_LAUNCH_CHECK("ladder_copy_plane_kernel");
    }
    return AG_OK;
}

}  // namespace ag

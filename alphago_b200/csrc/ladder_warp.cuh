// Warp-cooperative ladder read-out: the search of ladder.cuh (same predicates, same candidate order, same node budget --
// the build-owned spec of oracle/ladder_spec.py / oracle/go_oracle.c; the reference's
// preprocessing.py:147-152 raise NotImplementedError) with the 361-bit sets spread over the warp instead of held by one
// lane: lane l owns word l (points 32l .. 32l+31) of every set, lanes 12..31 hold zeros.  A dilation is four shuffles and
// a dozen ALU operations for the whole board, counts are warp reductions, and every decision is taken from a warp-wide
// value, so all 32 lanes follow one control path: one query per WARP at a time.  (The lane-per-query version ran 2.2
// of 32 lanes on average and spent most of its time waiting for instructions of divergent paths.)
// ladder.cuh stays as the scalar statement of the same state machine: tests/native/ladder_host_check.cpp runs it on the
// CPU against the oracle; this file is checked on the GPU against the same oracle (tests/test_encoder_gpu.py).
#pragma once
#include <stdint.h>

#include "ladder.cuh"  // kLadderDepth, kLadderNodes, kNoPoint, LadderFrame, LP_*

namespace ag {
namespace enc {

constexpr unsigned kFullWarp = 0xffffffffu;

// this lane's word of: board points, column y = 0, column y = size-1; and the next lane's column-0 word / the previous lane's
// last-column word (so that one pair of shuffles serves all four shifts of a dilation)
struct WMasks { uint32_t all, col_first, col_last, col_first_next, col_last_prev; };

__device__ __forceinline__ WMasks wmasks_make(int size, int lane) {
    WMasks m = {0u, 0u, 0u, 0u, 0u};
    for (int b = 0; b < 32; b++) {
        const int p = lane * 32 + b;
        if (p < size * size) {
            m.all |= 1u << b;
            const int y = p % size;
            if (y == 0) m.col_first |= 1u << b;
            if (y == size - 1) m.col_last |= 1u << b;
        }
    }
    m.col_first_next = __shfl_down_sync(0xffffffffu, m.col_first, 1);
    m.col_last_prev = __shfl_up_sync(0xffffffffu, m.col_last, 1);
    if (lane == 31) m.col_first_next = 0;
    if (lane == 0) m.col_last_prev = 0;
    return m;
}

__device__ __forceinline__ uint32_t wb_point(int p, int lane) { return lane == (p >> 5) ? 1u << (p & 31) : 0u; }
__device__ __forceinline__ bool wb_get(uint32_t b, int p) { return (__shfl_sync(kFullWarp, b, p >> 5) >> (p & 31)) & 1u; }
__device__ __forceinline__ bool wb_any(uint32_t b) { return __any_sync(kFullWarp, b != 0); }
__device__ __forceinline__ int wb_count(uint32_t b) { return (int)__reduce_add_sync(kFullWarp, (unsigned)__popc(b)); }
// smallest set point strictly above `after` (after = -1 for the first), or -1
__device__ __forceinline__ int wb_next(uint32_t b, int after, int lane) {
    const int i = after + 1;
    uint32_t m = b;
    if (lane == (i >> 5)) m &= 0xFFFFFFFFu << (i & 31);
    else if (lane < (i >> 5)) m = 0;
    const unsigned cand = m ? (unsigned)(lane * 32 + __ffs(m) - 1) : 0xFFFFu;
    const unsigned best = __reduce_min_sync(kFullWarp, cand);
    return best == 0xFFFFu ? -1 : (int)best;
}
// whole-set shifts by 0 < s < 32 (towards higher / lower point indices)
__device__ __forceinline__ uint32_t wb_shl(uint32_t b, int s, int lane) {
    uint32_t prev = __shfl_up_sync(kFullWarp, b, 1);
    if (lane == 0) prev = 0;
    return (b << s) | (prev >> (32 - s));
}
__device__ __forceinline__ uint32_t wb_shr(uint32_t b, int s, int lane) {
    uint32_t next = __shfl_down_sync(kFullWarp, b, 1);
    if (lane == 31) next = 0;
    return (b >> s) | (next << (32 - s));
}

struct WBoard {
    uint32_t blk, wht;   // this lane's word of the stones
    WMasks m;
    int size, player, ko, lane;
    uint32_t undo_cap[kLadderDepth + 2];   // per move on the path: this lane's word of the captured stones
    int16_t undo_pt[kLadderDepth + 2], undo_ko[kLadderDepth + 2];
    uint32_t replies[kLadderDepth + 2];    // per frame: this lane's word of the current prey's reply set
    uint32_t pgrp[kLadderDepth + 2];       // per frame: this lane's word of the prey group after the frame's move
    uint32_t plib[kLadderDepth + 2];       // per frame: ... and of that group's liberties
    int undo_top;
};

__device__ __forceinline__ uint32_t wl_dilate(const WBoard& B, uint32_t g) {
    // rotate instead of shift-with-fill: lane 0's "previous" word is lane 31's, which is zero like every lane >= 12, and what
    // lanes >= 12 compute is masked away by m.all; every shift across the word boundary is one funnel shift
    const uint32_t prev = __shfl_sync(kFullWarp, g, (B.lane + 31) & 31), next = __shfl_sync(kFullWarp, g, (B.lane + 1) & 31);
    const uint32_t up = __funnelshift_r(g, next, B.size);                                                   // x - 1
    const uint32_t down = __funnelshift_l(prev, g, B.size);                                                 // x + 1
    const uint32_t left = __funnelshift_r(g & ~B.m.col_first, next & ~B.m.col_first_next, 1);               // y - 1 (not from column 0)
    const uint32_t right = __funnelshift_l(prev & ~B.m.col_last_prev, g & ~B.m.col_last, 1);                // y + 1 (not from the last column)
    return (up | down | left | right) & B.m.all;
}
// stones / points with at least two members of `e` among their four neighbours (bit-sliced count of the four shifted copies)
__device__ __forceinline__ uint32_t wl_two_neighbours_in(const WBoard& B, uint32_t e) {
    const uint32_t prev = __shfl_sync(kFullWarp, e, (B.lane + 31) & 31), next = __shfl_sync(kFullWarp, e, (B.lane + 1) & 31);
    const uint32_t up = __funnelshift_r(e, next, B.size);
    const uint32_t down = __funnelshift_l(prev, e, B.size);
    const uint32_t left = __funnelshift_r(e & ~B.m.col_first, next & ~B.m.col_first_next, 1);
    const uint32_t right = __funnelshift_l(prev & ~B.m.col_last_prev, e & ~B.m.col_last, 1);
    return ((up & down) | (left & right) | ((up | down) & (left | right))) & B.m.all;
}
__device__ __forceinline__ int wl_colour(const WBoard& B, int p) {
    const uint32_t bw = __shfl_sync(kFullWarp, B.blk, p >> 5), ww = __shfl_sync(kFullWarp, B.wht, p >> 5);
    return ((bw >> (p & 31)) & 1u) ? 1 : (((ww >> (p & 31)) & 1u) ? -1 : 0);
}
// connected group of `colour` containing point p (which must hold that colour)
__device__ __forceinline__ uint32_t wl_group(const WBoard& B, int p, int colour) {
    const uint32_t own = colour > 0 ? B.blk : B.wht;
    uint32_t g = wb_point(p, B.lane);
    while (true) {
        const uint32_t n = g | (wl_dilate(B, g) & own);
        if (!__any_sync(kFullWarp, n != g)) return g;
        g = n;
    }
}
__device__ __forceinline__ uint32_t wl_liberties(const WBoard& B, uint32_t group) { return wl_dilate(B, group) & ~B.blk & ~B.wht; }
// Grows the group of `colour` at p and counts its liberties, giving up as soon as more than `limit` of them are certain
// (most groups a ladder search looks at are large and far from capture: one or two dilations settle the question).
// Returns the liberty count when it is <= limit -- *group and *libs are then complete -- else limit + 1.
__device__ __forceinline__ int wl_group_libs(const WBoard& B, int p, int colour, int limit, uint32_t* group, uint32_t* libs) {
    const uint32_t own = colour > 0 ? B.blk : B.wht, stones = B.blk | B.wht;
    uint32_t g = wb_point(p, B.lane);
    while (true) {
        const uint32_t d = wl_dilate(B, g);
        const uint32_t l = d & ~stones;
        const int c = wb_count(l);
        const uint32_t n = g | (d & own);
        if (c > limit) { *group = n; *libs = l; return limit + 1; }
        if (!__any_sync(kFullWarp, n != g)) { *group = g; *libs = l; return c; }
        g = n;
    }
}

__device__ __forceinline__ bool wl_legal(const WBoard& B, int a) {  // go.py:181-216 for the side to move
    if (wl_colour(B, a) != 0 || a == B.ko) return false;
    const uint32_t nb = wl_dilate(B, wb_point(a, B.lane));
    if (wb_any(nb & ~B.blk & ~B.wht)) return true;  // an empty neighbour
    for (int q = wb_next(nb, -1, B.lane); q >= 0; q = wb_next(nb, q, B.lane)) {  // every neighbour holds a stone
        const int c = wl_colour(B, q);
        uint32_t g, l;
        const bool other = wl_group_libs(B, q, c, 1, &g, &l) > 1;  // a itself is one of its liberties
        if (c == B.player && other) return true;
        if (c == -B.player && !other) return true;
    }
    return false;
}

// go.py:295-338 without bookkeeping; a must be legal.  Pushes an undo record.
__device__ __forceinline__ void wl_play(WBoard& B, int a) {
    const int color = B.player, n = B.size * B.size, y = a % B.size;
    const int old_ko = B.ko;
    uint32_t captured = 0;
    B.ko = -1;
    if (color > 0) B.blk |= wb_point(a, B.lane); else B.wht |= wb_point(a, B.lane);
    // go.py:96 neighbour order: (x-1,y), (x+1,y), (x,y-1), (x,y+1)
    const int cand[4] = {a - B.size >= 0 ? a - B.size : -1, a + B.size < n ? a + B.size : -1, y > 0 ? a - 1 : -1, y < B.size - 1 ? a + 1 : -1};
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int q = cand[j];
        if (q < 0 || wl_colour(B, q) != -color) continue;
        uint32_t g, l;
        if (wl_group_libs(B, q, -color, 0, &g, &l) > 0) continue;
        if (color > 0) B.wht &= ~g; else B.blk &= ~g;
        captured |= g;
        if (wb_count(g) == 1) {  // go.py:323-331, evaluated inside the loop
            uint32_t me, ml;
            if (wl_group_libs(B, a, color, 1, &me, &ml) == 1 && wb_count(me) == 1) B.ko = q;
        }
    }
    B.undo_cap[B.undo_top] = captured;
    B.undo_pt[B.undo_top] = (int16_t)a;
    B.undo_ko[B.undo_top] = (int16_t)old_ko;
    B.undo_top++;
    B.player = -color;
}
__device__ __forceinline__ void wl_undo(WBoard& B) {
    B.undo_top--;
    B.player = -B.player;  // the side that played
    const uint32_t cap = B.undo_cap[B.undo_top];
    const uint32_t pt = wb_point(B.undo_pt[B.undo_top], B.lane);
    if (B.player > 0) { B.wht |= cap; B.blk &= ~pt; }
    else { B.blk |= cap; B.wht &= ~pt; }
    B.ko = B.undo_ko[B.undo_top];
}

// is (kind ? ladder_escape : ladder_capture)(a) true on board B for the side to move?  B is left unchanged.
// fr: kLadderDepth + 2 frames in shared memory, private to this warp (every lane writes the same values).
// prey_mask != 0: the caller (the encoder) knows that a is legal and which of its neighbours (go.py:96 order: x-1, x+1, y-1, y+1)
// name the prey groups -- exactly what LP_ENTER would find -- so the root frame starts past it, with its node already counted.
__device__ __forceinline__ bool wl_query(WBoard& B, LadderFrame* fr, int kind, int a, int prey_mask = 0) {
    int nodes = kLadderNodes, sp = 0;
    bool ret = false;
    fr[0].a = (uint16_t)a; fr[0].prey = kNoPoint; fr[0].kind = (uint8_t)kind; fr[0].phase = LP_ENTER;
    if (prey_mask) {
        const int nbr[4] = {a - B.size, a + B.size, a - 1, a + 1};
        int np = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) if (prey_mask & (1 << j)) fr[0].preys[np++] = (uint16_t)nbr[j];
        fr[0].nprey = (uint8_t)np; fr[0].iprey = 0;
        fr[0].phase = LP_NEXT_PREY;
        nodes--;
    }
    sp = 1;
    while (sp > 0) {
        LadderFrame& f = fr[sp - 1];
        const int depth = kLadderDepth - (sp - 1);
        bool done = false, value = false;
        const int phase = f.phase, fkind = f.kind, fa = f.a;
        if (phase == LP_ENTER) {
            if (!wl_legal(B, fa)) { done = true; value = false; }
            else if (depth <= 0 || nodes <= 0) { done = true; value = fkind == 0; }
            else {
                nodes--;
                int nprey = 0;
                if (f.prey != kNoPoint) f.preys[nprey++] = f.prey;
                else {
                    const int want_col = fkind == 0 ? -B.player : B.player;
                    const int want_libs = fkind == 0 ? 2 : 1;
                    const int n = B.size * B.size, y = fa % B.size;
                    const int cand[4] = {fa - B.size >= 0 ? fa - B.size : -1, fa + B.size < n ? fa + B.size : -1, y > 0 ? fa - 1 : -1,
                                         y < B.size - 1 ? fa + 1 : -1};
                    uint32_t seen = 0;
#pragma unroll
                    for (int j = 0; j < 4; j++) {  // each adjacent group once, named by the first neighbour that touches it
                        const int q = cand[j];
                        if (q < 0 || wl_colour(B, q) != want_col || wb_get(seen, q)) continue;
                        uint32_t g, l;
                        const int libs = wl_group_libs(B, q, want_col, want_libs, &g, &l);
                        if (libs <= want_libs) seen |= g;  // complete group (a group given up on early is no prey from any side)
                        if (libs == want_libs) f.preys[nprey++] = (uint16_t)q;
                    }
                }
                f.nprey = (uint8_t)nprey; f.iprey = 0;
                f.phase = LP_NEXT_PREY;
            }
        } else if (phase == LP_NEXT_PREY) {
            if (f.iprey >= f.nprey) { done = true; value = false; }
            else {
                const int pr = f.preys[f.iprey];
                wl_play(B, fa);
                f.last = kNoPoint;
                // The prey group after the move and its liberties.  Below the root the parent frame holds the group as it was
                // before this move (the board between a frame's move and its children's moves does not change): a hunter's move
                // leaves it as it is, the prey's own move adds the stone -- if it touches the group; a reply may also capture
                // hunter stones elsewhere -- and whatever it connects to: one or two dilations instead of a flood fill along the
                // whole ladder.
                const int prey_col = fkind == 0 ? B.player : -B.player;   // wl_play has passed the turn
                const uint32_t own = prey_col > 0 ? B.blk : B.wht;
                uint32_t g = sp >= 2 ? B.pgrp[sp - 2] : wb_point(pr, B.lane);
                uint32_t d;
                while (true) {
                    d = wl_dilate(B, g);
                    const uint32_t nn = g | (d & own);
                    if (!__any_sync(kFullWarp, nn != g)) break;
                    g = nn;
                }
                const uint32_t libset = d & ~B.blk & ~B.wht;
                B.pgrp[sp - 1] = g;
                B.plib[sp - 1] = libset;
                bool go_on = true;
                if (fkind == 1) {
                    const int libs = wb_count(libset);
                    if (libs >= 3) { wl_undo(B); done = true; value = true; go_on = false; }
                    else if (libs <= 1) { wl_undo(B); f.iprey++; go_on = false; }
                }
                if (go_on) f.phase = LP_NEXT_CHILD;
            }
        } else if (phase == LP_NEXT_CHILD) {
            const int pr = f.preys[f.iprey];
            const int after = f.last == kNoPoint ? -1 : (int)f.last;
            uint32_t replies;
            if (f.last == kNoPoint) {  // first child of this prey: work the reply set out once, keep it for the siblings
                replies = B.plib[sp - 1];
                if (fkind == 0) {
                    // + the single liberty of every hunter group that touches the prey and is in atari.  A group in atari has no
                    // stone with two empty neighbours: those are dropped before the per-group check
                    const int prey_col = wl_colour(B, pr);
                    const uint32_t empty = ~B.blk & ~B.wht & B.m.all;
                    uint32_t touching = wl_dilate(B, B.pgrp[sp - 1]) & (prey_col > 0 ? B.wht : B.blk) & ~wl_two_neighbours_in(B, empty);
                    for (int s = wb_next(touching, -1, B.lane); s >= 0; s = wb_next(touching, -1, B.lane)) {
                        uint32_t h, hl;
                        const int nl = wl_group_libs(B, s, -prey_col, 1, &h, &hl);  // h is partial when given up early: it still holds s
                        touching &= ~h;
                        if (nl == 1) replies |= hl;
                    }
                }
                B.replies[sp - 1] = replies;
            } else {
                replies = B.replies[sp - 1];
            }
            const int e = wb_next(replies, after, B.lane);
            if (e < 0) {  // no reply left: the capture holds / nothing catches the escaper
                wl_undo(B);
                done = true; value = true;
            } else {
                f.last = (uint16_t)e;
                f.phase = LP_WAIT_CHILD;
                LadderFrame& c = fr[sp++];
                c.a = (uint16_t)e; c.prey = (uint16_t)pr; c.kind = (uint8_t)(fkind ^ 1); c.phase = LP_ENTER;
            }
        } else {  // LP_WAIT_CHILD
            if (ret) { wl_undo(B); f.iprey++; f.phase = LP_NEXT_PREY; }
            else f.phase = LP_NEXT_CHILD;
        }
        __syncwarp();
        if (done) { ret = value; sp--; }
    }
    return ret;
}

// load a position (int8 board) into the warp's bitboards: 32 points per ballot
// (B.m, B.size and B.lane are per-launch constants: set once by the caller with wmasks_make)
__device__ __forceinline__ void wl_load(WBoard& B, const int8_t* board, int size, int player, int ko, int lane) {
    const int n = size * size;
    uint32_t bw = 0, ww = 0;
    for (int w = 0; w * 32 < n; w++) {
        const int p = w * 32 + lane;
        const int c = p < n ? board[p] : 0;
        const uint32_t b = __ballot_sync(kFullWarp, c > 0), wt = __ballot_sync(kFullWarp, c < 0);
        if (lane == w) { bw = b; ww = wt; }
    }
    B.blk = bw; B.wht = ww;
    B.player = player; B.ko = ko; B.undo_top = 0;
}

}  // namespace enc
}  // namespace ag

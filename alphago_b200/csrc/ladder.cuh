// Ladder read-out for the ladder_capture / ladder_escape planes (AlphaGo/preprocessing/preprocessing.py:
// 147-152 declare them; the reference leaves both as NotImplementedError, so the semantics are the build-owned
// spec of oracle/ladder_spec.py / oracle/go_oracle.c -- SURVEY Appendix A).
//
// One lane searches one candidate move.  The mutual recursion
//   capture(a, prey, d): a illegal -> 0; d <= 0 or budget spent -> 1; for each prey (adjacent opponent group with
//       exactly 2 liberties, or the given one): play a; replies = prey's liberties U liberties of hunter groups
//       adjacent to the prey that are in atari; if no reply escapes -> 1.  -> 0
//   escape(a, prey, d):  a illegal -> 0; d <= 0 or budget spent -> 0; for each prey (adjacent own group in atari, or
//       the given one): play a; >= 3 liberties -> 1; 1 liberty -> next prey; if neither liberty captures -> 1.  -> 0
// runs as an explicit frame stack with play/undo; candidates are visited in ascending flat index and the node budget
// counts predicate entries, so results equal the CPU spec exactly.
//
// The position is two 361-bit sets (black, white) held in registers: a group is grown by repeated dilation
// (shift by +-1 within rows, +-size across rows) masked with its colour, its liberties are one more dilation masked
// with the empty points, counts are popcounts.  No per-stone loops and no memory traffic in the hot primitives;
// only the undo log and the frame stack live in local memory.
#pragma once
#include <stdint.h>

namespace ag {
namespace enc {

constexpr int kLadderDepth = 80;    // OG_LADDER_DEPTH
constexpr int kLadderNodes = 160;   // OG_LADDER_NODES
constexpr int kNoPoint = 0x1FF;

struct Bits { uint32_t w[LW]; };  // 361-bit set, bit p = point p (x*size+y)

__device__ __forceinline__ bool bit_get(const Bits& b, int i) {
    uint32_t v = 0;
#pragma unroll
    for (int w = 0; w < LW; w++) if (w == (i >> 5)) v = b.w[w];  // no dynamic register indexing
    return (v >> (i & 31)) & 1;
}
__device__ __forceinline__ void bit_set(Bits& b, int i) {
#pragma unroll
    for (int w = 0; w < LW; w++) if (w == (i >> 5)) b.w[w] |= 1u << (i & 31);
}
__device__ __forceinline__ void bit_clr(Bits& b, int i) {
#pragma unroll
    for (int w = 0; w < LW; w++) if (w == (i >> 5)) b.w[w] &= ~(1u << (i & 31));
}
__device__ __forceinline__ Bits bits_zero() {
    Bits r;
#pragma unroll
    for (int i = 0; i < LW; i++) r.w[i] = 0;
    return r;
}
__device__ __forceinline__ Bits bits_and(const Bits& a, const Bits& b) {
    Bits r;
#pragma unroll
    for (int i = 0; i < LW; i++) r.w[i] = a.w[i] & b.w[i];
    return r;
}
__device__ __forceinline__ Bits bits_or(const Bits& a, const Bits& b) {
    Bits r;
#pragma unroll
    for (int i = 0; i < LW; i++) r.w[i] = a.w[i] | b.w[i];
    return r;
}
__device__ __forceinline__ Bits bits_andnot(const Bits& a, const Bits& b) {
    Bits r;
#pragma unroll
    for (int i = 0; i < LW; i++) r.w[i] = a.w[i] & ~b.w[i];
    return r;
}
__device__ __forceinline__ bool bits_any(const Bits& a) {
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < LW; i++) o |= a.w[i];
    return o != 0;
}
__device__ __forceinline__ bool bits_eq(const Bits& a, const Bits& b) {
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < LW; i++) o |= a.w[i] ^ b.w[i];
    return o == 0;
}
__device__ __forceinline__ int bits_count(const Bits& a) {
    int c = 0;
#pragma unroll
    for (int i = 0; i < LW; i++) c += __popc(a.w[i]);
    return c;
}
// smallest set bit strictly above `after` (after = -1 for the first), or -1
__device__ __forceinline__ int bits_next(const Bits& b, int after) {
    const int i = after + 1;
    int best = -1;
#pragma unroll
    for (int w = LW - 1; w >= 0; w--) {
        uint32_t m = b.w[w];
        if (w == (i >> 5)) m &= 0xFFFFFFFFu << (i & 31);
        else if (w < (i >> 5)) m = 0;
        if (m) best = (w << 5) + __ffs(m) - 1;
    }
    return best;
}
// whole-set shifts by a run-time amount 0 < s < 32
__device__ __forceinline__ Bits bits_shl(const Bits& a, int s) {
    Bits r;
#pragma unroll
    for (int i = LW - 1; i > 0; i--) r.w[i] = __funnelshift_l(a.w[i - 1], a.w[i], s);
    r.w[0] = a.w[0] << s;
    return r;
}
__device__ __forceinline__ Bits bits_shr(const Bits& a, int s) {
    Bits r;
#pragma unroll
    for (int i = 0; i < LW - 1; i++) r.w[i] = __funnelshift_r(a.w[i], a.w[i + 1], s);
    r.w[LW - 1] = a.w[LW - 1] >> s;
    return r;
}

// per-board-size constants: the points of the board, of its first column (y = 0) and last column (y = size-1)
struct BoardMasks { Bits all, col_first, col_last; };
__device__ __forceinline__ void board_masks_fill(BoardMasks& M, int size) {  // by one thread
    for (int i = 0; i < LW; i++) { M.all.w[i] = 0; M.col_first.w[i] = 0; M.col_last.w[i] = 0; }
    for (int p = 0; p < size * size; p++) {
        M.all.w[p >> 5] |= 1u << (p & 31);
        int y = p % size;
        if (y == 0) M.col_first.w[p >> 5] |= 1u << (p & 31);
        if (y == size - 1) M.col_last.w[p >> 5] |= 1u << (p & 31);
    }
}

struct LadderBoard {
    Bits blk, wht;          // stones
    const BoardMasks* m;
    int size, player, ko;
    uint32_t undo[(kLadderDepth + 2) * (LW + 2)];  // per move on the path: captured stones (LW words), point, old ko
    int undo_top;
};

// the 4-neighbourhood of a set
__device__ __forceinline__ Bits lb_dilate(const LadderBoard& B, const Bits& g) {
    Bits up = bits_shr(g, B.size), down = bits_shl(g, B.size);                 // x -+ 1
    Bits left = bits_shr(bits_andnot(g, B.m->col_first), 1);                     // y - 1 (not from column 0)
    Bits right = bits_shl(bits_andnot(g, B.m->col_last), 1);                     // y + 1 (not from the last column)
    return bits_and(bits_or(bits_or(up, down), bits_or(left, right)), B.m->all);
}
__device__ __forceinline__ int lb_colour(const LadderBoard& B, int p) { return bit_get(B.blk, p) ? 1 : (bit_get(B.wht, p) ? -1 : 0); }
__device__ __forceinline__ Bits lb_point(int p) { Bits b = bits_zero(); bit_set(b, p); return b; }

// connected group of `colour` containing point p (which must hold that colour)
__device__ __forceinline__ Bits lb_group(const LadderBoard& B, int p, int colour) {
    const Bits own = colour > 0 ? B.blk : B.wht;
    Bits g = lb_point(p);
    while (true) {
        Bits n = bits_or(g, bits_and(lb_dilate(B, g), own));
        if (bits_eq(n, g)) return g;
        g = n;
    }
}
__device__ __forceinline__ Bits lb_liberties(const LadderBoard& B, const Bits& group) {
    return bits_andnot(bits_andnot(lb_dilate(B, group), B.blk), B.wht);
}
// go.py:96 neighbour order of point a: (x-1,y), (x+1,y), (x,y-1), (x,y+1); -1 where off the board
__device__ __forceinline__ void lb_nbrs(const LadderBoard& B, int a, int out[4]) {
    const int y = a % B.size, n = B.size * B.size;
    out[0] = a - B.size >= 0 ? a - B.size : -1;
    out[1] = a + B.size < n ? a + B.size : -1;
    out[2] = y > 0 ? a - 1 : -1;
    out[3] = y < B.size - 1 ? a + 1 : -1;
}

__device__ inline bool lb_legal(const LadderBoard& B, int a) {  // go.py:181-216 for the side to move
    if (lb_colour(B, a) != 0 || a == B.ko) return false;
    const Bits nb = lb_dilate(B, lb_point(a));
    if (bits_any(bits_andnot(bits_andnot(nb, B.blk), B.wht))) return true;  // an empty neighbour
    for (int q = bits_next(nb, -1); q >= 0; q = bits_next(nb, q)) {        // every neighbour holds a stone
        const int c = lb_colour(B, q);
        const bool other = bits_count(lb_liberties(B, lb_group(B, q, c))) > 1;  // a itself is one of its liberties
        if (c == B.player && other) return true;
        if (c == -B.player && !other) return true;
    }
    return false;
}

// go.py:295-338 without bookkeeping; a must be legal.  Pushes an undo record.
__device__ inline void lb_play(LadderBoard& B, int a) {
    const int color = B.player;
    const int old_ko = B.ko;
    Bits captured = bits_zero();
    B.ko = -1;
    if (color > 0) bit_set(B.blk, a); else bit_set(B.wht, a);
    int cand[4];
    lb_nbrs(B, a, cand);
    for (int j = 0; j < 4; j++) {
        const int q = cand[j];
        if (q < 0 || lb_colour(B, q) != -color) continue;
        const Bits g = lb_group(B, q, -color);
        if (bits_any(lb_liberties(B, g))) continue;
        if (color > 0) B.wht = bits_andnot(B.wht, g); else B.blk = bits_andnot(B.blk, g);
        captured = bits_or(captured, g);
        if (bits_count(g) == 1) {  // go.py:323-331, evaluated inside the loop
            const Bits me = lb_group(B, a, color);
            if (bits_count(me) == 1 && bits_count(lb_liberties(B, me)) == 1) B.ko = q;
        }
    }
    uint32_t* u = B.undo + B.undo_top;
#pragma unroll
    for (int i = 0; i < LW; i++) u[i] = captured.w[i];
    u[LW] = (uint32_t)a;
    u[LW + 1] = (uint32_t)(old_ko + 1);
    B.undo_top += LW + 2;
    B.player = -color;
}
__device__ inline void lb_undo(LadderBoard& B) {
    B.undo_top -= LW + 2;
    const uint32_t* u = B.undo + B.undo_top;
    B.player = -B.player;  // the side that played
    Bits cap;
#pragma unroll
    for (int i = 0; i < LW; i++) cap.w[i] = u[i];
    if (B.player > 0) { B.wht = bits_or(B.wht, cap); bit_clr(B.blk, (int)u[LW]); }
    else { B.blk = bits_or(B.blk, cap); bit_clr(B.wht, (int)u[LW]); }
    B.ko = (int)u[LW + 1] - 1;
}

struct LadderFrame {
    uint16_t a;        // the move this predicate is about
    uint16_t prey;     // given prey point or kNoPoint
    uint16_t preys[4];
    uint16_t last;     // last reply tried (kNoPoint before the first)
    uint8_t kind;      // 0 capture, 1 escape
    uint8_t phase;
    uint8_t nprey, iprey;
};
enum { LP_ENTER, LP_NEXT_PREY, LP_NEXT_CHILD, LP_WAIT_CHILD };

// is (kind ? ladder_escape : ladder_capture)(a) true on board B for the side to move?  B is left unchanged.
__device__ inline bool ladder_query(LadderBoard& B, LadderFrame* fr, int kind, int a) {
    int nodes = kLadderNodes, sp = 0;
    bool ret = false;
    fr[0].a = (uint16_t)a; fr[0].prey = kNoPoint; fr[0].kind = (uint8_t)kind; fr[0].phase = LP_ENTER;
    sp = 1;
    while (sp > 0) {
        LadderFrame& f = fr[sp - 1];
        const int depth = kLadderDepth - (sp - 1);
        bool done = false, value = false;
        switch (f.phase) {
            case LP_ENTER: {
                if (!lb_legal(B, f.a)) { done = true; value = false; break; }
                if (depth <= 0 || nodes <= 0) { done = true; value = f.kind == 0; break; }
                nodes--;
                f.nprey = 0; f.iprey = 0;
                if (f.prey != kNoPoint) f.preys[f.nprey++] = f.prey;
                else {
                    const int want_col = f.kind == 0 ? -B.player : B.player;
                    const int want_libs = f.kind == 0 ? 2 : 1;
                    int cand[4];
                    lb_nbrs(B, f.a, cand);
                    Bits seen = bits_zero();
                    for (int j = 0; j < 4; j++) {  // each adjacent group once, named by the first neighbour that touches it
                        const int q = cand[j];
                        if (q < 0 || lb_colour(B, q) != want_col || bit_get(seen, q)) continue;
                        const Bits g = lb_group(B, q, want_col);
                        seen = bits_or(seen, g);
                        if (bits_count(lb_liberties(B, g)) == want_libs) f.preys[f.nprey++] = (uint16_t)q;
                    }
                }
                f.phase = LP_NEXT_PREY;
                break;
            }
            case LP_NEXT_PREY: {
                if (f.iprey >= f.nprey) { done = true; value = false; break; }
                lb_play(B, f.a);
                f.last = kNoPoint;
                if (f.kind == 1) {
                    const int pr = f.preys[f.iprey];
                    const int libs = bits_count(lb_liberties(B, lb_group(B, pr, lb_colour(B, pr))));
                    if (libs >= 3) { lb_undo(B); done = true; value = true; break; }
                    if (libs <= 1) { lb_undo(B); f.iprey++; break; }
                }
                f.phase = LP_NEXT_CHILD;
                break;
            }
            case LP_NEXT_CHILD: {
                const int pr = f.preys[f.iprey];
                const int after = f.last == kNoPoint ? -1 : (int)f.last;
                const int prey_col = lb_colour(B, pr);
                const Bits prey = lb_group(B, pr, prey_col);
                Bits replies = lb_liberties(B, prey);
                if (f.kind == 0) {
                    // + the single liberty of every hunter group that touches the prey and is in atari
                    Bits touching = bits_and(lb_dilate(B, prey), prey_col > 0 ? B.wht : B.blk);
                    for (int s = bits_next(touching, -1); s >= 0; s = bits_next(touching, -1)) {
                        const Bits h = lb_group(B, s, -prey_col);
                        touching = bits_andnot(touching, h);
                        const Bits hl = lb_liberties(B, h);
                        if (bits_count(hl) == 1) replies = bits_or(replies, hl);
                    }
                }
                const int e = bits_next(replies, after);
                if (e < 0) {  // no reply left: the capture holds / nothing catches the escaper
                    lb_undo(B);
                    done = true; value = true;
                    break;
                }
                f.last = (uint16_t)e;
                f.phase = LP_WAIT_CHILD;
                LadderFrame& c = fr[sp++];
                c.a = (uint16_t)e; c.prey = (uint16_t)pr; c.kind = f.kind ^ 1; c.phase = LP_ENTER;
                break;
            }
            case LP_WAIT_CHILD: {
                if (ret) { lb_undo(B); f.iprey++; f.phase = LP_NEXT_PREY; }
                else f.phase = LP_NEXT_CHILD;
                break;
            }
        }
        if (done) { ret = value; sp--; }
    }
    return ret;
}

// load a position (int8 board) into the bitboards
__device__ __forceinline__ void lb_load(LadderBoard& B, const int8_t* board, int size, int player, int ko, const BoardMasks* m) {
    uint32_t bw[LW], ww[LW];
    for (int i = 0; i < LW; i++) { bw[i] = 0; ww[i] = 0; }
    for (int p = 0; p < size * size; p++) {
        const int c = board[p];
        if (c > 0) bw[p >> 5] |= 1u << (p & 31);
        else if (c < 0) ww[p >> 5] |= 1u << (p & 31);
    }
    for (int i = 0; i < LW; i++) { B.blk.w[i] = bw[i]; B.wht.w[i] = ww[i]; }
    B.m = m; B.size = size; B.player = player; B.ko = ko; B.undo_top = 0;
}

}  // namespace enc
}  // namespace ag

// Ladder read-out for the ladder_capture / ladder_escape planes (AlphaGo/preprocessing/preprocessing.py:
// 147-152 declare them; the reference leaves both as NotImplementedError, so the semantics are the build-owned
// spec of oracle/ladder_spec.py / oracle/go_oracle.c -- SURVEY Appendix A).
//
// One lane searches one candidate move on a lane-private board in local memory.  The mutual recursion
//   capture(a, prey, d): a illegal -> 0; d <= 0 or budget spent -> 1; for each prey (adjacent opponent group with
//       exactly 2 liberties, or the given one): play a; replies = prey's liberties U liberties of hunter groups
//       adjacent to the prey that are in atari; if no reply escapes -> 1.  -> 0
//   escape(a, prey, d):  a illegal -> 0; d <= 0 or budget spent -> 0; for each prey (adjacent own group in atari, or
//       the given one): play a; >= 3 liberties -> 1; 1 liberty -> next prey; if neither liberty captures -> 1.  -> 0
// runs as an explicit frame stack with play/undo instead of board copies; candidates are visited in ascending
// flat index and the node budget counts predicate entries, so results equal the CPU spec exactly.
#pragma once
#include <stdint.h>

namespace ag {
namespace enc {

constexpr int kLadderDepth = 80;    // OG_LADDER_DEPTH
constexpr int kLadderNodes = 160;   // OG_LADDER_NODES
constexpr int kNoPoint = 0x1FF;

// on-board neighbours of every point in go.py:96 order, -1 padded; built once per CTA in shared memory
struct NeighbourTable { int16_t nb[NPAD][4]; };
__device__ __forceinline__ void neighbour_table_fill(NeighbourTable& T, int size, int tid, int nthreads) {
    for (int p = tid; p < NPAD; p += nthreads) {
        int x = p / size, y = p - x * size, k = 0;
        int16_t v[4] = {-1, -1, -1, -1};
        if (p < size * size) {
            if (x > 0) v[k++] = (int16_t)(p - size);
            if (x < size - 1) v[k++] = (int16_t)(p + size);
            if (y > 0) v[k++] = (int16_t)(p - 1);
            if (y < size - 1) v[k++] = (int16_t)(p + 1);
        }
        for (int j = 0; j < 4; j++) T.nb[p][j] = v[j];
    }
}

struct LadderBoard {
    const NeighbourTable* nt;
    int8_t bd[NPAD];
    int size, n, player, ko;
    uint32_t vis[LW], lib[LW];  // last flood fill: stones / liberties
    uint16_t stk[MAXN];
    uint16_t undo[MAXN + 2 * (kLadderDepth + 2) + 3];  // captured points + (count, old ko) per move on the path
    int undo_top;
};

__device__ __forceinline__ int lb_nbrs(const LadderBoard& B, int p, int out[4]) {  // go.py:96 order
    const int2 v = *reinterpret_cast<const int2*>(B.nt->nb[p]);  // four int16 in one 8-byte load
    out[0] = (int16_t)(v.x & 0xFFFF); out[1] = v.x >> 16; out[2] = (int16_t)(v.y & 0xFFFF); out[3] = v.y >> 16;
    return (out[0] >= 0) + (out[1] >= 0) + (out[2] >= 0) + (out[3] >= 0);
}
__device__ __forceinline__ bool bit_get(const uint32_t* b, int i) { return (b[i >> 5] >> (i & 31)) & 1; }
__device__ __forceinline__ void bit_set(uint32_t* b, int i) { b[i >> 5] |= 1u << (i & 31); }
__device__ __forceinline__ int bits_count(const uint32_t* b) {
    int c = 0;
#pragma unroll
    for (int w = 0; w < LW; w++) c += __popc(b[w]);
    return c;
}
// smallest set bit strictly above `after` (after = -1 for the first), or -1
__device__ __forceinline__ int bits_next(const uint32_t* b, int after) {
    int i = after + 1;
    int w = i >> 5;
    if (w >= LW) return -1;
    uint32_t m = b[w] & (0xFFFFFFFFu << (i & 31));
    while (true) {
        if (m) return (w << 5) + __ffs(m) - 1;
        if (++w >= LW) return -1;
        m = b[w];
    }
}

// flood fill of the group at p into B.vis / B.lib; returns its stone count
__device__ int lb_fill(LadderBoard& B, int p) {
#pragma unroll
    for (int w = 0; w < LW; w++) { B.vis[w] = 0; B.lib[w] = 0; }
    const int col = B.bd[p];
    int sp = 0, count = 0;
    B.stk[sp++] = (uint16_t)p;
    bit_set(B.vis, p);
    while (sp) {
        int c = B.stk[--sp];
        count++;
        int nb[4], k = lb_nbrs(B, c, nb);
        for (int j = 0; j < k; j++) {
            int q = nb[j], v = B.bd[q];
            if (v == 0) bit_set(B.lib, q);
            else if (v == col && !bit_get(B.vis, q)) { bit_set(B.vis, q); B.stk[sp++] = (uint16_t)q; }
        }
    }
    return count;
}

__device__ bool lb_legal(LadderBoard& B, int a) {  // go.py:181-216 for the side to move
    if (B.bd[a] != 0 || a == B.ko) return false;
    int nb[4], k = lb_nbrs(B, a, nb);
    for (int j = 0; j < k; j++) if (B.bd[nb[j]] == 0) return true;
    for (int j = 0; j < k; j++) {
        int q = nb[j];
        lb_fill(B, q);
        bool other = bits_count(B.lib) > 1;
        if (B.bd[q] == B.player && other) return true;
        if (B.bd[q] == -B.player && !other) return true;
    }
    return false;
}

// go.py:295-338 without bookkeeping; a must be legal.  Pushes an undo record: captured points..., count, old ko.
__device__ void lb_play(LadderBoard& B, int a) {
    const int color = B.player;
    const int old_ko = B.ko;
    int captured = 0;
    B.ko = -1;
    B.bd[a] = (int8_t)color;
    int nb[4], k = lb_nbrs(B, a, nb);
    for (int j = 0; j < k; j++) {
        int q = nb[j];
        if (B.bd[q] != -color) continue;
        int cnt = lb_fill(B, q);
        if (bits_count(B.lib) != 0) continue;
        for (int t = bits_next(B.vis, -1); t >= 0; t = bits_next(B.vis, t)) {
            B.bd[t] = 0;
            B.undo[B.undo_top++] = (uint16_t)t;
        }
        captured += cnt;
        if (cnt == 1) {  // go.py:323-331, inside the loop
            int mine = lb_fill(B, a);
            if (bits_count(B.lib) == 1 && mine == 1) B.ko = q;
        }
    }
    B.undo[B.undo_top++] = (uint16_t)captured;
    B.undo[B.undo_top++] = (uint16_t)(old_ko + 1);
    B.player = -color;
}
__device__ void lb_undo(LadderBoard& B, int a) {
    B.player = -B.player;  // the side that played a
    B.ko = (int)B.undo[--B.undo_top] - 1;
    int captured = B.undo[--B.undo_top];
    for (int i = 0; i < captured; i++) B.bd[B.undo[--B.undo_top]] = (int8_t)(-B.player);
    B.bd[a] = 0;
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
__device__ bool ladder_query(LadderBoard& B, LadderFrame* fr, int kind, int a) {
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
                    int nb[4], k = lb_nbrs(B, f.a, nb);
                    bool dup[4] = {false, false, false, false};
                    for (int j = 0; j < k; j++) {
                        int q = nb[j];
                        if (B.bd[q] != want_col || dup[j]) continue;
                        lb_fill(B, q);
                        for (int t = j + 1; t < k; t++) if (bit_get(B.vis, nb[t])) dup[t] = true;
                        if (bits_count(B.lib) == want_libs) f.preys[f.nprey++] = (uint16_t)q;
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
                    lb_fill(B, f.preys[f.iprey]);
                    int libs = bits_count(B.lib);
                    if (libs >= 3) { lb_undo(B, f.a); done = true; value = true; break; }
                    if (libs <= 1) { lb_undo(B, f.a); f.iprey++; break; }
                }
                f.phase = LP_NEXT_CHILD;
                break;
            }
            case LP_NEXT_CHILD: {
                const int pr = f.preys[f.iprey];
                const int after = f.last == kNoPoint ? -1 : (int)f.last;
                int e = -1;
                if (f.kind == 1) {
                    lb_fill(B, pr);
                    e = bits_next(B.lib, after);
                } else {
                    // replies = prey liberties U liberties of adjacent hunter groups in atari: take the smallest
                    // point above `after` over all of those sets without materialising the union
                    lb_fill(B, pr);
                    e = bits_next(B.lib, after);
                    uint32_t stones[LW], seen[LW];  // prey stones; hunter stones whose group was already examined
#pragma unroll
                    for (int w = 0; w < LW; w++) { stones[w] = B.vis[w]; seen[w] = 0; }
                    const int hunter = -B.bd[pr];
                    for (int s = bits_next(stones, -1); s >= 0; s = bits_next(stones, s)) {
                        int nb[4], k = lb_nbrs(B, s, nb);
                        for (int j = 0; j < k; j++) {
                            int q = nb[j];
                            if (B.bd[q] != hunter || bit_get(seen, q)) continue;
                            lb_fill(B, q);  // each adjacent hunter group once
#pragma unroll
                            for (int w = 0; w < LW; w++) seen[w] |= B.vis[w];
                            if (bits_count(B.lib) != 1) continue;
                            int c = bits_next(B.lib, -1);
                            if (c > after && (e < 0 || c < e)) e = c;
                        }
                    }
                }
                if (e < 0) {  // no reply left: capture holds / nothing catches the escaper
                    lb_undo(B, f.a);
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
                if (ret) { lb_undo(B, f.a); f.iprey++; f.phase = LP_NEXT_PREY; }
                else f.phase = LP_NEXT_CHILD;
                break;
            }
        }
        if (done) { ret = value; sp--; }
    }
    return ret;
}

}  // namespace enc
}  // namespace ag

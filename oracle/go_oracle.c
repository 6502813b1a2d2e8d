/*
 * go_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * A plain-C restatement of the reference's Go rules engine and 48-plane feature
 * encoder, used only as the checker for the CUDA path (tests/, __graft_entry__.smoke(),
 * bench.py's cpu_baseline / --impl reference legs).  Nothing under alphago_b200/ may
 * call into this file.
 *
 * Citations are into the reference tree (not present on the GPU box):
 *   go  = AlphaGo/go.py
 *   pp  = AlphaGo/preprocessing/preprocessing.py
 *
 * Pinning: tests/test_oracle_*.py check this file against (i) the hand-coded 7x7
 * known-answer tests of the reference (tests/test_preprocessing.py:87-258,
 * tests/test_gamestate.py, tests/test_liberties.py) and (ii) golden planes produced by
 * importing the reference's own Preprocess in the build container
 * (oracle/gen_golden.py -> tests/golden/).  The two ladder planes have no reference
 * semantics (pp:147-152 raise NotImplementedError): for them this file *is* the spec
 * (see og_ladder_* below) and parity is "unpinned" against the reference.
 *
 * Conventions: points are (x,y), flat index x*size+y (AlphaGo/util.py:6-8); colours
 * BLACK=+1, WHITE=-1, EMPTY=0 (go:3-5); neighbour order (x-1,y),(x+1,y),(x,y-1),(x,y+1)
 * (go:96); diagonal order (x-1,y-1),(x+1,y+1),(x+1,y-1),(x-1,y+1) (go:102).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAXN 361
#define LW 12 /* 361 bits -> 12 u32 words */

typedef struct {
    int size, n;
    int8_t board[MAXN];
    int player;        /* go:18 */
    int ko;            /* flat idx or -1 (go:19) */
    int16_t *hist;     /* go:21; -1 == PASS_MOVE */
    int nhist, cap;
    int turns, prisoners_black, prisoners_white, passes_black, passes_white, end;
} og_state;

/* ---- derived structure: groups + liberties from a flood fill of the board.
 * The reference keeps these incrementally (go:104-157); they are a pure function of the
 * board (a captured group has no empty neighbour, so go:150-152 never leaves a stale
 * set) -- verified against the reference on random playouts by tests/test_oracle_ref.py. */
typedef struct {
    int16_t gid[MAXN];      /* group id per point, -1 for empty */
    int ng;
    int gsize[MAXN];        /* len(group_sets) */
    int glibs[MAXN];        /* len(liberty_sets) of a stone group */
    uint32_t glib[MAXN][LW];/* liberty bitset of each group */
} og_groups;

static int nbrs(int size, int p, int out[4]) { /* go:91-96 */
    int x = p / size, y = p % size, k = 0;
    if (x > 0) out[k++] = p - size;
    if (x < size - 1) out[k++] = p + size;
    if (y > 0) out[k++] = p - 1;
    if (y < size - 1) out[k++] = p + 1;
    return k;
}
static int diags(int size, int p, int out[4]) { /* go:98-102 */
    int x = p / size, y = p % size, k = 0;
    if (x > 0 && y > 0) out[k++] = p - size - 1;
    if (x < size - 1 && y < size - 1) out[k++] = p + size + 1;
    if (x < size - 1 && y > 0) out[k++] = p + size - 1;
    if (x > 0 && y < size - 1) out[k++] = p - size + 1;
    return k;
}
static inline void bs_set(uint32_t *b, int i) { b[i >> 5] |= 1u << (i & 31); }
static inline int bs_get(const uint32_t *b, int i) { return (b[i >> 5] >> (i & 31)) & 1; }
static inline int bs_count(const uint32_t *b) {
    int c = 0;
    for (int i = 0; i < LW; i++) c += __builtin_popcount(b[i]);
    return c;
}

static void analyse(int size, const int8_t *board, og_groups *g) {
    int n = size * size, stack[MAXN], nb[4];
    g->ng = 0;
    for (int i = 0; i < n; i++) g->gid[i] = -1;
    for (int s = 0; s < n; s++) {
        if (board[s] == 0 || g->gid[s] >= 0) continue;
        int id = g->ng++, sp = 0, col = board[s];
        memset(g->glib[id], 0, sizeof(g->glib[id]));
        g->gsize[id] = 0;
        g->gid[s] = id;
        stack[sp++] = s;
        while (sp) {
            int p = stack[--sp];
            g->gsize[id]++;
            int k = nbrs(size, p, nb);
            for (int j = 0; j < k; j++) {
                int q = nb[j];
                if (board[q] == 0) bs_set(g->glib[id], q);
                else if (board[q] == col && g->gid[q] < 0) { g->gid[q] = id; stack[sp++] = q; }
            }
        }
        g->glibs[id] = bs_count(g->glib[id]);
    }
}

/* ------------------------------------------------------------------ rules */

og_state *og_new(int size) { /* go:13-45 */
    og_state *s = (og_state *)calloc(1, sizeof(og_state));
    s->size = size; s->n = size * size; s->player = 1; s->ko = -1;
    s->cap = 64; s->hist = (int16_t *)malloc(sizeof(int16_t) * s->cap);
    return s;
}
void og_free(og_state *s) { if (s) { free(s->hist); free(s); } }
og_state *og_copy(const og_state *s) { /* go:159-179, but history is de-aliased */
    og_state *c = (og_state *)malloc(sizeof(og_state));
    *c = *s;
    c->hist = (int16_t *)malloc(sizeof(int16_t) * s->cap);
    memcpy(c->hist, s->hist, sizeof(int16_t) * s->nhist);
    return c;
}

/* go:181-202 on a board + analysis, for colour `me` */
static int suicide_b(int size, const int8_t *board, const og_groups *g, int p, int me) {
    int nb[4], k = nbrs(size, p, nb);
    for (int j = 0; j < k; j++) if (board[nb[j]] == 0) return 0; /* num_liberties_here > 0 */
    for (int j = 0; j < k; j++) {
        int q = nb[j];
        /* p is a liberty of every adjacent group, so "other liberties" <=> glibs > 1 */
        int other = g->glibs[g->gid[q]] > 1;
        if (board[q] == me && other) return 0;
        if (board[q] == -me && !other) return 0;
    }
    return 1;
}
/* go:204-216 (on-board, non-pass) */
static int legal_b(int size, const int8_t *board, const og_groups *g, int p, int me, int ko) {
    return board[p] == 0 && p != ko && !suicide_b(size, board, g, p, me);
}
static int eyeish_b(int size, const int8_t *board, int p, int owner) { /* go:218-228 */
    if (board[p] != 0) return 0;
    int nb[4], k = nbrs(size, p, nb);
    for (int j = 0; j < k; j++) if (board[nb[j]] != owner) return 0;
    return 1;
}
/* go:230-259 -- literal recursion incl. the path-dependent `stack` and the early exit */
static int eye_rec(int size, const int8_t *board, int p, int owner, int *stack, int *sp) {
    if (!eyeish_b(size, board, p, owner)) return 0;
    int nb[4], dg[4];
    int bad = 0, allow = (nbrs(size, p, nb) == 4) ? 1 : 0;
    int k = diags(size, p, dg);
    for (int j = 0; j < k; j++) {
        int d = dg[j];
        if (board[d] == -owner) bad++;
        else if (board[d] == 0) {
            int on = 0;
            for (int t = 0; t < *sp; t++) if (stack[t] == d) { on = 1; break; }
            if (!on) {
                stack[(*sp)++] = p;
                if (!eye_rec(size, board, d, owner, stack, sp)) bad++;
                (*sp)--;
            }
        }
        if (bad > allow) return 0;
    }
    return 1;
}
static int eye_b(int size, const int8_t *board, int p, int owner) {
    int stack[MAXN + 4], sp = 0;
    return eye_rec(size, board, p, owner, stack, &sp);
}

int og_is_legal(const og_state *s, int p) {
    if (p < 0) return 1; /* pass, go:210-211 */
    og_groups g; analyse(s->size, s->board, &g);
    return legal_b(s->size, s->board, &g, p, s->player, s->ko);
}
int og_is_eye(const og_state *s, int p, int owner) { return eye_b(s->size, s->board, p, owner); }
int og_is_eyeish(const og_state *s, int p, int owner) { return eyeish_b(s->size, s->board, p, owner); }

/* go:261-267; returns count, row-major order */
int og_legal_moves(const og_state *s, int include_eyes, int16_t *out) {
    og_groups g; analyse(s->size, s->board, &g);
    int c = 0;
    for (int p = 0; p < s->n; p++)
        if (legal_b(s->size, s->board, &g, p, s->player, s->ko) &&
            (include_eyes || !eye_b(s->size, s->board, p, s->player)))
            out[c++] = (int16_t)p;
    return c;
}

/* go:295-349.  color==0 means current player.  returns 0, or -1 for IllegalMove. */
int og_do_move(og_state *s, int p, int color) {
    if (color == 0) color = s->player;
    int reset = s->player;
    s->player = color;
    og_groups g;
    if (p >= 0) {
        analyse(s->size, s->board, &g);
        if (!legal_b(s->size, s->board, &g, p, color, s->ko)) { s->player = reset; return -1; }
    }
    s->ko = -1; /* go:306 */
    if (p >= 0) {
        s->board[p] = (int8_t)color;
        int nb[4], k = nbrs(s->size, p, nb);
        for (int j = 0; j < k; j++) { /* go:312-331: captures tested neighbour by neighbour */
            int q = nb[j];
            if (s->board[q] != -color) continue;
            analyse(s->size, s->board, &g);
            int id = g.gid[q];
            if (g.glibs[id] != 0) continue;
            int ncap = g.gsize[id];
            for (int t = 0; t < s->n; t++) if (g.gid[t] == id) s->board[t] = 0; /* go:138-157 */
            if (color == 1) s->prisoners_white += ncap; else s->prisoners_black += ncap;
            if (ncap == 1) { /* go:323-331: evaluated inside the loop; a later capture never clears it */
                analyse(s->size, s->board, &g);
                int me = g.gid[p];
                if (g.glibs[me] == 1 && g.gsize[me] == 1) s->ko = q;
            }
        }
    } else {
        if (color == 1) s->passes_black++; else s->passes_white++; /* go:333-336 */
    }
    s->player = -color;
    s->turns++;
    if (s->nhist == s->cap) { s->cap *= 2; s->hist = (int16_t *)realloc(s->hist, sizeof(int16_t) * s->cap); }
    s->hist[s->nhist++] = (int16_t)p;
    if (s->nhist > 1 && s->hist[s->nhist - 1] < 0 && s->hist[s->nhist - 2] < 0 && s->player == -1)
        s->end = 1; /* go:345-348 */
    return 0;
}

/* accessors for ctypes */
int og_size(const og_state *s) { return s->size; }
int og_player(const og_state *s) { return s->player; }
int og_ko(const og_state *s) { return s->ko; }
int og_nhist(const og_state *s) { return s->nhist; }
int og_hist(const og_state *s, int i) { return s->hist[i]; }
int og_is_end(const og_state *s) { return s->end; }
void og_set_player(og_state *s, int pl) { s->player = pl; }
void og_prisoners(const og_state *s, int *b, int *w) { *b = s->prisoners_black; *w = s->prisoners_white; }
void og_board(const og_state *s, int8_t *out) { memcpy(out, s->board, s->n); }
void og_liberty_counts(const og_state *s, int *out) { /* go:40-41: -1 on empty */
    og_groups g; analyse(s->size, s->board, &g);
    for (int p = 0; p < s->n; p++) out[p] = g.gid[p] < 0 ? -1 : g.glibs[g.gid[p]];
}
void og_group_sizes(const og_state *s, int *out) {
    og_groups g; analyse(s->size, s->board, &g);
    for (int p = 0; p < s->n; p++) out[p] = g.gid[p] < 0 ? 0 : g.gsize[g.gid[p]];
}

/* pp:20-43 literal walk -> per-point age in 0..7, 255 = no plane set.  `hist`/`nhist`
 * may describe moves never played on `board` (go:167 aliasing): those still advance depth. */
void og_age_from_history(int size, const int8_t *board, const int16_t *hist, int nhist, uint8_t *age) {
    int n = size * size, depth = 0;
    memset(age, 255, n);
    for (int i = nhist - 1; i >= 0; i--) {
        int m = hist[i];
        if (m >= 0 && board[m] != 0 && age[m] == 255) age[m] = (uint8_t)depth;
        if (depth < 7) depth++;
    }
}
void og_pack(const og_state *s, int8_t *board, uint8_t *age, int8_t *player, int16_t *ko) {
    memcpy(board, s->board, s->n);
    og_age_from_history(s->size, s->board, s->hist, s->nhist, age);
    *player = (int8_t)s->player; *ko = (int16_t)s->ko;
}

/* ------------------------------------------------------------------ ladders
 * BUILD-OWNED SPEC (no reference semantics: pp:147-152).  Mutual recursion on copies of
 * the position, following SURVEY.md Appendix A:
 *   capture(a, prey, d): a illegal -> 0; d<=0 or node budget spent -> 1.  Prey candidates:
 *     adjacent opponent groups with exactly 2 liberties (or the given prey).  Play a; the
 *     prey's replies are its remaining liberties plus the liberties of hunter groups
 *     adjacent to the prey that are now in atari; capture holds iff no reply escapes.
 *   escape(a, prey, d): a illegal -> 0; d<=0 or budget spent -> 0.  Prey candidates:
 *     adjacent own groups with exactly 1 liberty (or given).  Play a; >=3 liberties -> 1;
 *     1 liberty -> try next prey; 2 liberties -> escaped iff neither liberty is a capture.
 * Candidates are tried in ascending flat index; OG_LADDER_DEPTH / OG_LADDER_NODES are the
 * bounds shared with the CUDA kernel.  The node budget is per top-level query and counts
 * every capture()/escape() entry, so CPU and GPU (same DFS order) agree exactly. */
#define OG_LADDER_DEPTH 80
#define OG_LADDER_NODES 160

typedef struct { int size, n; int8_t board[MAXN]; int player, ko; } lpos;

static void lplay(lpos *s, int p) { /* go:295-338 without bookkeeping; p must be legal */
    int color = s->player;
    og_groups g;
    s->ko = -1;
    s->board[p] = (int8_t)color;
    int nb[4], k = nbrs(s->size, p, nb);
    for (int j = 0; j < k; j++) {
        int q = nb[j];
        if (s->board[q] != -color) continue;
        analyse(s->size, s->board, &g);
        int id = g.gid[q];
        if (g.glibs[id] != 0) continue;
        int ncap = g.gsize[id];
        for (int t = 0; t < s->n; t++) if (g.gid[t] == id) s->board[t] = 0;
        if (ncap == 1) {
            analyse(s->size, s->board, &g);
            int me = g.gid[p];
            if (g.glibs[me] == 1 && g.gsize[me] == 1) s->ko = q;
        }
    }
    s->player = -color;
}

static int lad_escape(const lpos *s, int a, int prey, int depth, int *nodes);

static int lad_capture(const lpos *s, int a, int prey, int depth, int *nodes) {
    og_groups g;
    analyse(s->size, s->board, &g);
    if (!legal_b(s->size, s->board, &g, a, s->player, s->ko)) return 0;
    if (depth <= 0 || *nodes <= 0) return 1;
    (*nodes)--;
    int hunter = s->player, cand[4], nc = 0;
    if (prey < 0) {
        int nb[4], k = nbrs(s->size, a, nb), seen[4], ns = 0;
        for (int j = 0; j < k; j++) {
            int q = nb[j];
            if (s->board[q] != -hunter) continue;
            int id = g.gid[q], dup = 0;
            for (int t = 0; t < ns; t++) if (seen[t] == id) dup = 1;
            if (dup) continue;
            seen[ns++] = id;
            if (g.glibs[id] == 2) cand[nc++] = q;
        }
    } else cand[nc++] = prey;
    for (int c = 0; c < nc; c++) {
        int pr = cand[c];
        lpos t = *s;
        lplay(&t, a);
        og_groups h;
        analyse(t.size, t.board, &h);
        int pid = h.gid[pr];
        if (pid < 0) continue; /* cannot happen: prey had 2 liberties */
        uint32_t esc[LW];
        memcpy(esc, h.glib[pid], sizeof(esc));
        for (int p = 0; p < t.n; p++) {
            if (h.gid[p] != pid) continue;
            int nb[4], k = nbrs(t.size, p, nb);
            for (int j = 0; j < k; j++) {
                int q = nb[j];
                if (t.board[q] == hunter && h.glibs[h.gid[q]] == 1)
                    for (int w = 0; w < LW; w++) esc[w] |= h.glib[h.gid[q]][w];
            }
        }
        int escaped = 0;
        for (int e = 0; e < t.n && !escaped; e++)
            if (bs_get(esc, e) && lad_escape(&t, e, pr, depth - 1, nodes)) escaped = 1;
        if (!escaped) return 1;
    }
    return 0;
}

static int lad_escape(const lpos *s, int a, int prey, int depth, int *nodes) {
    og_groups g;
    analyse(s->size, s->board, &g);
    if (!legal_b(s->size, s->board, &g, a, s->player, s->ko)) return 0;
    if (depth <= 0 || *nodes <= 0) return 0;
    (*nodes)--;
    int me = s->player, cand[4], nc = 0;
    if (prey < 0) {
        int nb[4], k = nbrs(s->size, a, nb), seen[4], ns = 0;
        for (int j = 0; j < k; j++) {
            int q = nb[j];
            if (s->board[q] != me) continue;
            int id = g.gid[q], dup = 0;
            for (int t = 0; t < ns; t++) if (seen[t] == id) dup = 1;
            if (dup) continue;
            seen[ns++] = id;
            if (g.glibs[id] == 1) cand[nc++] = q;
        }
    } else cand[nc++] = prey;
    for (int c = 0; c < nc; c++) {
        int pr = cand[c];
        lpos t = *s;
        lplay(&t, a);
        og_groups h;
        analyse(t.size, t.board, &h);
        int pid = h.gid[pr];
        if (pid < 0) continue;
        int libs = h.glibs[pid];
        if (libs >= 3) return 1;
        if (libs <= 1) continue;
        int caught = 0;
        for (int e = 0; e < t.n && !caught; e++)
            if (bs_get(h.glib[pid], e) && lad_capture(&t, e, pr, depth - 1, nodes)) caught = 1;
        if (!caught) return 1;
    }
    return 0;
}

static void lpos_from(lpos *l, int size, const int8_t *board, int player, int ko) {
    l->size = size; l->n = size * size; l->player = player; l->ko = ko;
    memcpy(l->board, board, l->n);
}
int og_is_ladder_capture(const og_state *s, int a) {
    lpos l; lpos_from(&l, s->size, s->board, s->player, s->ko);
    int nodes = OG_LADDER_NODES;
    return lad_capture(&l, a, -1, OG_LADDER_DEPTH, &nodes);
}
int og_is_ladder_escape(const og_state *s, int a) {
    lpos l; lpos_from(&l, s->size, s->board, s->player, s->ko);
    int nodes = OG_LADDER_NODES;
    return lad_escape(&l, a, -1, OG_LADDER_DEPTH, &nodes);
}

/* ------------------------------------------------------------------ encoder
 * Feature ids (order of pp:164-209 DEFAULT_FEATURES, plus the value net's colour plane). */
enum { F_BOARD, F_ONES, F_TURNS, F_LIBS, F_CAPTURE, F_SELFATARI, F_LIBSAFTER, F_LADCAP, F_LADESC,
       F_SENSIBLE, F_ZEROS, F_COLOR, F_COUNT };
static const int F_SIZE[F_COUNT] = {3, 1, 8, 8, 8, 8, 8, 1, 1, 1, 1, 1};
int og_feature_size(int f) { return (f >= 0 && f < F_COUNT) ? F_SIZE[f] : -1; }

/* Encode one position.  out is uint8 [planes][size*size], zero-filled here.  Returns the
 * number of planes written, or -1 on a bad feature id.  pp:238-245 concatenates in list order. */
int og_encode(const int8_t *board, const uint8_t *age, int player, int ko, int size,
              const int32_t *feats, int nfeat, uint8_t *out) {
    int n = size * size, planes = 0;
    for (int i = 0; i < nfeat; i++) {
        if (feats[i] < 0 || feats[i] >= F_COUNT) return -1;
        planes += F_SIZE[feats[i]];
    }
    memset(out, 0, (size_t)planes * n);
    og_groups g;
    analyse(size, board, &g);
    /* legal / sensible masks, once (go:261-267) */
    uint8_t legal[MAXN], sens[MAXN];
    int need_sens = 0;
    for (int i = 0; i < nfeat; i++) if (feats[i] == F_SENSIBLE) need_sens = 1;
    for (int p = 0; p < n; p++) {
        legal[p] = (uint8_t)legal_b(size, board, &g, p, player, ko);
        sens[p] = (need_sens && legal[p]) ? (uint8_t)!eye_b(size, board, p, player) : 0;
    }
    uint8_t *o = out;
    for (int i = 0; i < nfeat; i++) {
        int f = feats[i];
        switch (f) {
        case F_BOARD: /* pp:9-17 */
            for (int p = 0; p < n; p++) {
                o[0 * n + p] = board[p] == player;
                o[1 * n + p] = board[p] == -player;
                o[2 * n + p] = board[p] == 0;
            }
            break;
        case F_ONES: memset(o, 1, n); break; /* pp:169-172 */
        case F_ZEROS: break;                 /* pp:205-208 */
        case F_COLOR: memset(o, player == 1, n); break; /* build-owned */
        case F_TURNS: /* pp:20-43 via the packed age */
            for (int p = 0; p < n; p++) if (board[p] != 0 && age[p] < 8) o[age[p] * n + p] = 1;
            break;
        case F_LIBS: /* pp:46-61 */
            for (int p = 0; p < n; p++) if (g.gid[p] >= 0) {
                int l = g.glibs[g.gid[p]];
                if (l >= 1) o[(l >= 8 ? 7 : l - 1) * n + p] = 1;
            }
            break;
        case F_CAPTURE: /* pp:64-88 */
            for (int p = 0; p < n; p++) if (legal[p]) {
                int nb[4], k = nbrs(size, p, nb), seen[4], ns = 0, cap = 0;
                for (int j = 0; j < k; j++) {
                    int q = nb[j];
                    if (board[q] == 0) continue;
                    int id = g.gid[q], dup = 0;
                    for (int t = 0; t < ns; t++) if (seen[t] == id) dup = 1;
                    if (dup) continue;
                    seen[ns++] = id;
                    if (g.glibs[id] == 1 && board[q] != player) cap += g.gsize[id];
                }
                o[(cap > 7 ? 7 : cap) * n + p] = 1;
            }
            break;
        case F_SELFATARI: /* pp:91-115 */
        case F_LIBSAFTER: /* pp:118-144 */
            for (int p = 0; p < n; p++) if (legal[p]) {
                uint32_t L[LW] = {0};
                int nb[4], k = nbrs(size, p, nb), seen[4], ns = 0, gsz = 1;
                for (int j = 0; j < k; j++) if (board[nb[j]] == 0) bs_set(L, nb[j]); /* liberty_sets[x][y] of an empty point */
                for (int j = 0; j < k; j++) {
                    int q = nb[j];
                    if (board[q] != player) continue;
                    int id = g.gid[q], dup = 0;
                    for (int t = 0; t < ns; t++) if (seen[t] == id) dup = 1;
                    if (dup) continue;
                    seen[ns++] = id;
                    for (int w = 0; w < LW; w++) L[w] |= g.glib[id][w];
                    gsz += g.gsize[id];
                }
                L[p >> 5] &= ~(1u << (p & 31));
                int cnt = bs_count(L);
                if (f == F_SELFATARI) {
                    if (cnt == 1) o[(gsz - 1 > 7 ? 7 : gsz - 1) * n + p] = 1;
                } else {
                    /* pp:143: min(7, cnt-1); cnt==0 gives index -1 which numpy wraps to plane 7 */
                    int idx = cnt == 0 ? 7 : (cnt - 1 > 7 ? 7 : cnt - 1);
                    o[idx * n + p] = 1;
                }
            }
            break;
        case F_LADCAP:
        case F_LADESC: {
            lpos l; lpos_from(&l, size, board, player, ko);
            for (int p = 0; p < n; p++) if (legal[p]) {
                int nodes = OG_LADDER_NODES;
                o[p] = (uint8_t)(f == F_LADCAP ? lad_capture(&l, p, -1, OG_LADDER_DEPTH, &nodes)
                                               : lad_escape(&l, p, -1, OG_LADDER_DEPTH, &nodes));
            }
            break;
        }
        case F_SENSIBLE: /* pp:155-161 */
            for (int p = 0; p < n; p++) o[p] = sens[p];
            break;
        }
        o += (size_t)F_SIZE[f] * n;
    }
    return planes;
}

/* batch form: arrays are [n][size*size] etc.  nthreads>1 uses OpenMP when compiled with it. */
int og_encode_batch(const int8_t *board, const uint8_t *age, const int8_t *player, const int16_t *ko,
                    int npos, int size, const int32_t *feats, int nfeat, uint8_t *out, int nthreads) {
    int n = size * size, planes = 0, bad = 0;
    for (int i = 0; i < nfeat; i++) { if (feats[i] < 0 || feats[i] >= F_COUNT) return -1; planes += F_SIZE[feats[i]]; }
    (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int i = 0; i < npos; i++) {
        if (og_encode(board + (size_t)i * n, age + (size_t)i * n, player[i], ko[i], size, feats, nfeat,
                      out + (size_t)i * planes * n) < 0) bad = 1;
    }
    return bad ? -1 : planes;
}

/* ------------------------------------------------------------------ synthetic positions
 * random playouts, uniform over get_legal_moves(include_eyes=False); snapshots
 * packed as encoder inputs.  xorshift64* so fixtures are reproducible across platforms. */
static uint64_t rng_next(uint64_t *s) {
    uint64_t x = *s; x ^= x >> 12; x ^= x << 25; x ^= x >> 27; *s = x;
    return x * 2685821657736338717ULL;
}
/* Plays random games and records `per_game` snapshots per game at random move numbers in
 * [0, max_moves]; fills npos positions.  Returns npos. */
int og_playout_positions(uint64_t seed, int npos, int size, int max_moves, int per_game,
                         int8_t *board, uint8_t *age, int8_t *player, int16_t *ko) {
    int n = size * size, done = 0;
    uint64_t rs = seed * 0x9E3779B97F4A7C15ULL + 0x1234567ULL;
    if (!rs) rs = 1;
    int16_t moves[MAXN];
    while (done < npos) {
        int want[64], nw = per_game > 64 ? 64 : per_game;
        for (int i = 0; i < nw; i++) want[i] = (int)(rng_next(&rs) % (uint64_t)(max_moves + 1));
        og_state *s = og_new(size);
        for (int t = 0; t <= max_moves && done < npos; t++) {
            for (int i = 0; i < nw && done < npos; i++) if (want[i] == t) {
                og_pack(s, board + (size_t)done * n, age + (size_t)done * n, player + done, ko + done);
                done++;
            }
            int c = og_legal_moves(s, 0, moves);
            if (c == 0) { if (og_do_move(s, -1, 0)) break; if (s->end) break; continue; }
            og_do_move(s, moves[rng_next(&rs) % (uint64_t)c], 0);
        }
        og_free(s);
    }
    return done;
}

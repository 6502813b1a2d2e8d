// C ABI of libalphago_b200.so: library-level helpers, the encoder entry point and the host rules
// engine.  Declarations and reference citations live in include/alphago_b200.h.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <atomic>

#include "../../include/alphago_b200.h"
#include <stdlib.h>

#include "common.cuh"
#include "encode.cuh"
#include "host_go.hpp"

namespace ag {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return AG_ECUDA;
}
void count_launch(int n) { g_launches += n; }
bool pdl_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("AG_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}

bool chain_coop_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("AG_CHAIN_COOP"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}

int current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
    return dev < kMaxDevices ? dev : kMaxDevices - 1;
}

int sm_count() {
    static PerDevice<int> cache;
    int& n = cache.cur();
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, current_device()) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

static const int kFeaturePlanes[enc::F_COUNT] = {3, 1, 8, 8, 8, 8, 8, 1, 1, 1, 1, 1};

// feature list -> per-plane (code array, matching value)
static int build_plane_table(const int32_t* feats, int nf, enc::PlaneTable* t) {
    using namespace enc;
    memset(t, 0, sizeof(*t));
    int k = 0;
    for (int i = 0; i < nf; i++) {
        int f = feats[i];
        if (f < 0 || f >= F_COUNT) { set_error("unknown feature id %d", f); return AG_EINVAL; }
        int np = kFeaturePlanes[f];
        if (k + np > MAX_PLANES) { set_error("feature list exceeds %d planes", MAX_PLANES); return AG_EINVAL; }
        uint8_t arr = 0, base = 0;
        switch (f) {
            case F_BOARD: arr = C_BOARD; break;
            case F_ONES: arr = C_CONST; break;             // code 0 everywhere, val 0 -> ones
            case F_ZEROS: arr = C_CONST; base = 1; break;  // never matches -> zeros
            case F_TURNS: arr = C_TURNS; break;
            case F_LIBS: arr = C_LIBS; break;
            case F_CAPTURE: arr = C_CAPTURE; break;
            case F_SELFATARI: arr = C_SELFATARI; break;
            case F_LIBSAFTER: arr = C_LIBSAFTER; break;
            case F_LADCAP: arr = C_LADCAP; t->need_ladder = 1; break;
            case F_LADESC: arr = C_LADESC; t->need_ladder = 1; break;
            case F_SENSIBLE: arr = C_SENSIBLE; t->need_eye = 1; break;
            case F_COLOR: arr = C_COLOR; break;
        }
        for (int j = 0; j < np; j++) { t->arr[k] = arr; t->val[k] = (uint8_t)(base + j); k++; }
    }
    t->n_planes = k;
    return AG_OK;
}

}  // namespace ag

using ag::set_error;

struct ag_state { ag::HostBoard b; explicit ag_state(int size) : b(size) {} };

extern "C" {

const char* ag_last_error(void) { return ag::g_err; }
int ag_version(void) { return 1; }
long long ag_launch_count(void) { return ag::g_launches.load(); }
int ag_feature_planes(int f) { return (f >= 0 && f < ag::enc::F_COUNT) ? ag::kFeaturePlanes[f] : -1; }

int ag_encode(const int8_t* d_board, const uint8_t* d_age, const int8_t* d_player, const int16_t* d_ko, int n, int size,
              const int32_t* h_features, int n_features, uint8_t* d_out, void* stream) {
    if (n < 0 || size < 2 || size > 19) { set_error("ag_encode: bad n=%d or size=%d (2..19)", n, size); return AG_EINVAL; }
    if (n_features <= 0 || !h_features) { set_error("ag_encode: empty feature list"); return AG_EINVAL; }
    ag::enc::PlaneTable tab;
    int rc = ag::build_plane_table(h_features, n_features, &tab);
    if (rc) return rc;
    if (n == 0) return AG_OK;
    if (!d_board || !d_age || !d_player || !d_ko || !d_out) { set_error("ag_encode: null device pointer"); return AG_EINVAL; }
    return ag::encode_launch(d_board, d_age, d_player, d_ko, n, size, tab, d_out, (cudaStream_t)stream);
}

/* ---- host rules engine */
ag_state* ag_state_new(int size) {
    if (size < 2 || size > 19) { set_error("ag_state_new: size %d not in 2..19", size); return nullptr; }
    return new ag_state(size);
}
ag_state* ag_state_copy(const ag_state* s) { return s ? new ag_state(*s) : nullptr; }
void ag_state_free(ag_state* s) { delete s; }
int ag_state_do_move(ag_state* s, int point, int colour) {
    if (point >= s->b.n || point < -1 || (colour != 0 && colour != 1 && colour != -1)) { set_error("ag_state_do_move: bad argument"); return AG_EINVAL; }
    if (!s->b.play(point, colour)) { set_error("illegal move %d", point); return AG_ESTATE; }
    return AG_OK;
}
int ag_state_is_legal(ag_state* s, int point) { return s->b.legal(point, s->b.to_move) ? 1 : 0; }
int ag_state_is_eyeish(ag_state* s, int point, int owner) { return s->b.eyeish(point, owner) ? 1 : 0; }
int ag_state_is_eye(ag_state* s, int point, int owner) { return s->b.eye(point, owner) ? 1 : 0; }
int ag_state_legal_moves(ag_state* s, int include_eyes, int16_t* out) { return s->b.legal_moves(include_eyes != 0, out); }
int ag_state_winner(const ag_state* s, double komi) { return s->b.winner(komi); }
int ag_state_get(const ag_state* s, int field) {
    const ag::HostBoard& b = s->b;
    switch (field) {
        case 0: return b.size;
        case 1: return b.to_move;
        case 2: return b.ko;
        case 3: return (int)b.history.size();
        case 4: return b.finished ? 1 : 0;
        case 5: return b.turns;
        case 6: return b.prisoners_black;
        case 7: return b.prisoners_white;
        case 8: return b.passes_black;
        case 9: return b.passes_white;
    }
    return 0;
}
int ag_state_set_player(ag_state* s, int colour) {
    if (colour != 1 && colour != -1) { set_error("bad colour"); return AG_EINVAL; }
    s->b.to_move = colour;
    return AG_OK;
}
int ag_state_set_ko(ag_state* s, int point) { s->b.ko = point < 0 ? -1 : point; return AG_OK; }
void ag_state_board(const ag_state* s, int8_t* out) { memcpy(out, s->b.stone, s->b.n); }
void ag_state_history(const ag_state* s, int16_t* out) {
    if (!s->b.history.empty()) memcpy(out, s->b.history.data(), sizeof(int16_t) * s->b.history.size());
}
void ag_state_groups(ag_state* s, int16_t* gid, int16_t* stones, int16_t* libs) {
    ag::HostBoard& b = s->b;
    int16_t members[ag::HostBoard::kMax];
    for (int p = 0; p < b.n; p++) { gid[p] = -1; stones[p] = 0; libs[p] = -1; }
    int next = 0;
    for (int p = 0; p < b.n; p++) {
        if (b.stone[p] == 0 || gid[p] >= 0) continue;
        ag::HostBoard::Group g = b.probe(p, members);
        for (int t = 0; t < g.stones; t++) { gid[members[t]] = (int16_t)next; stones[members[t]] = (int16_t)g.stones; libs[members[t]] = (int16_t)g.liberties; }
        next++;
    }
}
void ag_age_from_history(int size, const int8_t* board, const int16_t* hist, int nhist, uint8_t* age) {
    ag::age_from_history(size, board, hist, nhist, age);
}
void ag_state_pack(const ag_state* s, int8_t* board, uint8_t* age, int8_t* player, int16_t* ko) {
    const ag::HostBoard& b = s->b;
    memcpy(board, b.stone, b.n);
    ag::age_from_history(b.size, b.stone, b.history.data(), (int)b.history.size(), age);
    *player = (int8_t)b.to_move;
    *ko = (int16_t)b.ko;
}

int ag_state_replay_pack(ag_state* s, int n_moves, const int16_t* points, const int8_t* colours, int8_t* boards, uint8_t* ages,
                         int8_t* players, int16_t* kos, int* n_packed) {
    int k = 0;
    const int pts = s->b.n;
    for (int i = 0; i < n_moves; i++) {
        const int point = points[i], colour = colours[i];
        if (point < -1 || (colour != 1 && colour != -1)) { *n_packed = k; set_error("ag_state_replay_pack: bad move %d", i); return AG_EINVAL; }
        if (point >= 0) {   // the position BEFORE a non-pass move is one training pair (game_converter.py:38-40)
            ag_state_pack(s, boards + (size_t)k * pts, ages + (size_t)k * pts, players + k, kos + k);
            k++;
        }
        if (point >= pts || !s->b.play(point, colour)) {   // off the board counts as illegal, as in GameState.do_move
            *n_packed = k;
            set_error("illegal move %d (entry %d)", point, i);
            return AG_ESTATE;
        }
    }
    *n_packed = k;
    return AG_OK;
}

int ag_playout_positions(uint64_t seed, int n, int size, int max_moves, int per_game, int8_t* board, uint8_t* age,
                         int8_t* player, int16_t* ko) {
    if (n < 0 || size < 2 || size > 19 || max_moves < 0 || per_game < 1) { set_error("ag_playout_positions: bad argument"); return AG_EINVAL; }
    const int pts = size * size;
    uint64_t rs = (seed + 1) * 0xD1B54A32D192ED03ULL;
    auto next = [&rs]() {  // splitmix64
        uint64_t z = (rs += 0x9E3779B97F4A7C15ULL);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        return z ^ (z >> 31);
    };
    if (per_game > 256) per_game = 256;
    std::vector<int> want(per_game);
    int16_t moves[ag::HostBoard::kMax];
    int done = 0;
    while (done < n) {
        for (int i = 0; i < per_game; i++) want[i] = (int)(next() % (uint64_t)(max_moves + 1));
        ag_state st(size);
        for (int t = 0; t <= max_moves && done < n; t++) {
            for (int i = 0; i < per_game && done < n; i++)
                if (want[i] == t) {
                    ag_state_pack(&st, board + (size_t)done * pts, age + (size_t)done * pts, player + done, ko + done);
                    done++;
                }
            int c = st.b.legal_moves(false, moves);
            st.b.play(c ? moves[next() % (uint64_t)c] : -1);
            if (st.b.finished) break;
        }
    }
    return AG_OK;
}

}  // extern "C"

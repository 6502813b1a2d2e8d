// Host-side Go rules engine of the product: backs alphago_b200.go.GameState and the synthetic
// random-playout position generator that feeds the encoder in bench.py.
//
// Mirrors the observable behaviour of the reference GameState (AlphaGo/go.py:9-349) including
// its quirks -- ko is set inside the per-neighbour capture loop (go.py:312-331), the eye test is
// the path-dependent recursion of go.py:230-259 -- but keeps no incremental group/liberty sets:
// group queries are stamp-marked flood fills over a flat board, which is all the bookkeeping a
// 19x19 board needs on a modern core.
#pragma once
#include <stdint.h>
#include <string.h>
#include <vector>

namespace ag {

class HostBoard {
  public:
    static constexpr int kMax = 361;
    int size, n;
    int8_t stone[kMax];
    int to_move = 1;   // +1 black, -1 white
    int ko = -1;       // flat index or -1
    std::vector<int16_t> history;  // -1 = pass
    int turns = 0, prisoners_black = 0, prisoners_white = 0, passes_black = 0, passes_white = 0;
    bool finished = false;

    explicit HostBoard(int sz) : size(sz), n(sz * sz) { memset(stone, 0, sizeof(stone)); memset(mark_, 0, sizeof(mark_)); }

    struct Group { int stones; int liberties; };

    // flood fill from p (must hold a stone).  Optionally collects members / liberty points.
    Group probe(int p, int16_t* members = nullptr, int16_t* libs = nullptr) {
        const int colour = stone[p];
        const uint32_t s = ++stamp_, l = ++stamp_;
        int head = 0, tail = 0, nlib = 0;
        queue_[tail++] = (int16_t)p;
        mark_[p] = s;
        while (head < tail) {
            int c = queue_[head++];
            int x = c / size, y = c - x * size;
            const int cand[4] = {x > 0 ? c - size : -1, x < size - 1 ? c + size : -1, y > 0 ? c - 1 : -1, y < size - 1 ? c + 1 : -1};
            for (int q : cand) {
                if (q < 0) continue;
                if (stone[q] == 0) {
                    if (mark_[q] != l) { mark_[q] = l; if (libs) libs[nlib] = (int16_t)q; nlib++; }
                } else if (stone[q] == colour && mark_[q] != s) {
                    mark_[q] = s;
                    queue_[tail++] = (int16_t)q;
                }
            }
        }
        if (members) memcpy(members, queue_, sizeof(int16_t) * tail);
        return {tail, nlib};
    }

    int neighbours(int p, int out[4]) const {  // go.py:96 order
        int x = p / size, y = p - x * size, k = 0;
        if (x > 0) out[k++] = p - size;
        if (x < size - 1) out[k++] = p + size;
        if (y > 0) out[k++] = p - 1;
        if (y < size - 1) out[k++] = p + 1;
        return k;
    }

    bool suicide(int p, int colour) {  // go.py:181-202
        int nb[4], k = neighbours(p, nb);
        for (int j = 0; j < k; j++) if (stone[nb[j]] == 0) return false;
        for (int j = 0; j < k; j++) {
            int q = nb[j];
            bool breathing = probe(q).liberties > 1;  // p itself is one of its liberties
            if (stone[q] == colour && breathing) return false;
            if (stone[q] == -colour && !breathing) return false;
        }
        return true;
    }
    bool legal(int p, int colour) {  // go.py:204-216; p < 0 is a pass
        if (p < 0) return true;
        return p < n && stone[p] == 0 && p != ko && !suicide(p, colour);
    }
    bool eyeish(int p, int owner) const {  // go.py:218-228
        if (stone[p] != 0) return false;
        int nb[4], k = neighbours(p, nb);
        for (int j = 0; j < k; j++) if (stone[nb[j]] != owner) return false;
        return true;
    }
    bool eye(int p, int owner) { path_len_ = 0; return eye_walk(p, owner); }

    // go.py:295-349.  colour 0 = side to move.  false = IllegalMove (state untouched).
    bool play(int p, int colour = 0) {
        if (colour == 0) colour = to_move;
        if (!legal(p, colour)) return false;
        ko = -1;
        if (p >= 0) {
            stone[p] = (int8_t)colour;
            int nb[4], k = neighbours(p, nb);
            int16_t dead[kMax];
            for (int j = 0; j < k; j++) {
                int q = nb[j];
                if (stone[q] != -colour) continue;
                Group g = probe(q, dead);
                if (g.liberties != 0) continue;
                for (int t = 0; t < g.stones; t++) stone[dead[t]] = 0;
                (colour == 1 ? prisoners_white : prisoners_black) += g.stones;
                if (g.stones == 1) {  // go.py:323-331, inside the loop: never cleared by a later capture
                    Group mine = probe(p);
                    if (mine.liberties == 1 && mine.stones == 1) ko = q;
                }
            }
        } else {
            (colour == 1 ? passes_black : passes_white)++;
        }
        to_move = -colour;
        turns++;
        history.push_back((int16_t)p);
        size_t h = history.size();
        if (h > 1 && history[h - 1] < 0 && history[h - 2] < 0 && to_move == -1) finished = true;  // go.py:345-348
        return true;
    }

    int legal_moves(bool include_eyes, int16_t* out) {  // go.py:261-267
        int c = 0;
        for (int p = 0; p < n; p++)
            if (legal(p, to_move) && (include_eyes || !eye(p, to_move))) out[c++] = (int16_t)p;
        return c;
    }

    int winner(double komi) const {  // go.py:269-293 (area scoring, eyeish empties)
        double sw = 0, sb = 0;
        for (int p = 0; p < n; p++) {
            if (stone[p] == 1) sb++;
            else if (stone[p] == -1) sw++;
            else if (eyeish(p, 1)) sb++;
            else if (eyeish(p, -1)) sw++;
        }
        sw += komi - passes_white;
        sb -= passes_black;
        return sb > sw ? 1 : (sw > sb ? -1 : 0);
    }

  private:
    uint32_t mark_[kMax];
    uint32_t stamp_ = 0;
    int16_t queue_[kMax];
    int16_t path_[kMax];
    int path_len_ = 0;

    bool eye_walk(int p, int owner) {  // go.py:230-259; path_ is the reference's `stack`
        if (!eyeish(p, owner)) return false;
        int x = p / size, y = p - x * size;
        const bool interior = x > 0 && y > 0 && x < size - 1 && y < size - 1;
        const int tolerated = interior ? 1 : 0;
        const int dx[4] = {-1, 1, 1, -1}, dy[4] = {-1, 1, -1, 1};  // go.py:102
        int bad = 0;
        for (int j = 0; j < 4; j++) {
            int xx = x + dx[j], yy = y + dy[j];
            if (xx < 0 || yy < 0 || xx >= size || yy >= size) continue;
            int d = xx * size + yy;
            if (stone[d] == -owner) bad++;
            else if (stone[d] == 0) {
                bool on_path = false;
                for (int t = 0; t < path_len_; t++) on_path |= (path_[t] == d);
                if (!on_path) {
                    path_[path_len_++] = (int16_t)p;
                    if (!eye_walk(d, owner)) bad++;
                    path_len_--;
                }
            }
            if (bad > tolerated) return false;
        }
        return true;
    }
};

// preprocessing.py:20-43 as data: age[p] = depth of the newest history entry at p whose point is
// still occupied (depth saturates at 7, passes and stale entries advance it too); 255 = none.
inline void age_from_history(int size, const int8_t* board, const int16_t* hist, int nhist, uint8_t* age) {
    memset(age, 255, (size_t)size * size);
    int depth = 0;
    for (int i = nhist - 1; i >= 0; i--) {
        int m = hist[i];
        if (m >= 0 && board[m] != 0 && age[m] == 255) age[m] = (uint8_t)depth;
        if (depth < 7) depth++;
    }
}

}  // namespace ag

// Device-resident self-play: a mini-batch of games lives in device memory and advances one ply per launch, so a ply is
// encode -> policy forward -> ag_selfplay_step with no host round trip.
//
// Replaces the per-ply host work of the reference's REINFORCE trainer
// (AlphaGo/training/reinforcement_policy_trainer.py:16-76: `get_moves` + `state.do_move` for every game of the mini-batch)
// and of ProbabilisticPolicyPlayer (AlphaGo/ai.py:51-68: distribution over the sensible moves raised to 1/temperature,
// one move sampled).  Rules are the reference's (AlphaGo/go.py:295-349): captures, the ko rule evaluated inside the
// per-neighbour capture loop, pass bookkeeping, end of game after two passes when White is to move -- played on bit
// boards by one warp per game with the ladder search's move generator (ladder_warp.cuh: wl_play), which the encoder
// tests already hold to the reference's rules.  tests/test_selfplay_gpu.py replays every move on the host engine and
// compares board, ko, side to move, turns_since ages and the end-of-game flag after every ply.
#include "common.cuh"
#include "encode.cuh"
#include "ladder.cuh"
#include "ladder_warp.cuh"

namespace ag {
namespace enc {

constexpr int kSelfplayWarps = 4;

// flags: bit 0 = game over, bit 1 = the previous move was a pass
__global__ void __launch_bounds__(kSelfplayWarps * 32)
selfplay_step_kernel(int8_t* __restrict__ g_board, uint8_t* __restrict__ g_age, int8_t* __restrict__ g_player, int16_t* __restrict__ g_ko,
                     uint8_t* __restrict__ g_flags, int32_t* __restrict__ g_passes, const float* __restrict__ g_probs,
                     const uint8_t* __restrict__ g_mask, long long mask_stride, const float* __restrict__ g_uniform, float beta,
                     int games, int size, int16_t* __restrict__ g_moves) {
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * kSelfplayWarps + (threadIdx.x >> 5);
    if (g >= games) return;
    const int n = size * size;
    if (g_flags[g] & 1) {   // go.py:297-349 is never reached for a finished game (rpt.py:46 `if not st.is_end_of_game`)
        if (lane == 0) g_moves[g] = -2;
        return;
    }
    int8_t* board = g_board + (size_t)g * n;
    uint8_t* age = g_age + (size_t)g * n;
    const uint8_t* mask = g_mask + (size_t)g * mask_stride;
    const float* probs = g_probs + (size_t)g * n;

    // ---- sample a move: weights p^beta over the masked points (ai.py:38-46), the first point whose running sum passes u * total
    float w[12];
    float total = 0.f;
#pragma unroll
    for (int k = 0; k < 12; k++) {
        const int p = k * 32 + lane;
        w[k] = (p < n && mask[p]) ? __powf(fmaxf(probs[p], 1e-30f), beta) : 0.f;
        total += w[k];
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) total += __shfl_xor_sync(0xffffffffu, total, d);
    int move = -1;
    if (total > 0.f) {
        const float target = g_uniform[g] * total;
        float base = 0.f;
        int last_pos = -1;   // the last point with weight: the fallback if rounding leaves the running sum short of the target
#pragma unroll
        for (int k = 0; k < 12; k++) {
            float incl = w[k];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const float v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
            const unsigned hit = __ballot_sync(0xffffffffu, w[k] > 0.f && base + incl > target);
            const unsigned any = __ballot_sync(0xffffffffu, w[k] > 0.f);
            if (any) last_pos = k * 32 + 31 - __clz(any);
            if (move < 0 && hit) move = k * 32 + __ffs(hit) - 1;
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (move < 0) move = last_pos;
    }

    // ---- play it (go.py:295-349)
    WBoard B;
    B.m = wmasks_make(size, lane);
    B.size = size; B.lane = lane;
    const int mover = g_player[g];
    wl_load(B, board, size, mover, g_ko[g], lane);
    const uint32_t before = B.blk | B.wht;
    if (move >= 0) {
        wl_play(B, move);               // sets B.ko, flips B.player
    } else {
        B.ko = -1;
        B.player = -mover;
        if (lane == 0) g_passes[2 * g + (mover > 0 ? 0 : 1)] += 1;
    }
    // board and turns_since ages back to memory: every surviving stone is one history entry older (preprocessing.py:20-43
    // counts passes too), the new stone is the newest
#pragma unroll
    for (int k = 0; k < 12; k++) {
        const int p = k * 32 + lane;
        const uint32_t bw = __shfl_sync(0xffffffffu, B.blk, k), ww = __shfl_sync(0xffffffffu, B.wht, k), old = __shfl_sync(0xffffffffu, before, k);
        if (p >= n) continue;
        const int c = ((bw >> lane) & 1u) ? 1 : (((ww >> lane) & 1u) ? -1 : 0);
        board[p] = (int8_t)c;
        uint8_t a = 255;
        if (c != 0) {
            if (p == move) a = 0;
            else { const uint8_t o = ((old >> lane) & 1u) ? age[p] : (uint8_t)255; a = o >= 7 ? (o == 255 ? 255 : 7) : (uint8_t)(o + 1); }
        }
        age[p] = a;
    }
    if (lane == 0) {
        const int flags = g_flags[g];
        const bool pass = move < 0;
        const bool over = pass && (flags & 2) && B.player == -1;   // go.py:345-348
        g_flags[g] = (uint8_t)((over ? 1 : 0) | (pass ? 2 : 0));
        g_player[g] = (int8_t)B.player;
        g_ko[g] = (int16_t)B.ko;
        g_moves[g] = (int16_t)move;
    }
}

// go.py:269-293: area scoring, an empty point counts for the colour that owns all its on-board neighbours; komi and passes
__global__ void selfplay_winner_kernel(const int8_t* __restrict__ g_board, const int32_t* __restrict__ g_passes, int games, int size,
                                       float komi, int8_t* __restrict__ g_winner) {
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= games) return;
    const int n = size * size;
    const int8_t* b = g_board + (size_t)g * n;
    int sb = 0, sw = 0;
    for (int p = lane; p < n; p += 32) {
        const int c = b[p];
        if (c == 1) sb++;
        else if (c == -1) sw++;
        else {
            const int x = p / size, y = p - x * size;
            bool allb = true, allw = true;
            if (x > 0) { const int q = b[p - size]; allb &= q == 1; allw &= q == -1; }
            if (x < size - 1) { const int q = b[p + size]; allb &= q == 1; allw &= q == -1; }
            if (y > 0) { const int q = b[p - 1]; allb &= q == 1; allw &= q == -1; }
            if (y < size - 1) { const int q = b[p + 1]; allb &= q == 1; allw &= q == -1; }
            if (allb) sb++;
            else if (allw) sw++;
        }
    }
    for (int d = 16; d > 0; d >>= 1) { sb += __shfl_xor_sync(0xffffffffu, sb, d); sw += __shfl_xor_sync(0xffffffffu, sw, d); }
    if (lane == 0) {
        const double white = (double)sw + (double)komi - (double)g_passes[2 * g + 1], black = (double)sb - (double)g_passes[2 * g];
        g_winner[g] = black > white ? 1 : (white > black ? -1 : 0);
    }
}

}  // namespace enc
}  // namespace ag

extern "C" {

int ag_selfplay_step(int8_t* d_board, uint8_t* d_age, int8_t* d_player, int16_t* d_ko, uint8_t* d_flags, int32_t* d_passes,
                     const float* d_probs, const uint8_t* d_mask, long long mask_stride, const float* d_uniform, float beta,
                     int games, int size, int16_t* d_moves, void* stream) {
    if (games < 0 || size < 2 || size > 19 || mask_stride < (long long)size * size) { ag::set_error("ag_selfplay_step: bad argument"); return AG_EINVAL; }
    if (games == 0) return AG_OK;
    if (!d_board || !d_age || !d_player || !d_ko || !d_flags || !d_passes || !d_probs || !d_mask || !d_uniform || !d_moves) {
        ag::set_error("ag_selfplay_step: null device pointer");
        return AG_EINVAL;
    }
    const int grid = (games + ag::enc::kSelfplayWarps - 1) / ag::enc::kSelfplayWarps;
    ag::enc::selfplay_step_kernel<<<grid, ag::enc::kSelfplayWarps * 32, 0, (cudaStream_t)stream>>>(
        d_board, d_age, d_player, d_ko, d_flags, d_passes, d_probs, d_mask, mask_stride, d_uniform, beta, games, size, d_moves);
    AG_LAUNCH_CHECK("selfplay_step_kernel");
    return AG_OK;
}

int ag_selfplay_winner(const int8_t* d_board, const int32_t* d_passes, int games, int size, float komi, int8_t* d_winner, void* stream) {
    if (games < 0 || size < 2 || size > 19) { ag::set_error("ag_selfplay_winner: bad argument"); return AG_EINVAL; }
    if (games == 0) return AG_OK;
    ag::enc::selfplay_winner_kernel<<<(games + 3) / 4, 128, 0, (cudaStream_t)stream>>>(d_board, d_passes, games, size, komi, d_winner);
    AG_LAUNCH_CHECK("selfplay_winner_kernel");
    return AG_OK;
}

}  // extern "C"

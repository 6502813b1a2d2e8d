// CPU harness for the device ladder search (alphago_b200/csrc/ladder.cuh): the header is plain C++ once the
// CUDA qualifiers and two intrinsics are stubbed, so the exact state machine the kernel runs can be checked
// against the oracle without a GPU.  Exposes one C function for tests/test_ladder_host.py.
#include <stdint.h>
#include <string.h>
#define __device__
#define __forceinline__ inline
static inline uint32_t __funnelshift_l(uint32_t lo, uint32_t hi, uint32_t s) { s &= 31; return s ? (hi << s) | (lo >> (32 - s)) : hi; }
static inline uint32_t __funnelshift_r(uint32_t lo, uint32_t hi, uint32_t s) { s &= 31; return s ? (lo >> s) | (hi << (32 - s)) : lo; }
static inline int __popc(uint32_t v) { return __builtin_popcount(v); }
static inline int __ffs(uint32_t v) { return __builtin_ffs((int)v); }
namespace ag { namespace enc {
constexpr int MAXN = 361, NPAD = 368, LW = 12;
} }
#include "../../alphago_b200/csrc/ladder.cuh"

extern "C" int ladder_planes_host(const int8_t* board, int size, int player, int ko, uint8_t* cap, uint8_t* esc) {
    using namespace ag::enc;
    static LadderBoard B;
    static LadderFrame fr[kLadderDepth + 2];
    static BoardMasks masks;
    board_masks_fill(masks, size);
    lb_load(B, board, size, player, ko, &masks);
    const Bits b0 = B.blk, w0 = B.wht;
    int n = size * size;
    for (int p = 0; p < n; p++) {
        cap[p] = esc[p] = 0;
        if (!lb_legal(B, p)) continue;
        cap[p] = ladder_query(B, fr, 0, p);
        esc[p] = ladder_query(B, fr, 1, p);
        if (B.undo_top != 0 || B.player != player || B.ko != ko) return -1;  // search must restore the board
        if (!bits_eq(B.blk, b0) || !bits_eq(B.wht, w0)) return -2;
    }
    return 0;
}

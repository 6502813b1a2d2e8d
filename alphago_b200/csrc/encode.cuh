// Encoder constants shared between the kernel and the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ag {
namespace enc {

constexpr int MAXN = 361;        // 19x19
constexpr int NPAD = 368;        // points padded to a multiple of 16
constexpr int LW = 12;           // 361-bit sets as 12 words
constexpr int WARPS = 10;        // positions per CTA (one warp each): 3 CTAs x 10 warps fill the 227 KB of an SM
constexpr int MAX_PLANES = 64;   // longest feature list the kernel accepts, in planes

// feature ids of the C ABI (order of DEFAULT_FEATURES, preprocessing.py:211-214, + value-net colour)
enum Feature { F_BOARD, F_ONES, F_TURNS, F_LIBS, F_CAPTURE, F_SELFATARI, F_LIBSAFTER, F_LADCAP, F_LADESC,
               F_SENSIBLE, F_ZEROS, F_COLOR, F_COUNT };

// per-point code arrays in shared memory: the plane index a point sets within a family (0xFF: none)
enum Code { C_BOARD, C_TURNS, C_LIBS, C_CAPTURE, C_SELFATARI, C_LIBSAFTER, C_SENSIBLE, C_LADCAP, C_LADESC,
            C_CONST, C_COLOR, N_CODE };
constexpr int CODE_STRIDE = (NPAD / 4 + 1) * 4;  // bytes per code array

// output plane j is 1 where code array arr[j] equals val[j]
struct PlaneTable {
    int n_planes;
    int need_eye;     // sensibleness requested
    int need_ladder;  // ladder_capture / ladder_escape requested
    uint8_t arr[MAX_PLANES];
    uint8_t val[MAX_PLANES];
};

}  // namespace enc

int encode_launch(const int8_t* board, const uint8_t* age, const int8_t* player, const int16_t* ko, int npos, int size,
                  const enc::PlaneTable& tab, uint8_t* out, cudaStream_t stream);
size_t encode_smem_bytes();

}  // namespace ag

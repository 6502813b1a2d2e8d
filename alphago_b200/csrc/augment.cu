// Minibatch assembly for SL training: gather samples from the uint8 dataset, apply one of the 8
// dihedral board symmetries per sample and turn the (x, y) action into a flat label.
//
// Replaces the reference's per-sample Python loop in shuffled_hdf5_batch_generator
// (AlphaGo/training/supervised_policy_trainer.py:18-40): F numpy rot90/flip calls per sample and a
// one-hot 19x19 target.  The transforms are BOARD_TRANSFORMATIONS (spt.py:71-80) on arrays indexed
// [x][y]; the label stays an integer (the one-hot never exists).
#include "common.cuh"

namespace ag {

// source point (sx, sy) read by output point (i, j) under transform t on an N x N array:
//  0 identity            1 rot90(f,1): f[j][N-1-i]   2 rot90(f,2): f[N-1-i][N-1-j]   3 rot90(f,3): f[N-1-j][i]
//  4 fliplr: f[i][N-1-j] 5 flipud: f[N-1-i][j]       6 transpose: f[j][i]            7 fliplr(rot90(f,1)): f[N-1-j][N-1-i]
__device__ __forceinline__ void d4_source(int t, int N, int i, int j, int& sx, int& sy) {
    const int M = N - 1;
    switch (t) {
        default: sx = i; sy = j; break;
        case 1: sx = j; sy = M - i; break;
        case 2: sx = M - i; sy = M - j; break;
        case 3: sx = M - j; sy = i; break;
        case 4: sx = i; sy = M - j; break;
        case 5: sx = M - i; sy = j; break;
        case 6: sx = j; sy = i; break;
        case 7: sx = M - j; sy = M - i; break;
    }
}
// where the action (x, y) lands: the inverse of d4_source
__device__ __forceinline__ void d4_dest(int t, int N, int x, int y, int& i, int& j) {
    const int M = N - 1;
    switch (t) {
        default: i = x; j = y; break;
        case 1: i = M - y; j = x; break;
        case 2: i = M - x; j = M - y; break;
        case 3: i = y; j = M - x; break;
        case 4: i = x; j = M - y; break;
        case 5: i = M - x; j = y; break;
        case 6: i = y; j = x; break;
        case 7: i = M - y; j = M - x; break;
    }
}

__global__ void __launch_bounds__(256)
augment_kernel(const uint8_t* __restrict__ states, const uint8_t* __restrict__ actions, const int64_t* __restrict__ index,
               const uint8_t* __restrict__ transform, int planes, int size, uint8_t* __restrict__ out, int32_t* __restrict__ labels) {
    pdl_trigger();
    pdl_wait();
    __shared__ uint16_t s_src[368];
    const int b = blockIdx.x, pts = size * size;
    const int t = transform[b] & 7;
    const int64_t src = index[b];
    for (int p = threadIdx.x; p < pts; p += blockDim.x) {
        int i = p / size, j = p - i * size, sx, sy;
        d4_source(t, size, i, j, sx, sy);
        s_src[p] = (uint16_t)(sx * size + sy);
    }
    if (threadIdx.x == 0) {
        int x = actions[src * 2], y = actions[src * 2 + 1], i, j;
        d4_dest(t, size, x, y, i, j);
        labels[b] = i * size + j;
    }
    __syncthreads();
    const uint8_t* in = states + (size_t)src * planes * pts;
    uint8_t* o = out + (size_t)b * planes * pts;
    const int total = planes * pts;
    // 4 output bytes per thread per iteration where the sample is 4-byte aligned, else byte by byte
    if (((reinterpret_cast<uintptr_t>(o) | (uintptr_t)total) & 3) == 0) {
        for (int e = threadIdx.x * 4; e < total; e += blockDim.x * 4) {
            uint32_t word = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int f = (e + k) / pts, p = (e + k) - f * pts;
                word |= (uint32_t)in[f * pts + s_src[p]] << (8 * k);
            }
            *reinterpret_cast<uint32_t*>(o + e) = word;
        }
    } else {
        for (int e = threadIdx.x; e < total; e += blockDim.x) {
            int f = e / pts, p = e - f * pts;
            o[e] = in[f * pts + s_src[p]];
        }
    }
}

}  // namespace ag

extern "C" int ag_augment(const uint8_t* d_states, const uint8_t* d_actions, const int64_t* d_index, const uint8_t* d_transform,
                          int batch, int planes, int size, uint8_t* d_out, int32_t* d_labels, void* stream) {
    if (batch < 0 || planes < 1 || size < 2 || size > 19) { ag::set_error("ag_augment: bad argument"); return AG_EINVAL; }
    if (batch == 0) return AG_OK;
    ag::launch_k(ag::augment_kernel, dim3(batch), dim3(256), 0, (cudaStream_t)stream, d_states, d_actions, d_index, d_transform, planes, size, d_out, d_labels);
    AG_LAUNCH_CHECK("augment_kernel");
    return AG_OK;
}

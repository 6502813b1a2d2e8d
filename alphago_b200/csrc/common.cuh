// Shared helpers for the alphago_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/alphago_b200.h"  // AG_OK / AG_E* codes

namespace ag {

// error reporting for the C ABI: every entry point returns 0 or a negative code and leaves a
// message for ag_last_error().

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
void count_launch(int n = 1);

#define AG_CUDA(expr)                                               \
    do {                                                            \
        cudaError_t _e = (expr);                                    \
        if (_e != cudaSuccess) return ::ag::cuda_fail(_e, #expr);   \
    } while (0)

// checked after every kernel launch (launch-configuration errors only; no sync)
#define AG_LAUNCH_CHECK(name)                                        \
    do {                                                             \
        ::ag::count_launch();                                        \
        cudaError_t _e = cudaGetLastError();                         \
        if (_e != cudaSuccess) return ::ag::cuda_fail(_e, name);     \
    } while (0)

int sm_count();

// Programmatic dependent launch: every kernel of the training step is launched with the stream-serialization
// attribute, raises `launch_dependents` on entry and does its set-up (barrier init, TMEM allocation, descriptor
// prefetch) BEFORE `griddepcontrol.wait`, so that part overlaps the tail of the kernel ahead of it in the stream.
// Nothing before pdl_wait() may touch global memory.  AG_PDL=0 in the environment turns the attribute off (the two
// instructions are then no-ops).
bool pdl_enabled();

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... P, typename... A>
inline cudaError_t launch_k(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}
#endif

}  // namespace ag

// Shared helpers for the alphago_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

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

int sm_count();   // of the CURRENT device

// One process may drive several GPUs (one thread or one net per device): everything the launchers cache about a device
// (shared-memory opt-ins, occupancy answers, symbol addresses) is kept per device ordinal, never process-wide.
constexpr int kMaxDevices = 64;
int current_device();   // cudaGetDevice(), clamped to [0, kMaxDevices)
template <typename T>
struct PerDevice {
    T v[kMaxDevices] = {};
    T& cur() { return v[current_device()]; }
};

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
inline cudaError_t launch_k_opt(bool early, void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = (early && pdl_enabled()) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}
template <typename... P, typename... A>
inline cudaError_t launch_k(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A... args) {
    return launch_k_opt(true, kernel, grid, block, smem, stream, args...);
}
// For kernels whose CTAs wait for each other (the layer chains): a COOPERATIVE launch.  The runtime refuses a grid that
// cannot be co-resident on the device (cudaErrorCooperativeLaunchTooLarge: the caller falls back to per-layer kernels) and
// the hardware starts the grid only as a whole, so CTAs never spin on CTAs that have no SM yet -- whatever else runs on
// the device (another net on a second stream, NCCL's CTAs, the encoder).  Never a programmatic-dependent-launch secondary:
// starting early behind the kernel in front is exactly the piecemeal residency this rules out.  AG_CHAIN_COOP=0 falls back to
// a plain launch (for A/B measurements only).
bool chain_coop_enabled();
template <typename... P, typename... A>
inline cudaError_t launch_k_whole(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = chain_coop_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}
#endif

}  // namespace ag

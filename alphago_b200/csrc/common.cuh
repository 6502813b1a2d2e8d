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

}  // namespace ag

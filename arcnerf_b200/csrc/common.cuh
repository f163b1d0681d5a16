// Shared host/device helpers for libarcb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#define ARCB200_OK 0
#define ARCB200_ERR_ARG (-1)      // bad argument (null pointer, negative size, unsupported shape)
#define ARCB200_ERR_CUDA (-2)     // a CUDA runtime call failed; see arcb200_last_error()
#define ARCB200_ERR_LIMIT (-3)    // size beyond a documented limit

// Per-thread last-error text (defined in abi.cu).
void arcb200_set_error(const char *fmt, ...);

#define ARC_REQUIRE(cond, ...)                          \
    do {                                                \
        if (!(cond)) {                                  \
            arcb200_set_error(__VA_ARGS__);             \
            return ARCB200_ERR_ARG;                     \
        }                                               \
    } while (0)

#define ARC_CUDA(call)                                                                       \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            arcb200_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return ARCB200_ERR_CUDA;                                                         \
        }                                                                                    \
    } while (0)

#define ARC_LAUNCH_CHECK()                                                                   \
    do {                                                                                     \
        cudaError_t e__ = cudaGetLastError();                                                \
        if (e__ != cudaSuccess) {                                                            \
            arcb200_set_error("%s:%d kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
            return ARCB200_ERR_CUDA;                                                         \
        }                                                                                    \
    } while (0)

static inline unsigned int arc_blocks(long long n, int threads) {
    return (unsigned int)((n + threads - 1) / threads);
}

// Number of SMs of the current device (cached per process; B200 = 148).
int arc_sm_count();

#define ARC_FULL_MASK 0xffffffffu

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

// ---- programmatic dependent launch (PDL) --------------------------------------------------------------
// Every kernel of the library is launched with cudaLaunchAttributeProgrammaticStreamSerialization and begins with
// arc_pdl_sync(): wait until the preceding kernel of the stream has completed (griddepcontrol.wait), then allow the
// following kernel to be scheduled (griddepcontrol.launch_dependents).  The next kernel's CTAs are placed on the SMs as
// this kernel's CTAs drain and sit in their own wait, so launch latency and CTA start-up no longer lie between two
// kernels of a step.  Because a kernel signals only AFTER its own wait, anything older than the direct predecessor is
// complete by the time a kernel starts.  Without the launch attribute both instructions are no-ops.
// ARCB200_PDL=0 in the environment turns the attribute off (plain stream order).
#ifdef __CUDACC__
__device__ __forceinline__ void arc_pdl_sync() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif

bool arc_pdl_enabled();   // abi.cu

template <typename... KArgs, typename... Args>
static inline void arc_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = arc_pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);   // errors surface through ARC_LAUNCH_CHECK()
}

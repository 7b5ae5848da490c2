// common.cuh -- shared helpers for libtsm_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/tsm_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libtsm_b200 is written for sm_100a (B200) only"
#endif

namespace tsm {

extern thread_local char g_err[512];

inline int fail(int code, const char* fmt, const char* a = "", long long b = 0, long long c = 0) {
    snprintf(g_err, sizeof(g_err), fmt, a, b, c);
    return code;
}

#define TSM_REQUIRE(cond, msg)                                                        \
    do {                                                                              \
        if (!(cond)) return ::tsm::fail(TSM_ERR_INVALID, "%s: requirement failed: " msg, __func__); \
    } while (0)

#define TSM_CUDA(expr)                                                                \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            snprintf(::tsm::g_err, sizeof(::tsm::g_err), "%s: CUDA error %s at %s:%d", __func__, \
                     cudaGetErrorString(_e), __FILE__, __LINE__);                     \
            return TSM_ERR_CUDA;                                                      \
        }                                                                             \
    } while (0)

#define TSM_LAUNCH_CHECK() TSM_CUDA(cudaGetLastError())

inline cudaStream_t as_stream(tsm_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

inline int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 148;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-stream-serialization attribute may start
// while its predecessor in the stream is still running; pdl_wait() blocks until the predecessor has completed and its
// writes are visible (no-op for a normal launch); pdl_launch_dependents() lets the successor start its prologue.
// Host side: the lock-step rollout loop (collect.cu) sets g_pdl_hint around the launches whose predecessor in the stream is the
// other kernel of its forward -> world step chain; launch sites that understand it (tc_net_forward, launch_step) then launch
// with the attribute.  Everything else launches normally (the device-side calls are no-ops then).
extern thread_local bool g_pdl_hint;
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(bool pdl, void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(grid); lc.blockDim = dim3(block); lc.dynamicSmemBytes = smem; lc.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&lc, kernel, static_cast<KArgs>(args)...);
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

}  // namespace tsm

// Shared helpers for libopb200 (error plumbing, exact-arithmetic wrappers, small device utilities).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include "../../include/opb200.h"

namespace opb
{
    // Thread-local error text behind opb200_last_error().
    void setError(const char* fmt, ...);
    const char* getError();

    inline opb200_status fail(opb200_status code, const char* what, const char* file, int line)
    {
        setError("%s (%s:%d)", what, file, line);
        return code;
    }

    constexpr int kNumSMsB200 = 148;
}

#define OPB_FAIL(code, msg) return ::opb::fail((code), (msg), __FILE__, __LINE__)
#define OPB_REQUIRE(cond, msg)                              \
    do                                                      \
    {                                                       \
        if (!(cond))                                        \
            OPB_FAIL(OPB200_ERR_INVALID, msg);              \
    } while (0)
#define OPB_CUDA(expr)                                                                       \
    do                                                                                       \
    {                                                                                        \
        const cudaError_t e__ = (expr);                                                      \
        if (e__ != cudaSuccess)                                                              \
        {                                                                                    \
            ::opb::setError("CUDA error %s: %s (%s:%d)", cudaGetErrorName(e__),             \
                            cudaGetErrorString(e__), __FILE__, __LINE__);                    \
            return OPB200_ERR_CUDA;                                                          \
        }                                                                                    \
    } while (0)
// After a kernel launch: same role as the reference's cudaCheck (src/openpose/gpu/cuda.cpp:18-39).
#define OPB_LAUNCH_CHECK() OPB_CUDA(cudaPeekAtLastError())

#ifdef __CUDACC__
namespace opb
{
    // The CPU reference is compiled without FMA contraction (x86-64 baseline). These wrappers pin
    // round-to-nearest single operations so nvcc cannot fuse a*b+c.
    __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    __device__ __forceinline__ float divf(float a, float b) { return __fdiv_rn(a, b); }

    // positiveIntRound (reference: include/openpose/utilities/fastMath.hpp:29-32): add 0.5f, truncate.
    __device__ __forceinline__ int posIntRound(float a) { return (int)add(a, 0.5f); }

    __device__ __forceinline__ unsigned laneId() { return threadIdx.x & 31u; }
}
#endif

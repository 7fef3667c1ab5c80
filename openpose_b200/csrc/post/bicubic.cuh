// Catmull-Rom bicubic pieces shared by the resize and pre-processing kernels: the evaluation order of
// include/openpose_private/gpu/cuda.hu:91-144 with explicit round-to-nearest multiplies/adds (no FMA contraction).
#pragma once
#include "../common.cuh"

namespace opb
{
    // cuda.hu:117-120, exact order, no FMA.
    __device__ __forceinline__ float cubicA(float v0, float v1, float v2, float v3)
    {
        return add(sub(add(mul(-0.5f, v0), mul(1.5f, v1)), mul(1.5f, v2)), mul(0.5f, v3));
    }
    __device__ __forceinline__ float cubicB(float v0, float v1, float v2, float v3)
    {
        return sub(add(sub(v0, mul(2.5f, v1)), mul(2.f, v2)), mul(0.5f, v3));
    }
    __device__ __forceinline__ float cubicC(float v0, float v2) { return mul(0.5f, sub(v0, v2)); }
    __device__ __forceinline__ float cubicEval(float A, float B, float C, float v1, float d)
    {
        const float t1 = mul(mul(mul(A, d), d), d);
        const float t2 = mul(mul(B, d), d);
        const float t3 = mul(C, d);
        return add(sub(add(t1, t2), t3), v1);
    }
    __device__ __forceinline__ float cubic1d(float v0, float v1, float v2, float v3, float d)
    {
        return cubicEval(cubicA(v0, v1, v2, v3), cubicB(v0, v1, v2, v3), cubicC(v0, v2), v1, d);
    }

    // cuda.hu:96-106 for one axis: clamped tap indices + (unclamped) fractional offset.
    __device__ __forceinline__ void taps(int idx[4], float& d, float s, int size)
    {
        idx[1] = min(size - 1, max(0, (int)floorf(s)));
        idx[0] = max(0, idx[1] - 1);
        idx[2] = min(size - 1, idx[1] + 1);
        idx[3] = min(size - 1, idx[2] + 1);
        d = sub(s, (float)idx[1]);
    }
    __host__ __device__ inline int align4(int v) { return (v + 3) & ~3; }
    __device__ __forceinline__ float srcCoord(int x, float scale) { return sub(divf(add((float)x, 0.5f), scale), 0.5f); }
    // (x + 0.5) / 8 - 0.5: the division by a power of two is exact, so a multiply gives the identical float.
    __device__ __forceinline__ float srcCoord8(int x) { return sub(mul(add((float)x, 0.5f), 0.125f), 0.5f); }
}

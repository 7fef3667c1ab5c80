// Catmull-Rom bicubic pieces shared by the resize and pre-processing kernels: cubicInterpolate of
// include/openpose_private/gpu/cuda.hu:109-122 evaluated exactly as the reference's CUDA build evaluates it. nvcc
// compiles the reference with its default -fmad=true and contracts the expression
//     (-0.5f*v0 + 1.5f*v1 - 1.5f*v2 + 0.5f*v3)*dx*dx*dx + (v0 - 2.5f*v1 + 2.f*v2 - 0.5f*v3)*dx*dx - 0.5f*(v0 - v2)*dx + v1
// left to right into the FMA chain below (read off the SASS of resize8TimesKernel<float> /
// resizeAndAddAndAverageKernel<float> / resizeAndPadKernel<float> in oracle/_ref/obj/cu_resizeAndMergeBase.o; every
// inlined instance has the same shape). The roundings are spelled out with __fmaf_rn / __fmul_rn so this file's result
// does not depend on what the compiler chooses to fuse: bit-identical to the reference kernels on the same GPU
// (tests/test_refcuda_gpu.py) and to the fmaf() evaluation of the C oracle (oracle/post_oracle.c, `contract` = 1).
// The non-contracted evaluation (cuda.hu compiled as host C++) differs from it by a few ulp; tests/test_oracle_vs_ref.py
// reports and bounds that difference.
#pragma once
#include "../common.cuh"

namespace opb
{
    __device__ __forceinline__ float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }
    // A = ((-0.5*v0 + 1.5*v1) - 1.5*v2) + 0.5*v3 with every multiply-add fused
    __device__ __forceinline__ float cubicA(float v0, float v1, float v2, float v3)
    {
        return fma_(v3, 0.5f, fma_(v2, -1.5f, fma_(v1, 1.5f, mul(v0, -0.5f))));
    }
    // B = ((v0 - 2.5*v1) + 2*v2) - 0.5*v3
    __device__ __forceinline__ float cubicB(float v0, float v1, float v2, float v3)
    {
        return fma_(v3, -0.5f, fma_(v2, 2.f, fma_(v1, -2.5f, v0)));
    }
    // C = -(0.5*(v0 - v2)) (the sign flip is exact)
    __device__ __forceinline__ float cubicC(float v0, float v2) { return mul(sub(v0, v2), -0.5f); }
    // ((A*d)*d)*d, then + (B*d)*d fused, then + C*d fused, then + v1
    __device__ __forceinline__ float cubicEval(float A, float B, float C, float v1, float d)
    {
        float t = mul(mul(mul(A, d), d), d);
        t = fma_(mul(B, d), d, t);
        t = fma_(C, d, t);
        return add(v1, t);
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

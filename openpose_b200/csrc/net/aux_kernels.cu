// Bandwidth-bound helpers of the net forward (everything that is not the implicit-GEMM conv):
//   inputPatchKernel : fp32 NCHW net input -> bf16 NHWC "patch" tensor (k x k im2col of the few input channels,
//                      so the first conv becomes a 1x1 GEMM with K = Cin*k*k <= 64; k = 1 is a plain layout change)
//   maxPoolKernel    : Caffe MAX pooling, ceil output size, window clipped to the image
//                      (reference: 3rdparty/caffe/src/caffe/layers/pooling_layer.cpp:91-97,141-175)
//   nhwcToNchwKernel : bf16 NHWC view -> fp32 NCHW (only for output blobs that no conv epilogue wrote directly)
#include "../common.cuh"
#include "net_kernels.h"
#include <cuda_bf16.h>

namespace opb
{
    __global__ void __launch_bounds__(256) inputPatchKernel(__nv_bfloat16* __restrict__ dst,
                                                            const float* __restrict__ src, int N, int C, int H, int W,
                                                            int k, int pad, int dstC)
    {
        // one thread per (pixel, 8-channel group of the patch vector)
        const int groups = dstC / 8;
        const long long total = (long long)N * H * W * groups;
        const int kk = k * k;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
             i += (long long)gridDim.x * blockDim.x)
        {
            const int g = (int)(i % groups);
            const long long pix = i / groups;
            const int x = (int)(pix % W);
            const int y = (int)((pix / W) % H);
            const int n = (int)(pix / ((long long)W * H));
            __align__(16) __nv_bfloat16 v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)
            {
                const int j = g * 8 + e;          // j = c*k*k + ky*k + kx  (Caffe weight order within one filter)
                float f = 0.f;
                if (j < C * kk)
                {
                    const int c = j / kk, r = j - c * kk;
                    const int yy = y + r / k - pad, xx = x + r % k - pad;
                    if (yy >= 0 && yy < H && xx >= 0 && xx < W)
                        f = __ldg(src + (((long long)n * C + c) * H + yy) * W + xx);
                }
                v[e] = __float2bfloat16_rn(f);
            }
            *reinterpret_cast<uint4*>(dst + pix * dstC + g * 8) = *reinterpret_cast<const uint4*>(v);
        }
    }

    opb200_status launchInputPatch(void* dst, const float* src, int N, int C, int H, int W, int k, int pad, int dstC,
                                   cudaStream_t st)
    {
        const long long total = (long long)N * H * W * (dstC / 8);
        const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
        inputPatchKernel<<<blocks > 0 ? blocks : 1, 256, 0, st>>>(static_cast<__nv_bfloat16*>(dst), src, N, C, H, W, k,
                                                                 pad, dstC);
        OPB_LAUNCH_CHECK();
        return OPB200_OK;
    }

    __device__ __forceinline__ uint32_t bmax2(uint32_t a, uint32_t b)
    {
        __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&a), y = *reinterpret_cast<__nv_bfloat162*>(&b);
        __nv_bfloat162 m = __hmax2(x, y);
        return *reinterpret_cast<uint32_t*>(&m);
    }

    // src view: pixel stride srcStride elements, channels [0, C) (C multiple of 8); dst likewise.
    __global__ void __launch_bounds__(256) maxPoolKernel(__nv_bfloat16* __restrict__ dst,
                                                         const __nv_bfloat16* __restrict__ src, int N, int C, int H,
                                                         int W, int Ho, int Wo, int k, int stride, int srcStride,
                                                         int dstStride)
    {
        const int groups = C / 8;
        const long long total = (long long)N * Ho * Wo * groups;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
             i += (long long)gridDim.x * blockDim.x)
        {
            const int g = (int)(i % groups);
            const long long pix = i / groups;
            const int xo = (int)(pix % Wo);
            const int yo = (int)((pix / Wo) % Ho);
            const int n = (int)(pix / ((long long)Wo * Ho));
            const int y0 = yo * stride, x0 = xo * stride;
            const int y1 = min(y0 + k, H), x1 = min(x0 + k, W);
            uint4 m = *reinterpret_cast<const uint4*>(src + (((long long)n * H + y0) * W + x0) * srcStride + g * 8);
            for (int y = y0; y < y1; ++y)
                for (int x = x0; x < x1; ++x)
                {
                    const uint4 v =
                        *reinterpret_cast<const uint4*>(src + (((long long)n * H + y) * W + x) * srcStride + g * 8);
                    m.x = bmax2(m.x, v.x);
                    m.y = bmax2(m.y, v.y);
                    m.z = bmax2(m.z, v.z);
                    m.w = bmax2(m.w, v.w);
                }
            *reinterpret_cast<uint4*>(dst + pix * dstStride + g * 8) = m;
        }
    }

    opb200_status launchMaxPool(void* dst, const void* src, int N, int C, int H, int W, int Ho, int Wo, int k,
                                int stride, int srcStride, int dstStride, cudaStream_t st)
    {
        OPB_REQUIRE(C % 8 == 0, "pool: channel count must be a multiple of 8");
        const long long total = (long long)N * Ho * Wo * (C / 8);
        const long long want = (total + 255) / 256;
        const int blocks = (int)(want < 148 * 16 ? want : 148 * 16);
        maxPoolKernel<<<blocks > 0 ? blocks : 1, 256, 0, st>>>(static_cast<__nv_bfloat16*>(dst),
                                                              static_cast<const __nv_bfloat16*>(src), N, C, H, W, Ho,
                                                              Wo, k, stride, srcStride, dstStride);
        OPB_LAUNCH_CHECK();
        return OPB200_OK;
    }

    __global__ void __launch_bounds__(256) nhwcToNchwKernel(float* __restrict__ dst,
                                                            const __nv_bfloat16* __restrict__ src, int N, int C, int H,
                                                            int W, int srcStride, int dstC, int dstOff)
    {
        const long long total = (long long)N * C * H * W;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
             i += (long long)gridDim.x * blockDim.x)
        {
            const int x = (int)(i % W);
            const int y = (int)((i / W) % H);
            const int c = (int)((i / ((long long)W * H)) % C);
            const int n = (int)(i / ((long long)W * H * C));
            dst[(((long long)n * dstC + dstOff + c) * H + y) * W + x] =
                __bfloat162float(src[(((long long)n * H + y) * W + x) * srcStride + c]);
        }
    }

    opb200_status launchNhwcToNchw(float* dst, const void* src, int N, int C, int H, int W, int srcStride, int dstC,
                                   int dstOff, cudaStream_t st)
    {
        const long long total = (long long)N * C * H * W;
        const long long want = (total + 255) / 256;
        const int blocks = (int)(want < 148 * 16 ? want : 148 * 16);
        nhwcToNchwKernel<<<blocks > 0 ? blocks : 1, 256, 0, st>>>(dst, static_cast<const __nv_bfloat16*>(src), N, C, H,
                                                                 W, srcStride, dstC, dstOff);
        OPB_LAUNCH_CHECK();
        return OPB200_OK;
    }
}

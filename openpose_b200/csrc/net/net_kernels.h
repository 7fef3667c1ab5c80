// Internal launch interface between the net runtime (net_runtime.cu) and its kernels.
#pragma once
#include "../common.cuh"
#include "conv_params.h"

namespace opb
{
    opb200_status launchConv(ConvParams& p, int bn, int gridCap, cudaStream_t st);   // fills p's pipeline fields
    opb200_status launchInputPatch(void* dstBf16Nhwc, const float* srcNchw, int N, int C, int H, int W, int k, int pad,
                                   int dstC, cudaStream_t st);
    // First layer straight from the fp32 NCHW input (Cin*k*k <= 32, Cout <= 64): mma.sync, fused bias/activation.
    opb200_status launchFirstConv(void* dstBf16Nhwc, const float* srcNchw, const uint32_t* wfrag, const float* bias,
                                  const float* slope, int N, int C, int H, int W, int k, int pad, int cout,
                                  int dstStride, bool reluOnly, cudaStream_t st);
    opb200_status launchMaxPool(void* dst, const void* src, int N, int C, int H, int W, int Ho, int Wo, int k,
                                int stride, int srcStride, int dstStride, cudaStream_t st);
    opb200_status launchNhwcToNchw(float* dst, const void* src, int N, int C, int H, int W, int srcStride, int dstC,
                                   int dstOff, cudaStream_t st, long long strideN = 0, long long strideY = 0);

    // cuTensorMapEncodeTiled through cudaGetDriverEntryPoint (no link-time dependency on libcuda).
    // Activation view: dims (C, W, H, N) of bf16, pixel stride `strideC` elements; box (boxC, TW, TH, 1).
    opb200_status makeActivationMap(CUtensorMap* m, const void* base, int C, int W, int H, int N, int strideC,
                                    int boxC, int TW, int TH, bool swizzle128);
    // Per-frame view of a frame-interleaved ("flat") level, memory [y][n][x of pitchW][c]: dims (C, W, N, H), box
    // (boxC, TW, 1, TH), SWIZZLE_64B. Used for stores: the separator column x = W of every frame is never written.
    opb200_status makeFlatStoreMap(CUtensorMap* m, const void* base, int C, int W, int H, int N, int pitchW, int strideC,
                                   int boxC, int TW, int TH);
    // Weight matrix: dims (K, rows) of bf16 row-major; box (64, boxRows), SWIZZLE_128B.
    opb200_status makeWeightMap(CUtensorMap* m, const void* base, int K, int rows, int boxRows, int boxBlocks);
}

// Launch descriptor of the implicit-GEMM convolution kernel (conv_sm100.cu), shared by host planner and device.
#pragma once
#include <cuda.h>
#include <cstdint>

namespace opb
{
    constexpr int kConvMaxSegments = 4;   // a conv input is a (virtual) concat of <= 4 blob views
    constexpr int kConvBlockM = 128;      // output pixels per tile = one TW x TH patch of one frame
    constexpr int kConvBlockK = 64;       // channels per k-block = one 128-byte swizzled smem row

    struct ConvParams
    {
        CUtensorMap aMap[kConvMaxSegments];   // input views: dims (C, W, H, N) bf16, box (64, TW, TH, 1), SWIZZLE_128B
        CUtensorMap bMap;                     // weights: dims (Ktotal, CoutPad) bf16, box (64, BN), SWIZZLE_128B
        CUtensorMap oMap;                     // output view: dims (Cout, W, H, N) bf16, box (32, TW, TH, 1), SWIZZLE_64B
        int segChunks[kConvMaxSegments];      // 64-channel chunks per segment (last chunk zero-filled by TMA)
        int nSeg;
        int chunksTotal;
        int kh, kw, pad;
        int H, W, N;                          // stride 1, "same" padding: output spatial == input spatial
        int TW, TH, tilesX, tilesY;
        int nTilesN;                          // Cout tiles of BN channels
        int cout;
        const float* bias;                    // [nTilesN * BN], zero padded
        const float* slope;                   // negative-side slope per channel: 0 = ReLU, a_c = PReLU, 1 = identity
        float* out32;                         // optional fp32 NCHW copy of the output (net_output members)
        int out32C, out32Off;                 // channels of that blob, channel offset of this layer in it
        int storeBf16;                        // 0: skip the bf16 NHWC store (nobody reads it)
    };
}

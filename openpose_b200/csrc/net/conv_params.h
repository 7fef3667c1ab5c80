// Launch descriptor of the implicit-GEMM convolution kernel (conv_sm100.cu), shared by host planner and device.
#pragma once
#include <cuda.h>
#include <cstdint>

namespace opb
{
    constexpr int kConvMaxSegments = 4;   // a conv input is a (virtual) concat of <= 4 blob views
    constexpr int kConvTileW = 8;         // output tile = 8 x 16 pixels of one frame = 128 GEMM rows;
    constexpr int kConvTileH = 16;        //   8 pixels = one 1024-byte swizzle group, so a filter-row shift is +1024 B
    constexpr int kConvBlockM = kConvTileW * kConvTileH;
    constexpr int kConvBlockK = 64;       // channels per k-block = one 128-byte swizzled smem row
    constexpr int kConvMaxKernel = 7;
    constexpr int kConvMaxChunks = 32;    // 64-channel chunks of one conv input (Cin <= 2048)

    struct ConvParams
    {
        // input views: dims (C, W, H, N) bf16; box (64, 8, 16 + k - 1, 1) = the tile plus its vertical halo, SWIZZLE_128B
        CUtensorMap aMap[kConvMaxSegments];
        CUtensorMap bMap;                     // weights: dims (Ktotal, CoutPad) bf16, box (64, BN), SWIZZLE_128B
        CUtensorMap oMap;                     // output view: dims (Cout, W, H, N) bf16, box (32, 8, 16, 1), SWIZZLE_64B
        int segChunks[kConvMaxSegments];      // 64-channel chunks per segment (last chunk zero-filled by TMA)
        uint8_t chunkSteps[kConvMaxChunks];               // per chunk (all segments in order): 16-channel k-steps with real channels, 1..4
        int nSeg;
        int chunksTotal;
        int k, pad;                           // square odd kernel, "same" padding
        int H, W, N;                          // stride 1: output spatial == input spatial
        int tilesX, tilesY, numMTiles;        // numMTiles = N * tilesY * tilesX; a CTA works on PAIRS of M tiles
        int nTilesN;                          // Cout tiles of BN channels
        int cout;
        int kxSpan;                           // filter columns served by one activation box: k (halo box) or 1
        int boxW;                             // box width in pixels = 8 + kxSpan - 1
        int aTileBytes;                       // boxW * (16 + k - 1) * 128, rounded up to 1024
        int aStages, bStages;                 // pipeline depths (A stage = 2 tiles x aTileBytes, B stage = BN x 128 B)
        int stagingBufs;                      // output staging buffers: 2 = one per tile of a work item
        int bGroup;                           // K blocks (filter rows) per weight stage / TMA box: 3 for 3x3, else 1
        int bResident;                        // 1: the whole weight tile fits in the B stages and is loaded once per CTA
        const float* bias;                    // [nTilesN * BN], zero padded
        const float* slope;                   // negative-side slope per channel: 0 = ReLU, a_c = PReLU, 1 = identity
        float* out32;                         // optional fp32 NCHW copy of the output (net_output members)
        int out32C, out32Off;                 // channels of that blob, channel offset of this layer in it
        int storeBf16;                        // 0: skip the bf16 NHWC store (nobody reads it)
        int poolFused;                        // 1: 2x2/2 MAX pooling in the epilogue; oMap is the pooled blob, box (32, 4, 8, 1)
        int actMode;                          // 1: every slope is 0 (ReLU: add + max); 0: general per-channel slope
        // Frame-interleaved ("flat") levels: memory [y][n][x of flatPitch = W + sep][c] with sep zero separator columns
        // (the widest horizontal padding of the level's convs: 1 for 3x3, 3 for 7x7) after every frame, so the batch is ONE image of width N * (W + sep) whose 8-pixel groups run across frame borders (no
        // per-frame rounding of W up to 8) and whose separators are the horizontal zero padding. The kernel then sees
        // N = 1, W = N * flatPitch, and its epilogue writes zeros into the separator columns; only the pooled store of the
        // level above goes through a per-frame map (C, W, N, H), which clips at each frame's right edge.
        int flatPitch;                        // 0: plain [n][y][x][c]; else W + sep of the real frames
        int flatW, flatN;                     // real frame width and count (fp32 NCHW output addressing)
        int poolOutFlat;                      // the pooled blob's level is flat: pooled store coordinates are (c, x, n, y)
        int reverse;                          // 1: M-tile groups are walked last to first (serpentine order across layers: the
                                              //    part of the input written last by the previous layer is still in L2)
        int debugMode;                        // experiments only (OPB200_CONV_DEBUG): 1 = epilogue skips math and stores, 2 = no MMAs
        int pairMode;                         // 1: 2-CTA clusters (cta_group::2); bMap box is then (64, BN/2)
        int tilesPerItem;                     // M tiles per CTA and work item: 2, or 1 when a small batch leaves too few tile
                                              //   pairs to occupy the SMs (the second half of an A stage then stays unused)
    };
}

// Implicit-GEMM convolution for sm_100a: TMA -> swizzled smem -> tcgen05.mma (bf16 x bf16 -> fp32 in TMEM)
// -> fused epilogue (bias + ReLU/PReLU/identity, bf16 NHWC store at a channel offset = free concat, optional
// fp32 NCHW copy for net_output members).
//
// Replaces, for the whole graph, what the reference runs per layer as cuDNN conv + bias + activation kernel +
// concat copy (3rdparty/caffe/src/caffe/layers/cudnn_conv_layer.cu:21-37, relu_layer.cu:9, prelu_layer.cu:11-18,
// concat_layer.cu:9) and on the CPU as im2col + sgemm (base_conv_layer.cpp:258-281).
//
// GEMM view: M = output pixels (tiles of 128 = TW x TH patch of one frame), N = Cout, K = taps x Cin.
// No im2col buffer exists: for every filter tap the producer issues one 4-D TMA box load of the input patch
// shifted by (kx - pad, ky - pad); TMA zero-fills out-of-image elements, which IS the conv's zero padding,
// and zero-fills channels beyond a view's extent, which pads Cin to a multiple of 64 for free.
//
// Roles (256 threads, persistent over tiles): warp 0 = TMA producer, warp 1 = MMA issuer (one thread),
// warp 2 = TMEM allocator, warps 4-7 = epilogue (thread t <-> TMEM lane t <-> pixel t of the tile).
// Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
#include "../common.cuh"
#include "conv_params.h"
#include "sm100.cuh"
#include <cuda_bf16.h>

namespace opb
{
    using namespace sm100;

    constexpr int kConvThreads = 256;
    constexpr int kABytes = kConvBlockM * kConvBlockK * 2;   // 16 KB

    template <int BN>
    struct ConvCfg
    {
        static constexpr int kBBytes = BN * kConvBlockK * 2;
        static constexpr int kStageBytes = kABytes + kBBytes;
        static constexpr int kStagingBytes = (BN / 32) * (kConvBlockM * 64);   // bf16 output tile, 32-channel pieces
        static constexpr int kBudget = 222 * 1024;   // stages + staging; + 5 KB align/aux = 227 KB max dynamic smem
        static constexpr int kStagesRaw = (kBudget - kStagingBytes) / kStageBytes;
        static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
        static constexpr int kAccStride = BN <= 32 ? 32 : (BN <= 64 ? 64 : (BN <= 128 ? 128 : 256));
        static constexpr int kTmemCols = 2 * kAccStride;
        static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 /*align*/ + 4096 /*aux*/;
    };

    struct ConvAux   // barriers and small tables, placed after the pipeline buffers
    {
        uint64_t full[8];
        uint64_t empty[8];
        uint64_t tmemFull[2];
        uint64_t tmemEmpty[2];
        uint32_t tmemBase;
        uint32_t pad;
        float bias[256];
        float slope[256];
    };
    static_assert(sizeof(ConvAux) <= 4096, "aux block too large");

    template <int BN>
    __global__ void __launch_bounds__(kConvThreads, 1) convIgemmKernel(const __grid_constant__ ConvParams p)
    {
        using Cfg = ConvCfg<BN>;
        extern __shared__ uint8_t smemRaw[];
        uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smemRaw) + 1023) & ~uintptr_t(1023));
        uint8_t* sStage = smem;                                        // kStages x (A 16 KB | B)
        uint8_t* sOut = smem + Cfg::kStages * Cfg::kStageBytes;        // staging for the TMA store
        ConvAux* aux = reinterpret_cast<ConvAux*>(sOut + Cfg::kStagingBytes);

        const int warp = threadIdx.x >> 5;
        const int lane = threadIdx.x & 31;
        const int tapsTotal = p.kh * p.kw;
        const int numK = tapsTotal * p.chunksTotal;
        const int tilesPerFrame = p.tilesX * p.tilesY;
        const int numTiles = p.N * tilesPerFrame * p.nTilesN;

        if (warp == 0 && lane == 0)
        {
            for (int s = 0; s < p.nSeg; ++s)
                prefetchTensorMap(&p.aMap[s]);
            prefetchTensorMap(&p.bMap);
            prefetchTensorMap(&p.oMap);
        }
        if (warp == 1 && lane == 0)
        {
            for (int i = 0; i < Cfg::kStages; ++i)
            {
                mbarInit(&aux->full[i], 1);
                mbarInit(&aux->empty[i], 1);
            }
            for (int i = 0; i < 2; ++i)
            {
                mbarInit(&aux->tmemFull[i], 1);
                mbarInit(&aux->tmemEmpty[i], 128);
            }
            fenceBarrierInit();
        }
        if (warp == 2)
            tmemAlloc<Cfg::kTmemCols>(&aux->tmemBase);
        tcFenceBefore();
        __syncthreads();
        tcFenceAfter();
        const uint32_t tmemBase = aux->tmemBase;

        if (warp == 0)
        {
            // ===================================== TMA producer ======================================
            if (lane == 0)
            {
                int stage = 0;
                uint32_t phase = 0;
                for (int tile = blockIdx.x; tile < numTiles; tile += gridDim.x)
                {
                    const int nt = tile % p.nTilesN;
                    const int m = tile / p.nTilesN;
                    const int x0 = (m % p.tilesX) * p.TW;
                    const int y0 = ((m / p.tilesX) % p.tilesY) * p.TH;
                    const int n = m / tilesPerFrame;
                    for (int tap = 0; tap < tapsTotal; ++tap)
                    {
                        const int ky = tap / p.kw, kx = tap - ky * p.kw;
                        int kb = tap * p.chunksTotal;
                        for (int s = 0; s < p.nSeg; ++s)
                            for (int c = 0; c < p.segChunks[s]; ++c, ++kb)
                            {
                                mbarWait(&aux->empty[stage], phase ^ 1);
                                uint8_t* sA = sStage + stage * Cfg::kStageBytes;
                                mbarExpectTx(&aux->full[stage], Cfg::kStageBytes);
                                tmaLoad4d(sA, &p.aMap[s], &aux->full[stage], c * kConvBlockK, x0 + kx - p.pad,
                                          y0 + ky - p.pad, n);
                                tmaLoad2d(sA + kABytes, &p.bMap, &aux->full[stage], kb * kConvBlockK, nt * BN);
                                if (++stage == Cfg::kStages)
                                {
                                    stage = 0;
                                    phase ^= 1;
                                }
                            }
                    }
                }
            }
        }
        else if (warp == 1)
        {
            // ===================================== MMA issuer ========================================
            if (lane == 0)
            {
                constexpr uint32_t idesc = instrDescBf16(kConvBlockM, BN);
                int stage = 0;
                uint32_t phase = 0;
                int acc = 0;
                uint32_t accPhase = 0;
                for (int tile = blockIdx.x; tile < numTiles; tile += gridDim.x)
                {
                    mbarWait(&aux->tmemEmpty[acc], accPhase ^ 1);
                    tcFenceAfter();
                    const uint32_t tmemD = tmemBase + acc * Cfg::kAccStride;
                    for (int kb = 0; kb < numK; ++kb)
                    {
                        mbarWait(&aux->full[stage], phase);
                        tcFenceAfter();
                        const uint32_t aAddr = smemAddr(sStage + stage * Cfg::kStageBytes);
                        const uint64_t aDesc = smemDescSw128(aAddr);
                        const uint64_t bDesc = smemDescSw128(aAddr + kABytes);
#pragma unroll
                        for (int k = 0; k < kConvBlockK / 16; ++k)   // 16 bf16 = 32 bytes per MMA along K
                            mmaBf16(tmemD, aDesc + 2 * k, bDesc + 2 * k, idesc, (kb | k) != 0);
                        mmaCommit(&aux->empty[stage]);                // smem slot free once these MMAs retire
                        if (++stage == Cfg::kStages)
                        {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                    mmaCommit(&aux->tmemFull[acc]);                   // accumulator complete -> epilogue
                    if (++acc == 2)
                    {
                        acc = 0;
                        accPhase ^= 1;
                    }
                }
            }
        }
        else if (warp >= 4)
        {
            // ===================================== epilogue ==========================================
            const int t = threadIdx.x - 128;          // 0..127 = TMEM lane = pixel of the tile
            const int quad = warp & 3;                // warp (id % 4) may only touch lanes [32*quad, 32*quad+32)
            const int tw = t % p.TW, th = t / p.TW;
            int acc = 0;
            uint32_t accPhase = 0;
            int lastNt = -1;
            for (int tile = blockIdx.x; tile < numTiles; tile += gridDim.x)
            {
                const int nt = tile % p.nTilesN;
                const int m = tile / p.nTilesN;
                const int x0 = (m % p.tilesX) * p.TW;
                const int y0 = ((m / p.tilesX) % p.tilesY) * p.TH;
                const int n = m / tilesPerFrame;
                // staging buffer and bias table are free once the previous tile's TMA store has read smem
                if (t == 0)
                    tmaStoreWaitRead0();
                namedBarrier(1, 128);
                if (nt != lastNt)
                {
                    for (int i = t; i < BN; i += 128)
                    {
                        aux->bias[i] = __ldg(p.bias + nt * BN + i);
                        aux->slope[i] = __ldg(p.slope + nt * BN + i);
                    }
                    lastNt = nt;
                    namedBarrier(1, 128);
                }
                mbarWait(&aux->tmemFull[acc], accPhase);
                tcFenceAfter();
                const uint32_t taddr = tmemBase + acc * Cfg::kAccStride + ((uint32_t)(quad * 32) << 16);
                const int x = x0 + tw, y = y0 + th;
                const bool inImage = (x < p.W) && (y < p.H);
#pragma unroll 1
                for (int j = 0; j < BN / 32; ++j)
                {
                    uint32_t v[32];
                    tmemLoad32(taddr + j * 32, v);
                    tmemLoadWait();
                    float f[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                    {
                        const float a = __uint_as_float(v[i]) + aux->bias[j * 32 + i];
                        f[i] = a > 0.f ? a : a * aux->slope[j * 32 + i];
                    }
                    if (p.out32 != nullptr && inImage)
                    {
                        const int cBase = nt * BN + j * 32;
                        float* o = p.out32 + (((size_t)n * p.out32C + p.out32Off + cBase) * p.H + y) * p.W + x;
                        const size_t plane = (size_t)p.H * p.W;
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (cBase + i < p.cout)
                                o[i * plane] = f[i];
                    }
                    if (p.storeBf16)
                    {
                        // row t of a 128 x 64-byte piece; 16-byte chunk c lands at c ^ ((t >> 1) & 3) (SWIZZLE_64B)
                        uint8_t* row = sOut + j * (kConvBlockM * 64) + t * 64;
                        const int sw = (t >> 1) & 3;
#pragma unroll
                        for (int c = 0; c < 4; ++c)
                        {
                            uint4 q;
                            __nv_bfloat162 b0 = __floats2bfloat162_rn(f[c * 8 + 0], f[c * 8 + 1]);
                            __nv_bfloat162 b1 = __floats2bfloat162_rn(f[c * 8 + 2], f[c * 8 + 3]);
                            __nv_bfloat162 b2 = __floats2bfloat162_rn(f[c * 8 + 4], f[c * 8 + 5]);
                            __nv_bfloat162 b3 = __floats2bfloat162_rn(f[c * 8 + 6], f[c * 8 + 7]);
                            q.x = *reinterpret_cast<uint32_t*>(&b0);
                            q.y = *reinterpret_cast<uint32_t*>(&b1);
                            q.z = *reinterpret_cast<uint32_t*>(&b2);
                            q.w = *reinterpret_cast<uint32_t*>(&b3);
                            *reinterpret_cast<uint4*>(row + ((c ^ sw) << 4)) = q;
                        }
                    }
                }
                // all TMEM reads of this thread are complete (wait::ld above): hand the accumulator back
                tcFenceBefore();
                mbarArrive(&aux->tmemEmpty[acc]);
                if (++acc == 2)
                {
                    acc = 0;
                    accPhase ^= 1;
                }
                if (p.storeBf16)
                {
                    fenceProxyAsyncShared();
                    namedBarrier(1, 128);
                    if (t == 0)
                    {
                        for (int j = 0; j < BN / 32; ++j)
                            if (nt * BN + j * 32 < p.cout)
                                tmaStore4d(&p.oMap, sOut + j * (kConvBlockM * 64), nt * BN + j * 32, x0, y0, n);
                        tmaStoreCommit();
                    }
                }
            }
            if (t == 0)
                tmaStoreWaitAll();
        }
        tcFenceBefore();
        __syncthreads();
        if (warp == 2)
            tmemFree<Cfg::kTmemCols>(tmemBase);
    }

    template <int BN>
    static opb200_status launchConvT(const ConvParams& p, int gridCap, cudaStream_t st)
    {
        using Cfg = ConvCfg<BN>;
        static bool configured = false;   // per process; the attribute is per function, set once per device below
        static int configuredDev[64] = {0};
        int dev = 0;
        OPB_CUDA(cudaGetDevice(&dev));
        if (!configured || !configuredDev[dev & 63])
        {
            OPB_CUDA(cudaFuncSetAttribute(convIgemmKernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          Cfg::kSmemBytes));
            configured = true;
            configuredDev[dev & 63] = 1;
        }
        const int numTiles = p.N * p.tilesX * p.tilesY * p.nTilesN;
        int grid = numTiles < gridCap ? numTiles : gridCap;
        if (grid < 1)
            grid = 1;
        convIgemmKernel<BN><<<grid, kConvThreads, Cfg::kSmemBytes, st>>>(p);
        OPB_LAUNCH_CHECK();
        return OPB200_OK;
    }

    opb200_status launchConv(const ConvParams& p, int bn, int gridCap, cudaStream_t st)
    {
        switch (bn)
        {
        case 32: return launchConvT<32>(p, gridCap, st);
        case 64: return launchConvT<64>(p, gridCap, st);
        case 96: return launchConvT<96>(p, gridCap, st);
        case 128: return launchConvT<128>(p, gridCap, st);
        case 256: return launchConvT<256>(p, gridCap, st);
        default: OPB_FAIL(OPB200_ERR_INVALID, "conv: unsupported BLOCK_N");
        }
    }
}
